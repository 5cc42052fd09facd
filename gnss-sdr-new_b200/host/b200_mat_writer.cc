/*
 * b200_mat_writer.cc — see the header. Layout after "MAT-File Format" (MathWorks), Level 5:
 * 128-byte header (116 text, 8 subsystem offset, version 0x0100, endian indicator "MI" as a 16-bit value), then one
 * miMATRIX element per variable = { array flags, dimensions, name, real part }, every sub-element padded to 8 bytes.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_mat_writer.h"
#include <cstring>
#include <ctime>

namespace
{
constexpr uint32_t miINT8 = 1, miINT32 = 5, miUINT32 = 6, miSINGLE = 7, miDOUBLE = 9, miUINT64 = 13, miMATRIX = 14;
constexpr uint32_t mxDOUBLE = 6, mxSINGLE = 7, mxINT32 = 12, mxUINT32 = 13, mxUINT64 = 15;

size_t padded(size_t n) { return (n + 7U) & ~static_cast<size_t>(7U); }

bool put(std::FILE* f, const void* p, size_t n) { return n == 0 || std::fwrite(p, 1, n, f) == n; }

bool put_tag(std::FILE* f, uint32_t type, uint32_t nbytes)
{
    const uint32_t tag[2] = {type, nbytes};
    return put(f, tag, sizeof(tag));
}

bool put_padding(std::FILE* f, size_t nbytes)
{
    static const char zeros[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    return put(f, zeros, padded(nbytes) - nbytes);
}
}  // namespace


bool B200_Mat_Writer::open(const std::string& filename)
{
    close();
    d_f = std::fopen(filename.c_str(), "wb");
    if (d_f == nullptr) return false;
    char header[128];
    std::memset(header, ' ', 116);
    char when[64] = "";
    const std::time_t now = std::time(nullptr);
    std::tm tm_buf{};
    if (gmtime_r(&now, &tm_buf) != nullptr) std::strftime(when, sizeof(when), "%a %b %d %H:%M:%S %Y", &tm_buf);
    char text[116];
    const int n = std::snprintf(text, sizeof(text), "MATLAB 5.0 MAT-file, Platform: GNSS-SDR B200, Created on: %s", when);
    std::memcpy(header, text, static_cast<size_t>(n < 116 ? n : 116));
    std::memset(header + 116, 0, 8);  // no subsystem-specific data
    const uint16_t version = 0x0100;
    const uint16_t endian = 0x4D49;  // bytes "IM" on a little-endian writer; the reader swaps if it sees "MI"
    std::memcpy(header + 124, &version, 2);
    std::memcpy(header + 126, &endian, 2);
    if (!put(d_f, header, sizeof(header)))
        {
            close();
            return false;
        }
    return true;
}


bool B200_Mat_Writer::close()
{
    if (d_f == nullptr) return true;
    const bool ok = std::fclose(d_f) == 0;
    d_f = nullptr;
    return ok;
}


bool B200_Mat_Writer::write_array(const std::string& name, uint32_t mx_class, uint32_t mi_type, size_t elem_size, const void* data, size_t rows, size_t cols)
{
    if (d_f == nullptr || data == nullptr || name.empty() || name.size() > 63) return false;
    const size_t data_bytes = rows * cols * elem_size;
    const size_t total = 16 + 16 + (8 + padded(name.size())) + (8 + padded(data_bytes));
    if (total > 0xFFFFFFFFULL || rows > 0x7FFFFFFFULL || cols > 0x7FFFFFFFULL) return false;  // Level 5 holds 32-bit sizes
    bool ok = put_tag(d_f, miMATRIX, static_cast<uint32_t>(total));
    const uint32_t flags[2] = {mx_class, 0U};  // no complex / global / logical flag
    ok = ok && put_tag(d_f, miUINT32, 8) && put(d_f, flags, 8);
    const int32_t dims[2] = {static_cast<int32_t>(rows), static_cast<int32_t>(cols)};
    ok = ok && put_tag(d_f, miINT32, 8) && put(d_f, dims, 8);
    ok = ok && put_tag(d_f, miINT8, static_cast<uint32_t>(name.size())) && put(d_f, name.data(), name.size()) && put_padding(d_f, name.size());
    ok = ok && put_tag(d_f, mi_type, static_cast<uint32_t>(data_bytes)) && put(d_f, data, data_bytes) && put_padding(d_f, data_bytes);
    return ok;
}


bool B200_Mat_Writer::write_single(const std::string& name, const float* data, size_t rows, size_t cols)
{
    return write_array(name, mxSINGLE, miSINGLE, sizeof(float), data, rows, cols);
}

bool B200_Mat_Writer::write_double(const std::string& name, const double* data, size_t rows, size_t cols)
{
    return write_array(name, mxDOUBLE, miDOUBLE, sizeof(double), data, rows, cols);
}

bool B200_Mat_Writer::write_int32(const std::string& name, const int32_t* data, size_t rows, size_t cols)
{
    return write_array(name, mxINT32, miINT32, sizeof(int32_t), data, rows, cols);
}

bool B200_Mat_Writer::write_uint32(const std::string& name, const uint32_t* data, size_t rows, size_t cols)
{
    return write_array(name, mxUINT32, miUINT32, sizeof(uint32_t), data, rows, cols);
}

bool B200_Mat_Writer::write_uint64(const std::string& name, const uint64_t* data, size_t rows, size_t cols)
{
    return write_array(name, mxUINT64, miUINT64, sizeof(uint64_t), data, rows, cols);
}
