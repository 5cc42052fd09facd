/*
 * b200_mat_writer.h — a minimal writer of MATLAB Level-5 MAT-files (uncompressed numeric arrays only).
 *
 * The reference writes its acquisition dump through matio (pcps_acquisition.cc:408-508: Mat_CreateVer(MAT_FT_MAT73),
 * Mat_VarCreate / Mat_VarWrite per variable). matio and HDF5 are not part of this build, so the dump is written in the
 * older Level-5 container instead: the same variable names, classes and shapes, readable by MATLAB/Octave `load`,
 * scipy.io.loadmat and by matio's Mat_Open (which the reference's Acquisition_Dump_Reader uses and which detects the
 * container version by itself, acquisition_dump_reader.cc:30-60).
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_B200_MAT_WRITER_H
#define GNSS_SDR_B200_MAT_WRITER_H

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <string>

class B200_Mat_Writer
{
public:
    B200_Mat_Writer() = default;
    ~B200_Mat_Writer() { close(); }
    B200_Mat_Writer(const B200_Mat_Writer&) = delete;
    B200_Mat_Writer& operator=(const B200_Mat_Writer&) = delete;

    bool open(const std::string& filename);
    bool close();
    bool is_open() const { return d_f != nullptr; }

    // column-major rows x cols arrays (arma::fmat::memptr() order in the reference)
    bool write_single(const std::string& name, const float* data, size_t rows, size_t cols);
    bool write_double(const std::string& name, const double* data, size_t rows, size_t cols);
    bool write_int32(const std::string& name, const int32_t* data, size_t rows, size_t cols);
    bool write_uint32(const std::string& name, const uint32_t* data, size_t rows, size_t cols);
    bool write_uint64(const std::string& name, const uint64_t* data, size_t rows, size_t cols);
    // 1 x 1 forms
    bool write_single(const std::string& name, float v) { return write_single(name, &v, 1, 1); }
    bool write_double(const std::string& name, double v) { return write_double(name, &v, 1, 1); }
    bool write_int32(const std::string& name, int32_t v) { return write_int32(name, &v, 1, 1); }
    bool write_uint32(const std::string& name, uint32_t v) { return write_uint32(name, &v, 1, 1); }
    bool write_uint64(const std::string& name, uint64_t v) { return write_uint64(name, &v, 1, 1); }

private:
    bool write_array(const std::string& name, uint32_t mx_class, uint32_t mi_type, size_t elem_size, const void* data, size_t rows, size_t cols);
    std::FILE* d_f{nullptr};
};

#endif  // GNSS_SDR_B200_MAT_WRITER_H
