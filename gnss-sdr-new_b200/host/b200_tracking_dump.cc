/*
 * b200_tracking_dump.cc — see the header. ":NNN" = line of dll_pll_veml_tracking.cc the statement follows.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_tracking_dump.h"
#include "b200_mat_writer.h"
#include <cmath>
#include <complex>
#include <cstring>
#include <filesystem>
#include <iostream>
#include <vector>

bool B200_Tracking_Dump::configure(const std::string& dump_filename, bool dump_mat)
{
    d_dump = true;
    d_dump_mat = dump_mat;  // :125
    d_dump_filename = dump_filename;
    std::string dump_path;
    if (d_dump_filename.find_last_of('/') != std::string::npos)
        {
            const std::string dump_filename_ = d_dump_filename.substr(d_dump_filename.find_last_of('/') + 1);
            dump_path = d_dump_filename.substr(0, d_dump_filename.find_last_of('/'));
            d_dump_filename = dump_filename_;
        }
    else
        {
            dump_path = std::string(".");
        }
    if (d_dump_filename.empty())
        {
            d_dump_filename = "trk_channel_";
        }
    if (d_dump_filename.substr(1).find_last_of('.') != std::string::npos)
        {
            d_dump_filename = d_dump_filename.substr(0, d_dump_filename.find_last_of('.'));
        }
    d_dump_filename = dump_path + '/' + d_dump_filename;
    std::error_code ec;
    std::filesystem::create_directories(dump_path, ec);
    if (ec)
        {
            std::cerr << "GNSS-SDR cannot create dump files for the tracking block. Wrong permissions?\n";
            d_dump = false;
        }
    return d_dump;
}


std::string B200_Tracking_Dump::dat_filename() const
{
    return d_dump_filename + std::to_string(d_channel) + ".dat";
}


bool B200_Tracking_Dump::set_channel(uint32_t channel)
{
    d_channel = channel;
    if (!d_dump) return false;
    if (!d_dump_file.is_open())
        {
            d_dump_file.open(dat_filename().c_str(), std::ios::out | std::ios::binary);
        }
    return d_dump_file.is_open();
}


void B200_Tracking_Dump::set_acquisition(uint32_t prn, double acq_code_phase_samples, double acq_carrier_doppler_hz)
{
    d_prn = prn;
    d_acq_code_phase_samples = acq_code_phase_samples;
    d_acq_carrier_doppler_hz = acq_carrier_doppler_hz;
}


bool B200_Tracking_Dump::log(const gnss_b200_trk_epoch* epochs, uint32_t n_epochs, bool veml)
{
    if (!d_dump || !d_dump_file.is_open() || epochs == nullptr) return false;
    auto put = [this](const auto& v) { d_dump_file.write(reinterpret_cast<const char*>(&v), sizeof(v)); };
    for (uint32_t i = 0; i < n_epochs; i++)
        {
            const gnss_b200_trk_epoch& e = epochs[i];
            if ((e.flags & GNSS_B200_TRK_VALID) == 0U) continue;
            auto mag = [&e](int tap) { return std::abs(std::complex<float>(e.corr[2 * tap], e.corr[2 * tap + 1])); };
            put(veml ? mag(0) : 0.0F);  // :1431-1443
            put(mag(1));
            put(mag(2));
            put(mag(3));
            put(veml ? mag(4) : 0.0F);
            put(e.corr[4]);  // prompt I, Q (:1421-1430; the loop tracks the data component)
            put(e.corr[5]);
            const uint64_t stamp = e.sample_counter + static_cast<uint64_t>(e.prn_length_samples);  // nitems_read(0) + d_current_prn_length_samples
            put(stamp);
            put(static_cast<float>(e.acc_carrier_phase_rad));
            put(static_cast<float>(e.carrier_doppler_hz));
            put(0.0F);  // carrier phase rate: zero without the high-dynamics option
            put(static_cast<float>(e.code_freq_chips));
            put(0.0F);  // code phase rate: the same
            put(static_cast<float>(e.carr_error_hz));
            put(static_cast<float>(e.carr_error_filt_hz));
            put(static_cast<float>(e.code_error_chips));
            put(static_cast<float>(e.code_error_filt_chips));
            put(static_cast<float>(e.cn0_db_hz));
            put(static_cast<float>(e.carrier_lock_test));
            put(static_cast<float>(e.rem_code_phase_samples));
            put(static_cast<double>(stamp));
            put(d_prn);
            put(static_cast<float>(d_acq_code_phase_samples));
            put(static_cast<float>(d_acq_carrier_doppler_hz));
            put(static_cast<float>(e.evm));
        }
    return static_cast<bool>(d_dump_file);
}


int32_t B200_Tracking_Dump::finish()
{
    if (d_dump_file.is_open()) d_dump_file.close();
    if (d_dump && d_dump_mat) return save_matfile();  // :897-903
    return 0;
}


int32_t B200_Tracking_Dump::save_matfile() const
{
    const std::string dump_filename_ = dat_filename();
    std::cout << "Generating .mat file for " << dump_filename_ << '\n';
    std::ifstream dump_file(dump_filename_.c_str(), std::ios::binary | std::ios::ate);
    if (!dump_file.is_open())
        {
            std::cerr << "Problem opening dump file:" << dump_filename_ << '\n';
            return 1;
        }
    const int64_t num_epoch = static_cast<int64_t>(dump_file.tellg()) / static_cast<int64_t>(epoch_size_bytes);
    dump_file.seekg(0, std::ios::beg);
    // the record's fields in file order: 'f' float, 'q' uint64, 'd' double, 'u' uint32
    static const struct
    {
        const char* name;
        char type;
    } fields[] = {{"abs_VE", 'f'}, {"abs_E", 'f'}, {"abs_P", 'f'}, {"abs_L", 'f'}, {"abs_VL", 'f'}, {"Prompt_I", 'f'}, {"Prompt_Q", 'f'},
        {"PRN_start_sample_count", 'q'}, {"acc_carrier_phase_rad", 'f'}, {"carrier_doppler_hz", 'f'}, {"carrier_doppler_rate_hz", 'f'},
        {"code_freq_chips", 'f'}, {"code_freq_rate_chips", 'f'}, {"carr_error_hz", 'f'}, {"carr_error_filt_hz", 'f'}, {"code_error_chips", 'f'},
        {"code_error_filt_chips", 'f'}, {"CN0_SNV_dB_Hz", 'f'}, {"carrier_lock_test", 'f'}, {"aux1", 'f'}, {"aux2", 'd'}, {"PRN", 'u'},
        {"acq_code_phase_samples", 'f'}, {"acq_carrier_doppler_hz", 'f'}, {"EVM", 'f'}};
    constexpr size_t n_fields = sizeof(fields) / sizeof(fields[0]);
    std::vector<char> raw(static_cast<size_t>(num_epoch) * epoch_size_bytes);
    dump_file.read(raw.data(), static_cast<std::streamsize>(raw.size()));
    if (!dump_file && num_epoch > 0)
        {
            std::cerr << "Problem reading dump file:" << dump_filename_ << '\n';
            return 1;
        }
    std::string filename = dump_filename_;
    filename.erase(filename.length() - 4, 4);
    filename.append(".mat");
    B200_Mat_Writer mat;
    if (!mat.open(filename)) return 1;
    const auto n = static_cast<size_t>(num_epoch);
    std::vector<float> col_f(n);
    std::vector<double> col_d(n);
    std::vector<uint64_t> col_q(n);
    std::vector<uint32_t> col_u(n);
    size_t offset = 0;
    bool ok = true;
    for (size_t k = 0; k < n_fields && ok; k++)
        {
            const char t = fields[k].type;
            const size_t width = (t == 'f' || t == 'u') ? 4 : 8;
            for (size_t i = 0; i < n; i++)
                {
                    const char* src = raw.data() + i * epoch_size_bytes + offset;
                    if (t == 'f') std::memcpy(&col_f[i], src, 4);
                    if (t == 'u') std::memcpy(&col_u[i], src, 4);
                    if (t == 'q') std::memcpy(&col_q[i], src, 8);
                    if (t == 'd') std::memcpy(&col_d[i], src, 8);
                }
            // dims {1, num_epoch} (:1621-1622)
            if (t == 'f') ok = mat.write_single(fields[k].name, col_f.data(), 1, n);
            if (t == 'u') ok = mat.write_uint32(fields[k].name, col_u.data(), 1, n);
            if (t == 'q') ok = mat.write_uint64(fields[k].name, col_q.data(), 1, n);
            if (t == 'd') ok = mat.write_double(fields[k].name, col_d.data(), 1, n);
            offset += width;
        }
    return (ok && mat.close()) ? 0 : 1;
}
