/*
 * b200_tracking_dump.h — the per-channel tracking dump of dll_pll_veml_tracking, fed from the device loop's epoch records.
 *
 * The reference appends one 108-byte record per code period to <dump_filename><channel>.dat (log_data(),
 * dll_pll_veml_tracking.cc:1403-1509: 22 floats, one uint64, one double, one uint32 in a fixed order) and, with
 * dump_mat, converts the file into <dump_filename><channel>.mat when the block is destroyed (save_matfile(), :1511-1735;
 * one 1 x num_epoch variable per field). This class produces both files from gnss_b200_trk_epoch records
 * (include/gnss_b200.h) so that the reference's readers (Tracking_Dump_Reader, src/utils/matlab/dll_pll_veml_read_*.m,
 * the .mat plotting scripts) work on the output of gnss_b200_trk_run. The .mat is a Level-5 container
 * (b200_mat_writer.h) where the reference writes v7.3 through matio.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_B200_TRACKING_DUMP_H
#define GNSS_SDR_B200_TRACKING_DUMP_H

#include "gnss_b200.h"
#include <cstdint>
#include <fstream>
#include <string>

class B200_Tracking_Dump
{
public:
    // dump_filename as configured (Dll_Pll_Conf::dump_filename): path split off, default "trk_channel_", extension
    // dropped, directory created (:565-597); false when the directory cannot be created.
    bool configure(const std::string& dump_filename, bool dump_mat);
    // set_channel (:1735-1760): opens <dump_filename><channel>.dat for writing
    bool set_channel(uint32_t channel);
    // the acquisition the channel was started from (set_gnss_synchro / start_tracking: PRN, d_acq_code_phase_samples,
    // d_acq_carrier_doppler_hz)
    void set_acquisition(uint32_t prn, double acq_code_phase_samples, double acq_carrier_doppler_hz);
    // log_data() for every valid record of a gnss_b200_trk_run output, in order. `veml` as in the loop's configuration.
    bool log(const gnss_b200_trk_epoch* epochs, uint32_t n_epochs, bool veml);
    // the block's destructor: closes the .dat and, with dump_mat, writes the .mat (returns save_matfile()'s 0 / 1)
    int32_t finish();
    std::string dat_filename() const;
    static constexpr int32_t epoch_size_bytes = static_cast<int32_t>(sizeof(uint64_t) + sizeof(double) + sizeof(float) * 22 + sizeof(uint32_t));

private:
    int32_t save_matfile() const;
    std::string d_dump_filename;
    std::ofstream d_dump_file;
    uint32_t d_channel{0};
    uint32_t d_prn{0};
    double d_acq_code_phase_samples{0.0};
    double d_acq_carrier_doppler_hz{0.0};
    bool d_dump{false};
    bool d_dump_mat{false};
};

#endif  // GNSS_SDR_B200_TRACKING_DUMP_H
