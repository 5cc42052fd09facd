/*
 * Compile-time stand-in for the reference's Acq_Conf (src/algorithms/acquisition/libs/acq_conf.h), used only when the host
 * mirror is built outside the GNSS-SDR tree (B200_HOST_STANDALONE). Inside the tree the real header is included instead, so
 * the member NAMES here have to be the reference's; everything else is this repository's own: only the members the mirror
 * and its test driver read are present, grouped by what they configure. Defaults follow acq_conf.h:40-80.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef B200_STUB_ACQ_CONF_H
#define B200_STUB_ACQ_CONF_H
#include <complex>
#include <cstddef>
#include <cstdint>
#include <string>

class ConfigurationInterface;

class Acq_Conf
{
public:
    // Reads the Acquisition_XX.* / GNSS-SDR.* keys and derives the per-rate members, as acq_conf.cc:24-119 does. Only the
    // integration sources call it; defined in integration/stubs/acq_conf_stub.cc for builds outside the GNSS-SDR tree.
    void SetFromConfiguration(const ConfigurationInterface* configuration, const std::string& role, double chip_rate, double opt_freq);
    uint32_t chips_per_second = 1023000;

    // --- input stream -------------------------------------------------------------------------------------
    int64_t fs_in = 4000000;                           // GNSS-SDR.internal_fs_sps: the UNDECIMATED rate, also with the resampler on [Hz]
    int64_t resampled_fs = 0;                          // rate of the items the block receives: fs_in, or fs_in / resampler_ratio
    std::string item_type = "gr_complex";              // "gr_complex" | "cshort" | "cbyte"
    size_t it_size = sizeof(std::complex<float>);      // bytes per item of item_type
    float samples_per_ms = 0.0F;                       // resampled_fs / 1000 (SetDerivedParams, acq_conf.cc:113-119)
    float samples_per_code = 0.0F;                     // samples_per_ms * ms_per_code
    uint32_t samples_per_chip = 2;                     // ceil(fs / chip rate), exclusion window of the 1st/2nd peak test

    // --- dwell geometry -----------------------------------------------------------------------------------
    uint32_t sampled_ms = 1;                           // coherent integration time [ms]
    uint32_t ms_per_code = 1;                          // code period [ms]
    uint32_t max_dwells = 1;                           // non-coherent dwells before a negative decision
    bool bit_transition_flag = false;                  // two code periods per dwell, half of the lags kept

    // --- Doppler search -----------------------------------------------------------------------------------
    int32_t doppler_max = 5000;                        // +- [Hz]
    float doppler_step = 250.0F;                       // coarse bin width [Hz]
    bool make_2_steps = false;                         // second, narrow search around the coarse answer
    uint32_t num_doppler_bins_step2 = 4;               //   its bin count (second_nbins)
    float doppler_step2 = 125.0F;                      //   its bin width (second_doppler_step) [Hz]
    bool make_repeat_steps = false;                    // the fork's mode: keep searching after the first detection

    // --- decision -----------------------------------------------------------------------------------------
    bool use_CFAR_algorithm_flag = true;               // max / input power, else first / second peak
    float pfa = 0.0F;                                  // > 0: threshold derived from the false-alarm probability
    float pfa2 = 0.0F;                                 //      the same for the second step (pfa_second_step)

    // --- scheduling ---------------------------------------------------------------------------------------
    bool blocking = true;                              // search inside general_work instead of a worker thread
    bool blocking_on_standby = false;                  // do not consume items while inactive
    bool enable_monitor_output = false;                // copy Gnss_Synchro of positive acquisitions to output 0

    // --- acquisition resampler (gnss_flowgraph.cc:1063-1118) ------------------------------------------------
    bool use_automatic_resampler = false;
    float resampler_ratio = 1.0F;                      // decimation
    uint32_t resampler_latency_samples = 0;            // (ntaps - 1) / 2 of the decimating FIR

    // --- dump ---------------------------------------------------------------------------------------------
    bool dump = false;
    uint32_t dump_channel = 0;                         // only this channel writes
    std::string dump_filename;                         // path prefix; "acquisition" if empty
};
#endif  // B200_STUB_ACQ_CONF_H
