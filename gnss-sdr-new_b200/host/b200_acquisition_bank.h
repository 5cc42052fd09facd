/*
 * b200_acquisition_bank.h — shared owner of B200 acquisition engines (include/gnss_b200.h).
 *
 * In the reference every channel owns a pcps_acquisition block with private FFT plans, wipe-off tables and grid
 * (pcps_acquisition.cc:49-166,249-295), and every block repeats the forward FFTs of the same samples. Here all blocks
 * with the same search geometry share one engine handle (one "bank"); a block is a SLOT of that bank. Dwells that
 * arrive for the same sample block (equal sample stamp) within a bounded window are issued as ONE batched engine
 * call, so the forward transforms are computed once for all of them.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_B200_ACQUISITION_BANK_H
#define GNSS_SDR_B200_ACQUISITION_BANK_H

#include "gnss_b200.h"
#include <complex>
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

class B200_Acquisition_Bank
{
public:
    struct Geometry
    {
        uint32_t fft_size, consumed_samples;
        int32_t doppler_max;
        uint32_t doppler_step;
        double fs;
        uint32_t samples_per_chip;
        bool bit_transition, use_cfar, keep_grid;
        int device;           // first CUDA device of the bank
        uint32_t n_devices;   // > 1: ONE search per dwell spread Doppler-major over devices [device, device + n_devices), the
                              // row records exchanged GPU to GPU (gnss_b200_acq_group_*); needs use_cfar, gr_complex items
        uint32_t step2_bins;  // make_two_steps: num_doppler_bins_step2 (0 = no second engine)
        float step2_hz;       //                 doppler_step2
        uint32_t source_id;   // which sample stream feeds the blocks (RF_channel_ID / signal-source index): banks are never
                              // shared across streams, because equal sample stamps only mean equal samples on ONE stream
        bool operator==(const Geometry& o) const;
    };

    // Banks are shared: the same geometry returns the same bank. nullptr if the engine cannot be created
    // (no CUDA device, unsupported FFT length) — there is no CPU fallback; `why` receives the engine's message.
    static std::shared_ptr<B200_Acquisition_Bank> get(const Geometry& g, std::string* why);
    static void set_default_capacity(uint32_t max_slots);           // slots per bank (default 64)
    static void set_coalescing(uint32_t window_us, uint32_t expected_callers);

    ~B200_Acquisition_Bank();
    int acquire_slot();  // -1 when full
    void release_slot(int slot);
    uint32_t num_doppler_bins() const { return d_num_doppler_bins; }

    bool set_local_code(int slot, const std::complex<float>* code);
    bool set_doppler_center(int slot, int32_t center_hz, int32_t bias_hz);
    bool reset(int slot);
    // One acquisition_core pass for `slot` on `samples` (consumed_samples items). Blocks until `result` is written.
    bool dwell(int slot, const void* samples, bool is_cshort, uint64_t sample_stamp, gnss_b200_acq_result* result);
    // The same with d_step_two set (pcps_acquisition.cc:715-775): the narrow grid around doppler_center_step_two, the
    // CFAR statistic over the input power the first step measured.
    bool dwell_step2(int slot, const void* samples, bool is_cshort, uint64_t sample_stamp, float doppler_center_step_two,
        float input_power_step_one, gnss_b200_acq_result* result);
    bool read_grid(int slot, float* out);        // [num_doppler_bins][effective size]; bank geometry with keep_grid
    bool read_grid_step2(int slot, float* out);  // [step2_bins][effective size]
    std::string last_error() const;
    uint64_t engine_calls() const;
    uint64_t dwells_served() const;

private:
    explicit B200_Acquisition_Bank(const Geometry& g) : d_geo(g) {}
    struct Request
    {
        uint32_t slot;
        const void* samples;
        bool is_cshort;
        uint64_t stamp;
        gnss_b200_acq_result* result;
        bool step2 = false;
        float center2 = 0.F, input_power1 = 0.F;
        bool done = false, ok = false;
    };
    bool submit(Request& req);
    void drain_locked(std::unique_lock<std::mutex>& lk);
    bool same_block(const void* a, const void* b, bool is_cshort) const;

    Geometry d_geo;
    gnss_b200_acq* d_engine{nullptr};
    gnss_b200_acq_group* d_group{nullptr};  // instead of d_engine when the bank spans several devices
    gnss_b200_acq* d_engine2{nullptr};  // second-step handle (gnss_b200_acq_create_step2), same slots
    uint32_t d_num_doppler_bins{0};
    mutable std::mutex d_mu;
    std::condition_variable d_cv_leader, d_cv_done;
    std::vector<uint8_t> d_used;
    std::vector<Request*> d_pending;
    bool d_leader_active{false};
    uint64_t d_engine_calls{0}, d_dwells{0};
    std::string d_error;
};

#endif  // GNSS_SDR_B200_ACQUISITION_BANK_H
