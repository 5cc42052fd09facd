/*
 * b200_correlator_pool.h — process-wide owner of the B200 multicorrelator engine (include/gnss_b200.h).
 *
 * The reference gives every tracking channel its own Cpu_Multicorrelator_Real_Codes object and calls it from that
 * channel's GNU Radio thread once per code period (dll_pll_veml_tracking.cc:1064-1089). One GPU launch per channel
 * per millisecond would be launch-latency bound, so the B200 correlator objects all draw a channel id from this pool
 * and their calls are COALESCED: the first caller to arrive becomes the leader, waits a bounded window for the other
 * channels' calls of the same epoch, and issues one gnss_b200_mc_correlate_batch for all of them. With a window of 0
 * every call launches alone (exact reference timing, lowest throughput).
 *
 * Plain host C++17; no CUDA or GNU Radio types.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_B200_CORRELATOR_POOL_H
#define GNSS_SDR_B200_CORRELATOR_POOL_H

#include "gnss_b200.h"
#include <complex>
#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

class B200_Correlator_Pool
{
public:
    static B200_Correlator_Pool& instance();

    // Limits of the shared engine; only effective before the first channel is acquired.
    void configure(int device, uint32_t max_channels, uint32_t max_samples, uint32_t max_taps);
    // Leader waits at most `window_us` for `expected_callers` requests before launching (0/0: no coalescing).
    void set_coalescing(uint32_t window_us, uint32_t expected_callers);

    int acquire_channel();  // -1 when the engine cannot be created (no CUDA device) or the pool is full
    void release_channel(int ch);
    bool set_code(int ch, int code_length, const float* code, bool is_complex);

    // One reference call (one channel, one code period). Blocks until out[0..n_taps) is written.
    bool correlate(int ch, const gnss_b200_mc_args& args, const float* shifts_chips, uint32_t n_taps, uint32_t flags, std::complex<float>* out);
    // Explicit batch for callers that already hold many channels (a multi-channel tracking block).
    bool correlate_many(uint32_t n, const uint32_t* ch, const gnss_b200_mc_args* args, const float* shifts_chips, uint32_t n_taps, uint32_t flags,
        std::complex<float>* out);

    std::string last_error() const;
    uint64_t launches() const;        // kernels launched so far
    uint64_t batches() const;         // engine calls issued
    uint64_t coalesced_calls() const; // reference-level calls served

    B200_Correlator_Pool(const B200_Correlator_Pool&) = delete;
    B200_Correlator_Pool& operator=(const B200_Correlator_Pool&) = delete;

private:
    B200_Correlator_Pool() = default;
    ~B200_Correlator_Pool();
    bool ensure_engine_locked();

    struct Request
    {
        uint32_t ch;
        gnss_b200_mc_args args;
        const float* shifts;
        uint32_t n_taps, flags;
        bool is_complex;  // local code type of the channel: the engine does not mix real and complex codes in one batch
        std::complex<float>* out;
        bool done = false, ok = false;
    };
    void run_batch_locked(std::unique_lock<std::mutex>& lk);

    mutable std::mutex d_mu;
    std::condition_variable d_cv_leader, d_cv_done;
    gnss_b200_mc* d_engine{nullptr};
    int d_device{0};
    uint32_t d_max_channels{4096}, d_max_samples{1u << 20}, d_max_taps{16};
    uint32_t d_window_us{0}, d_expected{0};
    std::vector<uint8_t> d_used;
    std::vector<uint8_t> d_is_complex;  // per channel, recorded by set_code
    std::vector<Request*> d_pending;
    bool d_leader_active{false};
    uint64_t d_batches{0}, d_calls{0};
    std::string d_error;
};

#endif  // GNSS_SDR_B200_CORRELATOR_POOL_H
