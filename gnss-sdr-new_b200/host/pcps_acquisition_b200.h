/*
 * pcps_acquisition_b200.h — B200 drop-in for the pcps_acquisition block
 * (reference: src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.h:93-270, .cc:49-1050).
 *
 * Same public member names, argument meaning and side effects (Gnss_Synchro write-back, ChannelFsm / "events"
 * notification, sample-buffering state machine); the search itself (acquisition_core's Doppler loop and the two
 * statistics) runs in the B200 engine through the C-ABI. The class is independent of GNU Radio so that it can be
 * compiled and tested here; INTEGRATION.md shows the 30-line gr::block shell that forwards general_work to it.
 *
 * make_2_steps is mirrored: the narrow second search runs on the bank's second-step engine handle.
 * dump = true (N4) writes the reference's per-acquisition .mat file (same name pattern, variables, classes and shapes,
 * :408-508) in the Level-5 container (b200_mat_writer.h) instead of matio's v7.3.
 * The fork's make_repeat_steps mode (the block re-arms itself after the first detection and keeps reporting) is mirrored too.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_PCPS_ACQUISITION_B200_H
#define GNSS_SDR_PCPS_ACQUISITION_B200_H

#if defined(B200_HOST_STANDALONE)
#include "stubs/acq_conf.h"
#include "stubs/channel_fsm.h"
#include "stubs/gnss_synchro.h"
#else
#include "acq_conf.h"
#include "channel_fsm.h"
#include "gnss_synchro.h"
#endif
#include "b200_acquisition_bank.h"
#include <complex>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <queue>
#include <string>
#include <vector>

class pcps_acquisition_b200;
using pcps_acquisition_b200_sptr = std::shared_ptr<pcps_acquisition_b200>;

pcps_acquisition_b200_sptr pcps_make_acquisition_b200(const Acq_Conf& conf_);

class pcps_acquisition_b200
{
public:
    ~pcps_acquisition_b200();

    void init();
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro);
    void set_local_code(std::complex<float>* code);
    void set_state(int32_t state);
    void set_resampler_latency(uint32_t latency_samples);
    uint32_t mag() const { return d_mag; }
    void set_active(bool active);
    void set_channel(uint32_t channel);
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) { d_channel_fsm = std::move(channel_fsm); }
    void set_threshold(float threshold);
    void set_doppler_max(uint32_t doppler_max);
    void set_doppler_step(uint32_t doppler_step);
    void set_doppler_center(int32_t doppler_center);
    bool start();  // gr::block::start(): resets the sample counter and derives the threshold from pfa

    // B200-only knobs (no counterpart in pcps_acquisition); both take effect at the next init().
    // Which sample stream feeds this block (the adapter passes the channel's RF_channel_ID / signal-source index): blocks on
    // different streams never share an engine bank, whatever their sample stamps.
    void set_source_id(uint32_t source_id) { d_source_id = source_id; }
    // CUDA device the block's bank lives on (Acquisition_XX.device, default 0).
    void set_device(int device) { d_device = device; }
    // Acquisition_XX.devices > 1: every dwell of the bank is ONE search spread over that many GPUs, first one = `device`
    // (Doppler rows split over the devices, row records exchanged GPU to GPU). CFAR statistic and gr_complex items only.
    void set_num_devices(uint32_t n_devices) { d_num_devices = n_devices; }

    // Same contract as gr::block::general_work (gr_vector_int = std::vector<int>, ...). Items consumed and events
    // published go through the two hooks below, which the gr::block shell binds to consume_each / message_port_pub.
    int general_work(int noutput_items, std::vector<int>& ninput_items, std::vector<const void*>& input_items, std::vector<void*>& output_items);
    std::function<void(int)> consume_each = [](int) {};
    std::function<void(long)> publish_event = [](long) {};  // 1 = ACQ_SUCCESS, 2 = ACQ_FAIL (pcps_acquisition.cc:347,386)

    // introspection for tests
    float test_statistics() const { return d_test_statistics; }
    float threshold() const { return d_threshold; }
    uint32_t num_doppler_bins() const { return d_num_doppler_bins; }
    bool step_two() const { return d_step_two; }
    uint32_t fft_size() const { return d_fft_size; }

private:
    friend pcps_acquisition_b200_sptr pcps_make_acquisition_b200(const Acq_Conf& conf_);
    explicit pcps_acquisition_b200(const Acq_Conf& conf_);

    void acquisition_core(uint64_t samp_count);
    void pass_items_while_idle(int n_items);
    void begin_search(int n_items);
    void fill_dwell_buffer(int n_items, const void* items);
    void start_search(std::unique_lock<std::recursive_mutex>& lk);
    void send_negative_acquisition();
    void send_positive_acquisition();
    bool is_fdma();
    void calculate_threshold();
    void refresh_doppler_grid();
    void forget_last_answer();
    void report_answer(uint32_t indext, int32_t doppler, uint64_t samp_count);
    void attach_engine();
    void dump_results(int32_t effective_fft_size);
    bool dumping() const { return d_dump && d_channel == d_dump_channel; }
    void require_engine(const char* where) const;

    std::recursive_mutex d_setlock;  // gr::block::d_setlock in the reference
    std::shared_ptr<B200_Acquisition_Bank> d_bank;
    int d_slot{-1};
    std::vector<std::complex<float>> d_data_buffer;
    std::vector<int16_t> d_data_buffer_sc;  // interleaved lv_16sc_t
    std::vector<std::complex<float>> d_code_copy;
    std::vector<float> d_grid;         // [num_doppler_bins][effective size] = arma::fmat(effective, bins) in memory
    std::vector<float> d_narrow_grid;  // [num_doppler_bins_step2][effective size]
    std::string d_dump_filename;
    int64_t d_dump_number{0};
    std::weak_ptr<ChannelFsm> d_channel_fsm;
    Acq_Conf d_acq_parameters;
    Gnss_Synchro* d_gnss_synchro{nullptr};
    std::queue<Gnss_Synchro> d_monitor_queue;

    uint64_t d_sample_counter{0};
    float d_threshold{0.0F};
    float d_mag{0.0F};
    float d_input_power{0.0F};
    float d_test_statistics{0.0F};
    float d_doppler_center_step_two{0.0F};
    int32_t d_state{0};
    int32_t d_positive_acq{0};
    int32_t d_doppler_center{0};
    int32_t d_doppler_bias{0};
    uint32_t d_channel{0};
    uint32_t d_source_id{0};
    int d_device{0};
    uint32_t d_num_devices{1};
    uint32_t d_dump_channel{0};
    uint32_t d_samplesPerChip{0};
    uint32_t d_doppler_step{0};
    uint32_t d_num_noncoherent_integrations_counter{0};
    uint32_t d_fft_size{0};
    uint32_t d_consumed_samples{0};
    uint32_t d_num_doppler_bins{0};
    uint32_t d_num_doppler_bins_step2{0};
    uint32_t d_buffer_count{0};
    bool d_active{false};
    bool d_step_two{false};
    bool d_step_repeat{false};
    bool d_dump{false};
    bool d_worker_active{false};
    bool d_cshort{false};
    bool d_use_CFAR_algorithm_flag{true};
    bool d_have_code{false};
};

#endif  // GNSS_SDR_PCPS_ACQUISITION_B200_H
