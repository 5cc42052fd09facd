/*
 * pcps_acquisition_b200_block.h — the GNU Radio face of the B200 PCPS acquisition.
 *
 * Goes to src/algorithms/acquisition/gnuradio_blocks/ next to pcps_acquisition.h and has the same public surface
 * (pcps_acquisition.h:92-236), so an adapter written for `pcps_acquisition` works with it unchanged. The block itself holds no
 * search arithmetic: sample buffering, the decision state machine and Gnss_Synchro write-back live in `pcps_acquisition_b200`
 * (gnss-sdr-new_b200/host/, plain C++), which talks to the GPU engine through the C-ABI of include/gnss_b200.h. This shell
 * only binds that object to the scheduler: consume_each, the "events" message port, the optional monitor output.
 *
 * SPDX-License-Identifier: GPL-3.0-or-later (it derives from GNU Radio's gr::block and GNSS-SDR's interfaces)
 */
#ifndef GNSS_SDR_PCPS_ACQUISITION_B200_BLOCK_H
#define GNSS_SDR_PCPS_ACQUISITION_B200_BLOCK_H

#include "acq_conf.h"
#include "channel_fsm.h"
#include "gnss_block_interface.h"  // gnss_shared_ptr
#include "pcps_acquisition_b200.h"
#include <gnuradio/block.h>
#include <cstdint>
#include <memory>
#include <utility>

class Gnss_Synchro;
class pcps_acquisition_b200_block;

using pcps_acquisition_b200_block_sptr = gnss_shared_ptr<pcps_acquisition_b200_block>;

pcps_acquisition_b200_block_sptr pcps_make_acquisition_b200_block(const Acq_Conf& conf_);

class pcps_acquisition_b200_block : public gr::block
{
public:
    ~pcps_acquisition_b200_block() override = default;

    // ---- what the adapters call (same names and meaning as on pcps_acquisition) ----
    void init() { d_core->init(); }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) { d_core->set_gnss_synchro(p_gnss_synchro); }
    void set_local_code(std::complex<float>* code) { d_core->set_local_code(code); }
    void set_state(int32_t state) { d_core->set_state(state); }
    void set_active(bool active) { d_core->set_active(active); }
    void set_channel(uint32_t channel) { d_core->set_channel(channel); }
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) { d_core->set_channel_fsm(std::move(channel_fsm)); }
    void set_threshold(float threshold) { d_core->set_threshold(threshold); }
    void set_doppler_max(uint32_t doppler_max) { d_core->set_doppler_max(doppler_max); }
    void set_doppler_step(uint32_t doppler_step) { d_core->set_doppler_step(doppler_step); }
    void set_doppler_center(int32_t doppler_center) { d_core->set_doppler_center(doppler_center); }
    void set_resampler_latency(uint32_t latency_samples) { d_core->set_resampler_latency(latency_samples); }
    uint32_t mag() const { return d_core->mag(); }

    // ---- B200-only placement knobs (Acquisition_XX.device / .devices, the channel's signal source) ----
    void set_device(int device) { d_core->set_device(device); }
    void set_num_devices(uint32_t n_devices) { d_core->set_num_devices(n_devices); }
    void set_source_id(uint32_t source_id) { d_core->set_source_id(source_id); }

    // ---- gr::block ----
    bool start() override { return d_core->start(); }
    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override;

private:
    friend pcps_acquisition_b200_block_sptr pcps_make_acquisition_b200_block(const Acq_Conf& conf_);
    explicit pcps_acquisition_b200_block(const Acq_Conf& conf_);

    pcps_acquisition_b200_sptr d_core;
};

#endif  // GNSS_SDR_PCPS_ACQUISITION_B200_BLOCK_H
