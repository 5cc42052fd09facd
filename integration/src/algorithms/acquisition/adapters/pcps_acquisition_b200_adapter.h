/*
 * pcps_acquisition_b200_adapter.h — what the three B200 PCPS acquisition adapters share.
 *
 * The reference has one adapter file pair per signal (gps_l1_ca_pcps_acquisition.{h,cc}, galileo_e1_pcps_ambiguous_acquisition,
 * beidou_b1i_pcps_acquisition, ...) that differ in three things only: the signal constants handed to
 * Acq_Conf::SetFromConfiguration, the rate the local code is sampled at, and the sampled-code generator. Here those three are a
 * small description object (B200_Acq_Signal) and everything else — the AcquisitionInterface / GNSSBlockInterface methods, the
 * "cbyte" conversion chain, the replication of the code over the coherent time — is written once.
 *
 * Configuration keys: everything Acq_Conf reads (Acquisition_XX.item_type, .doppler_max, .doppler_step,
 * .coherent_integration_time_ms, .max_dwells, .pfa, .make_two_steps, .dump, ... and GNSS-SDR.internal_fs_sps,
 * GNSS-SDR.use_acquisition_resampler), unchanged, plus
 *     Acquisition_XX.device   = 0    CUDA device of the engine bank the channel joins
 *     Acquisition_XX.devices  = 1    > 1: every dwell is ONE search spread over that many GPUs (CFAR statistic, gr_complex)
 *     Channel<N>.RF_channel_ID / SignalSource ID are passed by the factory as `source_id` so that channels fed by different
 *     front ends never share a sample block.
 *
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_PCPS_ACQUISITION_B200_ADAPTER_H
#define GNSS_SDR_PCPS_ACQUISITION_B200_ADAPTER_H

#include "acq_conf.h"
#include "acquisition_interface.h"
#include "channel_fsm.h"
#include "gnss_synchro.h"
#include "pcps_acquisition_b200_block.h"
#include <complex>
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>

class ConfigurationInterface;

struct B200_Acq_Signal
{
    std::string implementation;   // what implementation() answers, e.g. "GPS_L1_CA_PCPS_Acquisition_B200"
    uint32_t ms_per_code;         // code period [ms]
    double code_rate_cps;         // chip rate
    double code_length_chips;     // chips per period
    double opt_acq_fs_sps;        // target rate of the automatic acquisition resampler
    bool code_at_resampled_rate;  // false: the reference adapter always samples the code at fs_in (BeiDou B1I)
    // one code period of PRN `prn` sampled at `fs` into dest (dest.size() = samples per period)
    std::function<void(std::complex<float>* dest, size_t n, const Gnss_Synchro& synchro, int32_t fs, unsigned int channel)> sampled_code;
};

class PcpsAcquisitionB200Adapter : public AcquisitionInterface
{
public:
    PcpsAcquisitionB200Adapter(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
        B200_Acq_Signal signal, uint32_t source_id);
    ~PcpsAcquisitionB200Adapter() override = default;

    // ---- GNSSBlockInterface ----
    std::string role() override { return role_; }
    std::string implementation() override { return signal_.implementation; }
    size_t item_size() override { return acq_parameters_.it_size; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override;
    gr::basic_block_sptr get_right_block() override { return acquisition_; }

    // ---- AcquisitionInterface ----
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) override;
    void set_threshold(float threshold) override { acquisition_->set_threshold(threshold); }
    void set_doppler_max(unsigned int doppler_max) override { acquisition_->set_doppler_max(doppler_max); }
    void set_doppler_step(unsigned int doppler_step) override { acquisition_->set_doppler_step(doppler_step); }
    void set_doppler_center(int doppler_center) override { acquisition_->set_doppler_center(doppler_center); }
    void init() override { acquisition_->init(); }
    void set_local_code() override;
    void set_state(int state) override { acquisition_->set_state(state); }
    signed int mag() override { return static_cast<signed int>(acquisition_->mag()); }
    void reset() override { acquisition_->set_active(true); }
    void stop_acquisition() override { acquisition_->set_active(false); }
    void set_resampler_latency(uint32_t latency_samples) override { acquisition_->set_resampler_latency(latency_samples); }

protected:
    bool converts_cbyte() const { return acq_parameters_.item_type == "cbyte"; }
    bool known_item_type() const;

    B200_Acq_Signal signal_;
    Acq_Conf acq_parameters_;
    std::string role_;
    pcps_acquisition_b200_block_sptr acquisition_;
    gr::basic_block_sptr cbyte_to_float_x2_, float_to_complex_;  // only for "cbyte" items, as in the reference adapters
    Gnss_Synchro* gnss_synchro_{nullptr};
    std::vector<std::complex<float>> code_;  // the code period replicated over the coherent time
    unsigned int channel_{0};
    unsigned int code_length_{0};  // samples of one code period at the acquisition rate
    unsigned int vector_length_{0};
};

#endif  // GNSS_SDR_PCPS_ACQUISITION_B200_ADAPTER_H
