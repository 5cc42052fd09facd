/*
 * pcps_acquisition_b200_adapter.cc — see the header. The numbers (code_length_, vector_length_, how often the code period is
 * repeated) follow gps_l1_ca_pcps_acquisition.cc:40-75,151-170 and its siblings.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "pcps_acquisition_b200_adapter.h"
#include "configuration_interface.h"
#include "conversion_blocks.h"  // in the tree: complex_byte_to_float_x2.h + <gnuradio/blocks/float_to_complex.h>
#include "gnss_sdr_flags.h"
#include <glog/logging.h>
#include <algorithm>
#include <cmath>
#include <utility>


PcpsAcquisitionB200Adapter::PcpsAcquisitionB200Adapter(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, B200_Acq_Signal signal, uint32_t source_id)
    : signal_(std::move(signal)), role_(role)
{
    acq_parameters_.ms_per_code = signal_.ms_per_code;
    acq_parameters_.SetFromConfiguration(configuration, role, signal_.code_rate_cps, signal_.opt_acq_fs_sps);
    DLOG(INFO) << "role " << role;
    if (FLAGS_doppler_max != 0) acq_parameters_.doppler_max = FLAGS_doppler_max;

    const double code_fs = static_cast<double>(signal_.code_at_resampled_rate ? acq_parameters_.resampled_fs : acq_parameters_.fs_in);
    code_length_ = static_cast<unsigned int>(std::floor(code_fs / (signal_.code_rate_cps / signal_.code_length_chips)));
    vector_length_ = static_cast<unsigned int>(std::floor(static_cast<float>(acq_parameters_.sampled_ms) * acq_parameters_.samples_per_ms) *
                                               (acq_parameters_.bit_transition_flag ? 2.0 : 1.0));
    code_.assign(vector_length_, std::complex<float>(0.0F, 0.0F));

    acquisition_ = pcps_make_acquisition_b200_block(acq_parameters_);
    acquisition_->set_device(configuration->property(role + ".device", 0));
    acquisition_->set_num_devices(configuration->property(role + ".devices", 1U));
    acquisition_->set_source_id(source_id);
    DLOG(INFO) << "acquisition(" << acquisition_->unique_id() << ")";

    if (converts_cbyte())
        {
            cbyte_to_float_x2_ = make_complex_byte_to_float_x2();
            float_to_complex_ = gr::blocks::float_to_complex::make();
        }
    if (in_streams > 1) LOG(ERROR) << "This implementation only supports one input stream";
    if (out_streams > 0) LOG(ERROR) << "This implementation does not provide an output stream";
}


bool PcpsAcquisitionB200Adapter::known_item_type() const
{
    const std::string& t = acq_parameters_.item_type;
    return t == "gr_complex" || t == "cshort" || t == "cbyte";
}


void PcpsAcquisitionB200Adapter::set_gnss_synchro(Gnss_Synchro* p_gnss_synchro)
{
    gnss_synchro_ = p_gnss_synchro;
    acquisition_->set_gnss_synchro(gnss_synchro_);
}


void PcpsAcquisitionB200Adapter::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_->set_channel(channel_);
}


void PcpsAcquisitionB200Adapter::set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm)
{
    acquisition_->set_channel_fsm(std::move(channel_fsm));
}


// One code period from the signal's generator, repeated sampled_ms / ms_per_code times into the buffer the block reads
// consumed_samples items from (the last, possibly partial, repetition is cut at vector_length_).
void PcpsAcquisitionB200Adapter::set_local_code()
{
    std::vector<std::complex<float>> period(code_length_);
    const bool resampled = signal_.code_at_resampled_rate && acq_parameters_.use_automatic_resampler;
    const auto fs = static_cast<int32_t>(resampled ? acq_parameters_.resampled_fs : acq_parameters_.fs_in);
    signal_.sampled_code(period.data(), period.size(), *gnss_synchro_, fs, channel_);
    const unsigned int repetitions = std::max(1U, acq_parameters_.sampled_ms / signal_.ms_per_code);
    for (unsigned int i = 0; i < repetitions; i++)
        {
            const size_t at = static_cast<size_t>(i) * code_length_;
            if (at >= code_.size()) break;
            std::copy_n(period.begin(), std::min<size_t>(code_length_, code_.size() - at), code_.begin() + static_cast<std::ptrdiff_t>(at));
        }
    acquisition_->set_local_code(code_.data());
}


// "cbyte" items pass through the reference's own conversion chain in front of the block; the other item types go straight in.
void PcpsAcquisitionB200Adapter::connect(gr::top_block_sptr top_block)
{
    if (converts_cbyte())
        {
            top_block->connect(cbyte_to_float_x2_, 0, float_to_complex_, 0);
            top_block->connect(cbyte_to_float_x2_, 1, float_to_complex_, 1);
            top_block->connect(float_to_complex_, 0, acquisition_, 0);
        }
    else if (!known_item_type())
        {
            LOG(WARNING) << acq_parameters_.item_type << " unknown acquisition item type";
        }
}


void PcpsAcquisitionB200Adapter::disconnect(gr::top_block_sptr top_block)
{
    if (converts_cbyte())
        {
            top_block->disconnect(cbyte_to_float_x2_, 0, float_to_complex_, 0);
            top_block->disconnect(cbyte_to_float_x2_, 1, float_to_complex_, 1);
            top_block->disconnect(float_to_complex_, 0, acquisition_, 0);
        }
    else if (!known_item_type())
        {
            LOG(WARNING) << acq_parameters_.item_type << " unknown acquisition item type";
        }
}


gr::basic_block_sptr PcpsAcquisitionB200Adapter::get_left_block()
{
    if (converts_cbyte()) return cbyte_to_float_x2_;
    if (known_item_type()) return acquisition_;
    LOG(WARNING) << acq_parameters_.item_type << " unknown acquisition item type";
    return nullptr;
}
