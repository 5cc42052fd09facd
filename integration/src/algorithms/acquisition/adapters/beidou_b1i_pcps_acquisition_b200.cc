/*
 * beidou_b1i_pcps_acquisition_b200.cc — BeiDou B1I: 1 ms code, 2.046 Mcps; the reference adapter samples the code at fs_in
 * whatever the resampler does and asks for a 10 MS/s acquisition rate (beidou_b1i_pcps_acquisition.cc:50-51,66,140-153).
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "beidou_b1i_pcps_acquisition_b200.h"
#include "signal_replicas.h"  // in the tree: Beidou_B1I.h + beidou_b1i_signal_replica.h

namespace
{
B200_Acq_Signal beidou_b1i()
{
    B200_Acq_Signal s;
    s.implementation = "BEIDOU_B1I_PCPS_Acquisition_B200";
    s.ms_per_code = 1;
    s.code_rate_cps = BEIDOU_B1I_CODE_RATE_CPS;
    s.code_length_chips = BEIDOU_B1I_CODE_LENGTH_CHIPS;
    s.opt_acq_fs_sps = 10e6;
    s.code_at_resampled_rate = false;
    s.sampled_code = [](std::complex<float>* dest, size_t n, const Gnss_Synchro& synchro, int32_t fs, unsigned int) {
        beidou_b1i_code_gen_complex_sampled(own::span<std::complex<float>>(dest, n), synchro.PRN, fs, 0);
    };
    return s;
}
}  // namespace

BeidouB1iPcpsAcquisitionB200::BeidouB1iPcpsAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, uint32_t source_id)
    : PcpsAcquisitionB200Adapter(configuration, role, in_streams, out_streams, beidou_b1i(), source_id)
{
}
