/*
 * gps_l1_ca_pcps_acquisition_b200.cc — GPS L1 C/A: 1 ms code, 1.023 Mcps, code sampled at the (resampled) acquisition rate
 * by the reference's own generator (gps_l1_ca_pcps_acquisition.cc:52-53,151-170).
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "gps_l1_ca_pcps_acquisition_b200.h"
#include "signal_replicas.h"  // in the tree: GPS_L1_CA.h + gps_sdr_signal_replica.h

namespace
{
B200_Acq_Signal gps_l1_ca()
{
    B200_Acq_Signal s;
    s.implementation = "GPS_L1_CA_PCPS_Acquisition_B200";
    s.ms_per_code = 1;
    s.code_rate_cps = GPS_L1_CA_CODE_RATE_CPS;
    s.code_length_chips = GPS_L1_CA_CODE_LENGTH_CHIPS;
    s.opt_acq_fs_sps = GPS_L1_CA_OPT_ACQ_FS_SPS;
    s.code_at_resampled_rate = true;
    s.sampled_code = [](std::complex<float>* dest, size_t n, const Gnss_Synchro& synchro, int32_t fs, unsigned int) {
        gps_l1_ca_code_gen_complex_sampled(own::span<std::complex<float>>(dest, n), synchro.PRN, fs, 0);
    };
    return s;
}
}  // namespace

GpsL1CaPcpsAcquisitionB200::GpsL1CaPcpsAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams,
    unsigned int out_streams, uint32_t source_id)
    : PcpsAcquisitionB200Adapter(configuration, role, in_streams, out_streams, gps_l1_ca(), source_id)
{
}
