/*
 * galileo_e1_pcps_ambiguous_acquisition_b200.cc — Galileo E1: 4 ms code, sinBOC(1,1) or CBOC replica of the data (1B) or, with
 * Acquisition_1B.acquire_pilot, the pilot (1C) component (galileo_e1_pcps_ambiguous_acquisition.cc:50-51,62,150-197).
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "galileo_e1_pcps_ambiguous_acquisition_b200.h"
#include "configuration_interface.h"
#include "signal_replicas.h"  // in the tree: Galileo_E1.h + galileo_e1_signal_replica.h
#include <array>

namespace
{
B200_Acq_Signal galileo_e1(const ConfigurationInterface* configuration, const std::string& role)
{
    B200_Acq_Signal s;
    s.implementation = "Galileo_E1_PCPS_Ambiguous_Acquisition_B200";
    s.ms_per_code = 4;
    s.code_rate_cps = GALILEO_E1_CODE_CHIP_RATE_CPS;
    s.code_length_chips = GALILEO_E1_B_CODE_LENGTH_CHIPS;
    s.opt_acq_fs_sps = GALILEO_E1_OPT_ACQ_FS_SPS;
    s.code_at_resampled_rate = true;
    const bool acquire_pilot = configuration->property(role + ".acquire_pilot", false);
    s.sampled_code = [configuration, acquire_pilot](std::complex<float>* dest, size_t n, const Gnss_Synchro& synchro, int32_t fs, unsigned int channel) {
        // the reference reads the CBOC switch per channel, at set_local_code time (:152-153)
        const bool cboc = configuration->property("Acquisition" + std::to_string(channel) + ".cboc", false);
        const std::array<char, 3> component = acquire_pilot ? std::array<char, 3>{{'1', 'C', '\0'}} : std::array<char, 3>{{synchro.Signal[0], synchro.Signal[1], '\0'}};
        galileo_e1_code_gen_complex_sampled(own::span<std::complex<float>>(dest, n), component, cboc, synchro.PRN, fs, 0, false);
    };
    return s;
}
}  // namespace

GalileoE1PcpsAmbiguousAcquisitionB200::GalileoE1PcpsAmbiguousAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role,
    unsigned int in_streams, unsigned int out_streams, uint32_t source_id)
    : PcpsAcquisitionB200Adapter(configuration, role, in_streams, out_streams, galileo_e1(configuration, role), source_id)
{
}
