/*
 * gnss_block_factory_b200.cc — see the header.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "gnss_block_factory_b200.h"
#include "beidou_b1i_pcps_acquisition_b200.h"
#include "configuration_interface.h"
#include "galileo_e1_pcps_ambiguous_acquisition_b200.h"
#include "gps_l1_ca_pcps_acquisition_b200.h"
#include <cctype>
#include <cstdint>

namespace
{
// "Acquisition_1C" -> ("1C", no channel); "Acquisition_1C3" -> ("1C", channel 3)
uint32_t rf_channel_of(const ConfigurationInterface* configuration, const std::string& role)
{
    const std::string prefix = "Acquisition_";
    if (role.compare(0, prefix.size(), prefix) != 0 || role.size() < prefix.size() + 2) return 0U;
    const std::string signal = role.substr(prefix.size(), 2);
    int32_t id = configuration->property("Channels_" + signal + ".RF_channel_ID", 0);
    const std::string tail = role.substr(prefix.size() + 2);
    if (!tail.empty() && std::isdigit(static_cast<unsigned char>(tail[0])) != 0)
        {
            id = configuration->property("Channel" + tail + ".RF_channel_ID", id);
        }
    return id < 0 ? 0U : static_cast<uint32_t>(id);
}
}  // namespace


bool IsB200AcquisitionImplementation(const std::string& implementation)
{
    return implementation == "GPS_L1_CA_PCPS_Acquisition_B200" || implementation == "Galileo_E1_PCPS_Ambiguous_Acquisition_B200" ||
           implementation == "BEIDOU_B1I_PCPS_Acquisition_B200";
}


std::unique_ptr<AcquisitionInterface> GetB200AcquisitionBlock(const ConfigurationInterface* configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams)
{
    const uint32_t source_id = rf_channel_of(configuration, role);
    if (implementation == "GPS_L1_CA_PCPS_Acquisition_B200")
        {
            return std::make_unique<GpsL1CaPcpsAcquisitionB200>(configuration, role, in_streams, out_streams, source_id);
        }
    if (implementation == "Galileo_E1_PCPS_Ambiguous_Acquisition_B200")
        {
            return std::make_unique<GalileoE1PcpsAmbiguousAcquisitionB200>(configuration, role, in_streams, out_streams, source_id);
        }
    if (implementation == "BEIDOU_B1I_PCPS_Acquisition_B200")
        {
            return std::make_unique<BeidouB1iPcpsAcquisitionB200>(configuration, role, in_streams, out_streams, source_id);
        }
    return nullptr;
}
