/*
 * galileo_e1_pcps_ambiguous_acquisition_b200.h — AcquisitionInterface adapter: GalileoE1PcpsAmbiguousAcquisition on the B200 engine.
 * Selected by Acquisition_XX.implementation=Galileo_E1_PCPS_Ambiguous_Acquisition_B200 (gnss_block_factory_b200.cc).
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_GALILEO_E1_PCPS_AMBIGUOUS_ACQUISITION_B200_H
#define GNSS_SDR_GALILEO_E1_PCPS_AMBIGUOUS_ACQUISITION_B200_H

#include "pcps_acquisition_b200_adapter.h"
#include <string>

class ConfigurationInterface;

class GalileoE1PcpsAmbiguousAcquisitionB200 : public PcpsAcquisitionB200Adapter
{
public:
    GalileoE1PcpsAmbiguousAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
        uint32_t source_id = 0);
};

#endif  // GNSS_SDR_GALILEO_E1_PCPS_AMBIGUOUS_ACQUISITION_B200_H
