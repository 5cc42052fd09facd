/*
 * gnss_block_factory_b200.h — the B200 arms of GNSSBlockFactory::GetBlock / GetAcqBlock (gnss_block_factory.cc:908,1345), kept
 * in a file of their own so that the patch to the reference's factory is two `else if` lines (gnss_block_factory.patch).
 *
 *   Acquisition_1C.implementation = GPS_L1_CA_PCPS_Acquisition_B200
 *   Acquisition_1B.implementation = Galileo_E1_PCPS_Ambiguous_Acquisition_B200
 *   Acquisition_B1.implementation = BEIDOU_B1I_PCPS_Acquisition_B200
 *
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_GNSS_BLOCK_FACTORY_B200_H
#define GNSS_SDR_GNSS_BLOCK_FACTORY_B200_H

#include "acquisition_interface.h"
#include <memory>
#include <string>

class ConfigurationInterface;

// true for the implementation names above
bool IsB200AcquisitionImplementation(const std::string& implementation);

// The adapter for `implementation`, or nullptr when it is not a B200 acquisition. `role` is what GetAcqBlock received
// ("Acquisition_1C", or "Acquisition_1C3" for a per-channel section): the signal and, when present, the channel number in it
// select the RF_channel_ID (gnss_flowgraph.cc:1017-1018) that keeps channels of different front ends in different engine banks.
std::unique_ptr<AcquisitionInterface> GetB200AcquisitionBlock(const ConfigurationInterface* configuration, const std::string& role,
    const std::string& implementation, unsigned int in_streams, unsigned int out_streams);

#endif  // GNSS_SDR_GNSS_BLOCK_FACTORY_B200_H
