/* Stand-in for src/core/system_parameters/gnss_synchro.h (see README.md in this directory). SPDX-License-Identifier: GPL-3.0-or-later */
#ifndef B200_STUB_GNSS_SYNCHRO_H
#define B200_STUB_GNSS_SYNCHRO_H
#include <cstdint>

class Gnss_Synchro
{
public:
    char System{};
    char Signal[3]{};
    uint32_t PRN{};
    int32_t Channel_ID{};
    double Acq_delay_samples{};
    double Acq_doppler_hz{};
    uint64_t Acq_samplestamp_samples{};
    uint32_t Acq_doppler_step{};
    bool Flag_valid_acquisition{};
    bool Flag_valid_symbol_output{};
    bool Flag_valid_word{};
    bool Flag_valid_pseudorange{};
};
#endif
