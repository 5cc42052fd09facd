// Stand-in for src/algorithms/libs/gnss_sdr_flags.h (gflags): the one flag the acquisition adapters read.
#ifndef B200_STUB_GNSS_SDR_FLAGS_H
#define B200_STUB_GNSS_SDR_FLAGS_H
#include <cstdint>
extern int32_t FLAGS_doppler_max;  // --doppler_max command-line override (0 = not given)
#endif
