// Stand-in for the reference's sampled-code generators (declarations only; the adapters stay on the reference's own
// generators inside the tree, the test harness links doubles):
//   src/algorithms/libs/gps_sdr_signal_replica.h:48, galileo_e1_signal_replica.h:67-75, beidou_b1i_signal_replica.h:50
// and for the signal constants of src/core/system_parameters/{GPS_L1_CA.h, Galileo_E1.h, Beidou_B1I.h}.
#ifndef B200_STUB_SIGNAL_REPLICAS_H
#define B200_STUB_SIGNAL_REPLICAS_H
#include <array>
#include <complex>
#include <cstdint>
#include <span>
namespace own = std;

void gps_l1_ca_code_gen_complex_sampled(own::span<std::complex<float>> dest, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift);
void galileo_e1_code_gen_complex_sampled(own::span<std::complex<float>> dest, const std::array<char, 3>& signal_id, bool cboc, uint32_t prn,
    int32_t sampling_freq, uint32_t chip_shift, bool secondary_flag);
void beidou_b1i_code_gen_complex_sampled(own::span<std::complex<float>> dest, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift);

constexpr double GPS_L1_CA_CODE_RATE_CPS = 1.023e6;
constexpr double GPS_L1_CA_CODE_LENGTH_CHIPS = 1023.0;
constexpr double GPS_L1_CA_OPT_ACQ_FS_SPS = 2000000.0;
constexpr double GALILEO_E1_CODE_CHIP_RATE_CPS = 1.023e6;
constexpr double GALILEO_E1_B_CODE_LENGTH_CHIPS = 4092.0;
constexpr double GALILEO_E1_OPT_ACQ_FS_SPS = 2000000.0;
constexpr double BEIDOU_B1I_CODE_RATE_CPS = 2.046e6;
constexpr double BEIDOU_B1I_CODE_LENGTH_CHIPS = 2046.0;
constexpr double BEIDOU_B1I_OPT_ACQ_FS_SPS = 2000000.0;
#endif
