/* TEST INFRASTRUCTURE — CPU oracle, not product code. See oracle_codes.c. */
#ifndef ORACLE_CODES_H
#define ORACLE_CODES_H
#include <stdint.h>
#ifdef __cplusplus
extern "C"
{
#endif
    int oracle_samples_per_code(int fs, int chip_rate, int code_len);
    /* chips are +1/-1; out_c64 holds samples_per_code complex samples */
    void oracle_sample_code_lfsr_style(const int8_t* chips, int code_len, int chip_rate, int fs, int imag_code, int tc_via_double, float* out_c64);
    void oracle_sample_code_l5(const int8_t* chips, int code_len, int chip_rate, int fs, float* out_c64);
    void oracle_sample_code_e1_sinboc11(const int8_t* chips, int fs, float* out_c64);
#ifdef __cplusplus
}
#endif
#endif
