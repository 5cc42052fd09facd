/*
 * TEST INFRASTRUCTURE — CPU oracle, not product code.
 *
 * Sampling of 1-chip-per-sample PRN codes to the receiver rate, as the
 * acquisition adapters of the reference do before set_local_code()
 * (paths under src/algorithms/libs/):
 *   gps_l1_ca_code_gen_complex_sampled   gps_sdr_signal_replica.cc:136-176  (imaginary code)
 *   beidou_b1i_code_gen_complex_sampled  beidou_b1i_signal_replica.cc:130-170 (real code)
 *   gps_l5i_code_gen_complex_sampled     gps_l5_signal_replica.cc:193-227   (real code, ceil)
 *   galileo_e1_code_gen_complex_sampled  galileo_e1_signal_replica.cc:143-228 (sinBOC(1,1), cboc=false)
 *   resampler                            gnss_signal_replica.cc:257-272
 * The chips themselves (ICD data / LFSR outputs) are NOT regenerated here: they
 * come from tests/golden/prn_chips.npz, written by tools/make_golden_codes.py
 * from the reference's own generators. Tests check these samplers against the
 * reference generators bit for bit where /root/reference is present.
 */
#include "oracle_codes.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

static int32_t aux_ceil(float x) { return (int32_t)(int64_t)(x + 1); }

int oracle_samples_per_code(int fs, int chip_rate, int code_len)
{
    return (int)((double)fs / ((double)chip_rate / (double)code_len));
}

void oracle_sample_code_lfsr_style(const int8_t* chips, int code_len, int chip_rate, int fs, int imag_code, int tc_via_double, float* out_c64)
{
    /* GPS: `1.0F / float(rate)`; BeiDou: `1.0 / float(rate)` (double division, then narrowed) */
    const float tc = tc_via_double ? (float)(1.0 / (float)chip_rate) : 1.0f / (float)chip_rate;
    const int spc = oracle_samples_per_code(fs, chip_rate, code_len);
    const float ts = 1.0f / (float)fs;
    for (int i = 0; i < spc; i++)
        {
            const float aux = (ts * ((float)i + 1)) / tc;
            int idx = aux_ceil(aux) - 1;
            if (i == spc - 1) idx = code_len - 1;
            const float v = (float)chips[idx];
            out_c64[2 * i] = imag_code ? 0.0f : v;
            out_c64[2 * i + 1] = imag_code ? v : 0.0f;
        }
}

void oracle_sample_code_l5(const int8_t* chips, int code_len, int chip_rate, int fs, float* out_c64)
{
    const float tc = (float)(1.0 / (float)chip_rate);
    const int spc = oracle_samples_per_code(fs, chip_rate, code_len);
    const float ts = 1.0f / (float)fs;
    for (int i = 0; i < spc; i++)
        {
            int idx = (int32_t)ceilf(ts * (float)(i + 1.0f) / tc) - 1;
            if (i == spc - 1) idx = code_len - 1;
            out_c64[2 * i] = (float)chips[idx];
            out_c64[2 * i + 1] = 0.0f;
        }
}

void oracle_sample_code_e1_sinboc11(const int8_t* chips, int fs, float* out_c64)
{
    const int code_len = 4092, chip_rate = 1023000, spchip = 2;
    const int len2 = spchip * code_len;
    const int spc = (int)((double)fs / ((double)chip_rate / code_len));
    float* boc = (float*)malloc(sizeof(float) * len2);
    for (int i = 0; i < code_len; i++)
        {
            boc[2 * i] = (float)chips[i];
            boc[2 * i + 1] = (float)(-chips[i]);
        }
    const float* src = boc;
    float* res = NULL;
    if (fs != spchip * chip_rate)
        {
            res = (float*)malloc(sizeof(float) * spc);
            const float fs_in = (float)(spchip * chip_rate);
            const float t_out = 1.0f / (float)fs;
            for (int i = 0; i < spc - 1; i++)
                {
                    const float aux = (t_out * ((float)i + 1.0f)) * fs_in;
                    res[i] = boc[aux_ceil(aux) - 1];
                }
            res[spc - 1] = boc[len2 - 1];
            src = res;
        }
    /* chip_shift = 0 → delay 0 */
    for (int i = 0; i < spc; i++)
        {
            out_c64[2 * i] = src[i];
            out_c64[2 * i + 1] = 0.0f;
        }
    free(boc);
    free(res);
}
