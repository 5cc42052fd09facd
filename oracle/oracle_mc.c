/*
 * TEST INFRASTRUCTURE — CPU oracle, not product code.
 *
 * Restatement of the tracking multicorrelator of the reference:
 *   Cpu_Multicorrelator_Real_Codes::update_local_code                         cpu_multicorrelator_real_codes.cc:75-100
 *   Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler :103-126 (7-arg), :129-144 (6-arg)
 *   Cpu_Multicorrelator::Carrier_wipeoff_multicorrelator_resampler            cpu_multicorrelator.cc:87-101
 * (paths under src/algorithms/tracking/libs/) over the kernels of oracle_kernels.h.
 * The reference's own unit tests assert nothing numeric for these classes, so
 * this oracle is pinned by bit-identity with the reference's generic kernels
 * (oracle/_ref build) rather than by golden vectors.
 */
#include "oracle_mc.h"
#include "oracle_kernels.h"

void oracle_mc_real(const float* sig_c64, int n, const float* code, int code_len, const float* shifts, int ntaps,
    float rem_carr_rad, float carr_step_rad, float carr_rate_rad, float rem_code_chips, float code_step_chips,
    float code_rate_chips, int high_dyn_resampler, int six_arg_overload, float* out_c64)
{
    float** res = (float**)malloc(sizeof(float*) * ntaps);
    for (int t = 0; t < ntaps; t++) res[t] = (float*)malloc(sizeof(float) * (size_t)n);
    float* sh = (float*)malloc(sizeof(float) * ntaps);
    memcpy(sh, shifts, sizeof(float) * ntaps);
    /* update_local_code :75-100 */
    if (high_dyn_resampler)
        k_resampler_real_hd(res, code, rem_code_chips, code_step_chips, code_rate_chips, sh, (unsigned)code_len, ntaps, (unsigned)n);
    else
        k_resampler_real(res, code, rem_code_chips, code_step_chips, sh, (unsigned)code_len, ntaps, (unsigned)n);
    /* :114-115 */
    ocf phase = CMPLXF(cosf(rem_carr_rad), -sinf(rem_carr_rad));
    /* std::exp(lv_32fc_t(0, -step)) → cexpf */
    const ocf inc = cexpf(CMPLXF(0.0f, -carr_step_rad));
    if (high_dyn_resampler && !six_arg_overload)
        {
            const ocf inc_rate = cexpf(CMPLXF(0.0f, -carr_rate_rad));
            k_rotator_dot_real_hd((ocf*)out_c64, (const ocf*)sig_c64, inc, inc_rate, &phase, (const float**)res, ntaps, (unsigned)n);
        }
    else
        {
            k_rotator_dot_real((ocf*)out_c64, (const ocf*)sig_c64, inc, &phase, (const float**)res, ntaps, (unsigned)n);
        }
    for (int t = 0; t < ntaps; t++) free(res[t]);
    free(res);
    free(sh);
}

void oracle_mc_cplx(const float* sig_c64, int n, const float* code_c64, int code_len, const float* shifts, int ntaps,
    float rem_carr_rad, float carr_step_rad, float rem_code_chips, float code_step_chips, float* out_c64)
{
    ocf** res = (ocf**)malloc(sizeof(ocf*) * ntaps);
    for (int t = 0; t < ntaps; t++) res[t] = (ocf*)malloc(sizeof(ocf) * (size_t)n);
    float* sh = (float*)malloc(sizeof(float) * ntaps);
    memcpy(sh, shifts, sizeof(float) * ntaps);
    k_resampler_cplx(res, (const ocf*)code_c64, rem_code_chips, code_step_chips, sh, (unsigned)code_len, ntaps, (unsigned)n);
    ocf phase = CMPLXF(cosf(rem_carr_rad), -sinf(rem_carr_rad));
    const ocf inc = cexpf(CMPLXF(0.0f, -carr_step_rad));
    k_rotator_dot_cplx((ocf*)out_c64, (const ocf*)sig_c64, inc, &phase, (const ocf**)res, ntaps, (unsigned)n);
    for (int t = 0; t < ntaps; t++) free(res[t]);
    free(res);
    free(sh);
}

void oracle_mc_chip_indices(int n, int code_len, const float* shifts, int ntaps, float rem_code_chips,
    float code_step_chips, float code_rate_chips, int high_dyn_resampler, int32_t* idx_out)
{
    /* resample an identity table: the result IS the wrapped chip index */
    float* ident = (float*)malloc(sizeof(float) * (size_t)code_len);
    for (int i = 0; i < code_len; i++) ident[i] = (float)i;
    float** res = (float**)malloc(sizeof(float*) * ntaps);
    for (int t = 0; t < ntaps; t++) res[t] = (float*)malloc(sizeof(float) * (size_t)n);
    float* sh = (float*)malloc(sizeof(float) * ntaps);
    memcpy(sh, shifts, sizeof(float) * ntaps);
    if (high_dyn_resampler)
        k_resampler_real_hd(res, ident, rem_code_chips, code_step_chips, code_rate_chips, sh, (unsigned)code_len, ntaps, (unsigned)n);
    else
        k_resampler_real(res, ident, rem_code_chips, code_step_chips, sh, (unsigned)code_len, ntaps, (unsigned)n);
    for (int t = 0; t < ntaps; t++)
        for (int i = 0; i < n; i++) idx_out[(size_t)t * n + i] = (int32_t)res[t][i];
    for (int t = 0; t < ntaps; t++) free(res[t]);
    free(res);
    free(sh);
    free(ident);
}

void oracle_mc_batch(int n_ch, const float* const* sig_c64, const int* n_samples, const float* const* codes, const int* code_lens,
    const float* shifts, int ntaps, const float* args7 /* [n_ch][7]: rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate, unused */,
    int high_dyn, float* out_c64)
{
    for (int c = 0; c < n_ch; c++)
        {
            const float* a = args7 + (size_t)c * 7;
            oracle_mc_real(sig_c64[c], n_samples[c], codes[c], code_lens[c], shifts + (size_t)c * ntaps, ntaps, a[0], a[1], a[2], a[3], a[4], a[5],
                high_dyn, 0, out_c64 + (size_t)c * ntaps * 2);
        }
}
