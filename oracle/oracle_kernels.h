/*
 * TEST INFRASTRUCTURE — CPU oracle, not product code.
 *
 * Scalar restatement of the reference's `_generic` volk_gnsssdr kernels that sit
 * on the hot path, plus the four trivial upstream-VOLK element-wise ops the
 * reference calls (VOLK itself is third-party, not vendored, not installed).
 * Paths below are relative to
 *   /root/reference/src/algorithms/libs/volk_gnsssdr_module/volk_gnsssdr/kernels/volk_gnsssdr/
 *
 * When built with -DORACLE_USE_REF (the oracle/_ref build, only possible where
 * /root/reference exists) the k_* names map to the reference's own headers
 * instead, so the same glue runs over the real kernels; tests require the two
 * builds to agree bit for bit.
 *
 * All arithmetic is IEEE binary32, evaluated left to right with no FMA
 * contraction (build flags: -ffp-contract=off, no -mfma), as the reference's
 * generic kernels are built.
 */
#ifndef ORACLE_KERNELS_H
#define ORACLE_KERNELS_H

#include <complex.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef float complex ocf;

#ifdef ORACLE_USE_REF

#define LV_HAVE_GENERIC 1
#include <volk_gnsssdr/volk_gnsssdr.h> /* oracle/shim */
#include "volk_gnsssdr_16ic_convert_32fc.h"
#include "volk_gnsssdr_32f_index_max_32u.h"
#include "volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h"
#include "volk_gnsssdr_32f_xn_resampler_32f_xn.h"
#include "volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h"
#include "volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h"
#include "volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h"
#include "volk_gnsssdr_32fc_xn_resampler_32fc_xn.h"
#include "volk_gnsssdr_s32f_sincos_32fc.h"

#define k_sincos volk_gnsssdr_s32f_sincos_32fc_generic
#define k_index_max volk_gnsssdr_32f_index_max_32u_generic
#define k_convert_16ic volk_gnsssdr_16ic_convert_32fc_generic
#define k_resampler_real volk_gnsssdr_32f_xn_resampler_32f_xn_generic
#define k_resampler_real_hd volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_generic
#define k_resampler_cplx volk_gnsssdr_32fc_xn_resampler_32fc_xn_generic
#define k_rotator_dot_real volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_generic
#define k_rotator_dot_real_hd volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn_generic
#define k_rotator_dot_cplx volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_generic
#define ORACLE_KERNEL_FLAVOUR "reference-generic"

#else /* own restatement */

#define ORACLE_KERNEL_FLAVOUR "restated"

/* volk_gnsssdr_s32f_sincos_32fc.h:390-400 — running float32 phase sum */
static inline void k_sincos(ocf* out, const float phase_inc, float* phase, unsigned int n)
{
    float ph = *phase;
    for (unsigned int i = 0; i < n; i++)
        {
            out[i] = CMPLXF(cosf(ph), sinf(ph));
            ph += phase_inc;
        }
    *phase = ph;
}

/* volk_gnsssdr_32f_index_max_32u.h:446-465 — first index of the strict maximum */
static inline void k_index_max(uint32_t* target, const float* src, uint32_t n)
{
    if (n == 0) return;
    float best = src[0];
    uint32_t at = 0;
    for (uint32_t i = 1; i < n; i++)
        {
            if (src[i] > best)
                {
                    best = src[i];
                    at = i;
                }
        }
    *target = at;
}

/* volk_gnsssdr_16ic_convert_32fc.h:49-56 */
static inline void k_convert_16ic(ocf* out, const int16_t* in_iq, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) out[i] = CMPLXF((float)in_iq[2 * i], (float)in_iq[2 * i + 1]);
}

static inline int k_wrap_chip(int idx, unsigned int len)
{
    /* volk_gnsssdr_32f_xn_resampler_32f_xn.h:75-76 (same integer semantics: the
       division and the final modulo are unsigned because `len` is unsigned) */
    if (idx < 0) idx += (int)len * (abs(idx) / len + 1);
    return (int)((unsigned int)idx % len);
}

/* volk_gnsssdr_32f_xn_resampler_32f_xn.h:63-80 */
static inline void k_resampler_real(float** result, const float* code, float rem, float step, float* shifts, unsigned int len, int ntaps, unsigned int n_pts)
{
    for (int t = 0; t < ntaps; t++)
        for (unsigned int n = 0; n < n_pts; n++)
            {
                int idx = (int)floor(step * (float)n + shifts[t] - rem);
                result[t][n] = code[k_wrap_chip(idx, len)];
            }
}

/* volk_gnsssdr_32fc_xn_resampler_32fc_xn.h:60-80 */
static inline void k_resampler_cplx(ocf** result, const ocf* code, float rem, float step, float* shifts, unsigned int len, int ntaps, unsigned int n_pts)
{
    for (int t = 0; t < ntaps; t++)
        for (unsigned int n = 0; n < n_pts; n++)
            {
                int idx = (int)floor(step * (float)n + shifts[t] - rem);
                result[t][n] = code[k_wrap_chip(idx, len)];
            }
}

/* volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h:67-91 — tap 0 carries the
   rate term, taps 1.. are circular sample shifts of tap 0 */
static inline void k_resampler_real_hd(float** result, const float* code, float rem, float step, float rate, float* shifts, unsigned int len, int ntaps, unsigned int n_pts)
{
    for (unsigned int n = 0; n < n_pts; n++)
        {
            int idx = (int)floor(step * (float)n + rate * (float)(n * n) + shifts[0] - rem);
            result[0][n] = code[k_wrap_chip(idx, len)];
        }
    unsigned int shift_samples = 0;
    for (int t = 1; t < ntaps; t++)
        {
            shift_samples += (int)round((shifts[t] - shifts[t - 1]) / step);
            memcpy(&result[t][0], &result[0][shift_samples], (n_pts - shift_samples) * sizeof(float));
            memcpy(&result[t][n_pts - shift_samples], &result[0][0], shift_samples * sizeof(float));
        }
}

/* volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h:66-98 */
static inline void k_rotator_dot_real(ocf* result, const ocf* in, const ocf phase_inc, ocf* phase, const float** codes, int ntaps, unsigned int n_pts)
{
    for (int t = 0; t < ntaps; t++) result[t] = 0.0f;
    for (unsigned int n = 0; n < n_pts; n++)
        {
            const ocf wiped = in[n] * (*phase);
            if (n % 256 == 0) (*phase) /= hypotf(crealf(*phase), cimagf(*phase));
            (*phase) *= phase_inc;
            for (int t = 0; t < ntaps; t++)
                {
                    const ocf prod = wiped * codes[t][n];
                    result[t] += prod;
                }
        }
}

/* volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h:67-99 */
static inline void k_rotator_dot_cplx(ocf* result, const ocf* in, const ocf phase_inc, ocf* phase, const ocf** codes, int ntaps, unsigned int n_pts)
{
    for (int t = 0; t < ntaps; t++) result[t] = 0.0f;
    for (unsigned int n = 0; n < n_pts; n++)
        {
            const ocf wiped = in[n] * (*phase);
            if (n % 256 == 0) (*phase) /= hypotf(crealf(*phase), cimagf(*phase));
            (*phase) *= phase_inc;
            for (int t = 0; t < ntaps; t++)
                {
                    const ocf prod = wiped * codes[t][n];
                    result[t] += prod;
                }
        }
}

/* volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h:68-109 */
static inline void k_rotator_dot_real_hd(ocf* result, const ocf* in, const ocf phase_inc, const ocf phase_inc_rate, ocf* phase, const float** codes, int ntaps, unsigned int n_pts)
{
    ocf doppler = *phase;
    for (int t = 0; t < ntaps; t++) result[t] = 0.0f;
    for (unsigned int n = 0; n < n_pts; n++)
        {
            if (n % 256 == 0) (*phase) /= hypotf(crealf(*phase), cimagf(*phase));
            const ocf wiped = in[n] * (*phase);
            doppler *= phase_inc;
            ocf rate = cpowf(phase_inc_rate, CMPLXF((float)(n * n), 0.0f));
            rate /= hypotf(crealf(rate), cimagf(rate));
            *phase = doppler * rate;
            for (int t = 0; t < ntaps; t++) result[t] += (wiped * codes[t][n]);
        }
}

#endif /* ORACLE_USE_REF */

/* ---- upstream VOLK element-wise ops (third-party, unpinned version; call sites
 *      pcps_acquisition.cc:208,658,665,674,679) — restated from their definitions ---- */
static inline void v_multiply(ocf* out, const ocf* a, const ocf* b, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++)
        {
            const float ar = crealf(a[i]), ai = cimagf(a[i]), br = crealf(b[i]), bi = cimagf(b[i]);
            out[i] = CMPLXF(ar * br - ai * bi, ar * bi + ai * br);
        }
}
static inline void v_conjugate(ocf* out, const ocf* a, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) out[i] = conjf(a[i]);
}
static inline void v_mag_squared(float* out, const ocf* a, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++)
        {
            const float r = crealf(a[i]), q = cimagf(a[i]);
            out[i] = r * r + q * q;
        }
}
static inline void v_add(float* out, const float* a, const float* b, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) out[i] = a[i] + b[i];
}

#endif
