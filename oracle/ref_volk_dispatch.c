/*
 * TEST INFRASTRUCTURE — oracle/_ref only (needs /root/reference at build time).
 * C translation unit that includes the reference's volk_gnsssdr kernel headers
 * and exposes their `_generic` variants (a) under the refd_* names the C++ shim
 * forwards to and (b) under ref_k_* names for direct kernel-level checks from
 * Python. Also supplies the two allocator symbols the reference classes call
 * (plumbing only; the reference's lib/volk_gnsssdr_malloc.c needs generated
 * config headers).
 */
#define LV_HAVE_GENERIC 1
#ifdef REF_SIMD_AVX
/* the TIMED baseline flavour (_ref/libref_mc_avx.so): the reference's own AVX protokernels of the two kernels the real-code
   tracking correlator spends its time in — what volk_gnsssdr_profile selects on an AVX host */
#define LV_HAVE_AVX 1
#include <immintrin.h> /* volk_gnsssdr_common.h's bit256 union names __m256 when LV_HAVE_AVX is set */
#define REF_RESAMPLER_REAL volk_gnsssdr_32f_xn_resampler_32f_xn_u_avx
#define REF_ROTATOR_REAL volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_u_avx
#else
#define REF_RESAMPLER_REAL volk_gnsssdr_32f_xn_resampler_32f_xn_generic
#define REF_ROTATOR_REAL volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_generic
#endif
#include <volk_gnsssdr/volk_gnsssdr.h>
#include "volk_gnsssdr_16ic_convert_32fc.h"
#include "volk_gnsssdr_32f_index_max_32u.h"
#include "volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h"
#include "volk_gnsssdr_32f_xn_resampler_32f_xn.h"
#include "volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h"
#include "volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h"
#include "volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h"
#include "volk_gnsssdr_32fc_xn_resampler_32fc_xn.h"
#include "volk_gnsssdr_s32f_sincos_32fc.h"
#include <stdlib.h>

size_t volk_gnsssdr_get_alignment(void) { return 64; }
void* volk_gnsssdr_malloc(size_t size, size_t alignment)
{
    void* p = NULL;
    if (alignment < sizeof(void*)) alignment = sizeof(void*);
    if (posix_memalign(&p, alignment, size ? size : alignment) != 0) return NULL;
    return p;
}
void volk_gnsssdr_free(void* p) { free(p); }

void refd_resampler_real(float** result, const float* code, float rem, float step, float* shifts, unsigned int len, int ntaps, unsigned int n)
{
    REF_RESAMPLER_REAL(result, code, rem, step, shifts, len, ntaps, n);
}
void refd_resampler_real_hd(float** result, const float* code, float rem, float step, float rate, float* shifts, unsigned int len, int ntaps, unsigned int n)
{
    volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_generic(result, code, rem, step, rate, shifts, len, ntaps, n);
}
void refd_resampler_cplx(void** result, const void* code, float rem, float step, float* shifts, unsigned int len, int ntaps, unsigned int n)
{
    volk_gnsssdr_32fc_xn_resampler_32fc_xn_generic((lv_32fc_t**)result, (const lv_32fc_t*)code, rem, step, shifts, len, ntaps, n);
}
void refd_rotator_real(void* result, const void* in, const float* pi2, void* phase, const float** codes, int ntaps, unsigned int n)
{
    REF_ROTATOR_REAL((lv_32fc_t*)result, (const lv_32fc_t*)in, lv_cmake(pi2[0], pi2[1]), (lv_32fc_t*)phase, codes, ntaps, n);
}
void refd_rotator_real_hd(void* result, const void* in, const float* pi2, const float* pr2, void* phase, const float** codes, int ntaps, unsigned int n)
{
    volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn_generic((lv_32fc_t*)result, (const lv_32fc_t*)in, lv_cmake(pi2[0], pi2[1]), lv_cmake(pr2[0], pr2[1]), (lv_32fc_t*)phase, codes, ntaps, n);
}
void refd_rotator_cplx(void* result, const void* in, const float* pi2, void* phase, const void** codes, int ntaps, unsigned int n)
{
    volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_generic((lv_32fc_t*)result, (const lv_32fc_t*)in, lv_cmake(pi2[0], pi2[1]), (lv_32fc_t*)phase, (const lv_32fc_t**)codes, ntaps, n);
}

/* direct kernel access for Python */
void ref_k_sincos(float* out_c64, float phase_inc, float* phase, unsigned int n)
{
    volk_gnsssdr_s32f_sincos_32fc_generic((lv_32fc_t*)out_c64, phase_inc, phase, n);
}
uint32_t ref_k_index_max(const float* src, uint32_t n)
{
    uint32_t t = 0;
    volk_gnsssdr_32f_index_max_32u_generic(&t, src, n);
    return t;
}
