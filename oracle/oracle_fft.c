/*
 * TEST INFRASTRUCTURE — CPU oracle, not product code.
 * Instantiates the any-length DFT of oracle_fft_impl.inc in float32 (the
 * precision the reference's FFTW3f plans run in) and in float64 (error bounds).
 */
#include "oracle_fft.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define REAL float
#define SFX _f
#include "oracle_fft_impl.inc"
#undef REAL
#undef SFX

#define REAL double
#define SFX _d
#include "oracle_fft_impl.inc"
#undef REAL
#undef SFX

struct offt_f
{
    struct offt_plan_f* p;
};
struct offt_d
{
    struct offt_plan_d* p;
};

offt_f* offt_f_create(int n, int sign)
{
    offt_f* h = (offt_f*)malloc(sizeof(*h));
    h->p = offt_plan_create_f(n, sign);
    return h;
}
void offt_f_destroy(offt_f* h)
{
    if (!h) return;
    offt_plan_destroy_f(h->p);
    free(h);
}
void offt_f_execute(const offt_f* h, const float* in, float* out)
{
    offt_execute_f(h->p, (const ocpx_f*)in, (ocpx_f*)out);
}

offt_d* offt_d_create(int n, int sign)
{
    offt_d* h = (offt_d*)malloc(sizeof(*h));
    h->p = offt_plan_create_d(n, sign);
    return h;
}
void offt_d_destroy(offt_d* h)
{
    if (!h) return;
    offt_plan_destroy_d(h->p);
    free(h);
}
void offt_d_execute(const offt_d* h, const double* in, double* out)
{
    offt_execute_d(h->p, (const ocpx_d*)in, (ocpx_d*)out);
}

/* one-shot helpers for the Python harness */
void oracle_fft_c64(const float* in, float* out, int n, int sign)
{
    offt_f* h = offt_f_create(n, sign);
    offt_f_execute(h, in, out);
    offt_f_destroy(h);
}
void oracle_fft_c128(const double* in, double* out, int n, int sign)
{
    offt_d* h = offt_d_create(n, sign);
    offt_d_execute(h, in, out);
    offt_d_destroy(h);
}
