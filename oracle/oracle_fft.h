/* TEST INFRASTRUCTURE — CPU oracle, not product code. See oracle_fft_impl.inc. */
#ifndef ORACLE_FFT_H
#define ORACLE_FFT_H
#ifdef __cplusplus
extern "C"
{
#endif

    typedef struct offt_f offt_f;
    typedef struct offt_d offt_d;

    /* sign = -1 forward, +1 backward (unnormalised, FFTW convention). in != out.
       Buffers are interleaved {re,im}. A plan is NOT re-entrant (generic-radix scratch). */
    offt_f* offt_f_create(int n, int sign);
    void offt_f_destroy(offt_f* h);
    void offt_f_execute(const offt_f* h, const float* in, float* out);

    offt_d* offt_d_create(int n, int sign);
    void offt_d_destroy(offt_d* h);
    void offt_d_execute(const offt_d* h, const double* in, double* out);

    void oracle_fft_c64(const float* in, float* out, int n, int sign);
    void oracle_fft_c128(const double* in, double* out, int n, int sign);

#ifdef __cplusplus
}
#endif
#endif
