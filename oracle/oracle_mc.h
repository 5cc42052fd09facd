/* TEST INFRASTRUCTURE — CPU oracle, not product code. See oracle_mc.c. */
#ifndef ORACLE_MC_H
#define ORACLE_MC_H
#include <stdint.h>
#ifdef __cplusplus
extern "C"
{
#endif

    /* Cpu_Multicorrelator_Real_Codes: one Carrier_wipeoff_multicorrelator_resampler call.
       high_dyn_resampler = d_use_high_dynamics_resampler; six_arg_overload selects the
       overload without phase_rate_step_rad (always the plain rotator). */
    void oracle_mc_real(const float* sig_c64, int n, const float* code, int code_len, const float* shifts, int ntaps,
        float rem_carr_rad, float carr_step_rad, float carr_rate_rad, float rem_code_chips, float code_step_chips,
        float code_rate_chips, int high_dyn_resampler, int six_arg_overload, float* out_c64);

    /* Cpu_Multicorrelator (complex codes), 5-arg call */
    void oracle_mc_cplx(const float* sig_c64, int n, const float* code_c64, int code_len, const float* shifts, int ntaps,
        float rem_carr_rad, float carr_step_rad, float rem_code_chips, float code_step_chips, float* out_c64);

    /* wrapped chip index of every (tap, sample): idx_out[t*n + i] */
    void oracle_mc_chip_indices(int n, int code_len, const float* shifts, int ntaps, float rem_code_chips,
        float code_step_chips, float code_rate_chips, int high_dyn_resampler, int32_t* idx_out);

    /* serial loop over channels (timed CPU baseline: one thread per call) */
    void oracle_mc_batch(int n_ch, const float* const* sig_c64, const int* n_samples, const float* const* codes, const int* code_lens,
        const float* shifts, int ntaps, const float* args7, int high_dyn, float* out_c64);

#ifdef __cplusplus
}
#endif
#endif
