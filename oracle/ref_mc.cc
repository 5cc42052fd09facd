/*
 * TEST INFRASTRUCTURE — oracle/_ref only (needs /root/reference at build time).
 * Drives the reference's own Cpu_Multicorrelator_Real_Codes / Cpu_Multicorrelator
 * classes (src/algorithms/tracking/libs/cpu_multicorrelator{,_real_codes}.cc,
 * compiled unmodified) so that the oracle's restatement of the class glue can be
 * checked against the real thing. The volk_gnsssdr dispatcher names the classes
 * call are provided by ref_volk_dispatch.c (→ the reference's _generic kernels).
 */
#include "cpu_multicorrelator.h"
#include "cpu_multicorrelator_real_codes.h"
#include <complex>
#include <vector>

extern "C"
{
    void ref_mc_real(const float* sig_c64, int n, const float* code, int code_len, const float* shifts, int ntaps,
        float rem_carr, float carr_step, float carr_rate, float rem_code, float code_step, float code_rate,
        int high_dyn, int six_arg, float* out_c64)
    {
        Cpu_Multicorrelator_Real_Codes mc;
        std::vector<float> sh(shifts, shifts + ntaps);
        mc.init(2 * n, ntaps);
        mc.set_high_dynamics_resampler(high_dyn != 0);
        mc.set_local_code_and_taps(code_len, code, sh.data());
        mc.set_input_output_vectors(reinterpret_cast<std::complex<float>*>(out_c64), reinterpret_cast<const std::complex<float>*>(sig_c64));
        if (six_arg)
            mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr, carr_step, rem_code, code_step, code_rate, n);
        else
            mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate, n);
        mc.free();
    }

    void ref_mc_cplx(const float* sig_c64, int n, const float* code_c64, int code_len, const float* shifts, int ntaps,
        float rem_carr, float carr_step, float rem_code, float code_step, float* out_c64)
    {
        Cpu_Multicorrelator mc;
        std::vector<float> sh(shifts, shifts + ntaps);
        mc.init(2 * n, ntaps);
        mc.set_local_code_and_taps(code_len, reinterpret_cast<const std::complex<float>*>(code_c64), sh.data());
        mc.set_input_output_vectors(reinterpret_cast<std::complex<float>*>(out_c64), reinterpret_cast<const std::complex<float>*>(sig_c64));
        mc.Carrier_wipeoff_multicorrelator_resampler(rem_carr, carr_step, rem_code, code_step, n);
        mc.free();
    }
}
