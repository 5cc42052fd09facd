/*
 * b200_multicorrelator_real_codes.h — B200 drop-in for Cpu_Multicorrelator_Real_Codes
 * (reference: src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.h:37-61).
 *
 * Same method set, same borrowed-pointer conventions, same `bool` returns, so dll_pll_veml_tracking can hold it by
 * value in place of the CPU class (dll_pll_veml_tracking.h:90-91) with a one-line type change. The arithmetic runs in
 * the B200 engine (include/gnss_b200.h); there is no CPU fallback: without a CUDA device init() returns false.
 */
#ifndef GNSS_SDR_B200_MULTICORRELATOR_REAL_CODES_H
#define GNSS_SDR_B200_MULTICORRELATOR_REAL_CODES_H

#include <complex>

class B200_Multicorrelator_Real_Codes
{
public:
    B200_Multicorrelator_Real_Codes() = default;
    ~B200_Multicorrelator_Real_Codes();
    void set_high_dynamics_resampler(bool use_high_dynamics_resampler);
    bool init(int max_signal_length_samples, int n_correlators);
    bool set_local_code_and_taps(int code_length_chips, const float *local_code_in, float *shifts_chips);
    bool set_input_output_vectors(std::complex<float> *corr_out, const std::complex<float> *sig_in);
    // Kept for source compatibility: the resampled local codes are never materialised on the B200 (they are gathered
    // inside the correlation kernel), so this is a no-op that only validates its arguments.
    void update_local_code(int correlator_length_samples, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips = 0.0);
    bool Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad, float phase_rate_step_rad, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples);
    bool Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples);
    bool free();

    B200_Multicorrelator_Real_Codes(const B200_Multicorrelator_Real_Codes &) = delete;
    B200_Multicorrelator_Real_Codes &operator=(const B200_Multicorrelator_Real_Codes &) = delete;

private:
    bool run(float rem_carrier_phase_in_rad, float phase_step_rad, float phase_rate_step_rad, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples, unsigned flags);

    const std::complex<float> *d_sig_in{nullptr};
    const float *d_local_code_in{nullptr};
    std::complex<float> *d_corr_out{nullptr};
    float *d_shifts_chips{nullptr};
    int d_code_length_chips{0};
    int d_n_correlators{0};
    int d_max_signal_length_samples{0};
    int d_channel{-1};  // slot in B200_Correlator_Pool
    bool d_use_high_dynamics_resampler{true};
};

#endif  // GNSS_SDR_B200_MULTICORRELATOR_REAL_CODES_H
