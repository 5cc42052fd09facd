/*
 * b200_multicorrelator_real_codes.h — B200 drop-in for Cpu_Multicorrelator_Real_Codes
 * (reference: src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.h:37-61).
 *
 * Same method set, same borrowed-pointer conventions, same `bool` returns, so dll_pll_veml_tracking can hold it by
 * value in place of the CPU class (dll_pll_veml_tracking.h:90-91) with a one-line type change. The arithmetic runs in
 * the B200 engine (include/gnss_b200.h); there is no CPU fallback: without a CUDA device init() returns false.
 * SPDX-License-Identifier: GPL-3.0-or-later
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

    // Pointers the caller lends (the reference's classes keep raw pointers too and re-read them on every call): nothing here is owned.
    struct Borrowed
    {
        const std::complex<float> *samples = nullptr;  // input block of the current epoch
        const float *code = nullptr;                 // local code table, one period
        std::complex<float> *taps_out = nullptr;       // one complex output per correlator tap
        float *tap_shifts = nullptr;                   // tap offsets [chips]
    } m_io;
    int m_code_len = 0;   // entries of the code table
    int m_n_taps = 0;     // correlator taps
    int m_capacity = 0;   // longest block init() was asked for [samples]
    int m_pool_slot = -1; // channel slot in B200_Correlator_Pool, -1 = not initialised
    bool m_high_dynamics = true;  // set_high_dynamics_resampler(); the reference defaults to true as well
};

#endif  // GNSS_SDR_B200_MULTICORRELATOR_REAL_CODES_H
