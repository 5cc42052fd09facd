/*
 * b200_multicorrelator.h — B200 drop-in for Cpu_Multicorrelator (complex local codes; used by the GLONASS and
 * TCP-connector tracking blocks). Reference: src/algorithms/tracking/libs/cpu_multicorrelator.h:37-58.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_B200_MULTICORRELATOR_H
#define GNSS_SDR_B200_MULTICORRELATOR_H

#include <complex>

class B200_Multicorrelator
{
public:
    B200_Multicorrelator() = default;
    ~B200_Multicorrelator();
    bool init(int max_signal_length_samples, int n_correlators);
    bool set_local_code_and_taps(int code_length_chips, const std::complex<float> *local_code_in, float *shifts_chips);
    bool set_input_output_vectors(std::complex<float> *corr_out, const std::complex<float> *sig_in);
    void update_local_code(int correlator_length_samples, float rem_code_phase_chips, float code_phase_step_chips);  // no-op, see real-codes class
    bool Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips, int signal_length_samples);
    bool free();

    B200_Multicorrelator(const B200_Multicorrelator &) = delete;
    B200_Multicorrelator &operator=(const B200_Multicorrelator &) = delete;

private:
    // Pointers the caller lends (the reference's classes keep raw pointers too and re-read them on every call): nothing here is owned.
    struct Borrowed
    {
        const std::complex<float> *samples = nullptr;  // input block of the current epoch
        const std::complex<float> *code = nullptr;   // local code table, one period
        std::complex<float> *taps_out = nullptr;       // one complex output per correlator tap
        float *tap_shifts = nullptr;                   // tap offsets [chips]
    } m_io;
    int m_code_len = 0;   // entries of the code table
    int m_n_taps = 0;     // correlator taps
    int m_capacity = 0;   // longest block init() was asked for [samples]
    int m_pool_slot = -1; // channel slot in B200_Correlator_Pool, -1 = not initialised
};

#endif  // GNSS_SDR_B200_MULTICORRELATOR_H
