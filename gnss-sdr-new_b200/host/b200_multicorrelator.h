/*
 * b200_multicorrelator.h — B200 drop-in for Cpu_Multicorrelator (complex local codes; used by the GLONASS and
 * TCP-connector tracking blocks). Reference: src/algorithms/tracking/libs/cpu_multicorrelator.h:37-58.
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
    const std::complex<float> *d_sig_in{nullptr};
    const std::complex<float> *d_local_code_in{nullptr};
    std::complex<float> *d_corr_out{nullptr};
    float *d_shifts_chips{nullptr};
    int d_code_length_chips{0};
    int d_n_correlators{0};
    int d_max_signal_length_samples{0};
    int d_channel{-1};
};

#endif  // GNSS_SDR_B200_MULTICORRELATOR_H
