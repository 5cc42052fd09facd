/*
 * b200_multicorrelator.cc — see the header; mirrors cpu_multicorrelator.cc member by member.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_multicorrelator.h"
#include "b200_correlator_pool.h"

B200_Multicorrelator::~B200_Multicorrelator()
{
    if (m_pool_slot >= 0)
        {
            B200_Multicorrelator::free();
        }
}


// reference cpu_multicorrelator.cc:35-47
bool B200_Multicorrelator::init(int max_signal_length_samples, int n_correlators)
{
    if (max_signal_length_samples <= 0 || n_correlators <= 0) return false;
    if (m_pool_slot < 0)
        {
            m_pool_slot = B200_Correlator_Pool::instance().acquire_channel();
        }
    m_n_taps = n_correlators;
    m_capacity = max_signal_length_samples;
    return m_pool_slot >= 0;
}


// reference :50-60
bool B200_Multicorrelator::set_local_code_and_taps(int code_length_chips, const std::complex<float> *local_code_in, float *shifts_chips)
{
    m_io.code = local_code_in;
    m_io.tap_shifts = shifts_chips;
    m_code_len = code_length_chips;
    if (m_pool_slot < 0) return false;
    return B200_Correlator_Pool::instance().set_code(m_pool_slot, code_length_chips, reinterpret_cast<const float *>(local_code_in), true);
}


// reference :63-69
bool B200_Multicorrelator::set_input_output_vectors(std::complex<float> *corr_out, const std::complex<float> *sig_in)
{
    m_io.samples = sig_in;
    m_io.taps_out = corr_out;
    return true;
}


// reference :72-82
void B200_Multicorrelator::update_local_code(int correlator_length_samples, float rem_code_phase_chips, float code_phase_step_chips)
{
    (void)correlator_length_samples;
    (void)rem_code_phase_chips;
    (void)code_phase_step_chips;
}


// reference :85-99
bool B200_Multicorrelator::Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips,
    float code_phase_step_chips, int signal_length_samples)
{
    if (m_pool_slot < 0 || m_io.samples == nullptr || m_io.taps_out == nullptr || m_io.tap_shifts == nullptr || m_io.code == nullptr) return false;
    if (signal_length_samples <= 0 || signal_length_samples > m_capacity) return false;
    gnss_b200_mc_args a{};
    a.sig_in_c64 = reinterpret_cast<const float *>(m_io.samples);
    a.n_samples = static_cast<uint32_t>(signal_length_samples);
    a.rem_carr_phase_rad = rem_carrier_phase_in_rad;
    a.carr_step_rad = phase_step_rad;
    a.rem_code_phase_chips = rem_code_phase_chips;
    a.code_step_chips = code_phase_step_chips;
    return B200_Correlator_Pool::instance().correlate(m_pool_slot, a, m_io.tap_shifts, static_cast<uint32_t>(m_n_taps), 0U, m_io.taps_out);
}


// reference :102-116
bool B200_Multicorrelator::free()
{
    if (m_pool_slot >= 0)
        {
            B200_Correlator_Pool::instance().release_channel(m_pool_slot);
            m_pool_slot = -1;
        }
    return true;
}
