/*
 * b200_multicorrelator_real_codes.cc — see the header.
 * Each method notes the reference member it mirrors (cpu_multicorrelator_real_codes.cc).
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_multicorrelator_real_codes.h"
#include "b200_correlator_pool.h"

B200_Multicorrelator_Real_Codes::~B200_Multicorrelator_Real_Codes()
{
    // reference :27-33 — the destructor releases whatever init() took
    if (m_pool_slot >= 0)
        {
            B200_Multicorrelator_Real_Codes::free();
        }
}


// reference :63-66
void B200_Multicorrelator_Real_Codes::set_high_dynamics_resampler(bool use_high_dynamics_resampler)
{
    m_high_dynamics = use_high_dynamics_resampler;
}


// reference :36-49 — there the per-tap resampled code buffers are allocated; here a channel of the shared engine.
bool B200_Multicorrelator_Real_Codes::init(int max_signal_length_samples, int n_correlators)
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


// reference :52-60 — both pointers are borrowed; the code is additionally uploaded to the device once.
bool B200_Multicorrelator_Real_Codes::set_local_code_and_taps(int code_length_chips, const float *local_code_in, float *shifts_chips)
{
    m_io.code = local_code_in;
    m_io.tap_shifts = shifts_chips;
    m_code_len = code_length_chips;
    if (m_pool_slot < 0) return false;
    return B200_Correlator_Pool::instance().set_code(m_pool_slot, code_length_chips, local_code_in, false);
}


// reference :69-75
bool B200_Multicorrelator_Real_Codes::set_input_output_vectors(std::complex<float> *corr_out, const std::complex<float> *sig_in)
{
    m_io.samples = sig_in;
    m_io.taps_out = corr_out;
    return true;
}


// reference :78-100
void B200_Multicorrelator_Real_Codes::update_local_code(int correlator_length_samples, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips)
{
    (void)correlator_length_samples;
    (void)rem_code_phase_chips;
    (void)code_phase_step_chips;
    (void)code_phase_rate_step_chips;
}


bool B200_Multicorrelator_Real_Codes::run(float rem_carrier_phase_in_rad, float phase_step_rad, float phase_rate_step_rad, float rem_code_phase_chips,
    float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples, unsigned flags)
{
    if (m_pool_slot < 0 || m_io.samples == nullptr || m_io.taps_out == nullptr || m_io.tap_shifts == nullptr || m_io.code == nullptr) return false;
    if (signal_length_samples <= 0 || signal_length_samples > m_capacity) return false;
    gnss_b200_mc_args a{};
    a.sig_in_c64 = reinterpret_cast<const float *>(m_io.samples);
    a.stream_offset = 0;
    a.n_samples = static_cast<uint32_t>(signal_length_samples);
    a.rem_carr_phase_rad = rem_carrier_phase_in_rad;
    a.carr_step_rad = phase_step_rad;
    a.carr_rate_rad = phase_rate_step_rad;
    a.rem_code_phase_chips = rem_code_phase_chips;
    a.code_step_chips = code_phase_step_chips;
    a.code_rate_chips = code_phase_rate_step_chips;
    return B200_Correlator_Pool::instance().correlate(m_pool_slot, a, m_io.tap_shifts, static_cast<uint32_t>(m_n_taps), flags, m_io.taps_out);
}


// reference :103-126 — 7-argument form: high-dynamics rotator when the high-dynamics resampler is selected
bool B200_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad,
    float phase_rate_step_rad, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples)
{
    const unsigned flags = m_high_dynamics ? static_cast<unsigned>(GNSS_B200_MC_HIGH_DYN) : 0U;
    return run(rem_carrier_phase_in_rad, phase_step_rad, phase_rate_step_rad, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips,
        signal_length_samples, flags);
}


// reference :129-144 — 6-argument form: always the plain rotator
bool B200_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad,
    float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples)
{
    const unsigned flags = (m_high_dynamics ? static_cast<unsigned>(GNSS_B200_MC_HIGH_DYN) : 0U) | GNSS_B200_MC_SIX_ARG;
    return run(rem_carrier_phase_in_rad, phase_step_rad, 0.0F, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips,
        signal_length_samples, flags);
}


// reference :147-167
bool B200_Multicorrelator_Real_Codes::free()
{
    if (m_pool_slot >= 0)
        {
            B200_Correlator_Pool::instance().release_channel(m_pool_slot);
            m_pool_slot = -1;
        }
    return true;
}
