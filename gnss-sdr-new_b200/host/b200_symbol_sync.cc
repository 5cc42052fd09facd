/*
 * b200_symbol_sync.cc — see the header. ":NNN" = line of dll_pll_veml_tracking.cc the statement follows.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_symbol_sync.h"
#include <cstdlib>

void B200_Symbol_Sync::configure(bool secondary, const std::string& secondary_code_string, uint32_t secondary_code_length, uint32_t symbols_per_bit,
    const std::string& data_secondary_code_string, uint32_t data_secondary_code_length, bool interchange_iq)
{
    d_secondary = secondary;
    d_secondary_code_string = secondary_code_string;
    d_secondary_code_length = secondary_code_length;
    d_symbols_per_bit = symbols_per_bit;
    d_data_secondary_code_string = data_secondary_code_string;
    d_data_secondary_code_length = data_secondary_code_length;
    d_interchange_iq = interchange_iq;
    reset();
}


void B200_Symbol_Sync::reset()
{
    d_Prompt_circular_buffer.clear();
    d_P_data_accu = std::complex<float>(0.0F, 0.0F);
    d_current_symbol = 0;
    d_current_data_symbol = 0;
    d_locked = false;
    d_Flag_PLL_180_deg_phase_locked = false;
}


// :923-966
// acquire_secondary (:923-966): the signs of the buffered prompts against the expected symbols ('0' = +, anything else = -).
// A prompt at exactly 0 counts as positive. All agree -> locked; all disagree -> locked half a cycle off.
bool B200_Symbol_Sync::acquire_secondary()
{
    int32_t agree = 0;
    for (uint32_t i = 0; i < d_secondary_code_length; i++)
        {
            const bool prompt_negative = d_Prompt_circular_buffer[i].real() < 0.0F;
            const bool symbol_negative = d_secondary_code_string[i] == '0';  // the reference counts a NEGATIVE prompt on '0' as agreement
            agree += (prompt_negative == symbol_negative) ? 1 : -1;
        }
    if (std::abs(agree) != static_cast<int32_t>(d_secondary_code_length)) return false;
    d_Flag_PLL_180_deg_phase_locked = agree < 0;
    return true;
}


bool B200_Symbol_Sync::push(const gnss_b200_trk_epoch& e, Symbol* out)
{
    if ((e.flags & GNSS_B200_TRK_VALID) == 0U) return false;
    if ((e.flags & GNSS_B200_TRK_LOSS_OF_LOCK) != 0U)
        {
            reset();  // clear_tracking_vars() + d_state = 0
            return false;
        }
    const std::complex<float> prompt(e.corr[4], e.corr[5]);
    if (!d_locked)
        {
            // state 2 (:1890-1935): nothing is collected during the pull-in transitory
            if ((e.flags & GNSS_B200_TRK_PULL_IN) != 0U) return false;
            bool next_state = false;
            if (d_secondary || d_symbols_per_bit > 1)
                {
                    if (d_Prompt_circular_buffer.size() == d_secondary_code_length && !d_Prompt_circular_buffer.empty())
                        {
                            d_Prompt_circular_buffer.erase(d_Prompt_circular_buffer.begin());  // boost::circular_buffer::push_back on a full buffer
                        }
                    d_Prompt_circular_buffer.push_back(prompt);
                    if (d_Prompt_circular_buffer.size() == d_secondary_code_length)
                        {
                            next_state = acquire_secondary();
                        }
                }
            else
                {
                    next_state = true;
                }
            if (next_state)  // :1936-1946
                {
                    d_P_data_accu = std::complex<float>(0.0F, 0.0F);
                    d_Prompt_circular_buffer.clear();
                    d_current_symbol = 0;
                    d_current_data_symbol = 0;
                    d_locked = true;
                }
            return false;
        }

    // state 4 with one code period per integration: save_correlation_results() for the prompt (:1290-1400)
    if (d_secondary)
        {
            d_current_symbol++;
            d_current_symbol %= d_secondary_code_length;  // the sign applied to d_P_accu does not reach the outputs below
        }
    if (d_symbols_per_bit > 1)
        {
            // per-bit accumulation, with the data secondary code wiped off where the signal has one
            const bool wipe = d_data_secondary_code_length > 0;
            const bool minus = wipe && d_data_secondary_code_string[d_current_data_symbol] != '0';
            d_P_data_accu = minus ? d_P_data_accu - prompt : d_P_data_accu + prompt;
            d_current_data_symbol = (d_current_data_symbol + 1) % (wipe ? d_data_secondary_code_length : d_symbols_per_bit);
        }
    else
        {
            d_P_data_accu = prompt;
        }
    if (d_current_data_symbol != 0) return false;
    if (out != nullptr)  // :2043-2066
        {
            if (d_interchange_iq)
                {
                    out->Prompt_I = static_cast<double>(d_P_data_accu.imag());
                    out->Prompt_Q = static_cast<double>(d_P_data_accu.real());
                }
            else
                {
                    out->Prompt_I = static_cast<double>(d_P_data_accu.real());
                    out->Prompt_Q = static_cast<double>(d_P_data_accu.imag());
                }
            out->Code_phase_samples = e.rem_code_phase_samples;
            out->Carrier_phase_rads = e.acc_carrier_phase_rad;
            out->Carrier_Doppler_hz = e.carrier_doppler_hz;
            out->CN0_dB_hz = e.cn0_db_hz;
            out->EVM = e.evm;
            out->Tracking_sample_counter = e.sample_counter;  // this->nitems_read(0) at the end of general_work (:2124)
            out->Flag_valid_symbol_output = true;
            out->Flag_PLL_180_deg_phase_locked = d_Flag_PLL_180_deg_phase_locked;
        }
    d_P_data_accu = std::complex<float>(0.0F, 0.0F);
    return true;
}
