/*
 * b200_symbol_sync.h — the part of dll_pll_veml_tracking that turns per-code-period prompts into telemetry symbols,
 * fed from the device loop's epoch records (gnss_b200_trk_epoch).
 *
 * With one code period per integration (extend_correlation_symbols = 1) the device loop computes the NCO trajectory of
 * the reference's states 2 and 4; what the block does besides is bookkeeping on the prompt outputs,
 * mirrored here statement by statement:
 *   - secondary-code / bit synchronisation: the last d_secondary_code_length prompts in a circular buffer
 *     (dll_pll_veml_tracking.cc:1892-1920), acquire_secondary() (:923-966) on their signs, the 180-degree flag;
 *   - after lock (state 4): save_correlation_results() (:1290-1400) for the prompt — data secondary code wipe-off,
 *     accumulation of d_P_data_accu over d_symbols_per_bit code periods, one output symbol when d_current_data_symbol
 *     wraps (:2041-2066);
 *   - loss of lock (:1873-1879, :2035-2040) sends it back to the search.
 * The strings and lengths are the per-signal constants the constructor of the reference block selects (:180-450), e.g.
 * GPS L1 C/A: GPS_CA_PREAMBLE_SYMBOLS_STR / 160 with 20 symbols per bit; GPS L5I: GPS_L5I_NH_CODE_STR / 10, 10 symbols.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#ifndef GNSS_SDR_B200_SYMBOL_SYNC_H
#define GNSS_SDR_B200_SYMBOL_SYNC_H

#include "gnss_b200.h"
#include <complex>
#include <cstdint>
#include <string>
#include <vector>

class B200_Symbol_Sync
{
public:
    struct Symbol  // the Gnss_Synchro fields the block fills when a symbol is ready (:2043-2066)
    {
        double Prompt_I, Prompt_Q;
        double Code_phase_samples, Carrier_phase_rads, Carrier_Doppler_hz, CN0_dB_hz, EVM;
        uint64_t Tracking_sample_counter;
        bool Flag_valid_symbol_output;
        bool Flag_PLL_180_deg_phase_locked;
    };

    // d_secondary (there is a secondary code on the tracked component), d_secondary_code_string / _length (or the preamble
    // for signals without one), d_symbols_per_bit, d_data_secondary_code_string / _length (empty / 0 if none)
    void configure(bool secondary, const std::string& secondary_code_string, uint32_t secondary_code_length, uint32_t symbols_per_bit,
        const std::string& data_secondary_code_string, uint32_t data_secondary_code_length, bool interchange_iq);
    void reset();  // start_tracking()
    // One epoch record of the device loop. Returns true when `out` holds a symbol for the telemetry decoder.
    bool push(const gnss_b200_trk_epoch& e, Symbol* out);
    bool locked() const { return d_locked; }
    bool pll_180_deg() const { return d_Flag_PLL_180_deg_phase_locked; }
    uint32_t current_symbol() const { return d_current_symbol; }

private:
    bool acquire_secondary();
    std::string d_secondary_code_string, d_data_secondary_code_string;
    std::vector<std::complex<float>> d_Prompt_circular_buffer;  // oldest first
    std::complex<float> d_P_data_accu{0.0F, 0.0F};
    uint32_t d_secondary_code_length{0}, d_data_secondary_code_length{0}, d_symbols_per_bit{0};
    uint32_t d_current_symbol{0}, d_current_data_symbol{0};
    bool d_secondary{false}, d_interchange_iq{false}, d_locked{false}, d_Flag_PLL_180_deg_phase_locked{false};
};

#endif  // GNSS_SDR_B200_SYMBOL_SYNC_H
