/*
 * TEST INFRASTRUCTURE — oracle/_ref only (needs /root/reference at build time).
 * extern "C" entry points over the reference's own PRN code generators, which
 * are compiled unmodified from where they lie (see Makefile, target _ref).
 */
#include "beidou_b1i_signal_replica.h"
#include "galileo_e1_signal_replica.h"
#include "gps_l5_signal_replica.h"
#include "gps_sdr_signal_replica.h"
#include <array>
#include <complex>
#include <cstdint>
#include <vector>

extern "C"
{
    void ref_gps_l1_ca_chips(int32_t* dest, int prn) { gps_l1_ca_code_gen_int(own::span<int32_t>(dest, 1023), prn, 0); }
    void ref_gps_l1_ca_sampled(float* dest_c64, int prn, int fs, int n)
    {
        gps_l1_ca_code_gen_complex_sampled(own::span<std::complex<float>>(reinterpret_cast<std::complex<float>*>(dest_c64), n), prn, fs, 0);
    }
    void ref_beidou_b1i_chips(int32_t* dest, int prn) { beidou_b1i_code_gen_int(own::span<int32_t>(dest, 2046), prn, 0); }
    void ref_beidou_b1i_sampled(float* dest_c64, int prn, int fs, int n)
    {
        beidou_b1i_code_gen_complex_sampled(own::span<std::complex<float>>(reinterpret_cast<std::complex<float>*>(dest_c64), n), prn, fs, 0);
    }
    void ref_gps_l5i_chips(float* dest, int prn) { gps_l5i_code_gen_float(own::span<float>(dest, 10230), prn); }
    void ref_gps_l5q_chips(float* dest, int prn) { gps_l5q_code_gen_float(own::span<float>(dest, 10230), prn); }
    void ref_gps_l5i_sampled(float* dest_c64, int prn, int fs, int n)
    {
        gps_l5i_code_gen_complex_sampled(own::span<std::complex<float>>(reinterpret_cast<std::complex<float>*>(dest_c64), n), prn, fs);
    }
    /* 1 sample per chip E1B / E1C primary code, via the 2-samples-per-chip sinBOC(1,1) float generator */
    void ref_galileo_e1_chips(float* dest4092, int prn, int pilot)
    {
        std::array<char, 3> sig = {{'1', pilot ? 'C' : 'B', '\0'}};
        std::vector<float> boc(2 * 4092);
        galileo_e1_code_gen_sinboc11_float(own::span<float>(boc.data(), boc.size()), sig, prn);
        for (int i = 0; i < 4092; i++) dest4092[i] = boc[2 * i];
    }
    void ref_galileo_e1_sampled(float* dest_c64, int prn, int fs, int n, int pilot, int cboc)
    {
        std::array<char, 3> sig = {{'1', pilot ? 'C' : 'B', '\0'}};
        galileo_e1_code_gen_complex_sampled(own::span<std::complex<float>>(reinterpret_cast<std::complex<float>*>(dest_c64), n), sig, cboc != 0, prn, fs, 0, false);
    }
}
