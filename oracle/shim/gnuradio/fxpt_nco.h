/*
 * TEST INFRASTRUCTURE — build stub for oracle/_ref only.
 * gnss_signal_replica.cc includes <gnuradio/fxpt_nco.h> for complex_exp_gen(),
 * which is off the hot path and never called by the oracle. GNU Radio is absent,
 * so this stub only satisfies the compiler; calling it aborts.
 */
#ifndef ORACLE_SHIM_FXPT_NCO_H
#define ORACLE_SHIM_FXPT_NCO_H
#include <complex>
#include <cstdlib>
namespace gr
{
class fxpt_nco
{
public:
    void set_freq(float) {}
    void sincos(std::complex<float>*, std::size_t, double) { std::abort(); }
};
}  // namespace gr
#endif
