/*
 * TEST INFRASTRUCTURE — oracle/_ref only (needs /root/reference at build time).
 * Drives the reference's own tracking helper classes and functions, compiled unmodified from
 * src/algorithms/tracking/libs/{tracking_loop_filter,tracking_FLL_PLL_filter,exponential_smoother,
 * tracking_discriminators,lock_detectors}.cc, so that oracle_trk.c's restatement can be pinned on them bit for bit.
 * (gr_complex, gr::fast_atan2f and glog come from the stand-in headers under oracle/shim.)
 */
#include "exponential_smoother.h"
#include "lock_detectors.h"
#include "tracking_FLL_PLL_filter.h"
#include "tracking_discriminators.h"
#include "tracking_loop_filter.h"
#include <complex>

extern "C"
{
    /* y[i] = filter.apply(x[i]) after the block's constructor + initialize() sequence (dll_pll_veml_tracking.cc:462,864-868) */
    void ref_trk_loop_filter_run(float update_interval, float bw, int order, int include_last_integrator, float initial_output, const float* x, int n, float* y)
    {
        Tracking_loop_filter f(update_interval, bw, order, include_last_integrator != 0);
        f.set_noise_bandwidth(bw);
        f.set_update_interval(update_interval);
        f.initialize(initial_output);
        for (int i = 0; i < n; i++) y[i] = f.apply(x[i]);
    }

    void ref_trk_carrier_filter_run(float fll_bw, float pll_bw, int order, float acq_doppler, const float* fll, const float* pll, float T, int n, float* y)
    {
        Tracking_FLL_PLL_filter f;
        f.set_params(fll_bw, pll_bw, order);
        f.initialize(acq_doppler);
        for (int i = 0; i < n; i++) y[i] = f.get_carrier_error(fll[i], pll[i], T);
    }

    void ref_trk_smoother_run(float alpha, float min_value, float offset, int n_init, const float* x, int n, float* y)
    {
        Exponential_Smoother s;
        s.set_alpha(alpha);
        s.set_min_value(min_value);
        s.set_offset(offset);
        s.set_samples_for_initialization(n_init);
        s.reset();
        for (int i = 0; i < n; i++) y[i] = s.smooth(x[i]);
    }

    double ref_trk_fll_diff_atan(float a_re, float a_im, float b_re, float b_im, double t1, double t2)
    {
        return fll_diff_atan(gr_complex(a_re, a_im), gr_complex(b_re, b_im), t1, t2);
    }
    double ref_trk_pll_cloop(float re, float im) { return pll_cloop_two_quadrant_atan(gr_complex(re, im)); }
    double ref_trk_dll_e_minus_l(float e_re, float e_im, float l_re, float l_im, float spc, float slope, float y_intercept)
    {
        return dll_nc_e_minus_l_normalized(gr_complex(e_re, e_im), gr_complex(l_re, l_im), spc, slope, y_intercept);
    }
    double ref_trk_dll_vemlp(const float* ve, const float* e, const float* l, const float* vl)
    {
        return dll_nc_vemlp_normalized(gr_complex(ve[0], ve[1]), gr_complex(e[0], e[1]), gr_complex(l[0], l[1]), gr_complex(vl[0], vl[1]));
    }
    float ref_trk_cn0_m2m4(const float* p, int n, float T) { return cn0_m2m4_estimator(reinterpret_cast<const gr_complex*>(p), n, T); }
    float ref_trk_carrier_lock(const float* p, int n) { return carrier_lock_detector(const_cast<gr_complex*>(reinterpret_cast<const gr_complex*>(p)), n); }
}
