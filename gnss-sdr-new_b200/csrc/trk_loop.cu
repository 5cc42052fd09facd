// The dll_pll_veml tracking loop closed on the device.
//
// Per epoch and channel the reference block does, in state 2 (dll_pll_veml_tracking.cc:1846-1990):
//     do_correlation_step (:1064-1089) -> cn0_and_tracking_lock_status (:968-1013) -> run_dll_pll (:1092-1182)
//     -> update_tracking_vars (:1225-1297) -> consume_each(d_current_prn_length_samples) (:2118)
// and, once the secondary code / bit edge is found (:1888-1925), the state 3 / state 4 cycle of the extended correlator
// (:1989-2084) on accumulators fed by save_correlation_results (:1290-1400),
// with one host thread per channel. Here one launch pair (a third launch for the slave data correlator of pilot tracking)
// per epoch serves every channel:
//     multicorrelator kernel (multicorr.cu)  : correlator outputs of epoch e for all channels
//     trk_update_kernel (this file)          : discriminators, loop filters, lock detectors, NCO update, and the
//                                              multicorrelator parameter block of epoch e+1 — written in device memory,
// so a block of samples holding K code periods is processed by 2K back-to-back launches and ONE result copy.
// Scalar arithmetic keeps the reference's types (double where the block uses double, float where the filter
// classes use float), so that trajectories agree with the CPU loop to rounding.
#include "multicorr_internal.cuh"
#include <cmath>
#include <vector>

namespace b200
{
namespace
{
constexpr int kMaxCn0Samples = 64;
constexpr int kMaxSmootherInit = 256;
constexpr int kMaxSecondary = 208;
constexpr int kMaxSmootherLength = 64;
// the reference's own constants (MATH_CONSTANTS.h:47-49): pi "as defined for GNSS", not M_PI
constexpr double kPi = 3.1415926535898;
constexpr double kHalfPi = kPi / 2.0;
constexpr double kTwoPi = 2.0 * kPi;

// Exponential_Smoother (exponential_smoother.cc:63-101)
struct Smoother
{
    float alpha, one_minus_alpha, old_value, min_value, offset;
    int samples_for_initialization, init_counter, initializing;
    float init_buffer[kMaxSmootherInit];
};

// Tracking_loop_filter (tracking_loop_filter.cc:57-97) with coefficients from update_coefficients (:100-210)
struct LoopFilter
{
    float in_coef[4], out_coef[3];
    int n_in, n_out, idx;
    float inputs[4], outputs[4];
};

// Tracking_FLL_PLL_filter (tracking_FLL_PLL_filter.cc:23-97)
struct CarrierFilter
{
    int order;
    float pll_w, pll_x, pll_a2, pll_a3, pll_b3, pll_w0p, pll_w0p2, pll_w0p3, pll_w0f, pll_w0f2;
};

struct TrkChan
{
    double carrier_doppler_hz, carrier_phase_step_rad, code_freq_chips, code_phase_step_chips;
    double rem_code_phase_samples, rem_code_phase_chips, acc_carrier_phase_rad;
    double carr_phase_error_hz, carr_freq_error_hz, carr_error_filt_hz, code_error_chips, code_error_filt_chips;
    double cn0_db_hz, carrier_lock_test, evm;
    unsigned long long cursor, acq_sample_stamp;
    float rem_carr_phase_rad;
    float2 P_old;
    int current_prn_length_samples;
    int state;  // 0 standby, 2 wide tracking + symbol synchronisation, 3 coherent extension, 4 narrow tracking (state 1 is gnss_b200_trk_start)
    int pull_in;
    int carrier_lock_fail, code_lock_fail, cn0_counter;
    int correlated;  // the parameter block of the running epoch was live
    CarrierFilter cf;
    LoopFilter lf;
    Smoother cn0_s, lock_s;
    float2 prompt_buf[kMaxCn0Samples];
    const float* code;
    uint32_t code_len;
    // the symbol machine (symbol_sync), pilot tracking and high dynamics
    const float* data_code;
    double carrier_phase_rate_step_rad, code_phase_rate_step_chips, current_correlation_time_s;
    float spc;           // d_trk_parameters.spc: the E-L spacing the discriminator is scaled with
    int narrow;          // the narrow taps are in use (:1960-1975)
    float2 accu[5];      // d_VE_accu, d_E_accu, d_P_accu, d_L_accu, d_VL_accu
    float2 P_data_accu;
    int cloop, acc_carrier_phase_initialized, current_symbol, current_data_symbol, extend_count, pll_180;
    int circ_size, carr_hist_size, code_hist_size;
    float2 circ[kMaxSecondary];                                      // d_Prompt_circular_buffer, oldest first
    double carr_hist[4 * kMaxSmootherLength], code_hist[4 * kMaxSmootherLength];  // (value, samples) pairs, capacity 2 * smoother_length
};

struct TrkDevCfg
{
    double fs_in, code_chip_rate, signal_carrier_freq, code_period_s, carrier_lock_th;
    uint32_t code_length_chips, code_samples_per_chip, vector_length, n_taps;
    int veml, enable_fll_pull_in, enable_fll_steady_state, carrier_aiding;
    uint32_t pull_in_time_s;
    int cn0_samples, cn0_min, max_code_lock_fail, max_carrier_lock_fail;
    float spc, slope, y_intercept;
    // symbol machine
    int symbol_sync, secondary, ext, track_pilot, high_dyn;
    uint32_t secondary_code_length, symbols_per_bit, data_secondary_code_length, smoother_length, bit_synchronization_time_limit_s;
    float spc_narrow, T_ext;          // early_late_space_narrow_chips; (float)ext * (float)code_period
    float shifts_wide[5], shifts_narrow[5];
    LoopFilter lf_narrow;             // coefficients after set_update_interval(T_ext) + set_noise_bandwidth(dll_bw_narrow_hz)
    CarrierFilter cf_narrow;          // constants after set_params(fll_bw_hz, pll_bw_narrow_hz, order)
    char secondary_code_string[kMaxSecondary], data_secondary_code_string[112];
};

// gr::fast_atan2f (GNU Radio, gnuradio-runtime/lib/math/fast_atan2f.cc): table[i] = atan(i / 255), two guard entries
__constant__ float c_fast_atan_table[258];

__device__ float fast_atan2f_gr(float y, float x)
{
    const float y_abs = fabsf(y), x_abs = fabsf(x);
    if (!((y_abs > 0.0f) || (x_abs > 0.0f))) return 0.0f;
    const float z = (y_abs < x_abs) ? __fdiv_rn(y_abs, x_abs) : __fdiv_rn(x_abs, y_abs);
    float base_angle;
    if (z < 0.003921569f)
        base_angle = z;
    else
        {
            float alpha = __fmul_rn(z, 255.0f);
            const int index = ((int)alpha) & 0xff;
            alpha = __fsub_rn(alpha, (float)index);
            base_angle = __fadd_rn(c_fast_atan_table[index], __fmul_rn(__fsub_rn(c_fast_atan_table[index + 1], c_fast_atan_table[index]), alpha));
        }
    const float pi = 3.14159265358979323846f, half_pi = 1.57079632679489661923f;
    float angle;
    if (x_abs > y_abs)
        {
            if (x >= 0.0f)
                angle = (y >= 0.0f) ? base_angle : -base_angle;
            else
                angle = (y >= 0.0f) ? __fsub_rn(pi, base_angle) : __fsub_rn(base_angle, pi);
        }
    else
        {
            if (y >= 0.0f)
                angle = (x >= 0.0f) ? __fsub_rn(half_pi, base_angle) : __fadd_rn(half_pi, base_angle);
            else
                angle = (x >= 0.0f) ? __fadd_rn(-half_pi, base_angle) : __fsub_rn(-half_pi, base_angle);
        }
    return angle;
}

__device__ float smoother_smooth(Smoother& s, float raw)
{
    float v;
    if (s.initializing)
        {
            s.init_counter++;
            v = raw;
            if (s.init_counter <= kMaxSmootherInit) s.init_buffer[s.init_counter - 1] = v;
            if (s.init_counter == s.samples_for_initialization)
                {
                    float acc = 0.0f;  // std::accumulate(..., 0.0F)
                    for (int i = 0; i < s.init_counter; i++) acc += s.init_buffer[i];
                    s.old_value = acc / (float)s.init_counter;
                    if (s.old_value < (s.min_value + s.offset))
                        s.init_counter = 0;  // flush buffer and start again
                    else
                        s.initializing = 0;
                }
        }
    else
        {
            v = __fadd_rn(__fmul_rn(s.alpha, raw), __fmul_rn(s.one_minus_alpha, s.old_value));
            s.old_value = v;
        }
    return v;
}

__device__ float loop_filter_apply(LoopFilter& f, float x)
{
    float r = 0.0f;
    for (int i = 0; i < f.n_out; i++) r = __fadd_rn(r, __fmul_rn(f.out_coef[i], f.outputs[(f.idx + i) % 4]));
    f.idx--;
    if (f.idx < 0) f.idx += 4;
    f.inputs[f.idx] = x;
    for (int i = 0; i < f.n_in; i++) r = __fadd_rn(r, __fmul_rn(f.in_coef[i], f.inputs[(f.idx + i) % 4]));
    f.outputs[f.idx] = r;
    return r;
}

__device__ float carrier_filter(CarrierFilter& c, float fll, float pll, float T)
{
    // explicit round-to-nearest products and sums: the host compiler does not contract the reference's expressions
    float err;
    if (c.order == 3)
        {
            c.pll_w = __fadd_rn(c.pll_w, __fmul_rn(T, __fadd_rn(__fmul_rn(c.pll_w0p3, pll), __fmul_rn(c.pll_w0f2, fll))));
            const float t = __fadd_rn(__fadd_rn(__fmul_rn(0.5f, c.pll_w), __fmul_rn(__fmul_rn(c.pll_a2, c.pll_w0f), fll)), __fmul_rn(__fmul_rn(c.pll_a3, c.pll_w0p2), pll));
            c.pll_x = __fadd_rn(c.pll_x, __fmul_rn(T, t));
            err = __fadd_rn(__fmul_rn(0.5f, c.pll_x), __fmul_rn(__fmul_rn(c.pll_b3, c.pll_w0p), pll));
        }
    else
        {
            const float w_new = __fadd_rn(__fadd_rn(c.pll_w, __fmul_rn(__fmul_rn(pll, c.pll_w0p2), T)), __fmul_rn(__fmul_rn(fll, c.pll_w0f), T));
            err = __fadd_rn(__fmul_rn(0.5f, __fadd_rn(w_new, c.pll_w)), __fmul_rn(__fmul_rn(c.pll_a2, c.pll_w0p), pll));
            c.pll_w = w_new;
        }
    return err;
}

// cn0_m2m4_estimator (lock_detectors.cc:98-130)
__device__ float cn0_m2m4(const float2* buf, int length, float T)
{
    float Psig = 0.0f, m2 = 0.0f, m4 = 0.0f;
    for (int i = 0; i < length; i++)
        {
            Psig = __fadd_rn(Psig, fabsf(buf[i].x));
            const float aux = __fadd_rn(__fmul_rn(buf[i].y, buf[i].y), __fmul_rn(buf[i].x, buf[i].x));
            m2 = __fadd_rn(m2, aux);
            m4 = __fadd_rn(m4, __fmul_rn(aux, aux));
        }
    const float n = (float)length;
    Psig = __fdiv_rn(Psig, n);
    Psig = __fmul_rn(Psig, Psig);
    m2 = __fdiv_rn(m2, n);
    m4 = __fdiv_rn(m4, n);
    const float aux = sqrtf(__fsub_rn(__fmul_rn(__fmul_rn(2.0f, m2), m2), m4));
    float snr;
    if (isnan(aux))
        snr = __fdiv_rn(Psig, __fsub_rn(m2, Psig));
    else
        snr = __fdiv_rn(aux, __fsub_rn(m2, aux));
    return __fsub_rn(__fmul_rn(10.0f, log10f(snr)), __fmul_rn(10.0f, log10f(T)));
}

__device__ double phase_unwrap(double p)
{
    if (p >= kHalfPi) return p - kPi;
    if (p <= -kHalfPi) return p + kPi;
    return p;
}

// state -> multicorrelator parameter block of the next epoch (what do_correlation_step passes, :1064-1078), the channel's
// tap shifts (wide or narrow), the integer sample shifts of the high-dynamics resampler, and the block of the slave data correlator
__device__ void write_params(const TrkDevCfg& cfg, TrkChan& s, mc::ChanParams& p, mc::ChanParams* p_data, float* shifts, uint32_t* shift_samples,
    uint32_t ch_index, const float2* stream, unsigned long long stream_start, unsigned long long stream_end)
{
    const bool live = s.state >= 2 && s.cursor >= stream_start && s.cursor + 2ull * cfg.vector_length <= stream_end;  // forecast(): 2 * vector_length
    s.correlated = live ? 1 : 0;
    p.sig = stream + (live ? (s.cursor - stream_start) : 0ull);
    p.code = s.code;
    p.code_len = s.code_len;
    p.n = live ? cfg.vector_length : 0u;
    const float spc = (float)cfg.code_samples_per_chip;
    p.rem_code = __fmul_rn((float)s.rem_code_phase_chips, spc);
    p.code_step = __fmul_rn((float)s.code_phase_step_chips, spc);
    p.code_rate = __fmul_rn((float)s.code_phase_rate_step_chips, spc);
    // phase_offset_as_complex = (cos(rem), -sin(rem)) and the rounded per-sample phasor exp(-j*step) in binary32
    // (cpu_multicorrelator_real_codes.cc:114-119); sin/cos in double then rounded = correctly rounded binary32
    double s0, c0;
    sincos((double)s.rem_carr_phase_rad, &s0, &c0);
    p.p0_re = (float)c0;
    p.p0_im = -(float)s0;
    const float step = (float)s.carrier_phase_step_rad;
    double si, ci;
    sincos(-(double)step, &si, &ci);
    const double theta = atan2((double)(float)si, (double)(float)ci);
    double sb, cb;
    sincos(theta * (double)mc::kThreads, &sb, &cb);
    p.theta = theta;
    p.wb_re = (float)cb;
    p.wb_im = (float)sb;
    p.rate_arg = 0.0f;
    p.flags = mc::F_CODE_DOUBLED;  // gnss_b200_trk_set_code uploads the doubled layout
    p.shifts_off = ch_index * cfg.n_taps;
    const float* sh = s.narrow ? cfg.shifts_narrow : cfg.shifts_wide;
    for (uint32_t t = 0; t < cfg.n_taps; t++) shifts[p.shifts_off + t] = sh[t];
    if (cfg.high_dyn)
        {
            // set_high_dynamics_resampler(true): the rate-aware resampler and rotator, whatever the rates are (:1068-1075)
            p.flags |= mc::F_HIGH_DYN_RESAMPLER | mc::F_HIGH_DYN_ROTATOR;
            // arg of the binary32 phasor exp(-j * rate) the rotator raises to (n-1)^2 (cpowf = cexpf(m * clogf(z)))
            const float rate = (float)s.carrier_phase_rate_step_rad;
            double sr, cr;
            sincos(-(double)rate, &sr, &cr);
            p.rate_arg = (float)atan2((double)(float)sr, (double)(float)cr);
            // ...high_dynamics_resampler...:84-87: cumulative integer sample shifts of taps 1.. relative to tap 0
            uint32_t acc = 0;
            shift_samples[p.shifts_off] = 0;
            for (uint32_t t = 1; t < cfg.n_taps; t++)
                {
                    const float d = __fdiv_rn(__fsub_rn(sh[t], sh[t - 1]), p.code_step);
                    acc += (uint32_t)(int)rint((double)d);
                    shift_samples[p.shifts_off + t] = acc;
                }
        }
    if (p_data)
        {
            *p_data = p;
            p_data->code = s.data_code;
            p_data->shifts_off = 0;  // one tap at the prompt shift (0), shared by all channels
            if (s.data_code == nullptr) p_data->n = 0;
        }
}

// cn0_and_tracking_lock_status(coh_integration_time_s) (:968-1056); returns false on loss of lock
__device__ bool lock_status(const TrkDevCfg& cfg, TrkChan& s, const float2 P, const float coh_T)
{
    if (s.cn0_counter < cfg.cn0_samples)
        {
            s.prompt_buf[s.cn0_counter] = P;
            s.cn0_counter++;
            return true;
        }
    s.prompt_buf[s.cn0_counter % cfg.cn0_samples] = P;
    s.cn0_counter++;
    const float raw = cn0_m2m4(s.prompt_buf, cfg.cn0_samples, coh_T);
    s.cn0_db_hz = (double)smoother_smooth(s.cn0_s, raw);
    // carrier_lock_detector(buffer, 1) (lock_detectors.cc:133-148)
    const float I = s.prompt_buf[0].x, Q = s.prompt_buf[0].y;
    const float II = __fmul_rn(I, I), QQ = __fmul_rn(Q, Q);
    const float test = __fdiv_rn(__fsub_rn(II, QQ), __fadd_rn(II, QQ));
    s.carrier_lock_test = (double)smoother_smooth(s.lock_s, test);
    if (!s.pull_in)
        {
            if (s.carrier_lock_test < cfg.carrier_lock_th)
                s.carrier_lock_fail++;
            else if (s.carrier_lock_fail > 0)
                s.carrier_lock_fail--;
            if (s.cn0_db_hz < (double)cfg.cn0_min)
                s.code_lock_fail++;
            else if (s.code_lock_fail > 0)
                s.code_lock_fail--;
        }
    if (s.carrier_lock_fail > cfg.max_carrier_lock_fail || s.code_lock_fail > cfg.max_code_lock_fail)
        {
            s.carrier_lock_fail = 0;
            s.code_lock_fail = 0;
            return false;
        }
    // EVM indicator of the fork (:1027-1053): binary32 sums in buffer order
    float tmp_sum = 0.f;
    for (int k = 0; k < cfg.cn0_samples; k++) tmp_sum = __fadd_rn(tmp_sum, __fmul_rn(s.prompt_buf[k].x, s.prompt_buf[k].x));
    const float d = __fsqrt_rn(__fdiv_rn(tmp_sum, (float)cfg.cn0_samples));
    tmp_sum = 0.f;
    for (int k = 0; k < cfg.cn0_samples; k++)
        {
            const float a = __fsub_rn(fabsf(__fdiv_rn(s.prompt_buf[k].x, d)), 1.f);
            const float b = fabsf(__fdiv_rn(s.prompt_buf[k].y, d));
            tmp_sum = __fadd_rn(__fadd_rn(tmp_sum, __fmul_rn(a, a)), __fmul_rn(b, b));
        }
    s.evm = sqrt((double)__fdiv_rn(tmp_sum, (float)cfg.cn0_samples));
    return true;
}

// run_dll_pll (:1092-1160) on the accumulators
__device__ void run_dll_pll(const TrkDevCfg& cfg, TrkChan& s)
{
    const float2 VE = s.accu[0], E = s.accu[1], P = s.accu[2], L = s.accu[3], VL = s.accu[4];
    const float T_s = (float)s.current_correlation_time_s;
    double pll_rad = 0.0;
    if (s.cloop)
        {
            // pll_cloop_two_quadrant_atan (tracking_discriminators.cc:101-108)
            if (P.x != 0.0f) pll_rad = (double)atanf(__fdiv_rn(P.y, P.x));
        }
    else
        pll_rad = (double)fast_atan2f_gr(P.y, P.x);  // pll_four_quadrant_atan (:86-89)
    s.carr_phase_error_hz = pll_rad / kTwoPi;
    const bool fll_pull = s.pull_in && cfg.enable_fll_pull_in;
    if (fll_pull || cfg.enable_fll_steady_state)
        {
            // fll_diff_atan(P_old, P, 0, T) (:65-73): std::atan(float) - std::atan(float), the difference is formed in binary32
            double d = (double)__fsub_rn(atanf(__fdiv_rn(P.y, P.x)), atanf(__fdiv_rn(s.P_old.y, s.P_old.x)));
            if (isnan(d)) d = 0.0;
            s.carr_freq_error_hz = phase_unwrap(d) / (s.current_correlation_time_s - 0.0) / kTwoPi;
            s.P_old = P;
            if (fll_pull)
                s.carr_error_filt_hz = (double)carrier_filter(s.cf, (float)s.carr_freq_error_hz, 0.0f, T_s);
            else
                s.carr_error_filt_hz = (double)carrier_filter(s.cf, (float)s.carr_freq_error_hz, (float)s.carr_phase_error_hz, T_s);
        }
    else
        {
            s.carr_error_filt_hz = (double)carrier_filter(s.cf, 0.0f, (float)s.carr_phase_error_hz, T_s);
        }
    s.carrier_doppler_hz = s.carr_error_filt_hz;
    if (cfg.veml)
        {
            // dll_nc_vemlp_normalized (:140-149): float sums, sqrt in float promoted to double
            const float es = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(VE.x, VE.x), __fmul_rn(VE.y, VE.y)), __fmul_rn(E.x, E.x)), __fmul_rn(E.y, E.y));
            const float ls = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(L.x, L.x), __fmul_rn(L.y, L.y)), __fmul_rn(VL.x, VL.x)), __fmul_rn(VL.y, VL.y));
            const double Early = (double)sqrtf(es), Late = (double)sqrtf(ls);
            const double epl = Early + Late;
            s.code_error_chips = (epl == 0.0) ? 0.0 : (Early - Late) / epl;
        }
    else
        {
            // dll_nc_e_minus_l_normalized (:120-130): std::abs(complex<float>) = hypotf
            const double Pe = (double)hypotf(E.x, E.y), Pl = (double)hypotf(L.x, L.y);
            const double epl = Pe + Pl;
            const float k = __fdiv_rn(__fsub_rn(cfg.y_intercept, __fmul_rn(cfg.slope, s.spc)), cfg.slope);
            s.code_error_chips = (epl == 0.0) ? 0.0 : (double)k * (Pe - Pl) / epl;
        }
    s.code_error_filt_chips = (double)loop_filter_apply(s.lf, (float)s.code_error_chips);
    s.code_freq_chips = cfg.code_chip_rate - s.code_error_filt_chips;
    if (cfg.carrier_aiding) s.code_freq_chips += s.carrier_doppler_hz * cfg.code_chip_rate / cfg.signal_carrier_freq;
}

// push_back on a boost::circular_buffer of (value, samples) pairs with capacity cap, then the smoothed rate when it is full (:1232-1251)
__device__ void hist_push_rate(double* h, int& size, int cap, uint32_t L, double first, double second, double& rate)
{
    if (size == cap)
        {
            for (int k = 0; k < 2 * (cap - 1); k++) h[k] = h[k + 2];
            size--;
        }
    h[2 * size] = first;
    h[2 * size + 1] = second;
    size++;
    if (size == cap)
        {
            double tmp_cp1 = 0.0, tmp_cp2 = 0.0, tmp_samples = 0.0;
            for (uint32_t k = 0; k < L; k++)
                {
                    tmp_cp1 += h[2 * k];
                    tmp_cp2 += h[2 * (L * 2 - k - 1)];
                    tmp_samples += h[2 * (L * 2 - k - 1) + 1];
                }
            tmp_cp1 /= (double)L;
            tmp_cp2 /= (double)L;
            rate = (tmp_cp2 - tmp_cp1) / tmp_samples;
        }
}

// update_tracking_vars (:1216-1287)
__device__ void update_tracking_vars(const TrkDevCfg& cfg, TrkChan& s)
{
    const double T_chip = 1.0 / s.code_freq_chips;
    const double T_prn_s = T_chip * (double)cfg.code_length_chips;
    const double T_prn_samples = T_prn_s * cfg.fs_in;
    const double K_blk = T_prn_samples + s.rem_code_phase_samples;
    s.current_prn_length_samples = (int)floor(K_blk);
    const double n = (double)s.current_prn_length_samples;
    s.carrier_phase_step_rad = kTwoPi * s.carrier_doppler_hz / cfg.fs_in;
    if (cfg.high_dyn)
        hist_push_rate(s.carr_hist, s.carr_hist_size, 2 * (int)cfg.smoother_length, cfg.smoother_length, s.carrier_phase_step_rad, n, s.carrier_phase_rate_step_rad);
    const double adv = s.carrier_phase_step_rad * n + 0.5 * s.carrier_phase_rate_step_rad * n * n;
    s.rem_carr_phase_rad = __fadd_rn(s.rem_carr_phase_rad, (float)adv);
    s.rem_carr_phase_rad = (float)fmod((double)s.rem_carr_phase_rad, kTwoPi);
    s.acc_carrier_phase_rad -= adv;
    s.code_phase_step_chips = s.code_freq_chips / cfg.fs_in;
    if (cfg.high_dyn)
        hist_push_rate(s.code_hist, s.code_hist_size, 2 * (int)cfg.smoother_length, cfg.smoother_length, s.code_phase_step_chips, n, s.code_phase_rate_step_chips);
    s.rem_code_phase_samples = K_blk - n;
    s.rem_code_phase_chips = s.code_freq_chips * s.rem_code_phase_samples / cfg.fs_in;
}

// save_correlation_results (:1290-1400)
__device__ void save_correlation_results(const TrkDevCfg& cfg, TrkChan& s, const float2* o5, const float2 prompt_data)
{
    const bool minus = cfg.secondary && cfg.secondary_code_string[s.current_symbol] != '0';
    for (int k = 0; k < 5; k++)
        {
            if (!cfg.veml && (k == 0 || k == 4)) continue;
            if (minus)
                s.accu[k] = make_float2(__fsub_rn(s.accu[k].x, o5[k].x), __fsub_rn(s.accu[k].y, o5[k].y));
            else
                s.accu[k] = make_float2(__fadd_rn(s.accu[k].x, o5[k].x), __fadd_rn(s.accu[k].y, o5[k].y));
        }
    if (cfg.secondary)
        {
            s.current_symbol++;
            s.current_symbol %= (int)cfg.secondary_code_length;
        }
    const float2 src = cfg.track_pilot ? prompt_data : o5[2];
    if (cfg.symbols_per_bit > 1)
        {
            if (cfg.data_secondary_code_length > 0)
                {
                    if (cfg.data_secondary_code_string[s.current_data_symbol] == '0')
                        s.P_data_accu = make_float2(__fadd_rn(s.P_data_accu.x, src.x), __fadd_rn(s.P_data_accu.y, src.y));
                    else
                        s.P_data_accu = make_float2(__fsub_rn(s.P_data_accu.x, src.x), __fsub_rn(s.P_data_accu.y, src.y));
                    s.current_data_symbol++;
                    s.current_data_symbol %= (int)cfg.data_secondary_code_length;
                }
            else
                {
                    s.P_data_accu = make_float2(__fadd_rn(s.P_data_accu.x, src.x), __fadd_rn(s.P_data_accu.y, src.y));
                    s.current_data_symbol++;
                    s.current_data_symbol %= (int)cfg.symbols_per_bit;
                }
        }
    else
        s.P_data_accu = src;
    s.cloop = cfg.track_pilot ? 0 : 1;
}

// acquire_secondary (:923-966)
__device__ bool acquire_secondary(const TrkDevCfg& cfg, TrkChan& s)
{
    int corr_value = 0;
    for (uint32_t k = 0; k < cfg.secondary_code_length; k++)
        {
            const bool zero = cfg.secondary_code_string[k] == '0';
            if (s.circ[k].x < 0.0f)
                corr_value += zero ? 1 : -1;
            else
                corr_value += zero ? -1 : 1;
        }
    if (abs(corr_value) == (int)cfg.secondary_code_length)
        {
            s.pll_180 = corr_value < 0;
            return true;
        }
    return false;
}

__global__ void trk_update_kernel(const TrkDevCfg cfg, TrkChan* __restrict__ chans, const uint32_t* __restrict__ ch_list, uint32_t n_ch,
    const float2* __restrict__ corr, const float2* __restrict__ corr_data, mc::ChanParams* __restrict__ params, mc::ChanParams* __restrict__ params_data,
    float* __restrict__ shifts, uint32_t* __restrict__ shift_samples, gnss_b200_trk_epoch* __restrict__ log, const float2* __restrict__ stream,
    unsigned long long stream_start, unsigned long long stream_end, int first)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ch) return;
    TrkChan& s = chans[ch_list[i]];
    if (!first && log)
        {
            gnss_b200_trk_epoch& e = log[i];
            e.flags = 0;
            if (s.correlated)
                {
                    const unsigned long long nread = s.cursor;  // nitems_read(0) at this general_work call
                    const int state_in = s.state;
                    // pull-in transitory ends by elapsed whole seconds (:1795-1803)
                    if (s.pull_in && (unsigned long long)cfg.pull_in_time_s < (nread - s.acq_sample_stamp) / (unsigned long long)(int)cfg.fs_in)
                        {
                            s.pull_in = 0;
                            s.carrier_lock_fail = 0;
                            s.code_lock_fail = 0;
                        }
                    const uint32_t T = cfg.n_taps;
                    const float2* o = corr + (size_t)i * T;
                    float2 o5[5];
                    for (int k = 0; k < 5; k++) o5[k] = make_float2(0.f, 0.f);
                    for (uint32_t t = 0; t < T; t++) o5[(cfg.veml ? 0 : 1) + t] = o[t];
                    const float2 prompt_data = (cfg.track_pilot && corr_data) ? corr_data[i] : make_float2(0.f, 0.f);
                    bool lock_ok = true, symbol_out = false;
                    const bool extended = cfg.ext > 1;
                    if (state_in == 2)
                        {
                            for (int k = 0; k < 5; k++) s.accu[k] = o5[k];  // d_VE_accu = *d_Very_Early ... (:1849-1856)
                            s.spc = cfg.spc;
                            if (cfg.symbol_sync && (unsigned long long)cfg.bit_synchronization_time_limit_s < (nread - s.acq_sample_stamp) / (unsigned long long)(int)cfg.fs_in)
                                s.carrier_lock_fail = 300000;  // :1866-1871
                            lock_ok = lock_status(cfg, s, s.accu[2], (float)cfg.code_period_s);
                            for (int k = 0; k < 5; k++)
                                {
                                    e.accu[2 * k] = s.accu[k].x;
                                    e.accu[2 * k + 1] = s.accu[k].y;
                                }
                            if (lock_ok)
                                {
                                    run_dll_pll(cfg, s);
                                    update_tracking_vars(cfg, s);
                                    if (cfg.symbol_sync)
                                        {
                                            bool next_state = false;
                                            if (!s.pull_in)
                                                {
                                                    if (cfg.secondary || cfg.symbols_per_bit > 1)
                                                        {
                                                            const int cap = (int)cfg.secondary_code_length;
                                                            if (s.circ_size == cap)
                                                                {
                                                                    for (int k = 0; k < cap - 1; k++) s.circ[k] = s.circ[k + 1];
                                                                    s.circ_size--;
                                                                }
                                                            s.circ[s.circ_size++] = o5[2];
                                                            if (s.circ_size == cap) next_state = acquire_secondary(cfg, s);
                                                        }
                                                    else
                                                        next_state = true;
                                                }
                                            if (next_state)
                                                {
                                                    for (int k = 0; k < 5; k++) s.accu[k] = make_float2(0.f, 0.f);
                                                    s.P_data_accu = make_float2(0.f, 0.f);
                                                    s.circ_size = 0;
                                                    s.current_symbol = s.current_data_symbol = 0;
                                                    if (extended)
                                                        {
                                                            s.extend_count = 0;
                                                            s.current_correlation_time_s = (double)cfg.T_ext;
                                                            s.state = 3;
                                                            // new coefficients / constants, the filter memories stay (:1956-1958)
                                                            for (int k = 0; k < 4; k++) s.lf.in_coef[k] = cfg.lf_narrow.in_coef[k];
                                                            for (int k = 0; k < 3; k++) s.lf.out_coef[k] = cfg.lf_narrow.out_coef[k];
                                                            s.lf.n_in = cfg.lf_narrow.n_in;
                                                            s.lf.n_out = cfg.lf_narrow.n_out;
                                                            const float w = s.cf.pll_w, x = s.cf.pll_x;
                                                            s.cf = cfg.cf_narrow;
                                                            s.cf.pll_w = w;
                                                            s.cf.pll_x = x;
                                                            s.narrow = 1;
                                                            s.spc = cfg.spc_narrow;
                                                        }
                                                    else
                                                        s.state = 4;
                                                }
                                        }
                                }
                        }
                    else if (state_in == 3)
                        {
                            save_correlation_results(cfg, s, o5, prompt_data);
                            for (int k = 0; k < 5; k++)
                                {
                                    e.accu[2 * k] = s.accu[k].x;
                                    e.accu[2 * k + 1] = s.accu[k].y;
                                }
                            update_tracking_vars(cfg, s);
                            if (s.current_data_symbol == 0) symbol_out = true;
                            s.extend_count++;
                            if (s.extend_count == cfg.ext - 1)
                                {
                                    s.extend_count = 0;
                                    s.state = 4;
                                }
                        }
                    else
                        {
                            save_correlation_results(cfg, s, o5, prompt_data);
                            for (int k = 0; k < 5; k++)
                                {
                                    e.accu[2 * k] = s.accu[k].x;
                                    e.accu[2 * k + 1] = s.accu[k].y;
                                }
                            lock_ok = lock_status(cfg, s, s.accu[2], (float)(cfg.code_period_s * (double)(cfg.ext > 1 ? cfg.ext : 1)));
                            if (lock_ok)
                                {
                                    run_dll_pll(cfg, s);
                                    update_tracking_vars(cfg, s);
                                    if (!s.acc_carrier_phase_initialized)  // check_carrier_phase_coherent_initialization (:1184-1190)
                                        {
                                            s.acc_carrier_phase_rad = -(double)s.rem_carr_phase_rad;
                                            s.acc_carrier_phase_initialized = 1;
                                        }
                                    if (s.current_data_symbol == 0) symbol_out = true;
                                    for (int k = 0; k < 5; k++) s.accu[k] = make_float2(0.f, 0.f);
                                    if (extended) s.state = 3;
                                }
                        }
                    e.flags = GNSS_B200_TRK_VALID | (s.pull_in ? GNSS_B200_TRK_PULL_IN : 0u);
                    if (!lock_ok)
                        {
                            // clear_tracking_vars() + d_state = 0 (:1875-1881, :1192-1213); the old prn length is still consumed
                            s.state = 0;
                            s.P_data_accu = make_float2(0.f, 0.f);
                            s.P_old = make_float2(0.f, 0.f);
                            s.carr_phase_error_hz = s.carr_freq_error_hz = s.carr_error_filt_hz = 0.0;
                            s.code_error_chips = s.code_error_filt_chips = 0.0;
                            s.current_symbol = s.current_data_symbol = 0;
                            s.circ_size = 0;
                            s.carrier_phase_rate_step_rad = s.code_phase_rate_step_chips = 0.0;
                            s.carr_hist_size = s.code_hist_size = 0;
                            e.flags |= GNSS_B200_TRK_LOSS_OF_LOCK;
                        }
                    if (symbol_out)
                        {
                            e.flags |= GNSS_B200_TRK_SYMBOL;
                            e.prompt_data[0] = s.P_data_accu.x;
                            e.prompt_data[1] = s.P_data_accu.y;
                            s.P_data_accu = make_float2(0.f, 0.f);
                        }
                    if (s.pll_180) e.flags |= GNSS_B200_TRK_PLL_180;
                    // ---- log_data() / Gnss_Synchro fields ----
                    for (int k = 0; k < 5; k++)
                        {
                            e.corr[2 * k] = o5[k].x;
                            e.corr[2 * k + 1] = o5[k].y;
                        }
                    e.carrier_doppler_hz = s.carrier_doppler_hz;
                    e.code_freq_chips = s.code_freq_chips;
                    e.carr_error_hz = s.carr_phase_error_hz;
                    e.carr_error_filt_hz = s.carr_error_filt_hz;
                    e.code_error_chips = s.code_error_chips;
                    e.code_error_filt_chips = s.code_error_filt_chips;
                    e.rem_code_phase_samples = s.rem_code_phase_samples;
                    e.acc_carrier_phase_rad = s.acc_carrier_phase_rad;
                    e.cn0_db_hz = s.cn0_db_hz;
                    e.carrier_lock_test = s.carrier_lock_test;
                    e.evm = s.evm;
                    e.sample_counter = nread;
                    e.prn_length_samples = s.current_prn_length_samples;
                    e.state = state_in;
                    s.cursor += (unsigned long long)s.current_prn_length_samples;  // consume_each
                }
        }
    write_params(cfg, s, params[i], params_data ? params_data + i : nullptr, shifts, shift_samples, i, stream, stream_start, stream_end);
}

void loop_filter_coefficients(LoopFilter& f, float T, float bw, int order)
{
    // Tracking_loop_filter::update_coefficients with include_last_integrator = false (the block's constructor, :462)
    const float zeta = 1.0F / std::sqrt(2.0F);
    float g1, g2, g3, wn;
    std::memset(&f, 0, sizeof(f));
    switch (order)
        {
        case 1:
            wn = bw * 4.0F;
            g1 = wn;
            f.n_in = 1;
            f.in_coef[0] = g1;
            f.n_out = 0;
            break;
        case 2:
            wn = bw * (8.0F * zeta) / (4.0F * zeta * zeta + 1.0F);
            g1 = wn * wn;
            g2 = wn * 2.0F * zeta;
            f.n_in = 2;
            f.in_coef[0] = (float)(g1 * T / 2.0 + g2);
            f.in_coef[1] = (float)(g1 * T / 2.0 - g2);
            f.n_out = 1;
            f.out_coef[0] = 1.0f;
            break;
        default:
            {
                wn = bw / 0.7845F;
                const float a3 = 1.1;
                const float b3 = 2.4;
                g1 = wn * wn * wn;
                g2 = a3 * wn * wn;
                g3 = b3 * wn;
                f.n_in = 3;
                f.in_coef[0] = (float)(g3 + T / 2.0 * (g2 + T / 2.0 * g1));
                f.in_coef[1] = (float)(g1 * T * T / 2.0 - 2.0 * g3);
                f.in_coef[2] = (float)(g3 + T / 2.0 * (-g2 + T / 2.0 * g1));
                f.n_out = 2;
                f.out_coef[0] = 2.0f;
                f.out_coef[1] = -1.0f;
            }
        }
    // initialize(0.0) (:262-267)
    f.idx = 3;
}

void carrier_filter_params(CarrierFilter& c, float fll_bw, float pll_bw, int order)
{
    std::memset(&c, 0, sizeof(c));
    c.order = order;
    if (order == 3)
        {
            c.pll_b3 = 2.400;
            c.pll_a3 = 1.100;
            c.pll_a2 = 1.414;
            c.pll_w0p = pll_bw / 0.7845F;
            c.pll_w0p2 = c.pll_w0p * c.pll_w0p;
            c.pll_w0p3 = c.pll_w0p2 * c.pll_w0p;
            c.pll_w0f = fll_bw / 0.53F;
            c.pll_w0f2 = c.pll_w0f * c.pll_w0f;
        }
    else
        {
            c.pll_a2 = 1.414;
            c.pll_w0p = pll_bw / 0.53F;
            c.pll_w0p2 = c.pll_w0p * c.pll_w0p;
            c.pll_w0f = fll_bw / 0.25F;
        }
}
}  // namespace
}  // namespace b200

using namespace b200;

struct gnss_b200_trk
{
    std::mutex mu;
    gnss_b200_trk_cfg cfg;
    TrkDevCfg dcfg;
    int device = 0;
    uint32_t n_taps = 3;
    cudaStream_t stream = nullptr;
    std::vector<DevBuf> codes, data_codes;
    std::vector<uint32_t> code_len;
    std::vector<uint8_t> started;
    DevBuf d_chans, d_list, d_params, d_shifts, d_corr, d_partials, d_tickets, d_log, d_stream;
    DevBuf d_params_data, d_corr_data, d_shifts_data, d_partials_data, d_tickets_data;  // slave data correlator (track_pilot)
    PinBuf h_log, h_chan;
    uint32_t splits = 1, smem_elems = 32;
    bool fast = false;
    uint64_t launches = 0;
};

extern "C"
{
    int gnss_b200_trk_create(const gnss_b200_trk_cfg* cfg, gnss_b200_trk** out)
    {
        B200_REQUIRE(cfg && out, GNSS_B200_ERR_INVALID, "trk_create: null argument");
        B200_REQUIRE(cfg->max_channels > 0 && cfg->vector_length > 0 && cfg->fs_in > 0 && cfg->code_chip_rate > 0 && cfg->code_length_chips > 0 &&
                         cfg->code_samples_per_chip >= 1 && cfg->code_period_s > 0,
            GNSS_B200_ERR_INVALID, "trk_create: bad geometry");
        B200_REQUIRE(cfg->pll_filter_order == 2 || cfg->pll_filter_order == 3, GNSS_B200_ERR_INVALID, "trk_create: pll_filter_order must be 2 or 3 (dll_pll_conf.cc:77-86)");
        B200_REQUIRE(cfg->dll_filter_order >= 1 && cfg->dll_filter_order <= 3, GNSS_B200_ERR_INVALID, "trk_create: dll_filter_order must be 1..3 (dll_pll_conf.cc:67-76)");
        B200_REQUIRE(cfg->cn0_samples >= 1 && cfg->cn0_samples <= kMaxCn0Samples, GNSS_B200_ERR_INVALID, "trk_create: cn0_samples must be 1..%d", kMaxCn0Samples);
        const int ms = (int)(cfg->code_period_s * 1000.0);
        const int cn0_init = ms > 0 ? cfg->cn0_smoother_samples / ms : cfg->cn0_smoother_samples;
        B200_REQUIRE(cn0_init <= kMaxSmootherInit && cfg->carrier_lock_test_smoother_samples <= kMaxSmootherInit, GNSS_B200_ERR_INVALID,
            "trk_create: smoother initialisation longer than %d samples", kMaxSmootherInit);
        B200_REQUIRE(cfg->slope != 0.0f || cfg->veml, GNSS_B200_ERR_INVALID, "trk_create: slope must not be 0");
        if (cfg->symbol_sync)
            {
                B200_REQUIRE(cfg->secondary_code_length <= (uint32_t)kMaxSecondary && cfg->data_secondary_code_length <= 111, GNSS_B200_ERR_INVALID,
                    "trk_create: secondary code longer than %d / 111 symbols", kMaxSecondary);
                B200_REQUIRE(!(cfg->secondary || cfg->symbols_per_bit > 1) || cfg->secondary_code_length > 0, GNSS_B200_ERR_INVALID,
                    "trk_create: symbol synchronisation needs secondary_code_length > 0");
                B200_REQUIRE(std::strlen(cfg->secondary_code_string) >= cfg->secondary_code_length && std::strlen(cfg->data_secondary_code_string) >= cfg->data_secondary_code_length,
                    GNSS_B200_ERR_INVALID, "trk_create: secondary code string shorter than its length");
            }
        else
            B200_REQUIRE(cfg->extend_correlation_symbols <= 1 && !cfg->track_pilot, GNSS_B200_ERR_INVALID,
                "trk_create: extended integration and pilot tracking need symbol_sync = 1");
        B200_REQUIRE(!cfg->high_dyn || (cfg->smoother_length >= 1 && cfg->smoother_length <= (uint32_t)kMaxSmootherLength), GNSS_B200_ERR_INVALID,
            "trk_create: smoother_length must be 1..%d", kMaxSmootherLength);
        int rc = select_device(cfg->device);
        if (rc) return rc;
        auto* h = new gnss_b200_trk();
        h->cfg = *cfg;
        h->device = cfg->device;
        h->n_taps = cfg->veml ? 5 : 3;
        TrkDevCfg& d = h->dcfg;
        d.fs_in = cfg->fs_in;
        d.code_chip_rate = cfg->code_chip_rate;
        d.signal_carrier_freq = cfg->signal_carrier_freq;
        d.code_period_s = cfg->code_period_s;
        d.carrier_lock_th = cfg->carrier_lock_th;
        d.code_length_chips = cfg->code_length_chips;
        d.code_samples_per_chip = cfg->code_samples_per_chip;
        d.vector_length = cfg->vector_length;
        d.n_taps = h->n_taps;
        d.veml = cfg->veml;
        d.enable_fll_pull_in = cfg->enable_fll_pull_in;
        d.enable_fll_steady_state = cfg->enable_fll_steady_state;
        d.carrier_aiding = cfg->carrier_aiding;
        d.pull_in_time_s = cfg->pull_in_time_s;
        d.cn0_samples = cfg->cn0_samples;
        d.cn0_min = cfg->cn0_min;
        d.max_code_lock_fail = cfg->max_code_lock_fail;
        d.max_carrier_lock_fail = cfg->max_carrier_lock_fail;
        d.spc = cfg->early_late_space_chips;  // d_trk_parameters.spc = early_late_space_chips in state 2 (:1858)
        d.slope = cfg->slope;
        d.y_intercept = cfg->y_intercept;
        d.symbol_sync = cfg->symbol_sync;
        d.secondary = cfg->secondary;
        d.ext = cfg->extend_correlation_symbols > 1 ? cfg->extend_correlation_symbols : 1;  // :512-520
        d.track_pilot = cfg->track_pilot;
        d.high_dyn = cfg->high_dyn;
        d.secondary_code_length = cfg->secondary_code_length;
        d.symbols_per_bit = cfg->symbols_per_bit;
        d.data_secondary_code_length = cfg->data_secondary_code_length;
        d.smoother_length = cfg->smoother_length;
        d.bit_synchronization_time_limit_s = cfg->bit_synchronization_time_limit_s;
        d.spc_narrow = cfg->early_late_space_narrow_chips;
        d.T_ext = (float)d.ext * (float)cfg->code_period_s;  // :1949
        {
            const float spc = (float)cfg->code_samples_per_chip;
            for (int k = 0; k < 5; k++) d.shifts_wide[k] = d.shifts_narrow[k] = 0.0f;
            if (cfg->veml)
                {
                    const float w[5] = {-cfg->very_early_late_space_chips * spc, -cfg->early_late_space_chips * spc, 0.0f, cfg->early_late_space_chips * spc,
                        cfg->very_early_late_space_chips * spc};
                    const float n[5] = {-cfg->very_early_late_space_narrow_chips * spc, -cfg->early_late_space_narrow_chips * spc, 0.0f,
                        cfg->early_late_space_narrow_chips * spc, cfg->very_early_late_space_narrow_chips * spc};
                    for (int k = 0; k < 5; k++)
                        {
                            d.shifts_wide[k] = w[k];
                            d.shifts_narrow[k] = n[k];
                        }
                }
            else
                {
                    d.shifts_wide[0] = -cfg->early_late_space_chips * spc;
                    d.shifts_wide[2] = cfg->early_late_space_chips * spc;
                    d.shifts_narrow[0] = -cfg->early_late_space_narrow_chips * spc;
                    d.shifts_narrow[2] = cfg->early_late_space_narrow_chips * spc;
                }
        }
        loop_filter_coefficients(d.lf_narrow, d.T_ext, cfg->dll_bw_narrow_hz, cfg->dll_filter_order);
        carrier_filter_params(d.cf_narrow, cfg->fll_bw_hz, cfg->pll_bw_narrow_hz, cfg->pll_filter_order);
        std::memset(d.secondary_code_string, 0, sizeof(d.secondary_code_string));
        std::memset(d.data_secondary_code_string, 0, sizeof(d.data_secondary_code_string));
        std::memcpy(d.secondary_code_string, cfg->secondary_code_string, std::min(sizeof(d.secondary_code_string), sizeof(cfg->secondary_code_string)));
        std::memcpy(d.data_secondary_code_string, cfg->data_secondary_code_string, std::min(sizeof(d.data_secondary_code_string), sizeof(cfg->data_secondary_code_string)));
        {
            float table[258];
            for (int k = 0; k < 256; k++) table[k] = (float)std::atan((double)k / 255.0);
            table[256] = table[257] = (float)std::atan(1.0);
            cudaError_t te = cudaMemcpyToSymbol(c_fast_atan_table, table, sizeof(table));
            if (te != cudaSuccess)
                {
                    delete h;
                    return fail(GNSS_B200_ERR_CUDA, "trk_create: %s", cudaGetErrorString(te));
                }
        }
        h->codes.resize(cfg->max_channels);
        h->data_codes.resize(cfg->max_channels);
        h->code_len.assign(cfg->max_channels, 0);
        h->started.assign(cfg->max_channels, 0);

        // launch shape of the correlator, fixed for the life of the handle: the chip window of a CTA is bounded by the
        // nominal code step (loop corrections are ppm-level) plus the tap span, with a generous margin
        const double step = cfg->code_chip_rate / cfg->fs_in * cfg->code_samples_per_chip;
        const double span_taps = 2.0 * std::max(std::max(cfg->early_late_space_chips, cfg->veml ? cfg->very_early_late_space_chips : 0.0f),
                                           std::max(cfg->early_late_space_narrow_chips, cfg->veml ? cfg->very_early_late_space_narrow_chips : 0.0f)) * cfg->code_samples_per_chip;
        uint32_t splits = 1;
        const uint32_t target = kNumSMs * 4;
        if (cfg->max_channels < target) splits = std::max(1u, std::min((target + cfg->max_channels - 1) / cfg->max_channels, std::max(1u, cfg->vector_length / 2048)));
        while (step * 1.01 * cfg->vector_length / splits + span_taps + 64.0 > 10800.0 && (uint64_t)splits * 2048 < cfg->vector_length) splits++;
        const double window = step * 1.01 * ((cfg->vector_length + splits - 1) / splits) + span_taps + 64.0;
        h->splits = splits;
        h->fast = !cfg->high_dyn && window + 8 <= (double)(mc::kSmemCapBytes - mc::kWinRingBytes) / 4 && step * cfg->vector_length + 16.0 < 4.0e6 &&
                  cfg->vector_length < (1u << 24) && (cfg->vector_length + splits - 1) / splits <= mc::kStMaxSegment;
        if (h->fast)
            h->smem_elems = (uint32_t)window;
        else
            {
                const uint32_t L = cfg->code_samples_per_chip * cfg->code_length_chips;
                h->smem_elems = std::min<uint32_t>(std::max<uint32_t>(L, 32u), mc::kSmemCapBytes / 4);
            }

        const uint32_t M = cfg->max_channels;
        cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = h->d_chans.reserve(sizeof(TrkChan) * M);
        if (e == cudaSuccess) e = cudaMemsetAsync(h->d_chans.p, 0, sizeof(TrkChan) * M, h->stream);
        if (e == cudaSuccess) e = h->d_list.reserve(sizeof(uint32_t) * M);
        if (e == cudaSuccess) e = h->d_params.reserve(sizeof(mc::ChanParams) * M);
        if (e == cudaSuccess) e = h->d_shifts.reserve(sizeof(float) * M * h->n_taps * 2);
        if (e == cudaSuccess && cfg->track_pilot)
            {
                e = h->d_params_data.reserve(sizeof(mc::ChanParams) * M);
                if (e == cudaSuccess) e = h->d_corr_data.reserve(sizeof(float2) * M);
                if (e == cudaSuccess) e = h->d_shifts_data.reserve(sizeof(float) * 8);
                if (e == cudaSuccess) e = cudaMemsetAsync(h->d_shifts_data.p, 0, sizeof(float) * 8, h->stream);  // the prompt shift, and shift_samples[0]
                if (e == cudaSuccess) e = h->d_partials_data.reserve(sizeof(float2) * (size_t)M * splits);
                if (e == cudaSuccess) e = h->d_tickets_data.reserve(sizeof(uint32_t) * M);
                if (e == cudaSuccess) e = cudaMemsetAsync(h->d_tickets_data.p, 0, sizeof(uint32_t) * M, h->stream);
            }
        if (e == cudaSuccess) e = h->d_corr.reserve(sizeof(float2) * M * h->n_taps);
        if (e == cudaSuccess) e = h->d_partials.reserve(sizeof(float2) * (size_t)M * splits * h->n_taps);
        if (e == cudaSuccess) e = h->d_tickets.reserve(sizeof(uint32_t) * M);
        if (e == cudaSuccess) e = cudaMemsetAsync(h->d_tickets.p, 0, sizeof(uint32_t) * M, h->stream);
        if (e == cudaSuccess) e = h->h_chan.reserve(sizeof(TrkChan));
        if (e == cudaSuccess)
            {
                // d_local_code_shift_chips (:830-841) for every channel slot; the second half is the (unused) shift_samples array
                std::vector<float> sh((size_t)M * h->n_taps * 2, 0.0f);
                const float spc = (float)cfg->code_samples_per_chip;
                for (uint32_t c = 0; c < M; c++)
                    {
                        float* s = sh.data() + (size_t)c * h->n_taps;
                        if (cfg->veml)
                            {
                                s[0] = -cfg->very_early_late_space_chips * spc;
                                s[1] = -cfg->early_late_space_chips * spc;
                                s[2] = 0.0f;
                                s[3] = cfg->early_late_space_chips * spc;
                                s[4] = cfg->very_early_late_space_chips * spc;
                            }
                        else
                            {
                                s[0] = -cfg->early_late_space_chips * spc;
                                s[1] = 0.0f;
                                s[2] = cfg->early_late_space_chips * spc;
                            }
                    }
                e = cudaMemcpyAsync(h->d_shifts.p, sh.data(), sh.size() * sizeof(float), cudaMemcpyHostToDevice, h->stream);
                if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
            }
        if (e != cudaSuccess)
            {
                gnss_b200_trk_destroy(h);
                return fail(GNSS_B200_ERR_CUDA, "trk_create: %s", cudaGetErrorString(e));
            }
        *out = h;
        return GNSS_B200_OK;
    }

    int gnss_b200_trk_destroy(gnss_b200_trk* h)
    {
        if (!h) return GNSS_B200_OK;
        cudaSetDevice(h->device);
        if (h->stream) cudaStreamSynchronize(h->stream);
        for (auto& c : h->codes) c.release();
        for (auto& c : h->data_codes) c.release();
        for (DevBuf* b : {&h->d_chans, &h->d_list, &h->d_params, &h->d_shifts, &h->d_corr, &h->d_partials, &h->d_tickets, &h->d_log, &h->d_stream,
                 &h->d_params_data, &h->d_corr_data, &h->d_shifts_data, &h->d_partials_data, &h->d_tickets_data})
            b->release();
        h->h_log.release();
        h->h_chan.release();
        if (h->stream) cudaStreamDestroy(h->stream);
        delete h;
        return GNSS_B200_OK;
    }

    int gnss_b200_trk_set_code(gnss_b200_trk* h, uint32_t ch, uint32_t code_len, const float* code)
    {
        B200_REQUIRE(h && code && code_len > 0, GNSS_B200_ERR_INVALID, "trk_set_code: bad arguments");
        B200_REQUIRE(ch < h->cfg.max_channels, GNSS_B200_ERR_INVALID, "trk_set_code: channel %u out of range", ch);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        B200_CUDA(h->codes[ch].reserve(mc::doubled_code_floats(code_len) * 4));
        B200_CUDA(mc::upload_doubled_code(h->codes[ch].p, code, code_len, h->stream));
        h->code_len[ch] = code_len;
        return GNSS_B200_OK;
    }

    int gnss_b200_trk_set_data_code(gnss_b200_trk* h, uint32_t ch, uint32_t code_len, const float* code)
    {
        B200_REQUIRE(h && code && code_len > 0, GNSS_B200_ERR_INVALID, "trk_set_data_code: bad arguments");
        B200_REQUIRE(ch < h->cfg.max_channels, GNSS_B200_ERR_INVALID, "trk_set_data_code: channel %u out of range", ch);
        B200_REQUIRE(h->cfg.track_pilot, GNSS_B200_ERR_STATE, "trk_set_data_code: the handle was not created with track_pilot");
        B200_REQUIRE(h->code_len[ch] == 0 || h->code_len[ch] == code_len, GNSS_B200_ERR_INVALID, "trk_set_data_code: length differs from the tracking code (%u)", h->code_len[ch]);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        B200_CUDA(h->data_codes[ch].reserve(mc::doubled_code_floats(code_len) * 4));
        B200_CUDA(mc::upload_doubled_code(h->data_codes[ch].p, code, code_len, h->stream));
        return GNSS_B200_OK;
    }

    int gnss_b200_trk_start(gnss_b200_trk* h, uint32_t ch, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now_sample)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "trk_start: null handle");
        B200_REQUIRE(ch < h->cfg.max_channels, GNSS_B200_ERR_INVALID, "trk_start: channel %u out of range", ch);
        B200_REQUIRE(h->codes[ch].p != nullptr, GNSS_B200_ERR_STATE, "trk_start: channel %u has no tracking code (call gnss_b200_trk_set_code)", ch);
        B200_REQUIRE(now_sample >= acq_samplestamp, GNSS_B200_ERR_INVALID, "trk_start: now_sample precedes the acquisition sample stamp");
        B200_REQUIRE(!h->cfg.track_pilot || h->data_codes[ch].p != nullptr, GNSS_B200_ERR_STATE, "trk_start: channel %u has no data code (call gnss_b200_trk_set_data_code)", ch);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        const gnss_b200_trk_cfg& c = h->cfg;
        TrkChan& s = *h->h_chan.as<TrkChan>();
        std::memset(&s, 0, sizeof(s));
        // start_tracking() (:640-660, :820-880)
        s.carrier_doppler_hz = acq_doppler_hz;
        s.carrier_phase_step_rad = kTwoPi * s.carrier_doppler_hz / c.fs_in;
        s.acq_sample_stamp = acq_samplestamp;
        s.carrier_lock_test = 1.0;
        carrier_filter_params(s.cf, c.fll_bw_hz, c.pll_bw_hz, c.pll_filter_order);
        loop_filter_coefficients(s.lf, (float)c.code_period_s, c.dll_bw_hz, c.dll_filter_order);
        if (c.pll_filter_order == 3)  // Tracking_FLL_PLL_filter::initialize (:58-70)
            {
                s.cf.pll_x = 2.0F * (float)acq_doppler_hz;
                s.cf.pll_w = 0;
            }
        else
            {
                s.cf.pll_w = (float)acq_doppler_hz;
                s.cf.pll_x = 0;
            }
        s.pull_in = 1;
        // state 1, pull-in alignment (:1812-1845)
        const int64_t acq_trk_diff_samples = (int64_t)now_sample - (int64_t)acq_samplestamp;
        const double delta = (double)acq_trk_diff_samples - acq_delay_samples;
        s.code_freq_chips = c.code_chip_rate;
        s.code_phase_step_chips = s.code_freq_chips / c.fs_in;
        const double T_chip_mod = 1.0 / s.code_freq_chips;
        const double T_prn_mod_s = T_chip_mod * (double)c.code_length_chips;
        const double T_prn_mod_samples = T_prn_mod_s * c.fs_in;
        const double acq_code_phase_samples = T_prn_mod_samples - std::fmod(delta, T_prn_mod_samples);
        s.current_prn_length_samples = (int)std::round(T_prn_mod_samples);
        const int32_t samples_offset = (int32_t)std::round(acq_code_phase_samples);
        s.acc_carrier_phase_rad -= s.carrier_phase_step_rad * (double)samples_offset;
        s.state = 2;
        s.cursor = now_sample + (uint64_t)samples_offset;  // consume_each(samples_offset)
        // smoothers (:538-550) reset at the state change (:1835-1836)
        const int ms = (int)(c.code_period_s * 1000.0);
        auto init_sm = [](Smoother& m, float alpha, float min_value, float offset, int n) {
            m.alpha = alpha < 0 ? 0 : (alpha > 1 ? 1 : alpha);
            m.one_minus_alpha = 1.0F - m.alpha;
            m.old_value = 0.0f;
            m.min_value = min_value;
            m.offset = offset;
            m.samples_for_initialization = n <= 0 ? 1 : n;
            m.init_counter = 0;
            m.initializing = 1;
        };
        init_sm(s.cn0_s, c.cn0_smoother_alpha, 25.0f, 12.0f, ms > 0 ? c.cn0_smoother_samples / ms : 200);
        init_sm(s.lock_s, c.carrier_lock_test_smoother_alpha, -1.0f, 0.0f, c.carrier_lock_test_smoother_samples);
        s.code = h->codes[ch].as<float>();
        s.code_len = h->code_len[ch];
        s.data_code = h->cfg.track_pilot ? h->data_codes[ch].as<float>() : nullptr;
        s.current_correlation_time_s = c.code_period_s;  // :842
        s.spc = c.early_late_space_chips;
        s.cloop = 1;                                     // :872
        B200_CUDA(cudaMemcpyAsync(h->d_chans.as<TrkChan>() + ch, &s, sizeof(TrkChan), cudaMemcpyHostToDevice, h->stream));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        h->started[ch] = 1;
        return GNSS_B200_OK;
    }

    int gnss_b200_trk_run(gnss_b200_trk* h, const float* samples_c64, int samples_is_device, uint64_t stream_start, uint64_t n_samples, uint32_t n_ch,
        const uint32_t* ch, uint32_t max_epochs, gnss_b200_trk_epoch* out)
    {
        B200_REQUIRE(h && samples_c64 && ch && out, GNSS_B200_ERR_INVALID, "trk_run: null argument");
        B200_REQUIRE(n_ch > 0 && n_ch <= h->cfg.max_channels && max_epochs > 0 && n_samples > 0, GNSS_B200_ERR_INVALID, "trk_run: bad sizes");
        for (uint32_t i = 0; i < n_ch; i++)
            {
                B200_REQUIRE(ch[i] < h->cfg.max_channels, GNSS_B200_ERR_INVALID, "trk_run: channel %u out of range", ch[i]);
                B200_REQUIRE(h->started[ch[i]], GNSS_B200_ERR_STATE, "trk_run: channel %u was not started (call gnss_b200_trk_start)", ch[i]);
            }
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        const float2* d_stream = reinterpret_cast<const float2*>(samples_c64);
        if (!samples_is_device)
            {
                B200_CUDA(h->d_stream.reserve(n_samples * 8));
                B200_CUDA(cudaMemcpyAsync(h->d_stream.p, samples_c64, n_samples * 8, cudaMemcpyHostToDevice, h->stream));
                d_stream = h->d_stream.as<float2>();
            }
        const size_t log_bytes = sizeof(gnss_b200_trk_epoch) * (size_t)max_epochs * n_ch;
        B200_CUDA(h->d_log.reserve(log_bytes));
        B200_CUDA(h->h_log.reserve(log_bytes));
        B200_CUDA(cudaMemcpyAsync(h->d_list.p, ch, sizeof(uint32_t) * n_ch, cudaMemcpyHostToDevice, h->stream));
        B200_CUDA(cudaMemsetAsync(h->d_log.p, 0, log_bytes, h->stream));
        // one thread per channel, mostly FP64 (the reference's loop arithmetic is double): an SM has few FP64 lanes, so the channels are
        // spread over as many SMs as there are (2048 channels: 128 CTAs of 16 threads, 23 -> 10 us per epoch against 16 CTAs of 128)
        const uint32_t tb = n_ch >= 128u * kNumSMs ? 128u : (n_ch >= 32u * kNumSMs ? 32u : 16u);
        const uint32_t blocks = (n_ch + tb - 1) / tb;
        const unsigned long long s0 = stream_start, s1 = stream_start + n_samples;
        mc::ChanParams* d_cp = h->d_params.as<mc::ChanParams>();
        const float* d_shifts = h->d_shifts.as<float>();
        float* d_shifts_w = h->d_shifts.as<float>();
        uint32_t* d_ss = reinterpret_cast<uint32_t*>(d_shifts_w + (size_t)h->cfg.max_channels * h->n_taps);
        const bool pilot = h->cfg.track_pilot != 0;
        mc::ChanParams* d_cp_data = pilot ? h->d_params_data.as<mc::ChanParams>() : nullptr;
        const float2* d_corr_data = pilot ? h->d_corr_data.as<float2>() : nullptr;
        trk_update_kernel<<<blocks, tb, 0, h->stream>>>(h->dcfg, h->d_chans.as<TrkChan>(), h->d_list.as<uint32_t>(), n_ch, nullptr, nullptr, d_cp, d_cp_data, d_shifts_w, d_ss,
            nullptr, d_stream, s0, s1, 1);
        h->launches++;
        for (uint32_t e = 0; e < max_epochs; e++)
            {
                B200_CUDA(mc::launch_real_plain(h->fast, n_ch, h->splits, h->smem_elems, h->stream, d_cp, d_shifts, d_ss, h->d_partials.as<float2>(),
                    h->d_tickets.as<uint32_t>(), h->d_corr.as<float2>(), h->n_taps));
                if (pilot)
                    {
                        // d_correlator_data_cpu (:1078-1088): one tap at the prompt shift on the data code, same NCO arguments
                        B200_CUDA(mc::launch_real_plain(h->fast, n_ch, h->splits, h->smem_elems, h->stream, d_cp_data, h->d_shifts_data.as<float>(),
                            reinterpret_cast<const uint32_t*>(h->d_shifts_data.as<float>() + 4), h->d_partials_data.as<float2>(), h->d_tickets_data.as<uint32_t>(),
                            h->d_corr_data.as<float2>(), 1));
                        h->launches++;
                    }
                trk_update_kernel<<<blocks, tb, 0, h->stream>>>(h->dcfg, h->d_chans.as<TrkChan>(), h->d_list.as<uint32_t>(), n_ch, h->d_corr.as<float2>(), d_corr_data, d_cp,
                    d_cp_data, d_shifts_w, d_ss, h->d_log.as<gnss_b200_trk_epoch>() + (size_t)e * n_ch, d_stream, s0, s1, 0);
                h->launches += 2;
            }
        B200_CUDA(cudaGetLastError());
        B200_CUDA(cudaMemcpyAsync(h->h_log.p, h->d_log.p, log_bytes, cudaMemcpyDeviceToHost, h->stream));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        std::memcpy(out, h->h_log.p, log_bytes);
        return GNSS_B200_OK;
    }

    void* gnss_b200_trk_stream(gnss_b200_trk* h) { return h ? (void*)h->stream : nullptr; }
    uint64_t gnss_b200_trk_launch_count(const gnss_b200_trk* h) { return h ? h->launches : 0; }
}
