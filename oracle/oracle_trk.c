/*
 * TEST INFRASTRUCTURE — CPU oracle, not product code. Nothing under gnss-sdr-new_b200/ may include or link this.
 *
 * Plain-C restatement of the per-epoch loop of the reference's dll_pll_veml tracking block
 * (src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc): state 2 (one code period per integration, Costas PLL) and,
 * with cfg.symbol_sync, the whole symbol machine — secondary-code / bit synchronisation (:1888-1925, acquire_secondary :923-966),
 * state 3 coherent extension (:1989-2027), state 4 narrow tracking (:2028-2084), save_correlation_results (:1290-1400), the
 * pilot + data correlator pair (:1078-1088) and the high-dynamics rate smoothing (:1232-1284):
 *   start_tracking            :640-880      state 1 alignment      :1812-1845
 *   do_correlation_step       :1064-1089    (oracle_mc_real = Cpu_Multicorrelator_Real_Codes)
 *   cn0_and_tracking_lock_status :968-1013  run_dll_pll            :1092-1182
 *   update_tracking_vars      :1225-1297    consume_each           :2118
 * and of the helper classes it uses:
 *   Tracking_loop_filter      src/algorithms/tracking/libs/tracking_loop_filter.cc:57-210,262-267
 *   Tracking_FLL_PLL_filter   tracking_FLL_PLL_filter.cc:23-97
 *   Exponential_Smoother      exponential_smoother.cc:63-101
 *   discriminators            tracking_discriminators.cc:27-149
 *   cn0_m2m4_estimator, carrier_lock_detector  lock_detectors.cc:98-148
 * Pinned bit-for-bit against those reference files compiled unmodified into oracle/_ref/libref_trk.so
 * (tests/test_oracle_trk.py). The reference's unit tests hold no numeric expectations for this loop
 * (SURVEY.md §4): beyond that component pin, parity of the closed loop is "unpinned".
 * pll_four_quadrant_atan (pilot tracking) calls gr::fast_atan2f, which lives in GNU Radio (gnuradio-runtime/lib/math/
 * fast_atan2f.cc, GNU Radio 3.10; not vendored under /root/reference): restated here from its published algorithm (a 255-step
 * arctangent table with linear interpolation, octant folding) — parity of that one function is "unpinned".
 */
#include "oracle_trk.h"
#include "oracle_mc.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* MATH_CONSTANTS.h:47-49 — the reference's own (truncated) pi */
#define GNSS_PI 3.1415926535898
#define HALF_PI (GNSS_PI / 2.0)
#define TWO_PI (2.0 * GNSS_PI)

/* ---------------- Tracking_loop_filter ---------------- */
void oracle_loop_filter_init(oracle_loop_filter* f, float update_interval, float noise_bandwidth, int order, int include_last_integrator)
{
    float g1, g2, g3, wn;
    const float T = update_interval;
    const float zeta = 1.0F / sqrtf(2.0F);
    memset(f, 0, sizeof(*f));
    switch (order)
        {
        case 1:
            wn = noise_bandwidth * 4.0F;
            g1 = wn;
            if (include_last_integrator)
                {
                    f->n_in = 2;
                    f->in_coef[0] = g1 * T / 2.0;
                    f->in_coef[1] = g1 * T / 2.0;
                    f->n_out = 1;
                    f->out_coef[0] = 1.0;
                }
            else
                {
                    f->n_in = 1;
                    f->in_coef[0] = g1;
                    f->n_out = 0;
                }
            break;
        case 2:
            wn = noise_bandwidth * (8.0F * zeta) / (4.0F * zeta * zeta + 1.0F);
            g1 = wn * wn;
            g2 = wn * 2.0F * zeta;
            if (include_last_integrator)
                {
                    f->n_in = 3;
                    f->in_coef[0] = T / 2.0 * (g1 * T / 2.0 + g2);
                    f->in_coef[1] = T * T / 2.0 * g1;
                    f->in_coef[2] = T / 2.0 * (g1 * T / 2.0 - g2);
                    f->n_out = 2;
                    f->out_coef[0] = 2.0;
                    f->out_coef[1] = -1.0;
                }
            else
                {
                    f->n_in = 2;
                    f->in_coef[0] = (g1 * T / 2.0 + g2);
                    f->in_coef[1] = g1 * T / 2.0 - g2;
                    f->n_out = 1;
                    f->out_coef[0] = 1.0;
                }
            break;
        default:
            {
                const float a3 = 1.1;
                const float b3 = 2.4;
                wn = noise_bandwidth / 0.7845F;
                g1 = wn * wn * wn;
                g2 = a3 * wn * wn;
                g3 = b3 * wn;
                if (include_last_integrator)
                    {
                        f->n_in = 4;
                        f->in_coef[0] = T / 2.0 * (g3 + T / 2.0 * (g2 + T / 2.0 * g1));
                        f->in_coef[1] = T / 2.0 * (-g3 + T / 2.0 * (g2 + 3.0 * T / 2.0 * g1));
                        f->in_coef[2] = T / 2.0 * (-g3 - T / 2.0 * (g2 - 3.0 * T / 2.0 * g1));
                        f->in_coef[3] = T / 2.0 * (g3 - T / 2.0 * (g2 - T / 2.0 * g1));
                        f->n_out = 3;
                        f->out_coef[0] = 3.0;
                        f->out_coef[1] = -3.0;
                        f->out_coef[2] = 1.0;
                    }
                else
                    {
                        f->n_in = 3;
                        f->in_coef[0] = g3 + T / 2.0 * (g2 + T / 2.0 * g1);
                        f->in_coef[1] = g1 * T * T / 2.0 - 2.0 * g3;
                        f->in_coef[2] = g3 + T / 2.0 * (-g2 + T / 2.0 * g1);
                        f->n_out = 2;
                        f->out_coef[0] = 2.0;
                        f->out_coef[1] = -1.0;
                    }
            }
        }
    f->idx = 0; /* constructor state; initialize() below is what start_tracking calls */
}

void oracle_loop_filter_initialize(oracle_loop_filter* f, float initial_output)
{
    for (int i = 0; i < 4; i++)
        {
            f->inputs[i] = 0.0f;
            f->outputs[i] = initial_output;
        }
    f->idx = 3;
}

float oracle_loop_filter_apply(oracle_loop_filter* f, float x)
{
    float result = 0.0f;
    for (int i = 0; i < f->n_out; i++) result += f->out_coef[i] * f->outputs[(f->idx + i) % 4];
    f->idx--;
    if (f->idx < 0) f->idx += 4;
    f->inputs[f->idx] = x;
    for (int i = 0; i < f->n_in; i++) result += f->in_coef[i] * f->inputs[(f->idx + i) % 4];
    f->outputs[f->idx] = result;
    return result;
}

/* ---------------- Tracking_FLL_PLL_filter ---------------- */
void oracle_carrier_filter_set_params(oracle_carrier_filter* c, float fll_bw_hz, float pll_bw_hz, int order)
{
    c->order = order;
    if (order == 3)
        {
            c->pll_b3 = 2.400;
            c->pll_a3 = 1.100;
            c->pll_a2 = 1.414;
            c->pll_w0p = pll_bw_hz / 0.7845F;
            c->pll_w0p2 = c->pll_w0p * c->pll_w0p;
            c->pll_w0p3 = c->pll_w0p2 * c->pll_w0p;
            c->pll_w0f = fll_bw_hz / 0.53F;
            c->pll_w0f2 = c->pll_w0f * c->pll_w0f;
        }
    else
        {
            c->pll_a2 = 1.414;
            c->pll_w0p = pll_bw_hz / 0.53F;
            c->pll_w0p2 = c->pll_w0p * c->pll_w0p;
            c->pll_w0f = fll_bw_hz / 0.25F;
        }
}

void oracle_carrier_filter_initialize(oracle_carrier_filter* c, float acq_doppler_hz)
{
    if (c->order == 3)
        {
            c->pll_x = 2.0F * acq_doppler_hz;
            c->pll_w = 0;
        }
    else
        {
            c->pll_w = acq_doppler_hz;
            c->pll_x = 0;
        }
}

float oracle_carrier_filter_get_carrier_error(oracle_carrier_filter* c, float FLL, float PLL, float T)
{
    float err;
    if (c->order == 3)
        {
            c->pll_w = c->pll_w + T * (c->pll_w0p3 * PLL + c->pll_w0f2 * FLL);
            c->pll_x = c->pll_x + T * (0.5F * c->pll_w + c->pll_a2 * c->pll_w0f * FLL + c->pll_a3 * c->pll_w0p2 * PLL);
            err = 0.5F * c->pll_x + c->pll_b3 * c->pll_w0p * PLL;
        }
    else
        {
            const float w_new = c->pll_w + PLL * c->pll_w0p2 * T + FLL * c->pll_w0f * T;
            err = 0.5F * (w_new + c->pll_w) + c->pll_a2 * c->pll_w0p * PLL;
            c->pll_w = w_new;
        }
    return err;
}

/* ---------------- Exponential_Smoother ---------------- */
void oracle_smoother_init(oracle_smoother* s, float alpha, float min_value, float offset, int samples_for_initialization)
{
    memset(s, 0, sizeof(*s));
    s->alpha = alpha < 0 ? 0 : (alpha > 1 ? 1 : alpha);
    s->one_minus_alpha = 1.0F - s->alpha;
    s->min_value = min_value;
    s->offset = offset;
    s->samples_for_initialization = samples_for_initialization <= 0 ? 1 : samples_for_initialization;
    s->initializing = 1;
}

float oracle_smoother_smooth(oracle_smoother* s, float raw)
{
    float v;
    if (s->initializing)
        {
            s->init_counter++;
            v = raw;
            if (s->init_counter <= ORACLE_TRK_MAX_SMOOTHER_INIT) s->init_buffer[s->init_counter - 1] = v;
            if (s->init_counter == s->samples_for_initialization)
                {
                    float acc = 0.0F;
                    for (int i = 0; i < s->init_counter; i++) acc += s->init_buffer[i];
                    s->old_value = acc / (float)s->init_counter;
                    if (s->old_value < (s->min_value + s->offset))
                        s->init_counter = 0;
                    else
                        s->initializing = 0;
                }
        }
    else
        {
            v = s->alpha * raw + s->one_minus_alpha * s->old_value;
            s->old_value = v;
        }
    return v;
}

/* ---------------- discriminators / lock detectors ---------------- */
static double phase_unwrap(double p)
{
    if (p >= HALF_PI) return p - GNSS_PI;
    if (p <= -HALF_PI) return p + GNSS_PI;
    return p;
}

double oracle_fll_diff_atan(float s1_re, float s1_im, float s2_re, float s2_im, double t1, double t2)
{
    double diff_atan = atanf(s2_im / s2_re) - atanf(s1_im / s1_re); /* float - float, then widened */
    if (isnan(diff_atan)) diff_atan = 0;
    return phase_unwrap(diff_atan) / (t2 - t1);
}

double oracle_pll_cloop_two_quadrant_atan(float re, float im)
{
    if (re != 0.0) return (double)atanf(im / re);
    return 0.0;
}

/* gr::fast_atan2f: |error| < 1e-5 rad; table[i] = atan(i / 255), two guard entries at pi/4 */
float oracle_fast_atan2f(float y, float x)
{
    static float table[258];
    static int ready = 0;
    if (!ready)
        {
            for (int i = 0; i < 256; i++) table[i] = (float)atan((double)i / 255.0);
            table[256] = table[257] = (float)atan(1.0);
            ready = 1;
        }
    const float TAN_MAP_RES = 0.003921569f; /* 1 / 255 */
    const float y_abs = fabsf(y), x_abs = fabsf(x);
    if (!((y_abs > 0.0f) || (x_abs > 0.0f))) return 0.0f;
    const float z = (y_abs < x_abs) ? y_abs / x_abs : x_abs / y_abs;
    float base_angle;
    if (z < TAN_MAP_RES)
        base_angle = z;
    else
        {
            float alpha = z * 255.0f;
            const int index = ((int)alpha) & 0xff;
            alpha -= (float)index;
            base_angle = table[index] + (table[index + 1] - table[index]) * alpha;
        }
    float angle;
    if (x_abs > y_abs)
        {
            if (x >= 0.0f)
                angle = (y >= 0.0f) ? base_angle : -base_angle;
            else
                {
                    angle = (float)M_PI;
                    angle = (y >= 0.0f) ? angle - base_angle : base_angle - angle;
                }
        }
    else
        {
            if (y >= 0.0f)
                {
                    angle = (float)(M_PI / 2.0);
                    angle = (x >= 0.0f) ? angle - base_angle : angle + base_angle;
                }
            else
                {
                    angle = -(float)(M_PI / 2.0);
                    angle = (x >= 0.0f) ? angle + base_angle : angle - base_angle;
                }
        }
    return angle;
}

/* tracking_discriminators.cc:86-89 */
double oracle_pll_four_quadrant_atan(float re, float im) { return (double)oracle_fast_atan2f(im, re); }

double oracle_dll_nc_e_minus_l_normalized(float e_re, float e_im, float l_re, float l_im, float spc, float slope, float y_intercept)
{
    const double P_early = hypotf(e_re, e_im); /* std::abs(std::complex<float>) */
    const double P_late = hypotf(l_re, l_im);
    const double E_plus_L = P_early + P_late;
    if (E_plus_L == 0.0) return 0.0;
    return ((y_intercept - slope * spc) / slope) * (P_early - P_late) / E_plus_L;
}

double oracle_dll_nc_vemlp_normalized(const float* ve, const float* e, const float* l, const float* vl)
{
    const double Early = sqrtf(ve[0] * ve[0] + ve[1] * ve[1] + e[0] * e[0] + e[1] * e[1]);
    const double Late = sqrtf(l[0] * l[0] + l[1] * l[1] + vl[0] * vl[0] + vl[1] * vl[1]);
    const double E_plus_L = Early + Late;
    if (E_plus_L == 0.0) return 0.0;
    return (Early - Late) / E_plus_L;
}

float oracle_cn0_m2m4_estimator(const float* prompt_c64, int length, float coh_integration_time_s)
{
    float SNR_aux, Psig = 0.0, m_2 = 0.0, m_4 = 0.0, aux;
    const float n = (float)length;
    for (int i = 0; i < length; i++)
        {
            Psig += fabsf(prompt_c64[2 * i]);
            aux = prompt_c64[2 * i + 1] * prompt_c64[2 * i + 1] + prompt_c64[2 * i] * prompt_c64[2 * i];
            m_2 += aux;
            m_4 += (aux * aux);
        }
    Psig /= n;
    Psig = Psig * Psig;
    m_2 /= n;
    m_4 /= n;
    aux = sqrtf(2.0F * m_2 * m_2 - m_4);
    if (isnan(aux))
        SNR_aux = Psig / (m_2 - Psig);
    else
        SNR_aux = aux / (m_2 - aux);
    return 10.0F * log10f(SNR_aux) - 10.0F * log10f(coh_integration_time_s);
}

float oracle_carrier_lock_detector(const float* prompt_c64, int length)
{
    float I = 0.0, Q = 0.0;
    for (int i = 0; i < length; i++)
        {
            I += prompt_c64[2 * i];
            Q += prompt_c64[2 * i + 1];
        }
    return (I * I - Q * Q) / (I * I + Q * Q);
}

/* ---------------- the block's loop ---------------- */
#define ORACLE_TRK_MAX_SECONDARY 208
#define ORACLE_TRK_MAX_HISTORY 64
struct oracle_trk
{
    oracle_trk_cfg cfg;
    int n_taps;
    float shifts[5];
    float* code;
    int code_len;
    float* data_code;
    int data_code_len;
    /* members of dll_pll_veml_tracking (dll_pll_veml_tracking.h:118-204) */
    double carrier_doppler_hz, carrier_phase_step_rad, code_freq_chips, code_phase_step_chips;
    double carrier_phase_rate_step_rad, code_phase_rate_step_chips;
    double rem_code_phase_samples, rem_code_phase_chips, acc_carrier_phase_rad;
    double carr_phase_error_hz, carr_freq_error_hz, carr_error_filt_hz, code_error_chips, code_error_filt_chips;
    double cn0_db_hz, carrier_lock_test, evm, current_correlation_time_s;
    uint64_t cursor, acq_sample_stamp;
    float rem_carr_phase_rad, spc;
    float P_old[2];
    float accu[10];      /* d_VE_accu, d_E_accu, d_P_accu, d_L_accu, d_VL_accu */
    float P_data_accu[2];
    int current_prn_length_samples, state, pull_in, cloop, acc_carrier_phase_initialized;
    int carrier_lock_fail, code_lock_fail, cn0_counter;
    int current_symbol, current_data_symbol, extend_count, pll_180;
    /* boost::circular_buffer<gr_complex> d_Prompt_circular_buffer: oldest first, push_back on a full buffer drops the oldest */
    float circ[2 * ORACLE_TRK_MAX_SECONDARY];
    int circ_size;
    /* boost::circular_buffer<std::pair<double,double>> d_carr_ph_history / d_code_ph_history, capacity 2 * smoother_length */
    double carr_hist[2 * ORACLE_TRK_MAX_HISTORY], code_hist[2 * ORACLE_TRK_MAX_HISTORY];
    int carr_hist_size, code_hist_size;
    oracle_carrier_filter cf;
    oracle_loop_filter lf;
    float lf_update_interval, lf_noise_bandwidth;
    oracle_smoother cn0_s, lock_s;
    float prompt_buf[2 * ORACLE_TRK_MAX_CN0_SAMPLES];
};

oracle_trk* oracle_trk_create(const oracle_trk_cfg* cfg)
{
    oracle_trk* t = (oracle_trk*)calloc(1, sizeof(oracle_trk));
    t->cfg = *cfg;
    t->n_taps = cfg->veml ? 5 : 3;
    if (t->cfg.extend_correlation_symbols < 1) t->cfg.extend_correlation_symbols = 1; /* :512-520 */
    return t;
}

void oracle_trk_destroy(oracle_trk* t)
{
    if (!t) return;
    free(t->code);
    free(t->data_code);
    free(t);
}

void oracle_trk_set_code(oracle_trk* t, const float* code, int code_len)
{
    free(t->code);
    t->code = (float*)malloc(sizeof(float) * (size_t)code_len);
    memcpy(t->code, code, sizeof(float) * (size_t)code_len);
    t->code_len = code_len;
}

void oracle_trk_set_data_code(oracle_trk* t, const float* code, int code_len)
{
    free(t->data_code);
    t->data_code = (float*)malloc(sizeof(float) * (size_t)code_len);
    memcpy(t->data_code, code, sizeof(float) * (size_t)code_len);
    t->data_code_len = code_len;
}

static void wide_shifts(oracle_trk* t)
{
    const oracle_trk_cfg* c = &t->cfg;
    const float spc = (float)c->code_samples_per_chip;
    if (c->veml)
        {
            t->shifts[0] = -c->very_early_late_space_chips * spc;
            t->shifts[1] = -c->early_late_space_chips * spc;
            t->shifts[2] = 0.0f;
            t->shifts[3] = c->early_late_space_chips * spc;
            t->shifts[4] = c->very_early_late_space_chips * spc;
        }
    else
        {
            t->shifts[0] = -c->early_late_space_chips * spc;
            t->shifts[1] = 0.0f;
            t->shifts[2] = c->early_late_space_chips * spc;
        }
}

/* clear_tracking_vars (:1192-1213) except the correlator outputs, which the next correlation overwrites */
static void clear_tracking_vars(oracle_trk* t)
{
    t->P_data_accu[0] = t->P_data_accu[1] = 0.0f;
    t->P_old[0] = t->P_old[1] = 0.0f;
    t->carr_phase_error_hz = t->carr_freq_error_hz = t->carr_error_filt_hz = 0.0;
    t->code_error_chips = t->code_error_filt_chips = 0.0;
    t->current_symbol = t->current_data_symbol = 0;
    t->circ_size = 0;
    t->carrier_phase_rate_step_rad = 0.0;
    t->code_phase_rate_step_chips = 0.0;
    t->carr_hist_size = t->code_hist_size = 0;
}

void oracle_trk_start(oracle_trk* t, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now_sample)
{
    const oracle_trk_cfg* c = &t->cfg;
    /* start_tracking() */
    t->carrier_doppler_hz = acq_doppler_hz;
    t->carrier_phase_step_rad = TWO_PI * t->carrier_doppler_hz / c->fs_in;
    t->carrier_phase_rate_step_rad = 0.0;
    t->code_phase_rate_step_chips = 0.0; /* constructor :135; clear_tracking_vars afterwards */
    t->carr_hist_size = t->code_hist_size = 0;
    t->acq_sample_stamp = acq_samplestamp;
    t->carrier_lock_fail = t->code_lock_fail = 0;
    t->rem_code_phase_samples = 0.0;
    t->rem_carr_phase_rad = 0.0f;
    t->rem_code_phase_chips = 0.0;
    t->acc_carrier_phase_rad = 0.0;
    t->cn0_counter = 0;
    t->carrier_lock_test = 1.0;
    t->evm = 0.0; /* :845 */
    t->cn0_db_hz = 0.0;
    t->P_old[0] = t->P_old[1] = 0.0f;
    t->carr_phase_error_hz = t->carr_freq_error_hz = t->carr_error_filt_hz = t->code_error_chips = t->code_error_filt_chips = 0.0;
    memset(t->accu, 0, sizeof(t->accu));
    t->P_data_accu[0] = t->P_data_accu[1] = 0.0f;
    t->current_symbol = t->current_data_symbol = t->extend_count = 0;
    t->pll_180 = 0;
    wide_shifts(t);
    t->spc = c->early_late_space_chips;
    t->current_correlation_time_s = c->code_period_s;
    oracle_carrier_filter_set_params(&t->cf, c->fll_bw_hz, c->pll_bw_hz, c->pll_filter_order);
    t->lf_update_interval = (float)c->code_period_s;
    t->lf_noise_bandwidth = c->dll_bw_hz;
    oracle_loop_filter_init(&t->lf, t->lf_update_interval, t->lf_noise_bandwidth, c->dll_filter_order, 0);
    oracle_carrier_filter_initialize(&t->cf, (float)acq_doppler_hz);
    oracle_loop_filter_initialize(&t->lf, 0.0f);
    t->pull_in = 1;
    t->cloop = 1;
    t->circ_size = 0;
    t->acc_carrier_phase_initialized = 0;
    /* state 1: alignment with the local replica */
    {
        const int64_t acq_trk_diff_samples = (int64_t)now_sample - (int64_t)acq_samplestamp;
        const double delta_trk_to_acq_prn_start_samples = (double)acq_trk_diff_samples - acq_delay_samples;
        t->code_freq_chips = c->code_chip_rate;
        t->code_phase_step_chips = t->code_freq_chips / c->fs_in;
        t->code_phase_rate_step_chips = 0.0;
        const double T_chip_mod_seconds = 1.0 / t->code_freq_chips;
        const double T_prn_mod_seconds = T_chip_mod_seconds * (double)c->code_length_chips;
        const double T_prn_mod_samples = T_prn_mod_seconds * c->fs_in;
        const double acq_code_phase_samples = T_prn_mod_samples - fmod(delta_trk_to_acq_prn_start_samples, T_prn_mod_samples);
        t->current_prn_length_samples = (int)round(T_prn_mod_samples);
        const int32_t samples_offset = (int32_t)round(acq_code_phase_samples);
        t->acc_carrier_phase_rad -= t->carrier_phase_step_rad * (double)samples_offset;
        t->state = 2;
        t->cursor = now_sample + (uint64_t)samples_offset;
    }
    {
        const int ms = (int)(c->code_period_s * 1000.0);
        oracle_smoother_init(&t->cn0_s, c->cn0_smoother_alpha, 25.0f, 12.0f, c->code_period_s > 0.0 ? c->cn0_smoother_samples / ms : 200);
        oracle_smoother_init(&t->lock_s, c->carrier_lock_test_smoother_alpha, -1.0f, 0.0f, c->carrier_lock_test_smoother_samples);
    }
}

/* cn0_and_tracking_lock_status(coh_integration_time_s) (:968-1056): the prompt buffer takes d_P_accu */
static int cn0_and_tracking_lock_status(oracle_trk* t, const float* P, double coh_integration_time_s)
{
    const oracle_trk_cfg* c = &t->cfg;
    if (t->cn0_counter < c->cn0_samples)
        {
            t->prompt_buf[2 * t->cn0_counter] = P[0];
            t->prompt_buf[2 * t->cn0_counter + 1] = P[1];
            t->cn0_counter++;
            return 1;
        }
    t->prompt_buf[2 * (t->cn0_counter % c->cn0_samples)] = P[0];
    t->prompt_buf[2 * (t->cn0_counter % c->cn0_samples) + 1] = P[1];
    t->cn0_counter++;
    const float raw = oracle_cn0_m2m4_estimator(t->prompt_buf, c->cn0_samples, (float)coh_integration_time_s);
    t->cn0_db_hz = (double)oracle_smoother_smooth(&t->cn0_s, raw);
    t->carrier_lock_test = (double)oracle_smoother_smooth(&t->lock_s, oracle_carrier_lock_detector(t->prompt_buf, 1));
    if (!t->pull_in)
        {
            if (t->carrier_lock_test < c->carrier_lock_th)
                t->carrier_lock_fail++;
            else if (t->carrier_lock_fail > 0)
                t->carrier_lock_fail--;
            if (t->cn0_db_hz < c->cn0_min)
                t->code_lock_fail++;
            else if (t->code_lock_fail > 0)
                t->code_lock_fail--;
        }
    if (t->carrier_lock_fail > c->max_carrier_lock_fail || t->code_lock_fail > c->max_code_lock_fail)
        {
            t->carrier_lock_fail = 0;
            t->code_lock_fail = 0;
            return 0;
        }
    /* EVM indicator of the fork (dll_pll_veml_tracking.cc:1027-1053): prompt buffer scaled by the RMS of its real parts,
     * distance of (|I|, |Q|) from (1, 0); binary32 sums in buffer order, reached only when the lock test passed */
    {
        const float Prompt_I_ref = 1, Prompt_Q_ref = 0;
        float tmp_sum = 0;
        for (int i = 0; i < c->cn0_samples; i++) tmp_sum = tmp_sum + t->prompt_buf[2 * i] * t->prompt_buf[2 * i];
        float d = tmp_sum / (float)c->cn0_samples;
        d = sqrtf(d);
        tmp_sum = 0;
        for (int i = 0; i < c->cn0_samples; i++)
            {
                const float a = fabsf(t->prompt_buf[2 * i] / d) - Prompt_I_ref;
                const float b = fabsf(t->prompt_buf[2 * i + 1] / d) - Prompt_Q_ref;
                const float aa = a * a, bb = b * b;
                tmp_sum = tmp_sum + aa + bb;
            }
        t->evm = (double)(tmp_sum / (float)c->cn0_samples / (Prompt_I_ref * Prompt_I_ref + Prompt_Q_ref * Prompt_Q_ref));
        t->evm = sqrt(t->evm);
    }
    return 1;
}

/* run_dll_pll (:1092-1160) on the accumulators; the experimental Doppler correction (:1162-1182) is off by default and not restated */
static void run_dll_pll(oracle_trk* t)
{
    const oracle_trk_cfg* c = &t->cfg;
    const float* VE = t->accu;
    const float* E = t->accu + 2;
    const float* P = t->accu + 4;
    const float* L = t->accu + 6;
    const float* VL = t->accu + 8;
    const double T_s = t->current_correlation_time_s;
    if (t->cloop)
        t->carr_phase_error_hz = oracle_pll_cloop_two_quadrant_atan(P[0], P[1]) / TWO_PI;
    else
        t->carr_phase_error_hz = oracle_pll_four_quadrant_atan(P[0], P[1]) / TWO_PI;
    if ((t->pull_in && c->enable_fll_pull_in) || c->enable_fll_steady_state)
        {
            t->carr_freq_error_hz = oracle_fll_diff_atan(t->P_old[0], t->P_old[1], P[0], P[1], 0, T_s) / TWO_PI;
            t->P_old[0] = P[0];
            t->P_old[1] = P[1];
            if (t->pull_in && c->enable_fll_pull_in)
                t->carr_error_filt_hz = oracle_carrier_filter_get_carrier_error(&t->cf, (float)t->carr_freq_error_hz, 0.0F, (float)T_s);
            else
                t->carr_error_filt_hz = oracle_carrier_filter_get_carrier_error(&t->cf, (float)t->carr_freq_error_hz, (float)t->carr_phase_error_hz, (float)T_s);
        }
    else
        {
            t->carr_error_filt_hz = oracle_carrier_filter_get_carrier_error(&t->cf, 0, (float)t->carr_phase_error_hz, (float)T_s);
        }
    t->carrier_doppler_hz = t->carr_error_filt_hz;
    if (c->veml)
        t->code_error_chips = oracle_dll_nc_vemlp_normalized(VE, E, L, VL);
    else
        t->code_error_chips = oracle_dll_nc_e_minus_l_normalized(E[0], E[1], L[0], L[1], t->spc, c->slope, c->y_intercept);
    t->code_error_filt_chips = oracle_loop_filter_apply(&t->lf, (float)t->code_error_chips);
    t->code_freq_chips = c->code_chip_rate - t->code_error_filt_chips;
    if (c->carrier_aiding) t->code_freq_chips += t->carrier_doppler_hz * c->code_chip_rate / c->signal_carrier_freq;
}

/* push_back on boost::circular_buffer<std::pair<double,double>> of capacity cap (pairs stored flat) */
static void hist_push(double* h, int* size, int cap, double first, double second)
{
    if (*size == cap)
        {
            memmove(h, h + 2, sizeof(double) * 2 * (size_t)(cap - 1));
            (*size)--;
        }
    h[2 * *size] = first;
    h[2 * *size + 1] = second;
    (*size)++;
}

/* the smoothed rate of a history (:1236-1251, :1268-1283) */
static double hist_rate(const double* h, unsigned int L)
{
    double tmp_cp1 = 0.0, tmp_cp2 = 0.0, tmp_samples = 0.0;
    for (unsigned int k = 0; k < L; k++)
        {
            tmp_cp1 += h[2 * k];
            tmp_cp2 += h[2 * (L * 2 - k - 1)];
            tmp_samples += h[2 * (L * 2 - k - 1) + 1];
        }
    tmp_cp1 /= (double)L;
    tmp_cp2 /= (double)L;
    return (tmp_cp2 - tmp_cp1) / tmp_samples;
}

/* update_tracking_vars (:1216-1287) */
static void update_tracking_vars(oracle_trk* t)
{
    const oracle_trk_cfg* c = &t->cfg;
    const int hist_cap = c->high_dyn ? (int)c->smoother_length * 2 : 1;
    const double T_chip_seconds = 1.0 / t->code_freq_chips;
    const double T_prn_seconds = T_chip_seconds * (double)c->code_length_chips;
    const double T_prn_samples = T_prn_seconds * c->fs_in;
    const double K_blk_samples = T_prn_samples + t->rem_code_phase_samples;
    t->current_prn_length_samples = (int32_t)floor(K_blk_samples);
    const double n = (double)t->current_prn_length_samples;
    t->carrier_phase_step_rad = TWO_PI * t->carrier_doppler_hz / c->fs_in;
    if (c->high_dyn)
        {
            hist_push(t->carr_hist, &t->carr_hist_size, hist_cap, t->carrier_phase_step_rad, n);
            if (t->carr_hist_size == hist_cap) t->carrier_phase_rate_step_rad = hist_rate(t->carr_hist, c->smoother_length);
        }
    t->rem_carr_phase_rad += (float)(t->carrier_phase_step_rad * n + 0.5 * t->carrier_phase_rate_step_rad * n * n);
    t->rem_carr_phase_rad = fmod(t->rem_carr_phase_rad, TWO_PI);
    t->acc_carrier_phase_rad -= (t->carrier_phase_step_rad * n + 0.5 * t->carrier_phase_rate_step_rad * n * n);
    t->code_phase_step_chips = t->code_freq_chips / c->fs_in;
    if (c->high_dyn)
        {
            hist_push(t->code_hist, &t->code_hist_size, hist_cap, t->code_phase_step_chips, n);
            if (t->code_hist_size == hist_cap) t->code_phase_rate_step_chips = hist_rate(t->code_hist, c->smoother_length);
        }
    t->rem_code_phase_samples = K_blk_samples - n;
    t->rem_code_phase_chips = t->code_freq_chips * t->rem_code_phase_samples / c->fs_in;
}

/* acquire_secondary (:923-966) */
static int acquire_secondary(oracle_trk* t)
{
    const oracle_trk_cfg* c = &t->cfg;
    int32_t corr_value = 0;
    for (uint32_t i = 0; i < c->secondary_code_length; i++)
        {
            if (t->circ[2 * i] < 0.0)
                corr_value += (c->secondary_code_string[i] == '0') ? 1 : -1;
            else
                corr_value += (c->secondary_code_string[i] == '0') ? -1 : 1;
        }
    if (abs(corr_value) == (int32_t)c->secondary_code_length)
        {
            t->pll_180 = corr_value < 0;
            return 1;
        }
    return 0;
}

/* save_correlation_results (:1290-1400): outs = this epoch's VE,E,P,L,VL; prompt_data = d_Prompt_Data[0] */
static void save_correlation_results(oracle_trk* t, const float* outs5, const float* prompt_data)
{
    const oracle_trk_cfg* c = &t->cfg;
    const float sign = (c->secondary && c->secondary_code_string[t->current_symbol] != '0') ? -1.0f : 1.0f;
    for (int k = 0; k < 10; k++)
        {
            if (!c->veml && (k < 2 || k >= 8)) continue;
            if (sign > 0)
                t->accu[k] += outs5[k];
            else
                t->accu[k] -= outs5[k];
        }
    if (c->secondary)
        {
            t->current_symbol++;
            t->current_symbol %= (int)c->secondary_code_length;
        }
    const float* src = c->track_pilot ? prompt_data : outs5 + 4;
    if (c->symbols_per_bit > 1)
        {
            if (c->data_secondary_code_length > 0)
                {
                    if (c->data_secondary_code_string[t->current_data_symbol] == '0')
                        {
                            t->P_data_accu[0] += src[0];
                            t->P_data_accu[1] += src[1];
                        }
                    else
                        {
                            t->P_data_accu[0] -= src[0];
                            t->P_data_accu[1] -= src[1];
                        }
                    t->current_data_symbol++;
                    t->current_data_symbol %= (int)c->data_secondary_code_length;
                }
            else
                {
                    t->P_data_accu[0] += src[0];
                    t->P_data_accu[1] += src[1];
                    t->current_data_symbol++;
                    t->current_data_symbol %= (int)c->symbols_per_bit;
                }
        }
    else
        {
            t->P_data_accu[0] = src[0];
            t->P_data_accu[1] = src[1];
        }
    t->cloop = c->track_pilot ? 0 : 1;
}

int oracle_trk_run(oracle_trk* t, const float* samples_c64, uint64_t stream_start, uint64_t n_samples, int max_epochs, oracle_trk_epoch* out)
{
    return oracle_trk_run_traced(t, samples_c64, stream_start, n_samples, max_epochs, out, NULL);
}

/* trace (optional): per epoch {rem_carr_phase_rad, carrier step, rem code phase, code step (both in code samples), first sample
   of the epoch, 1 if the epoch ran, carrier rate step, code rate step (code samples)} exactly as handed to the correlator
   (do_correlation_step, :1064-1089): what an open-loop parity test feeds to the correlator under test. */
int oracle_trk_run_traced(oracle_trk* t, const float* samples_c64, uint64_t stream_start, uint64_t n_samples, int max_epochs, oracle_trk_epoch* out,
    double* trace)
{
    const oracle_trk_cfg* c = &t->cfg;
    const uint64_t stream_end = stream_start + n_samples;
    const float spc = (float)c->code_samples_per_chip;
    const int ext = c->extend_correlation_symbols;
    const int extended = ext > 1; /* d_enable_extended_integration */
    int done = 0;
    for (int e = 0; e < max_epochs; e++)
        {
            oracle_trk_epoch* o = &out[e];
            memset(o, 0, sizeof(*o));
            if (!(t->state >= 2 && t->cursor >= stream_start && t->cursor + 2ull * c->vector_length <= stream_end)) continue;
            const uint64_t nread = t->cursor;
            const int state_in = t->state;
            if (t->pull_in && (uint64_t)c->pull_in_time_s < (nread - t->acq_sample_stamp) / (uint64_t)(int)c->fs_in)
                {
                    t->pull_in = 0;
                    t->carrier_lock_fail = 0;
                    t->code_lock_fail = 0;
                }
            /* do_correlation_step */
            float outs[10], outs5[10], prompt_data[2] = {0.0f, 0.0f};
            const float a_rem_code = (float)t->rem_code_phase_chips * spc, a_code_step = (float)t->code_phase_step_chips * spc;
            const float a_code_rate = (float)t->code_phase_rate_step_chips * spc;
            if (trace)
                {
                    double* tr = trace + 8 * (size_t)e;
                    tr[0] = (double)t->rem_carr_phase_rad;
                    tr[1] = (double)(float)t->carrier_phase_step_rad;
                    tr[2] = (double)a_rem_code;
                    tr[3] = (double)a_code_step;
                    tr[4] = (double)nread;
                    tr[5] = 1.0;
                    tr[6] = (double)(float)t->carrier_phase_rate_step_rad;
                    tr[7] = (double)a_code_rate;
                }
            oracle_mc_real(samples_c64 + 2 * (nread - stream_start), (int)c->vector_length, t->code, t->code_len, t->shifts, t->n_taps,
                t->rem_carr_phase_rad, (float)t->carrier_phase_step_rad, (float)t->carrier_phase_rate_step_rad, a_rem_code, a_code_step, a_code_rate,
                c->high_dyn, 0, outs);
            if (c->track_pilot && t->data_code)
                {
                    const float prompt_shift = t->shifts[c->veml ? 2 : 1]; /* d_prompt_data_shift */
                    oracle_mc_real(samples_c64 + 2 * (nread - stream_start), (int)c->vector_length, t->data_code, t->data_code_len, &prompt_shift, 1,
                        t->rem_carr_phase_rad, (float)t->carrier_phase_step_rad, (float)t->carrier_phase_rate_step_rad, a_rem_code, a_code_step, a_code_rate,
                        c->high_dyn, 0, prompt_data);
                }
            memset(outs5, 0, sizeof(outs5));
            memcpy(outs5 + (c->veml ? 0 : 2), outs, sizeof(float) * 2 * (size_t)t->n_taps);
            o->flags = 1u;
            int lock_ok = 1, symbol_out = 0;
            if (state_in == 2)
                {
                    memcpy(t->accu, outs5, sizeof(outs5)); /* d_VE_accu = *d_Very_Early ... (:1849-1856) */
                    t->spc = c->early_late_space_chips;
                    if (c->symbol_sync && (uint64_t)c->bit_synchronization_time_limit_s < (nread - t->acq_sample_stamp) / (uint64_t)(int)c->fs_in)
                        t->carrier_lock_fail = 300000;
                    lock_ok = cn0_and_tracking_lock_status(t, t->accu + 4, c->code_period_s);
                    memcpy(o->accu, t->accu, sizeof(t->accu));
                    if (lock_ok)
                        {
                            run_dll_pll(t);
                            update_tracking_vars(t);
                            if (c->symbol_sync)
                                {
                                    int next_state = 0;
                                    if (!t->pull_in)
                                        {
                                            if (c->secondary || c->symbols_per_bit > 1)
                                                {
                                                    const int cap = (int)c->secondary_code_length;
                                                    if (t->circ_size == cap)
                                                        {
                                                            memmove(t->circ, t->circ + 2, sizeof(float) * 2 * (size_t)(cap - 1));
                                                            t->circ_size--;
                                                        }
                                                    t->circ[2 * t->circ_size] = outs5[4];
                                                    t->circ[2 * t->circ_size + 1] = outs5[5];
                                                    t->circ_size++;
                                                    if (t->circ_size == cap) next_state = acquire_secondary(t);
                                                }
                                            else
                                                next_state = 1;
                                        }
                                    if (next_state)
                                        {
                                            memset(t->accu, 0, sizeof(t->accu));
                                            t->P_data_accu[0] = t->P_data_accu[1] = 0.0f;
                                            t->circ_size = 0;
                                            t->current_symbol = t->current_data_symbol = 0;
                                            if (extended)
                                                {
                                                    t->extend_count = 0;
                                                    t->current_correlation_time_s = (double)((float)ext * (float)c->code_period_s);
                                                    t->state = 3;
                                                    /* set_update_interval + set_noise_bandwidth: new coefficients, the filter memory stays */
                                                    oracle_loop_filter lf_new;
                                                    oracle_loop_filter_init(&lf_new, (float)t->current_correlation_time_s, c->dll_bw_narrow_hz, c->dll_filter_order, 0);
                                                    memcpy(t->lf.in_coef, lf_new.in_coef, sizeof(lf_new.in_coef));
                                                    memcpy(t->lf.out_coef, lf_new.out_coef, sizeof(lf_new.out_coef));
                                                    t->lf.n_in = lf_new.n_in;
                                                    t->lf.n_out = lf_new.n_out;
                                                    {
                                                        const float w = t->cf.pll_w, x = t->cf.pll_x;
                                                        oracle_carrier_filter_set_params(&t->cf, c->fll_bw_hz, c->pll_bw_narrow_hz, c->pll_filter_order);
                                                        t->cf.pll_w = w;
                                                        t->cf.pll_x = x;
                                                    }
                                                    if (c->veml)
                                                        {
                                                            t->shifts[0] = -c->very_early_late_space_narrow_chips * spc;
                                                            t->shifts[1] = -c->early_late_space_narrow_chips * spc;
                                                            t->shifts[3] = c->early_late_space_narrow_chips * spc;
                                                            t->shifts[4] = c->very_early_late_space_narrow_chips * spc;
                                                        }
                                                    else
                                                        {
                                                            t->shifts[0] = -c->early_late_space_narrow_chips * spc;
                                                            t->shifts[2] = c->early_late_space_narrow_chips * spc;
                                                        }
                                                    t->spc = c->early_late_space_narrow_chips;
                                                }
                                            else
                                                t->state = 4;
                                        }
                                }
                        }
                }
            else if (state_in == 3)
                {
                    save_correlation_results(t, outs5, prompt_data);
                    memcpy(o->accu, t->accu, sizeof(t->accu));
                    update_tracking_vars(t);
                    if (t->current_data_symbol == 0) symbol_out = 1;
                    t->extend_count++;
                    if (t->extend_count == ext - 1)
                        {
                            t->extend_count = 0;
                            t->state = 4;
                        }
                }
            else /* 4 */
                {
                    save_correlation_results(t, outs5, prompt_data);
                    memcpy(o->accu, t->accu, sizeof(t->accu));
                    lock_ok = cn0_and_tracking_lock_status(t, t->accu + 4, c->code_period_s * (double)ext);
                    if (lock_ok)
                        {
                            run_dll_pll(t);
                            update_tracking_vars(t);
                            if (!t->acc_carrier_phase_initialized) /* check_carrier_phase_coherent_initialization (:1184-1190) */
                                {
                                    t->acc_carrier_phase_rad = -t->rem_carr_phase_rad;
                                    t->acc_carrier_phase_initialized = 1;
                                }
                            if (t->current_data_symbol == 0) symbol_out = 1;
                            memset(t->accu, 0, sizeof(t->accu));
                            if (extended) t->state = 3;
                        }
                }
            if (t->pull_in) o->flags |= 2u;
            if (!lock_ok)
                {
                    clear_tracking_vars(t);
                    t->state = 0;
                    o->flags |= 4u;
                }
            if (symbol_out)
                {
                    o->flags |= 8u;
                    o->prompt_data[0] = t->P_data_accu[0];
                    o->prompt_data[1] = t->P_data_accu[1];
                    t->P_data_accu[0] = t->P_data_accu[1] = 0.0f;
                }
            if (t->pll_180) o->flags |= 16u;
            memcpy(o->corr, outs5, sizeof(outs5));
            o->carrier_doppler_hz = t->carrier_doppler_hz;
            o->code_freq_chips = t->code_freq_chips;
            o->carr_error_hz = t->carr_phase_error_hz;
            o->carr_error_filt_hz = t->carr_error_filt_hz;
            o->code_error_chips = t->code_error_chips;
            o->code_error_filt_chips = t->code_error_filt_chips;
            o->rem_code_phase_samples = t->rem_code_phase_samples;
            o->acc_carrier_phase_rad = t->acc_carrier_phase_rad;
            o->cn0_db_hz = t->cn0_db_hz;
            o->carrier_lock_test = t->carrier_lock_test;
            o->evm = t->evm;
            o->sample_counter = nread;
            o->prn_length_samples = t->current_prn_length_samples;
            o->state = state_in;
            t->cursor += (uint64_t)t->current_prn_length_samples; /* consume_each */
            done++;
        }
    return done;
}
