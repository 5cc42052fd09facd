/*
 * TEST INFRASTRUCTURE — CPU oracle, not product code.
 *
 * Restatement of the PCPS acquisition hot path of the reference
 * (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc):
 *   set_local_code                 :176-209
 *   update_local_carrier           :233-246
 *   update_grid_doppler_wipeoffs   :298-305
 *   max_to_input_power_statistic   :511-543
 *   first_vs_second_peak_statistic :546-612
 *   acquisition_core (grid loop)   :615-713
 *   calculate_threshold            :894-909
 * over the kernels of oracle_kernels.h and the DFT of oracle_fft.c.
 * The GNU Radio block itself cannot be compiled here (no GNU Radio / VOLK /
 * FFTW / Boost), so this file is the oracle; it is pinned against the
 * reference's two in-tree captures and their expected answers in tests/.
 */
#include "oracle_acq.h"
#include "oracle_fft.h"
#include "oracle_kernels.h"
#include <stdio.h>

struct oracle_acq
{
    oracle_acq_cfg c;
    uint32_t n_doppler;
    ocf* fft_codes;   /* conj(FFT(padded code)), [N] */
    ocf* input;       /* samples then zeros, [N] */
    ocf* fft_in;      /* work */
    ocf* fft_out;     /* work */
    ocf* ifft_out;    /* work */
    float* tmp;       /* [N] */
    ocf* wipeoffs;    /* [nD][N] */
    float* grid;      /* [nD][N] (row stride N, like the reference's vector-of-vectors) */
    offt_f *fwd, *bwd;
    uint32_t dwell_count;
    float input_power;
};

const char* oracle_kernel_flavour(void) { return ORACLE_KERNEL_FLAVOUR; }

uint32_t oracle_acq_num_doppler_bins(int32_t doppler_max, int32_t doppler_step)
{
    /* pcps_acquisition.cc:264 */
    return (uint32_t)ceil((double)((int32_t)doppler_max - (int32_t)(-doppler_max)) / (double)doppler_step);
}

static void build_wipeoffs(oracle_acq* h)
{
    const uint32_t N = h->c.fft_size;
    for (uint32_t k = 0; k < h->n_doppler; k++)
        {
            /* :302-303 */
            const int32_t doppler = -(int32_t)h->c.doppler_max + h->c.doppler_center + h->c.doppler_step * (int32_t)k;
            const float freq = (float)(h->c.doppler_bias + doppler);
            /* :242 */
            const float phase_step_rad = (float)(2.0 * M_PI) * freq / (float)h->c.fs;
            float phase = 0.0f;
            k_sincos(h->wipeoffs + (size_t)k * N, -phase_step_rad, &phase, N);
        }
}

oracle_acq* oracle_acq_create(const oracle_acq_cfg* cfg)
{
    oracle_acq* h = (oracle_acq*)calloc(1, sizeof(*h));
    h->c = *cfg;
    const uint32_t N = cfg->fft_size;
    h->n_doppler = oracle_acq_num_doppler_bins(cfg->doppler_max, cfg->doppler_step);
    h->fft_codes = (ocf*)calloc(N, sizeof(ocf));
    h->input = (ocf*)calloc(N, sizeof(ocf));
    h->fft_in = (ocf*)calloc(N, sizeof(ocf));
    h->fft_out = (ocf*)calloc(N, sizeof(ocf));
    h->ifft_out = (ocf*)calloc(N, sizeof(ocf));
    h->tmp = (float*)calloc(N, sizeof(float));
    h->wipeoffs = (ocf*)calloc((size_t)N * h->n_doppler, sizeof(ocf));
    h->grid = (float*)calloc((size_t)N * h->n_doppler, sizeof(float));
    h->fwd = offt_f_create((int)N, -1);
    h->bwd = offt_f_create((int)N, +1);
    build_wipeoffs(h);
    return h;
}

void oracle_acq_destroy(oracle_acq* h)
{
    if (!h) return;
    free(h->fft_codes);
    free(h->input);
    free(h->fft_in);
    free(h->fft_out);
    free(h->ifft_out);
    free(h->tmp);
    free(h->wipeoffs);
    free(h->grid);
    offt_f_destroy(h->fwd);
    offt_f_destroy(h->bwd);
    free(h);
}

uint32_t oracle_acq_n_doppler(const oracle_acq* h) { return h->n_doppler; }
const float* oracle_acq_grid(const oracle_acq* h) { return h->grid; }
const float* oracle_acq_fft_codes(const oracle_acq* h) { return (const float*)h->fft_codes; }
const float* oracle_acq_wipeoffs(const oracle_acq* h) { return (const float*)h->wipeoffs; }

void oracle_acq_set_doppler_center(oracle_acq* h, int32_t center_hz, int32_t bias_hz)
{
    /* pcps_acquisition.h:199-208 → update_grid_doppler_wipeoffs */
    h->c.doppler_center = center_hz;
    h->c.doppler_bias = bias_hz;
    build_wipeoffs(h);
}

void oracle_acq_set_local_code(oracle_acq* h, const float* code_c64)
{
    const ocf* code = (const ocf*)code_c64;
    const uint32_t N = h->c.fft_size, consumed = h->c.consumed_samples;
    memset(h->fft_in, 0, sizeof(ocf) * N);
    if (h->c.bit_transition)
        {
            /* :188-193  [0 … 0 | code[0..N/2)] */
            const uint32_t off = N / 2;
            memcpy(h->fft_in + off, code, sizeof(ocf) * off);
        }
    else if (N == consumed)
        {
            /* :196-199 */
            memcpy(h->fft_in, code, sizeof(ocf) * consumed);
        }
    else
        {
            /* :200-204  zeros in the lower N-consumed, code in the upper part */
            memcpy(h->fft_in + consumed, code, sizeof(ocf) * consumed);
        }
    offt_f_execute(h->fwd, (const float*)h->fft_in, (float*)h->fft_out);
    v_conjugate(h->fft_codes, h->fft_out, N); /* :207-208 */
}

void oracle_acq_reset(oracle_acq* h)
{
    /* init() :281-284 zero-fills the grid; the dwell counter is cleared at :879 */
    memset(h->grid, 0, sizeof(float) * (size_t)h->c.fft_size * h->n_doppler);
    h->dwell_count = 0;
}

static float stat_max_to_input_power(oracle_acq* h, oracle_acq_result* r)
{
    /* :511-543 */
    const uint32_t N = h->c.fft_size;
    const uint32_t eff = h->c.bit_transition ? N / 2 : N;
    float grid_maximum = 0.0f;
    uint32_t index_doppler = 0, index_time = 0, tmp_t = 0;
    for (uint32_t i = 0; i < h->n_doppler; i++)
        {
            const float* row = h->grid + (size_t)i * N;
            k_index_max(&tmp_t, row, eff);
            if (row[tmp_t] > grid_maximum)
                {
                    grid_maximum = row[tmp_t];
                    index_doppler = i;
                    index_time = tmp_t;
                }
        }
    const uint32_t index_opp = (index_doppler + h->n_doppler / 2) % h->n_doppler;
    const float* opp = h->grid + (size_t)index_opp * N;
    float acc = 0.0f; /* std::accumulate with a float init: sequential float32 sum */
    for (uint32_t i = 0; i < eff; i++) acc = acc + opp[i];
    /* float / int32 → float; then / 2.0 and / counter are done in double; cast to float (:534) */
    h->input_power = (float)((acc / (int32_t)eff) / 2.0 / h->dwell_count);
    r->peak = grid_maximum;
    r->code_phase = index_time;
    r->doppler_index = index_doppler;
    r->doppler_hz = -(int32_t)h->c.doppler_max + h->c.doppler_center + h->c.doppler_step * (int32_t)index_doppler;
    r->input_power = h->input_power;
    r->second_peak = 0.0f;
    return grid_maximum / h->input_power;
}

static float stat_first_vs_second(oracle_acq* h, oracle_acq_result* r)
{
    /* :546-612 — note the argmax runs over the full row length N here */
    const uint32_t N = h->c.fft_size;
    float first = 0.0f;
    uint32_t index_doppler = 0, index_time = 0, tmp_t = 0;
    for (uint32_t i = 0; i < h->n_doppler; i++)
        {
            const float* row = h->grid + (size_t)i * N;
            k_index_max(&tmp_t, row, N);
            if (row[tmp_t] > first)
                {
                    first = row[tmp_t];
                    index_doppler = i;
                    index_time = tmp_t;
                }
        }
    int32_t ex1 = (int32_t)index_time - (int32_t)h->c.samples_per_chip;
    int32_t ex2 = (int32_t)index_time + (int32_t)h->c.samples_per_chip;
    if (ex1 < 0)
        ex1 = (int32_t)N + ex1;
    else if (ex2 >= (int32_t)N)
        ex2 = ex2 - (int32_t)N;
    memcpy(h->tmp, h->grid + (size_t)index_doppler * N, sizeof(float) * N);
    int32_t idx = ex1;
    do
        {
            h->tmp[idx] = 0.0f;
            idx++;
            if (idx == (int32_t)N) idx = 0;
        }
    while (idx != ex2);
    k_index_max(&tmp_t, h->tmp, N);
    const float second = h->tmp[tmp_t];
    r->peak = first;
    r->code_phase = index_time;
    r->doppler_index = index_doppler;
    r->doppler_hz = -(int32_t)h->c.doppler_max + h->c.doppler_center + h->c.doppler_step * (int32_t)index_doppler;
    r->input_power = h->input_power;
    r->second_peak = second;
    return first / second;
}

void oracle_acq_dwell(oracle_acq* h, const void* samples, int is_cshort, oracle_acq_result* r)
{
    /* acquisition_core :615-713 (first-step grid) */
    const uint32_t N = h->c.fft_size, consumed = h->c.consumed_samples;
    const uint32_t eff = h->c.bit_transition ? N / 2 : N;
    if (is_cshort)
        k_convert_16ic(h->input, (const void*)samples, consumed); /* :623-626 */
    else
        memcpy(h->input, samples, sizeof(ocf) * consumed); /* :627 */
    for (uint32_t i = consumed; i < N; i++) h->input[i] = 0.0f; /* :628-634 */
    h->dwell_count++;                                            /* :638 */
    const size_t offset = h->c.bit_transition ? eff : 0;         /* :671 */
    for (uint32_t k = 0; k < h->n_doppler; k++)
        {
            float* row = h->grid + (size_t)k * N;
            v_multiply(h->fft_in, h->input, h->wipeoffs + (size_t)k * N, N);        /* :658 */
            offt_f_execute(h->fwd, (const float*)h->fft_in, (float*)h->fft_out);    /* :662 */
            v_multiply(h->fft_in, h->fft_out, h->fft_codes, N);                     /* :665 */
            offt_f_execute(h->bwd, (const float*)h->fft_in, (float*)h->ifft_out);   /* :668 */
            if (h->dwell_count == 1)
                {
                    v_mag_squared(row, h->ifft_out + offset, eff); /* :674 */
                }
            else
                {
                    v_mag_squared(h->tmp, h->ifft_out + offset, eff); /* :678 */
                    v_add(row, row, h->tmp, eff);                     /* :679 */
                }
        }
    memset(r, 0, sizeof(*r));
    if (h->c.use_cfar)
        r->test_statistic = stat_max_to_input_power(h, r); /* :691 */
    else
        r->test_statistic = stat_first_vs_second(h, r); /* :695 */
    /* :709 */
    r->acq_delay_samples = (double)fmodf((float)r->code_phase, h->c.samples_per_code);
    r->dwell_count = h->dwell_count;
}

void oracle_acq_dwell_step2(oracle_acq* h, const void* samples, int is_cshort, uint32_t num_doppler_bins_step2, float doppler_step2,
    float doppler_center_step_two, oracle_acq_result* r)
{
    /* acquisition_core with d_step_two == true (:715-775): the narrow grid of make_two_steps around the coarse Doppler. The
     * block zeroes d_num_noncoherent_integrations_counter when it enters step two (:796-799), d_magnitude_grid rows
     * 0..n2-1 are reused, and d_input_power keeps the value step one left (:531-540 only refresh it when !d_step_two). */
    const uint32_t N = h->c.fft_size, consumed = h->c.consumed_samples, n2 = num_doppler_bins_step2;
    const uint32_t eff = h->c.bit_transition ? N / 2 : N;
    ocf* wipe2 = (ocf*)malloc(sizeof(ocf) * (size_t)n2 * N);
    float* grid2 = (float*)calloc((size_t)n2 * N, sizeof(float));
    for (uint32_t k = 0; k < n2; k++)
        {
            /* update_grid_doppler_wipeoffs_step2 :308-315 + update_local_carrier :233-246 */
            const float doppler = ((float)k - (float)floor(n2 / 2.0)) * doppler_step2;
            const float freq = doppler_center_step_two + doppler;
            const float phase_step_rad = (float)(2.0 * M_PI) * freq / (float)h->c.fs;
            float phase = 0.0f;
            k_sincos(wipe2 + (size_t)k * N, -phase_step_rad, &phase, N);
        }
    if (is_cshort)
        k_convert_16ic(h->input, (const void*)samples, consumed);
    else
        memcpy(h->input, samples, sizeof(ocf) * consumed);
    for (uint32_t i = consumed; i < N; i++) h->input[i] = 0.0f;
    const size_t offset = h->c.bit_transition ? eff : 0;
    for (uint32_t k = 0; k < n2; k++)
        {
            v_multiply(h->fft_in, h->input, wipe2 + (size_t)k * N, N);             /* :719 */
            offt_f_execute(h->fwd, (const float*)h->fft_in, (float*)h->fft_out);   /* :723 */
            v_multiply(h->fft_in, h->fft_out, h->fft_codes, N);                    /* :727 */
            offt_f_execute(h->bwd, (const float*)h->fft_in, (float*)h->ifft_out);  /* :730 */
            v_mag_squared(grid2 + (size_t)k * N, h->ifft_out + offset, eff);       /* :735 (counter == 1) */
        }
    memset(r, 0, sizeof(*r));
    /* the two statistics with d_step_two (:511-543, :546-612) */
    const uint32_t len = h->c.use_cfar ? eff : N;
    float best = 0.0f;
    uint32_t index_doppler = 0, index_time = 0, tmp_t = 0;
    for (uint32_t i = 0; i < n2; i++)
        {
            const float* row = grid2 + (size_t)i * N;
            k_index_max(&tmp_t, row, len);
            if (row[tmp_t] > best)
                {
                    best = row[tmp_t];
                    index_doppler = i;
                    index_time = tmp_t;
                }
        }
    r->peak = best;
    r->code_phase = index_time;
    r->doppler_index = index_doppler;
    r->doppler_hz = (int32_t)(doppler_center_step_two + ((float)index_doppler - (float)floor(n2 / 2.0)) * doppler_step2); /* :539 */
    r->input_power = h->input_power;
    if (h->c.use_cfar)
        {
            r->test_statistic = best / h->input_power; /* :542 */
        }
    else
        {
            int32_t ex1 = (int32_t)index_time - (int32_t)h->c.samples_per_chip;
            int32_t ex2 = (int32_t)index_time + (int32_t)h->c.samples_per_chip;
            if (ex1 < 0)
                ex1 = (int32_t)N + ex1;
            else if (ex2 >= (int32_t)N)
                ex2 = ex2 - (int32_t)N;
            memcpy(h->tmp, grid2 + (size_t)index_doppler * N, sizeof(float) * N);
            int32_t idx = ex1;
            do
                {
                    h->tmp[idx] = 0.0f;
                    idx++;
                    if (idx == (int32_t)N) idx = 0;
                }
            while (idx != ex2);
            k_index_max(&tmp_t, h->tmp, N);
            r->second_peak = h->tmp[tmp_t];
            r->test_statistic = best / r->second_peak;
        }
    r->acq_delay_samples = (double)fmodf((float)r->code_phase, h->c.samples_per_code); /* :769 */
    r->dwell_count = 1;
    free(wipe2);
    free(grid2);
}

/* ---- threshold: 2 * gamma_p_inv(a, p) (:894-909). boost::math is third-party
 * and absent; P(a,x) is evaluated by its series / continued fraction and inverted
 * by bracketed Newton. Cross-checked against scipy.special.gammaincinv in tests. ---- */
static double reg_lower_gamma(double a, double x)
{
    if (x <= 0.0) return 0.0;
    const double lg = lgamma(a);
    if (x < a + 1.0)
        {
            double term = 1.0 / a, sum = term;
            for (int n = 1; n < 100000; n++)
                {
                    term *= x / (a + n);
                    sum += term;
                    if (fabs(term) < fabs(sum) * 1e-17) break;
                }
            return sum * exp(-x + a * log(x) - lg);
        }
    /* Lentz continued fraction for Q(a,x) */
    const double tiny = 1e-300;
    double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, f = d;
    for (int i = 1; i < 100000; i++)
        {
            const double an = -i * (i - a);
            b += 2.0;
            d = an * d + b;
            if (fabs(d) < tiny) d = tiny;
            c = b + an / c;
            if (fabs(c) < tiny) c = tiny;
            d = 1.0 / d;
            const double del = d * c;
            f *= del;
            if (fabs(del - 1.0) < 1e-16) break;
        }
    return 1.0 - exp(-x + a * log(x) - lg) * f;
}

double oracle_gamma_p_inv(double a, double p)
{
    if (p <= 0.0) return 0.0;
    if (p >= 1.0) return INFINITY;
    double lo = 0.0, hi = a + 10.0;
    while (reg_lower_gamma(a, hi) < p) hi *= 2.0;
    double x = 0.5 * (lo + hi);
    for (int it = 0; it < 400; it++)
        {
            const double f = reg_lower_gamma(a, x) - p;
            if (f > 0)
                hi = x;
            else
                lo = x;
            /* Newton step with dP/dx = x^(a-1) e^-x / Gamma(a) */
            const double dpdx = exp((a - 1.0) * log(x) - x - lgamma(a));
            double xn = (dpdx > 0) ? x - f / dpdx : 0.5 * (lo + hi);
            if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
            if (fabs(xn - x) <= 1e-15 * fabs(x))
                {
                    x = xn;
                    break;
                }
            x = xn;
        }
    return x;
}

float oracle_acq_threshold(float pfa, uint32_t effective_fft_size, uint32_t num_doppler_bins, uint32_t max_dwells, int bit_transition)
{
    /* :903-908; the exponent is 1.0 / float(num_bins), evaluated in double */
    const int num_bins = (int)effective_fft_size * (int)num_doppler_bins;
    const double p = pow(1.0 - (double)pfa, 1.0 / (double)(float)num_bins);
    return (float)(2.0 * oracle_gamma_p_inv(2.0 * (bit_transition ? 1 : (double)max_dwells), p));
}
