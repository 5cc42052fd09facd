/* TEST INFRASTRUCTURE — CPU oracle, not product code. See oracle_trk.c. */
#ifndef ORACLE_TRK_H
#define ORACLE_TRK_H
#include <stdint.h>
#ifdef __cplusplus
extern "C"
{
#endif

#define ORACLE_TRK_MAX_CN0_SAMPLES 64
#define ORACLE_TRK_MAX_SMOOTHER_INIT 256

    /* same fields, order and meaning as gnss_b200_trk_cfg (include/gnss_b200.h) so that one ctypes struct serves both */
    typedef struct
    {
        double fs_in, code_chip_rate, signal_carrier_freq, code_period_s;
        uint32_t code_length_chips, code_samples_per_chip, vector_length;
        int32_t veml;
        float early_late_space_chips, very_early_late_space_chips, slope, y_intercept;
        float pll_bw_hz, dll_bw_hz, fll_bw_hz;
        int32_t pll_filter_order, dll_filter_order;
        int32_t enable_fll_pull_in, enable_fll_steady_state, carrier_aiding;
        uint32_t pull_in_time_s;
        int32_t cn0_samples, cn0_min, max_code_lock_fail, max_carrier_lock_fail;
        double carrier_lock_th;
        float cn0_smoother_alpha, carrier_lock_test_smoother_alpha;
        int32_t cn0_smoother_samples, carrier_lock_test_smoother_samples;
        uint32_t max_channels;
        int32_t device;
        /* symbol synchronisation, extended integration, pilot tracking, high dynamics (all 0: the state-2 loop only) */
        int32_t symbol_sync, secondary;
        uint32_t secondary_code_length, symbols_per_bit, data_secondary_code_length;
        int32_t extend_correlation_symbols, track_pilot;
        float pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_narrow_chips, very_early_late_space_narrow_chips;
        int32_t high_dyn;
        uint32_t smoother_length, bit_synchronization_time_limit_s;
        char secondary_code_string[208], data_secondary_code_string[112];
    } oracle_trk_cfg;

    /* same layout as gnss_b200_trk_epoch */
    typedef struct
    {
        float corr[10];
        double carrier_doppler_hz, code_freq_chips, carr_error_hz, carr_error_filt_hz, code_error_chips, code_error_filt_chips;
        double rem_code_phase_samples, acc_carrier_phase_rad, cn0_db_hz, carrier_lock_test;
        double evm;
        uint64_t sample_counter;
        int32_t prn_length_samples;
        uint32_t flags;
        float accu[10];       /* d_VE_accu .. d_VL_accu as the loop of this epoch saw them */
        float prompt_data[2]; /* d_P_data_accu handed to the telemetry decoder (flag 8) */
        int32_t state;        /* d_state the epoch ran in */
        int32_t reserved;
    } oracle_trk_epoch;

    typedef struct
    {
        float in_coef[4], out_coef[3];
        int n_in, n_out, idx;
        float inputs[4], outputs[4];
    } oracle_loop_filter;

    typedef struct
    {
        int order;
        float pll_w, pll_x, pll_a2, pll_a3, pll_b3, pll_w0p, pll_w0p2, pll_w0p3, pll_w0f, pll_w0f2;
    } oracle_carrier_filter;

    typedef struct
    {
        float alpha, one_minus_alpha, old_value, min_value, offset;
        int samples_for_initialization, init_counter, initializing;
        float init_buffer[ORACLE_TRK_MAX_SMOOTHER_INIT];
    } oracle_smoother;

    void oracle_loop_filter_init(oracle_loop_filter* f, float update_interval, float noise_bandwidth, int order, int include_last_integrator);
    void oracle_loop_filter_initialize(oracle_loop_filter* f, float initial_output);
    float oracle_loop_filter_apply(oracle_loop_filter* f, float x);
    void oracle_carrier_filter_set_params(oracle_carrier_filter* c, float fll_bw_hz, float pll_bw_hz, int order);
    void oracle_carrier_filter_initialize(oracle_carrier_filter* c, float acq_doppler_hz);
    float oracle_carrier_filter_get_carrier_error(oracle_carrier_filter* c, float FLL, float PLL, float T);
    void oracle_smoother_init(oracle_smoother* s, float alpha, float min_value, float offset, int samples_for_initialization);
    float oracle_smoother_smooth(oracle_smoother* s, float raw);
    double oracle_fll_diff_atan(float s1_re, float s1_im, float s2_re, float s2_im, double t1, double t2);
    double oracle_pll_cloop_two_quadrant_atan(float re, float im);
    double oracle_pll_four_quadrant_atan(float re, float im);
    float oracle_fast_atan2f(float y, float x);
    double oracle_dll_nc_e_minus_l_normalized(float e_re, float e_im, float l_re, float l_im, float spc, float slope, float y_intercept);
    double oracle_dll_nc_vemlp_normalized(const float* ve, const float* e, const float* l, const float* vl);
    float oracle_cn0_m2m4_estimator(const float* prompt_c64, int length, float coh_integration_time_s);
    float oracle_carrier_lock_detector(const float* prompt_c64, int length);

    typedef struct oracle_trk oracle_trk;
    oracle_trk* oracle_trk_create(const oracle_trk_cfg* cfg);
    void oracle_trk_destroy(oracle_trk* t);
    void oracle_trk_set_code(oracle_trk* t, const float* code, int code_len);
    void oracle_trk_set_data_code(oracle_trk* t, const float* code, int code_len); /* d_data_code (track_pilot) */
    void oracle_trk_start(oracle_trk* t, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now_sample);
    /* out[max_epochs]; returns the number of epochs that ran */
    int oracle_trk_run(oracle_trk* t, const float* samples_c64, uint64_t stream_start, uint64_t n_samples, int max_epochs, oracle_trk_epoch* out);
    /* same, and trace[max_epochs][8] = the correlator arguments of every epoch (NULL: no trace) */
    int oracle_trk_run_traced(oracle_trk* t, const float* samples_c64, uint64_t stream_start, uint64_t n_samples, int max_epochs, oracle_trk_epoch* out,
        double* trace);

#ifdef __cplusplus
}
#endif
#endif
