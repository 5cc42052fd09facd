/* TEST INFRASTRUCTURE — CPU oracle, not product code. See oracle_acq.c. */
#ifndef ORACLE_ACQ_H
#define ORACLE_ACQ_H
#include <stdint.h>
#ifdef __cplusplus
extern "C"
{
#endif

    typedef struct
    {
        uint32_t fft_size;         /* d_fft_size */
        uint32_t consumed_samples; /* d_consumed_samples */
        int32_t bit_transition;    /* Acq_Conf::bit_transition_flag */
        int32_t doppler_max;       /* Acq_Conf::doppler_max */
        int32_t doppler_step;      /* d_doppler_step */
        int32_t doppler_center;    /* d_doppler_center */
        int32_t doppler_bias;      /* d_doppler_bias (FDMA) */
        double fs;                 /* Acq_Conf::fs_in (or resampled_fs) */
        uint32_t samples_per_chip; /* d_samplesPerChip */
        float samples_per_code;    /* Acq_Conf::samples_per_code */
        int32_t use_cfar;          /* d_use_CFAR_algorithm_flag */
    } oracle_acq_cfg;

    typedef struct
    {
        float test_statistic;
        float peak;
        float input_power;
        float second_peak;
        uint32_t code_phase; /* indext */
        uint32_t doppler_index;
        int32_t doppler_hz;
        uint32_t dwell_count;
        double acq_delay_samples;
    } oracle_acq_result;

    typedef struct oracle_acq oracle_acq;

    const char* oracle_kernel_flavour(void);
    uint32_t oracle_acq_num_doppler_bins(int32_t doppler_max, int32_t doppler_step);
    oracle_acq* oracle_acq_create(const oracle_acq_cfg* cfg);
    void oracle_acq_destroy(oracle_acq* h);
    uint32_t oracle_acq_n_doppler(const oracle_acq* h);
    void oracle_acq_set_doppler_center(oracle_acq* h, int32_t center_hz, int32_t bias_hz);
    void oracle_acq_set_local_code(oracle_acq* h, const float* code_c64);
    void oracle_acq_reset(oracle_acq* h);
    void oracle_acq_dwell(oracle_acq* h, const void* samples, int is_cshort, oracle_acq_result* r);
    /* make_two_steps: the narrow second search (acquisition_core with d_step_two, :715-775) after a first-step dwell */
    void oracle_acq_dwell_step2(oracle_acq* h, const void* samples, int is_cshort, uint32_t num_doppler_bins_step2, float doppler_step2,
        float doppler_center_step_two, oracle_acq_result* r);
    const float* oracle_acq_grid(const oracle_acq* h);      /* [nD][fft_size] */
    const float* oracle_acq_fft_codes(const oracle_acq* h); /* [fft_size] c64 */
    const float* oracle_acq_wipeoffs(const oracle_acq* h);  /* [nD][fft_size] c64 */

    double oracle_gamma_p_inv(double a, double p);
    float oracle_acq_threshold(float pfa, uint32_t effective_fft_size, uint32_t num_doppler_bins, uint32_t max_dwells, int bit_transition);

#ifdef __cplusplus
}
#endif
#endif
