/*
 * gnss_b200.h — C-ABI of libgnss_b200.so, the B200-native GNSS correlation engine.
 *
 * Two data-parallel hot paths of GNSS-SDR (bobzhangbo2012/gnss-sdr-new) live behind
 * this boundary, each implemented as hand-written sm_100a CUDA kernels:
 *
 *   (a) PCPS acquisition: the Doppler-grid loop of pcps_acquisition::acquisition_core
 *       (reference: src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc:615-713)
 *       batched over PRN x Doppler x code phase;
 *   (b) the tracking multicorrelator: Cpu_Multicorrelator_Real_Codes /
 *       Cpu_Multicorrelator::Carrier_wipeoff_multicorrelator_resampler
 *       (reference: src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:103-144,
 *        cpu_multicorrelator.cc:87-101) batched over channels.
 *
 * Conventions
 *   - every call returns 0 on success, a negative gnss_b200_status otherwise;
 *     gnss_b200_last_error() returns a thread-local message for the last failure;
 *   - no C++ or torch types cross the boundary: opaque handles, plain pointers, sizes;
 *   - complex data is interleaved {re, im} float32 (gr_complex / lv_32fc_t layout);
 *   - the caller owns all host buffers; the engine owns all device memory and its stream;
 *   - a handle is serialised internally; distinct handles may be used concurrently;
 *   - calls are synchronous unless named *_async (then pair with *_sync);
 *   - there is NO CPU fallback: without a usable CUDA device create() fails.
 */
#ifndef GNSS_B200_H
#define GNSS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C"
{
#endif

#if defined(__GNUC__)
#define GNSS_B200_API __attribute__((visibility("default")))
#else
#define GNSS_B200_API
#endif

    typedef enum
    {
        GNSS_B200_OK = 0,
        GNSS_B200_ERR_INVALID = -1,     /* bad argument */
        GNSS_B200_ERR_CUDA = -2,        /* CUDA runtime error (message in last_error) */
        GNSS_B200_ERR_UNSUPPORTED = -3, /* e.g. a statistic that cannot be decided from row records */
        GNSS_B200_ERR_NO_DEVICE = -4,   /* no CUDA device: the engine has no CPU path */
        GNSS_B200_ERR_STATE = -5        /* call order violated (e.g. dwell before set_local_code) */
    } gnss_b200_status;

    GNSS_B200_API const char* gnss_b200_last_error(void);
    GNSS_B200_API const char* gnss_b200_version(void);
    GNSS_B200_API int gnss_b200_device_count(void);
    /* Pin / unpin a caller-owned host buffer so copies run at full PCIe rate (optional). */
    GNSS_B200_API int gnss_b200_host_register(void* ptr, size_t bytes);
    GNSS_B200_API int gnss_b200_host_unregister(void* ptr);

    /* ------------------------------------------------------------------ */
    /* (a) PCPS acquisition                                                */
    /* ------------------------------------------------------------------ */

    /* Replaces the state set up by pcps_acquisition::pcps_acquisition (pcps_acquisition.cc:49-166)
       and pcps_acquisition::init (:249-295) for a whole batch of PRN "slots" that share one
       sample stream and one search geometry (one Acq_Conf, acq_conf.h:33-82). */
    typedef struct
    {
        uint32_t fft_size;         /* d_fft_size (:85-92) — any length (see gnss_b200_acq_transform_length for lengths with a prime factor >= 31) */
        uint32_t consumed_samples; /* d_consumed_samples (:71) */
        uint32_t bit_transition;   /* Acq_Conf::bit_transition_flag */
        int32_t doppler_max;       /* Acq_Conf::doppler_max [Hz] */
        uint32_t doppler_step;     /* d_doppler_step [Hz]; bins k = 0 .. ceil(2*max/step)-1 (:264) */
        double fs;                 /* Acq_Conf::fs_in, or resampled_fs with the automatic resampler */
        uint32_t samples_per_chip; /* d_samplesPerChip (exclusion half-width of the 2nd-peak test, :580-581) */
        uint32_t use_cfar;         /* d_use_CFAR_algorithm_flag: 1 = max/input-power (:511-543), 0 = 1st/2nd peak (:546-612) */
        uint32_t keep_grid;        /* 1 = persistent magnitude grid per slot (needed for max_dwells > 1 or dump);
                                      0 = fused search, the grid never reaches HBM (single-dwell searches) */
        uint32_t max_slots;        /* number of PRN slots (each = one reference pcps_acquisition instance) */
        int32_t device;            /* CUDA device ordinal */
        /* Doppler rows [row_begin, row_end) searched by THIS handle; 0,0 = all rows. Lets one PRN's grid be
           sharded across GPUs; merge with gnss_b200_acq_read_rows + gnss_b200_acq_finalize_rows. */
        uint32_t row_begin, row_end;
    } gnss_b200_acq_cfg;

    /* What acquisition_core leaves behind for one PRN after one dwell (:688-713). */
    typedef struct
    {
        float test_statistic;   /* d_test_statistics */
        float peak;             /* grid_maximum / firstPeak */
        float input_power;      /* d_input_power (CFAR only) */
        float second_peak;      /* secondPeak (1st/2nd-peak test only) */
        uint32_t code_phase;    /* indext */
        uint32_t doppler_index; /* index_doppler */
        int32_t doppler_hz;     /* -doppler_max + doppler_center + step*index (:535) */
        uint32_t dwell_count;   /* d_num_noncoherent_integrations_counter after this dwell */
    } gnss_b200_acq_result;

    /* Per-Doppler-row reduction record (exchange unit when rows are sharded over GPUs). */
    typedef struct
    {
        float max_val;    /* max over code phase of the (accumulated) |.|^2 row */
        uint32_t max_idx; /* lowest code-phase index attaining it (volk_gnsssdr_32f_index_max_32u tie rule) */
        float row_sum;    /* sum over the effective row (input-power estimate, :534) */
        uint32_t valid;   /* 1 if this handle computed the row */
    } gnss_b200_acq_row;

    enum
    {
        GNSS_B200_SAMPLES_GR_COMPLEX = 0,
        GNSS_B200_SAMPLES_CSHORT = 1,
        GNSS_B200_SAMPLES_CBYTE = 2
    };

    typedef struct gnss_b200_acq gnss_b200_acq;

    GNSS_B200_API int gnss_b200_acq_create(const gnss_b200_acq_cfg* cfg, gnss_b200_acq** out);
    GNSS_B200_API int gnss_b200_acq_destroy(gnss_b200_acq* h);
    GNSS_B200_API uint32_t gnss_b200_acq_num_doppler_bins(const gnss_b200_acq* h);

    /* make_two_steps (Acq_Conf::make_2_steps, num_doppler_bins_step2, doppler_step2, acq_conf.h:51,62,77):
       a handle for the SECOND, narrow search of acquisition_core (:715-775) — num_doppler_bins_step2 bins spaced
       doppler_step2 Hz around each slot's own d_doppler_center_step_two (update_grid_doppler_wipeoffs_step2, :308-315).
       Same geometry fields as the first-step handle (doppler_max / doppler_step / row range are ignored). Codes, dwell and
       results work as for a first-step handle; doppler_hz is static_cast<int32_t>(centre + (k - floor(n2/2)) * step2) (:539)
       and the CFAR statistic divides by the input power the FIRST step left behind (:531-540 do not refresh it). At most
       64 slots per dwell call (every slot has its own Doppler grid). */
    GNSS_B200_API int gnss_b200_acq_create_step2(const gnss_b200_acq_cfg* cfg, uint32_t num_doppler_bins_step2, float doppler_step2, gnss_b200_acq** out);
    /* d_doppler_center_step_two = Acq_doppler_hz of the first step (:937-940), and that step's d_input_power */
    GNSS_B200_API int gnss_b200_acq_set_step2_center(gnss_b200_acq* h, uint32_t slot, float doppler_center_step_two_hz, float input_power_step_one);

    /* pcps_acquisition::set_local_code (:176-209). `code_c64` is the adapter's time-domain code
       (consumed_samples complex values; fft_size/2 are read when bit_transition). The zero-pad layout,
       forward FFT and conjugation run on the device. */
    GNSS_B200_API int gnss_b200_acq_set_local_code(gnss_b200_acq* h, uint32_t slot, const float* code_c64);
    /* Batch form: codes_c64 = [n_slots][consumed_samples] for slots[0..n_slots). */
    GNSS_B200_API int gnss_b200_acq_set_local_codes(gnss_b200_acq* h, uint32_t n_slots, const uint32_t* slots, const float* codes_c64);

    /* pcps_acquisition::set_doppler_center (pcps_acquisition.h:199-208) plus the FDMA bias of is_fdma() (:212-230). */
    GNSS_B200_API int gnss_b200_acq_set_doppler_center(gnss_b200_acq* h, uint32_t slot, int32_t center_hz, int32_t bias_hz);

    /* Zero the slot's grid and dwell counter (init() :281-284, :879). */
    GNSS_B200_API int gnss_b200_acq_reset(gnss_b200_acq* h, uint32_t slot);

    /* One coherent dwell (one acquisition_core pass, :615-713) for a batch of slots on ONE sample block.
       samples: consumed_samples items, gr_complex (is_cshort = 0) or lv_16sc_t (is_cshort = 1, :623-626); is_cshort = 2
       (GNSS_B200_SAMPLES_CBYTE) takes interleaved int8 "cbyte" items straight from the signal conditioner and
       does the adapter's cbyte -> gr_complex conversion (gps_l1_ca_pcps_acquisition.cc:76-80,191-198) on the device.
       results[i] belongs to slots[i]. */
    GNSS_B200_API int gnss_b200_acq_dwell(gnss_b200_acq* h, const void* samples, int is_cshort,
        uint32_t n_slots, const uint32_t* slots, gnss_b200_acq_result* results);

    /* Same, with the sample block already resident in device memory (gr_complex only) and no host
       synchronisation: results land in the handle's pinned result buffer; call _sync to wait and fetch. */
    GNSS_B200_API int gnss_b200_acq_dwell_device_async(gnss_b200_acq* h, const void* d_samples,
        uint32_t n_slots, const uint32_t* slots);
    GNSS_B200_API int gnss_b200_acq_sync(gnss_b200_acq* h, uint32_t n_slots, gnss_b200_acq_result* results);

    /* ---- sharded search: Doppler rows of ONE search spread over several handles / GPUs / processes ------------------------
       No counterpart in the reference (its only multi-GPU logic is `rand() % num_devices`, cuda_multicorrelator.cu:153); the
       unit that is exchanged is the per-row record the reference reduces its grid to (:511-543).
       With `world` shards, shard r is a handle created with row_begin = r * chunk, row_end = min(n_doppler, (r + 1) * chunk),
       chunk = ceil(n_doppler / world). gnss_b200_acq_exchange_buffers allocates, on the handle's device, `d_send`
       ([max_slots][chunk] records: what this shard contributes) and `d_recv` ([world][max_slots][chunk]: everybody's).
       From then on every dwell leaves this shard's rows of the call's n_slots slots in d_send as [n_slots][chunk], on the
       handle's stream. The caller moves them — NCCL all-gather of n_slots * chunk records per rank straight between the
       device buffers, or cudaMemcpyPeerAsync into the root's d_recv + r * n_slots * chunk — and then
       gnss_b200_acq_finalize_gathered_async decides every slot from d_recv on the device (CFAR statistic; same arithmetic
       and tie rule as an unsharded handle) and queues the result copy; gnss_b200_acq_sync fetches it. */
    GNSS_B200_API int gnss_b200_acq_exchange_buffers(gnss_b200_acq* h, uint32_t world, void** d_send, void** d_recv,
        uint64_t* bytes_per_shard, uint32_t* rows_per_shard);
    GNSS_B200_API int gnss_b200_acq_finalize_gathered_async(gnss_b200_acq* h, uint32_t n_slots);

    /* The exchange itself over peer memory, fused behind the search — no collective library between the kernels:
       every shard owns a region (flags + double-buffered [world][max_slots][chunk] records) that its peers can store into;
       gnss_b200_acq_peer_exchange_async queues, on the handle's stream, (1) a kernel that folds this shard's per-tile records
       into row records, stores them into EVERY shard's region over NVLink and then raises this shard's flag there (release,
       system scope), (2) the decision kernel, which waits for all `world` flags of the current exchange (acquire, system
       scope; ~2 s time-out -> NaN statistics) and decides from the records, (3) the result copy. All of it may sit inside
       gnss_b200_acq_capture_begin / _end, so a rank's whole step is one graph launch. Every shard must call it once per
       search, in the same order. Once connected, a dwell leaves the shard-local merge / decision / pack launches out (their
       work is kernel (1)): the search's results exist only after this call, and gnss_b200_acq_sync asked for results before
       it returns GNSS_B200_ERR_STATE.
       Setup, after gnss_b200_acq_exchange_buffers: _peer_setup(rank) allocates the region and returns its CUDA IPC handle (64
       opaque bytes, for peers in OTHER processes: exchange them with any host-side all-gather) and its device pointer (for
       handles of the SAME process, also on other devices); _peer_connect takes either all `world` IPC handles or all `world`
       region pointers (the entry of the own rank is ignored). One shard per GPU: shards that share a GPU could wait on each
       other's unscheduled kernels. */
    GNSS_B200_API int gnss_b200_acq_peer_setup(gnss_b200_acq* h, uint32_t rank, void* ipc_handle_out /* 64 bytes or NULL */, void** region_out);
    GNSS_B200_API int gnss_b200_acq_peer_connect(gnss_b200_acq* h, const void* ipc_handles /* [world][64] or NULL */, void* const* regions /* [world] or NULL */);
    GNSS_B200_API int gnss_b200_acq_peer_exchange_async(gnss_b200_acq* h, uint32_t n_slots);
    /* Rendezvous of all connected shards on the device, queued on the handle's stream: a one-warp kernel that announces itself in
       every region and waits for all `world` shards (~2 s time-out). Shards leave it within an NVLink store latency of each
       other - put it in front of a search that has to start on all GPUs together (a collective's ranks finish tens of
       microseconds apart). Every shard must call it, in the same order relative to its exchanges. */
    GNSS_B200_API int gnss_b200_acq_peer_barrier_async(gnss_b200_acq* h);

    /* The same inside ONE process that drives several GPUs (the reference receiver is one process): a group owns one engine
       per listed device, splits the Doppler rows of every slot over them (Doppler-major: no forward transform is computed
       twice and every device carries an equal share of every PRN), broadcasts each sample block host -> devices[0] -> the
       others, lets every device reduce its rows, moves the row records device-to-device into devices[0]
       (cudaMemcpyPeerAsync, NVLink) and decides there. Same slots / codes / Doppler-centre / reset / dwell semantics as a
       single handle (gnss_b200_acq_*); cfg->device and the row range are ignored; CFAR statistic and gr_complex blocks only
       (the first-vs-second-peak test needs the winning row's magnitudes on one device). Listing one device twice is legal. */
    typedef struct gnss_b200_acq_group gnss_b200_acq_group;
    GNSS_B200_API int gnss_b200_acq_group_create(const gnss_b200_acq_cfg* cfg, uint32_t n_devices, const int32_t* devices, gnss_b200_acq_group** out);
    GNSS_B200_API int gnss_b200_acq_group_destroy(gnss_b200_acq_group* g);
    GNSS_B200_API uint32_t gnss_b200_acq_group_num_doppler_bins(const gnss_b200_acq_group* g);
    GNSS_B200_API uint32_t gnss_b200_acq_group_num_shards(const gnss_b200_acq_group* g);   /* min(n_devices, Doppler rows) */
    GNSS_B200_API int gnss_b200_acq_group_set_local_codes(gnss_b200_acq_group* g, uint32_t n_slots, const uint32_t* slots, const float* codes_c64);
    GNSS_B200_API int gnss_b200_acq_group_set_doppler_center(gnss_b200_acq_group* g, uint32_t slot, int32_t center_hz, int32_t bias_hz);
    GNSS_B200_API int gnss_b200_acq_group_reset(gnss_b200_acq_group* g, uint32_t slot);
    GNSS_B200_API int gnss_b200_acq_group_dwell(gnss_b200_acq_group* g, const void* samples, int is_cshort, uint32_t n_slots, const uint32_t* slots,
        gnss_b200_acq_result* results);
    GNSS_B200_API int gnss_b200_acq_group_read_grid(gnss_b200_acq_group* g, uint32_t slot, float* out);  /* [n_doppler][effective], keep_grid */
    GNSS_B200_API uint64_t gnss_b200_acq_group_launch_count(const gnss_b200_acq_group* g);

    /* A dwell as a CUDA graph: between _capture_begin and _capture_end exactly one gnss_b200_acq_dwell_device_async is
       recorded instead of executed (after at least one ordinary dwell with the same slots, so that nothing allocates);
       gnss_b200_acq_replay_async then re-issues all its launches with ONE call on the handle's stream. The device sample
       pointer and the slot list are baked in; handles with a persistent grid are refused (dwell counters live on the host).
       For small grids (BASELINE configs[0]) and sharded searches the per-launch overhead is a large part of a dwell. */
    GNSS_B200_API int gnss_b200_acq_capture_begin(gnss_b200_acq* h);
    GNSS_B200_API int gnss_b200_acq_capture_end(gnss_b200_acq* h);
    GNSS_B200_API int gnss_b200_acq_replay_async(gnss_b200_acq* h);

    /* The accumulated magnitude grid of one slot, [n_doppler][effective_size] float32 (the content of the
       reference's d_magnitude_grid / MAT dump `acq_grid`, :433-434,:684). Requires keep_grid = 1. */
    GNSS_B200_API int gnss_b200_acq_read_grid(gnss_b200_acq* h, uint32_t slot, float* out);

    /* Row records of the last dwell for one slot: rows[n_doppler]. */
    GNSS_B200_API int gnss_b200_acq_read_rows(gnss_b200_acq* h, uint32_t slot, gnss_b200_acq_row* rows);
    /* Pure host function: cross-row decision of max_to_input_power_statistic (:511-543) from (merged) row
       records. Only the CFAR statistic can be finalised from records alone. */
    GNSS_B200_API int gnss_b200_acq_finalize_rows(const gnss_b200_acq_row* rows, uint32_t n_doppler, uint32_t effective_size,
        int32_t doppler_max, uint32_t doppler_step, int32_t doppler_center, uint32_t dwell_count, gnss_b200_acq_result* out);

    /* calculate_threshold (:894-909): 2 * gamma_p_inv(2*dof_dwells, (1-pfa)^(1/num_bins)). Host only. */
    GNSS_B200_API float gnss_b200_acq_threshold(float pfa, uint32_t effective_size, uint32_t n_doppler, uint32_t max_dwells, int bit_transition);

    /* Per-kernel timing with CUDA events on the handle's stream. enable = 1 starts recording around every kernel
       launch; the call synchronises, returns the accumulated milliseconds and launch counts per kernel class and
       (reset != 0) clears them. Kernel classes of the acquisition handle: 0 forward FFT stages, 1 inverse first
       stage (two-stage plans), 2 inverse last stage + |.|^2 + row reductions, 3 code spectra, 4 decision kernels.
       Multicorrelator handle: 0 the fused correlator kernel. */
#define GNSS_B200_PROFILE_SLOTS 8
    GNSS_B200_API int gnss_b200_acq_profile(gnss_b200_acq* h, int enable, double* ms_by_kernel, uint64_t* launches_by_kernel, int reset);

    /* The CUDA stream (cudaStream_t) the handle launches on, and the number of kernel launches it has issued
       (for event timing and launch accounting by the caller). */
    GNSS_B200_API void* gnss_b200_acq_stream(gnss_b200_acq* h);
    GNSS_B200_API uint64_t gnss_b200_acq_launch_count(const gnss_b200_acq* h);

    /* Length of the transforms the handle runs: fft_size, or the longer 2^a 3^b 5^c length of the padded cyclic correlation
       the engine chose because fft_size has a prime factor its codelets do not cover (> 127) or cover only by a slow direct
       pass (31 ... 127). The reference's FFTW plan takes any d_fft_size (pcps_acquisition.cc:85-92,114-115); results, indices
       and magnitudes are those of the fft_size-point search either way. */
    GNSS_B200_API uint32_t gnss_b200_acq_transform_length(const gnss_b200_acq* h);

    /* ------------------------------------------------------------------ */
    /* (b) tracking multicorrelator                                        */
    /* ------------------------------------------------------------------ */

    /* Arguments of one Carrier_wipeoff_multicorrelator_resampler call
       (cpu_multicorrelator_real_codes.cc:103-110), for one channel of a batch. */
    typedef struct
    {
        const float* sig_in_c64;     /* d_sig_in (host pointer), or NULL to window the shared stream */
        uint64_t stream_offset;      /* first sample in the shared stream when sig_in_c64 == NULL */
        uint32_t n_samples;          /* signal_length_samples */
        float rem_carr_phase_rad;    /* rem_carrier_phase_in_rad */
        float carr_step_rad;         /* phase_step_rad */
        float carr_rate_rad;         /* phase_rate_step_rad (high-dynamics rotator only) */
        float rem_code_phase_chips;  /* rem_code_phase_chips */
        float code_step_chips;       /* code_phase_step_chips */
        float code_rate_chips;       /* code_phase_rate_step_chips (high-dynamics resampler only) */
    } gnss_b200_mc_args;

    enum
    {
        GNSS_B200_MC_HIGH_DYN = 1, /* d_use_high_dynamics_resampler = true (set_high_dynamics_resampler) */
        GNSS_B200_MC_SIX_ARG = 2   /* the 6-argument overload (:129-144): plain rotator even when HIGH_DYN */
    };

    typedef struct gnss_b200_mc gnss_b200_mc;

    /* Cpu_Multicorrelator_Real_Codes::init (:36-49) for `max_channels` correlator objects at once. */
    GNSS_B200_API int gnss_b200_mc_create(uint32_t max_channels, uint32_t max_samples, uint32_t max_taps, int32_t device, gnss_b200_mc** out);
    GNSS_B200_API int gnss_b200_mc_destroy(gnss_b200_mc* h);

    /* set_local_code_and_taps (:52-62), code part: copies the local code of channel `ch` to the device.
       is_complex = 0: `code` holds code_len float32 (Cpu_Multicorrelator_Real_Codes);
       is_complex = 1: code_len interleaved complex (Cpu_Multicorrelator, cpu_multicorrelator.cc:50-60). */
    GNSS_B200_API int gnss_b200_mc_set_code(gnss_b200_mc* h, uint32_t ch, uint32_t code_len, const float* code, int is_complex);

    /* Upload a sample stream shared by all channels (the signal-conditioner output every tracking block
       reads, gnss_flowgraph.cc:1129-1135); channels then window it through stream_offset. */
    GNSS_B200_API int gnss_b200_mc_set_stream(gnss_b200_mc* h, const float* samples_c64, uint64_t n_samples);

    /* One correlation epoch for n_ch channels. shifts = [n_ch][n_taps] (d_shifts_chips, re-read on every
       call: the reference mutates the borrowed array in place, dll_pll_veml_tracking.cc:1961-1979).
       out_c64 = [n_ch][n_taps] (d_corr_out). */
    GNSS_B200_API int gnss_b200_mc_correlate_batch(gnss_b200_mc* h, uint32_t n_ch, const uint32_t* ch,
        const gnss_b200_mc_args* args, const float* shifts, uint32_t n_taps, uint32_t flags, float* out_c64);

    /* Device-resident form for throughput measurement: d_sig is [n_ch][max_samples] complex in device memory
       (independent per-channel streams), no host synchronisation; fetch with _sync. */
    GNSS_B200_API int gnss_b200_mc_correlate_device_async(gnss_b200_mc* h, uint32_t n_ch, const uint32_t* ch,
        const gnss_b200_mc_args* args, const float* shifts, uint32_t n_taps, uint32_t flags, const void* d_sig, uint64_t ch_stride_samples);
    GNSS_B200_API int gnss_b200_mc_sync(gnss_b200_mc* h, uint32_t n_ch, uint32_t n_taps, float* out_c64);

    /* Debug / parity aid: wrapped chip index of every (tap, sample) of one channel, as the kernel computes it. */
    GNSS_B200_API int gnss_b200_mc_chip_indices(gnss_b200_mc* h, uint32_t ch, const gnss_b200_mc_args* args,
        const float* shifts, uint32_t n_taps, uint32_t flags, int32_t* idx_out /* [n_taps][n_samples] */);

    GNSS_B200_API int gnss_b200_mc_profile(gnss_b200_mc* h, int enable, double* ms_by_kernel, uint64_t* launches_by_kernel, int reset);
    GNSS_B200_API void* gnss_b200_mc_stream(gnss_b200_mc* h);
    GNSS_B200_API uint64_t gnss_b200_mc_launch_count(const gnss_b200_mc* h);


    /* ------------------------------------------------------------------ */
    /* (f) N1: the dll_pll_veml tracking loop closed on the device          */
    /* ------------------------------------------------------------------ */

    /* Per-epoch loop of dll_pll_veml_tracking::general_work: do_correlation_step (dll_pll_veml_tracking.cc:1064-1089) ->
       cn0_and_tracking_lock_status (:968-1013) -> run_dll_pll (:1092-1182) -> update_tracking_vars (:1225-1297) ->
       consume_each(d_current_prn_length_samples) (:2118), for many channels at once and for as many epochs as the
       uploaded stream holds, WITHOUT a host round trip between epochs: a loop-update kernel turns the correlator
       outputs of epoch e into the multicorrelator arguments of epoch e+1 in device memory.
       symbol_sync == 0: state 2 only (one code period per integration, Costas PLL), symbol bookkeeping on the host
       (host/b200_symbol_sync). symbol_sync == 1: the block's whole symbol machine runs in the loop — secondary-code /
       bit synchronisation in state 2 (:1888-1925, acquire_secondary :923-966), state 3 coherent extension
       (:1989-2027), state 4 narrow tracking with the narrow bandwidths and taps (:1945-1975, :2028-2084),
       save_correlation_results (:1290-1400), the slave data correlator of pilot tracking (:1078-1088) and the
       high-dynamics rate smoothing (:1232-1284). Not covered: the experimental Doppler correction (:1162-1182), dumps.
       Field names follow Dll_Pll_Conf (dll_pll_conf.h:37-84) and the block's members (dll_pll_veml_tracking.h). */
    typedef struct
    {
        double fs_in;                 /* Dll_Pll_Conf::fs_in */
        double code_chip_rate;        /* d_code_chip_rate [chips/s] */
        double signal_carrier_freq;   /* d_signal_carrier_freq [Hz] (carrier aiding) */
        double code_period_s;         /* d_code_period */
        uint32_t code_length_chips;   /* d_code_length_chips */
        uint32_t code_samples_per_chip; /* d_code_samples_per_chip: 1, or 2 for Galileo E1 (:270) */
        uint32_t vector_length;       /* Dll_Pll_Conf::vector_length: samples correlated per epoch */
        int32_t veml;                 /* 1: five taps VE,E,P,L,VL (d_veml), 0: three taps E,P,L */
        float early_late_space_chips, very_early_late_space_chips;
        float slope, y_intercept;     /* dll_nc_e_minus_l_normalized parameters */
        float pll_bw_hz, dll_bw_hz, fll_bw_hz;
        int32_t pll_filter_order;     /* 2 or 3 */
        int32_t dll_filter_order;     /* 1, 2 or 3 */
        int32_t enable_fll_pull_in, enable_fll_steady_state, carrier_aiding;
        uint32_t pull_in_time_s;
        int32_t cn0_samples;          /* FLAGS_cn0_samples (<= 64) */
        int32_t cn0_min, max_code_lock_fail, max_carrier_lock_fail;
        double carrier_lock_th;
        float cn0_smoother_alpha, carrier_lock_test_smoother_alpha;
        int32_t cn0_smoother_samples, carrier_lock_test_smoother_samples;  /* smoother initialisation lengths (<= 256) */
        uint32_t max_channels;
        int32_t device;
        /* symbol synchronisation, extended integration, pilot tracking, high dynamics (all 0: the state-2 loop only) */
        int32_t symbol_sync;                 /* 1: states 2 -> 3 -> 4 on the device */
        int32_t secondary;                   /* d_secondary: the tracked component has a secondary code */
        uint32_t secondary_code_length;      /* d_secondary_code_length (or the preamble length), <= 200 */
        uint32_t symbols_per_bit;            /* d_symbols_per_bit */
        uint32_t data_secondary_code_length; /* d_data_secondary_code_length, <= 100 */
        int32_t extend_correlation_symbols;  /* Dll_Pll_Conf::extend_correlation_symbols (<= 1: no extension) */
        int32_t track_pilot;                 /* Dll_Pll_Conf::track_pilot: second, 1-tap correlator on the data code */
        float pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_narrow_chips, very_early_late_space_narrow_chips;
        int32_t high_dyn;                    /* Dll_Pll_Conf::high_dyn */
        uint32_t smoother_length;            /* Dll_Pll_Conf::smoother_length (<= 64) */
        uint32_t bit_synchronization_time_limit_s;
        char secondary_code_string[208];     /* d_secondary_code_string: '0' adds, anything else subtracts */
        char data_secondary_code_string[112];
    } gnss_b200_trk_cfg;

    /* What one epoch leaves behind per channel: the quantities log_data() dumps (:1403-1530) and the Gnss_Synchro
       tracking fields. */
    typedef struct
    {
        float corr[10];               /* VE,E,P,L,VL as (re,im) pairs; E,P,L in slots 1..3 when veml == 0 too */
        double carrier_doppler_hz;    /* d_carrier_doppler_hz after this epoch's PLL update */
        double code_freq_chips;       /* d_code_freq_chips */
        double carr_error_hz, carr_error_filt_hz, code_error_chips, code_error_filt_chips;
        double rem_code_phase_samples, acc_carrier_phase_rad;
        double cn0_db_hz, carrier_lock_test;
        double evm;                   /* d_EVM, the fork's error-vector-magnitude indicator over the prompt buffer (:1027-1053) */
        uint64_t sample_counter;      /* absolute index of the first sample correlated in this epoch */
        int32_t prn_length_samples;   /* samples consumed after this epoch (d_current_prn_length_samples) */
        uint32_t flags;               /* GNSS_B200_TRK_* */
        float accu[10];               /* d_VE_accu .. d_VL_accu as the loop of this epoch saw them */
        float prompt_data[2];         /* d_P_data_accu handed to the telemetry decoder when GNSS_B200_TRK_SYMBOL is set */
        int32_t state;                /* d_state the epoch ran in (2, 3 or 4) */
        int32_t reserved;
    } gnss_b200_trk_epoch;

    enum
    {
        GNSS_B200_TRK_VALID = 1,        /* the channel correlated in this epoch */
        GNSS_B200_TRK_PULL_IN = 2,      /* d_pull_in_transitory still set */
        GNSS_B200_TRK_LOSS_OF_LOCK = 4, /* cn0_and_tracking_lock_status() returned false: channel went to standby */
        GNSS_B200_TRK_SYMBOL = 8,       /* Flag_valid_symbol_output: prompt_data holds a symbol (:1997-2018, :2049-2070) */
        GNSS_B200_TRK_PLL_180 = 16      /* d_Flag_PLL_180_deg_phase_locked */
    };

    typedef struct gnss_b200_trk gnss_b200_trk;

    GNSS_B200_API int gnss_b200_trk_create(const gnss_b200_trk_cfg* cfg, gnss_b200_trk** out);
    GNSS_B200_API int gnss_b200_trk_destroy(gnss_b200_trk* h);
    /* the tracking code of channel `ch` (d_tracking_code: code_samples_per_chip * code_length_chips floats) */
    GNSS_B200_API int gnss_b200_trk_set_code(gnss_b200_trk* h, uint32_t ch, uint32_t code_len, const float* code);
    /* the data-component code of channel `ch` for pilot tracking (d_data_code, :670-700); same length as the tracking code */
    GNSS_B200_API int gnss_b200_trk_set_data_code(gnss_b200_trk* h, uint32_t ch, uint32_t code_len, const float* code);
    /* start_tracking() (:640-880) + the state-1 alignment (:1812-1845): `now_sample` is the absolute index of the next
       unread stream sample (nitems_read(0)); the channel's read cursor is moved to the next code-period start. */
    GNSS_B200_API int gnss_b200_trk_start(gnss_b200_trk* h, uint32_t ch, double acq_delay_samples, double acq_doppler_hz,
        uint64_t acq_samplestamp, uint64_t now_sample);
    /* Runs every started channel over the stream block samples_c64[0 .. n_samples) whose first sample has absolute index
       stream_start: up to max_epochs epochs, each needing 2 * vector_length samples ahead of the channel's cursor
       (the block's forecast(), :609). out = [max_epochs][n_ch] records for the listed channels (flags == 0 where a channel
       did not run). samples_is_device != 0: samples_c64 is a device pointer. */
    GNSS_B200_API int gnss_b200_trk_run(gnss_b200_trk* h, const float* samples_c64, int samples_is_device, uint64_t stream_start,
        uint64_t n_samples, uint32_t n_ch, const uint32_t* ch, uint32_t max_epochs, gnss_b200_trk_epoch* out);
    GNSS_B200_API void* gnss_b200_trk_stream(gnss_b200_trk* h);
    GNSS_B200_API uint64_t gnss_b200_trk_launch_count(const gnss_b200_trk* h);

    /* ------------------------------------------------------------------ */
    /* parity aids (pure host, no device needed)                           */
    /* ------------------------------------------------------------------ */

    /* ---------------------------------------------------------------------------------------------
     * Multicorrelator without allocation. NO COUNTERPART IN THE REFERENCE: added for
     * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
     * The same computation as gnss_b200_mc_correlate_batch for real codes and the plain resampler, without a handle: inputs
     * and outputs are the caller's DEVICE buffers, and the scratch (parameter blocks, tap shifts, split partials, tickets)
     * is a fixed device array inside the library (gnss_b200_mc_noalloc_scratch_bytes(), ~141 KB, held by every GPU that
     * loads the library until the process exits). The call allocates nothing (no cudaMalloc / cudaFree / allocating
     * library routine), enqueues its work on `cuda_stream` (a cudaStream_t; NULL = the default stream) and returns without
     * synchronising: d_out_c64 is valid once that stream has caught up.
     * Limits (GNSS_B200_ERR_INVALID with a message beyond them): GNSS_B200_MC_NOALLOC_MAX_CHANNELS channels,
     * GNSS_B200_MC_NOALLOC_MAX_TAPS taps, GNSS_B200_MC_NOALLOC_MAX_SAMPLES samples per channel and call.
     * All calls on one GPU share the scratch: calls whose GPU work overlaps (two streams, two threads) RACE — serialise them.
     * The call issues kernel launches only — parameters and shifts travel as by-value kernel arguments, no host pointer
     * outlives the call — so it may be recorded under stream capture and the graph replayed (the replay re-runs the same
     * parameters on whatever the caller's device buffers hold then). Make the first call on a device before capturing (it
     * resolves the scratch address). The caller's current device is restored on return.
     * ------------------------------------------------------------------------------------------- */
#define GNSS_B200_MC_NOALLOC_MAX_CHANNELS 64u
#define GNSS_B200_MC_NOALLOC_MAX_TAPS 16u
#define GNSS_B200_MC_NOALLOC_MAX_SAMPLES (1u << 22)
    typedef struct
    {
        const void* d_sig_c64;      /* DEVICE pointer: n_samples gr_complex (interleaved float re, im) */
        const float* d_code;        /* DEVICE pointer: code_len floats (d_local_code_in) */
        uint32_t code_len;          /* code_length_chips (x samples per chip) */
        uint32_t n_samples;         /* signal_length_samples */
        float rem_carr_phase_rad, carr_step_rad, rem_code_phase_chips, code_step_chips;
    } gnss_b200_mc_noalloc_chan;
    /* shifts: HOST [n_ch][n_taps] (d_shifts_chips); d_out_c64: DEVICE [n_ch][n_taps] gr_complex (d_corr_out) */
    GNSS_B200_API int gnss_b200_mc_correlate_noalloc(int32_t device, void* cuda_stream, uint32_t n_ch, const gnss_b200_mc_noalloc_chan* chans,
        const float* shifts, uint32_t n_taps, void* d_out_c64);
    GNSS_B200_API uint64_t gnss_b200_mc_noalloc_scratch_bytes(void);

    /* ---------------------------------------------------------------------------------------------
     * Acquisition resampler: the decimating low-pass FIR GNSSFlowgraph connects in front of the
     * acquisition blocks with GNSS-SDR.use_acquisition_resampler (gnss_flowgraph.cc:1063-1118; decimation from
     * Acq_Conf::ConfigureAutomaticResampler, acq_conf.cc:93-111). Replaces gr::filter::firdes::low_pass and
     * gr::filter::fir_filter_ccf (GNU Radio, not part of the reference tree: published algorithm restated).
     * ------------------------------------------------------------------------------------------- */
    typedef struct gnss_b200_resampler gnss_b200_resampler;
    /* firdes::low_pass(gain, fs, cutoff, transition) with the default Hamming window: odd tap count
       int(53 fs / (22 transition)), windowed sinc, unity DC gain times `gain`. The flow graph calls it as
       low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 2) (:1084-1087). */
    GNSS_B200_API int gnss_b200_firdes_low_pass(double gain, double sampling_freq, double cutoff_freq, double transition_width,
        uint32_t max_taps, float* taps, uint32_t* n_taps);
    /* fir_filter_ccf::make(decimation, taps) (:1089); max_input_items bounds one work() call */
    GNSS_B200_API int gnss_b200_resampler_create(uint32_t decimation, uint32_t n_taps, const float* taps, uint32_t max_input_items,
        int32_t device, gnss_b200_resampler** out);
    GNSS_B200_API int gnss_b200_resampler_destroy(gnss_b200_resampler* h);
    /* forget the history: the next item is the first of a stream (ntaps - 1 zeros before it) */
    GNSS_B200_API int gnss_b200_resampler_reset(gnss_b200_resampler* h);
    /* one work() call: n_input_items (a multiple of the decimation) gr_complex items from the host, y[m] = sum_j taps[j] x[m D - j]
       with the previous calls' items as history. The n_input_items / decimation output items stay in device memory
       (*out_device, valid until the next call: pass it to gnss_b200_acq_dwell_device_async) and are copied to out_c64 if
       that is not NULL. The filter delay the block has to report is (ntaps - 1) / 2 input items
       (set_resampler_latency, :1111). */
    GNSS_B200_API int gnss_b200_resampler_work(gnss_b200_resampler* h, const float* in_c64, uint32_t n_input_items, float* out_c64,
        const void** out_device);

    /* The binary32 running-sum phase of the reference's wipe-off row for one Doppler bin
       (volk_gnsssdr_s32f_sincos_32fc_generic, volk_gnsssdr_s32f_sincos_32fc.h:390-400), as reproduced in closed
       form by the engine: phase_out[n], n < n_samples. n_pieces (optional) = number of linear pieces used. */
    GNSS_B200_API int gnss_b200_debug_wipeoff_phase(int32_t doppler_hz_with_bias, double fs, uint32_t n_samples, float* phase_out, uint32_t* n_pieces);
    /* The FFT plan for length n: number of stages, radices (lowest stride first) and the position -> frequency
       permutation of the forward transform's output (any of the out pointers may be NULL). */
    /* Measured FP32 FMA throughput of the device [TFLOP/s] (dependent-chain FFMA loop over all SMs): the
       denominator for the FP32 roofline of the acquisition kernels, which no driver-written file provides. */
    GNSS_B200_API int gnss_b200_debug_fp32_peak(int32_t device, double* tflops);
    GNSS_B200_API int gnss_b200_debug_fft_plan(uint32_t n, uint32_t* n_stages, uint32_t* radices, uint32_t* n_radices, uint32_t* perm);
    /* The transform length gnss_b200_acq_create would choose for this search (gnss_b200_acq_transform_length of the handle);
       needs no device. */
    GNSS_B200_API int gnss_b200_debug_transform_length(uint32_t fft_size, uint32_t consumed_samples, uint32_t* length);

#ifdef __cplusplus
}
#endif
#endif /* GNSS_B200_H */
