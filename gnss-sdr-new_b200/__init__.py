"""gnss-sdr-new_b200 — B200-native GNSS correlation engine (PCPS acquisition + tracking multicorrelator).

Python is only the test / bench driver: this module is a thin ctypes layer over the C-ABI of
libgnss_b200.so (include/gnss_b200.h). All arithmetic runs in hand-written sm_100a CUDA kernels; there is
no CPU fallback and this package never imports anything from oracle/.
"""
import ctypes as C
import math
import os

import numpy as np

from . import sharding  # noqa: F401  (multi-GPU partition / gather host logic)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgnss_b200.so")

c_u32 = C.c_uint32
c_i32 = C.c_int32
_fp = C.POINTER(C.c_float)
_u32p = C.POINTER(c_u32)


class B200Error(RuntimeError):
    pass


class AcqCfg(C.Structure):
    _fields_ = [
        ("fft_size", c_u32), ("consumed_samples", c_u32), ("bit_transition", c_u32), ("doppler_max", c_i32),
        ("doppler_step", c_u32), ("fs", C.c_double), ("samples_per_chip", c_u32), ("use_cfar", c_u32),
        ("keep_grid", c_u32), ("max_slots", c_u32), ("device", c_i32), ("row_begin", c_u32), ("row_end", c_u32),
    ]


class AcqResult(C.Structure):
    _fields_ = [
        ("test_statistic", C.c_float), ("peak", C.c_float), ("input_power", C.c_float), ("second_peak", C.c_float),
        ("code_phase", c_u32), ("doppler_index", c_u32), ("doppler_hz", c_i32), ("dwell_count", c_u32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class AcqRow(C.Structure):
    _fields_ = [("max_val", C.c_float), ("max_idx", c_u32), ("row_sum", C.c_float), ("valid", c_u32)]


ACQ_ROW_DTYPE = np.dtype([("max_val", np.float32), ("max_idx", np.uint32), ("row_sum", np.float32), ("valid", np.uint32)])
ACQ_RESULT_DTYPE = np.dtype([("test_statistic", np.float32), ("peak", np.float32), ("input_power", np.float32),
                             ("second_peak", np.float32), ("code_phase", np.uint32), ("doppler_index", np.uint32),
                             ("doppler_hz", np.int32), ("dwell_count", np.uint32)])


class McArgs(C.Structure):
    _fields_ = [
        ("sig_in_c64", C.c_void_p), ("stream_offset", C.c_uint64), ("n_samples", c_u32),
        ("rem_carr_phase_rad", C.c_float), ("carr_step_rad", C.c_float), ("carr_rate_rad", C.c_float),
        ("rem_code_phase_chips", C.c_float), ("code_step_chips", C.c_float), ("code_rate_chips", C.c_float),
    ]


MC_HIGH_DYN = 1
MC_SIX_ARG = 2


class McNoAllocChan(C.Structure):
    """gnss_b200_mc_noalloc_chan: one channel of gnss_b200_mc_correlate_noalloc (device pointers, no handle)."""
    _fields_ = [
        ("d_sig_c64", C.c_void_p), ("d_code", C.c_void_p), ("code_len", c_u32), ("n_samples", c_u32),
        ("rem_carr_phase_rad", C.c_float), ("carr_step_rad", C.c_float), ("rem_code_phase_chips", C.c_float), ("code_step_chips", C.c_float),
    ]


MC_NOALLOC_MAX_CHANNELS, MC_NOALLOC_MAX_TAPS, MC_NOALLOC_MAX_SAMPLES = 64, 16, 1 << 22


class TrkCfg(C.Structure):
    """gnss_b200_trk_cfg: Dll_Pll_Conf (dll_pll_conf.h:37-84) + the signal constants the block derives (dll_pll_veml_tracking.cc:105-460)."""
    _fields_ = [
        ("fs_in", C.c_double), ("code_chip_rate", C.c_double), ("signal_carrier_freq", C.c_double), ("code_period_s", C.c_double),
        ("code_length_chips", c_u32), ("code_samples_per_chip", c_u32), ("vector_length", c_u32), ("veml", c_i32),
        ("early_late_space_chips", C.c_float), ("very_early_late_space_chips", C.c_float), ("slope", C.c_float), ("y_intercept", C.c_float),
        ("pll_bw_hz", C.c_float), ("dll_bw_hz", C.c_float), ("fll_bw_hz", C.c_float),
        ("pll_filter_order", c_i32), ("dll_filter_order", c_i32),
        ("enable_fll_pull_in", c_i32), ("enable_fll_steady_state", c_i32), ("carrier_aiding", c_i32),
        ("pull_in_time_s", c_u32),
        ("cn0_samples", c_i32), ("cn0_min", c_i32), ("max_code_lock_fail", c_i32), ("max_carrier_lock_fail", c_i32),
        ("carrier_lock_th", C.c_double),
        ("cn0_smoother_alpha", C.c_float), ("carrier_lock_test_smoother_alpha", C.c_float),
        ("cn0_smoother_samples", c_i32), ("carrier_lock_test_smoother_samples", c_i32),
        ("max_channels", c_u32), ("device", c_i32),
        # symbol synchronisation, extended integration, pilot tracking, high dynamics (all 0: the state-2 loop only)
        ("symbol_sync", c_i32), ("secondary", c_i32),
        ("secondary_code_length", c_u32), ("symbols_per_bit", c_u32), ("data_secondary_code_length", c_u32),
        ("extend_correlation_symbols", c_i32), ("track_pilot", c_i32),
        ("pll_bw_narrow_hz", C.c_float), ("dll_bw_narrow_hz", C.c_float), ("early_late_space_narrow_chips", C.c_float), ("very_early_late_space_narrow_chips", C.c_float),
        ("high_dyn", c_i32), ("smoother_length", c_u32), ("bit_synchronization_time_limit_s", c_u32),
        ("secondary_code_string", C.c_char * 208), ("data_secondary_code_string", C.c_char * 112),
    ]


def trk_config(fs_in, code_chip_rate, code_length_chips, signal_carrier_freq, max_channels, code_samples_per_chip=1, veml=False, device=0, **kw):
    """A gnss_b200_trk_cfg with the reference's defaults (dll_pll_conf.h:41-84, gnss_sdr_flags.cc:45-54); keyword
    arguments override single fields."""
    code_period_s = code_length_chips / code_chip_rate
    c = TrkCfg(fs_in=fs_in, code_chip_rate=code_chip_rate, signal_carrier_freq=signal_carrier_freq, code_period_s=code_period_s,
               code_length_chips=code_length_chips, code_samples_per_chip=code_samples_per_chip,
               vector_length=int(round(fs_in / (code_chip_rate / code_length_chips))), veml=int(veml),
               early_late_space_chips=0.25 if not veml else 0.15, very_early_late_space_chips=0.5, slope=1.0, y_intercept=1.0,
               pll_bw_hz=35.0, dll_bw_hz=2.0, fll_bw_hz=35.0, pll_filter_order=3, dll_filter_order=2,
               enable_fll_pull_in=0, enable_fll_steady_state=0, carrier_aiding=1, pull_in_time_s=10,
               cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.7,
               cn0_smoother_alpha=0.002, carrier_lock_test_smoother_alpha=0.002, cn0_smoother_samples=200,
               carrier_lock_test_smoother_samples=25, max_channels=max_channels, device=device)
    for k, v in kw.items():
        if not hasattr(c, k):
            raise B200Error(f"trk_config: unknown field {k}")
        setattr(c, k, v)
    return c


TRK_EPOCH_DTYPE = np.dtype([("corr", np.float32, (10,)), ("carrier_doppler_hz", np.float64), ("code_freq_chips", np.float64),
                            ("carr_error_hz", np.float64), ("carr_error_filt_hz", np.float64), ("code_error_chips", np.float64),
                            ("code_error_filt_chips", np.float64), ("rem_code_phase_samples", np.float64),
                            ("acc_carrier_phase_rad", np.float64), ("cn0_db_hz", np.float64), ("carrier_lock_test", np.float64),
                            ("evm", np.float64), ("sample_counter", np.uint64), ("prn_length_samples", np.int32), ("flags", np.uint32),
                            ("accu", np.float32, (10,)), ("prompt_data", np.float32, (2,)), ("state", np.int32), ("reserved", np.int32)], align=True)
TRK_VALID, TRK_PULL_IN, TRK_LOSS_OF_LOCK, TRK_SYMBOL, TRK_PLL_180 = 1, 2, 4, 8, 16

# every symbol include/gnss_b200.h declares (tests check the library exports all of them)
EXPORTS = [
    "gnss_b200_last_error", "gnss_b200_version", "gnss_b200_device_count", "gnss_b200_host_register",
    "gnss_b200_host_unregister", "gnss_b200_acq_create", "gnss_b200_acq_destroy", "gnss_b200_acq_num_doppler_bins",
    "gnss_b200_acq_set_local_code", "gnss_b200_acq_set_local_codes", "gnss_b200_acq_set_doppler_center",
    "gnss_b200_acq_reset", "gnss_b200_acq_dwell", "gnss_b200_acq_dwell_device_async", "gnss_b200_acq_sync",
    "gnss_b200_acq_read_grid", "gnss_b200_acq_read_rows", "gnss_b200_acq_finalize_rows", "gnss_b200_acq_threshold",
    "gnss_b200_acq_stream", "gnss_b200_acq_launch_count", "gnss_b200_acq_transform_length", "gnss_b200_mc_create", "gnss_b200_mc_destroy",
    "gnss_b200_mc_set_code", "gnss_b200_mc_set_stream", "gnss_b200_mc_correlate_batch",
    "gnss_b200_mc_correlate_device_async", "gnss_b200_mc_sync", "gnss_b200_mc_chip_indices", "gnss_b200_mc_stream",
    "gnss_b200_mc_launch_count", "gnss_b200_debug_wipeoff_phase", "gnss_b200_debug_fft_plan", "gnss_b200_debug_transform_length",
    "gnss_b200_acq_profile", "gnss_b200_mc_profile", "gnss_b200_debug_fp32_peak",
    "gnss_b200_acq_create_step2", "gnss_b200_acq_set_step2_center",
    "gnss_b200_mc_correlate_noalloc", "gnss_b200_mc_noalloc_scratch_bytes",
    "gnss_b200_firdes_low_pass", "gnss_b200_resampler_create", "gnss_b200_resampler_destroy", "gnss_b200_resampler_reset",
    "gnss_b200_resampler_work",
    "gnss_b200_trk_create", "gnss_b200_trk_destroy", "gnss_b200_trk_set_code", "gnss_b200_trk_set_data_code", "gnss_b200_trk_start", "gnss_b200_trk_run",
    "gnss_b200_trk_stream", "gnss_b200_trk_launch_count",
    "gnss_b200_acq_exchange_buffers", "gnss_b200_acq_finalize_gathered_async",
    "gnss_b200_acq_capture_begin", "gnss_b200_acq_capture_end", "gnss_b200_acq_replay_async",
    "gnss_b200_acq_peer_setup", "gnss_b200_acq_peer_connect", "gnss_b200_acq_peer_exchange_async", "gnss_b200_acq_peer_barrier_async",
    "gnss_b200_acq_group_create", "gnss_b200_acq_group_destroy", "gnss_b200_acq_group_num_doppler_bins", "gnss_b200_acq_group_num_shards",
    "gnss_b200_acq_group_set_local_codes", "gnss_b200_acq_group_set_doppler_center", "gnss_b200_acq_group_reset", "gnss_b200_acq_group_dwell",
    "gnss_b200_acq_group_read_grid", "gnss_b200_acq_group_launch_count",
]

_lib = None


def lib():
    """Load libgnss_b200.so. Fails loudly if it was not built (python gnss-sdr-new_b200/build.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(f"{LIB_PATH} is missing: build it with `python gnss-sdr-new_b200/build.py` (needs nvcc). "
                        "There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    L.gnss_b200_last_error.restype = C.c_char_p
    L.gnss_b200_version.restype = C.c_char_p
    L.gnss_b200_acq_create.argtypes = [C.POINTER(AcqCfg), C.POINTER(C.c_void_p)]
    L.gnss_b200_mc_correlate_noalloc.argtypes = [c_i32, C.c_void_p, c_u32, C.POINTER(McNoAllocChan), C.POINTER(C.c_float), c_u32, C.c_void_p]
    L.gnss_b200_mc_noalloc_scratch_bytes.restype = C.c_uint64
    L.gnss_b200_mc_noalloc_scratch_bytes.argtypes = []
    L.gnss_b200_firdes_low_pass.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, c_u32, C.POINTER(C.c_float), C.POINTER(c_u32)]
    L.gnss_b200_resampler_create.argtypes = [c_u32, c_u32, C.POINTER(C.c_float), c_u32, c_i32, C.POINTER(C.c_void_p)]
    L.gnss_b200_resampler_destroy.argtypes = [C.c_void_p]
    L.gnss_b200_resampler_reset.argtypes = [C.c_void_p]
    L.gnss_b200_resampler_work.argtypes = [C.c_void_p, C.c_void_p, c_u32, C.c_void_p, C.POINTER(C.c_void_p)]
    L.gnss_b200_acq_destroy.argtypes = [C.c_void_p]
    L.gnss_b200_acq_create_step2.argtypes = [C.POINTER(AcqCfg), c_u32, C.c_float, C.POINTER(C.c_void_p)]
    L.gnss_b200_acq_set_step2_center.argtypes = [C.c_void_p, c_u32, C.c_float, C.c_float]
    L.gnss_b200_acq_num_doppler_bins.argtypes = [C.c_void_p]
    L.gnss_b200_acq_num_doppler_bins.restype = c_u32
    L.gnss_b200_acq_set_local_code.argtypes = [C.c_void_p, c_u32, _fp]
    L.gnss_b200_acq_set_local_codes.argtypes = [C.c_void_p, c_u32, _u32p, _fp]
    L.gnss_b200_acq_set_doppler_center.argtypes = [C.c_void_p, c_u32, c_i32, c_i32]
    L.gnss_b200_acq_reset.argtypes = [C.c_void_p, c_u32]
    L.gnss_b200_acq_dwell.argtypes = [C.c_void_p, C.c_void_p, C.c_int, c_u32, _u32p, C.c_void_p]
    L.gnss_b200_acq_dwell_device_async.argtypes = [C.c_void_p, C.c_void_p, c_u32, _u32p]
    L.gnss_b200_acq_sync.argtypes = [C.c_void_p, c_u32, C.c_void_p]
    L.gnss_b200_acq_exchange_buffers.argtypes = [C.c_void_p, c_u32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), _u32p]
    L.gnss_b200_acq_finalize_gathered_async.argtypes = [C.c_void_p, c_u32]
    L.gnss_b200_acq_capture_begin.argtypes = [C.c_void_p]
    L.gnss_b200_acq_capture_end.argtypes = [C.c_void_p]
    L.gnss_b200_acq_replay_async.argtypes = [C.c_void_p]
    L.gnss_b200_acq_peer_setup.argtypes = [C.c_void_p, c_u32, C.c_void_p, C.POINTER(C.c_void_p)]
    L.gnss_b200_acq_peer_connect.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
    L.gnss_b200_acq_peer_exchange_async.argtypes = [C.c_void_p, c_u32]
    L.gnss_b200_acq_peer_barrier_async.argtypes = [C.c_void_p]
    L.gnss_b200_acq_group_create.argtypes = [C.POINTER(AcqCfg), c_u32, C.POINTER(c_i32), C.POINTER(C.c_void_p)]
    L.gnss_b200_acq_group_destroy.argtypes = [C.c_void_p]
    L.gnss_b200_acq_group_num_doppler_bins.argtypes = [C.c_void_p]
    L.gnss_b200_acq_group_num_doppler_bins.restype = c_u32
    L.gnss_b200_acq_group_num_shards.argtypes = [C.c_void_p]
    L.gnss_b200_acq_group_num_shards.restype = c_u32
    L.gnss_b200_acq_group_set_local_codes.argtypes = [C.c_void_p, c_u32, _u32p, _fp]
    L.gnss_b200_acq_group_set_doppler_center.argtypes = [C.c_void_p, c_u32, c_i32, c_i32]
    L.gnss_b200_acq_group_reset.argtypes = [C.c_void_p, c_u32]
    L.gnss_b200_acq_group_dwell.argtypes = [C.c_void_p, C.c_void_p, C.c_int, c_u32, _u32p, C.c_void_p]
    L.gnss_b200_acq_group_read_grid.argtypes = [C.c_void_p, c_u32, _fp]
    L.gnss_b200_acq_group_launch_count.argtypes = [C.c_void_p]
    L.gnss_b200_acq_group_launch_count.restype = C.c_uint64
    L.gnss_b200_acq_read_grid.argtypes = [C.c_void_p, c_u32, _fp]
    L.gnss_b200_acq_read_rows.argtypes = [C.c_void_p, c_u32, C.c_void_p]
    L.gnss_b200_acq_finalize_rows.argtypes = [C.c_void_p, c_u32, c_u32, c_i32, c_u32, c_i32, c_u32, C.c_void_p]
    L.gnss_b200_acq_threshold.argtypes = [C.c_float, c_u32, c_u32, c_u32, C.c_int]
    L.gnss_b200_acq_threshold.restype = C.c_float
    L.gnss_b200_acq_stream.argtypes = [C.c_void_p]
    L.gnss_b200_acq_stream.restype = C.c_void_p
    L.gnss_b200_acq_launch_count.argtypes = [C.c_void_p]
    L.gnss_b200_acq_launch_count.restype = C.c_uint64
    L.gnss_b200_acq_transform_length.argtypes = [C.c_void_p]
    L.gnss_b200_acq_transform_length.restype = C.c_uint32
    L.gnss_b200_host_register.argtypes = [C.c_void_p, C.c_size_t]
    L.gnss_b200_host_unregister.argtypes = [C.c_void_p]
    L.gnss_b200_mc_create.argtypes = [c_u32, c_u32, c_u32, c_i32, C.POINTER(C.c_void_p)]
    L.gnss_b200_mc_destroy.argtypes = [C.c_void_p]
    L.gnss_b200_mc_set_code.argtypes = [C.c_void_p, c_u32, c_u32, _fp, C.c_int]
    L.gnss_b200_mc_set_stream.argtypes = [C.c_void_p, _fp, C.c_uint64]
    L.gnss_b200_mc_correlate_batch.argtypes = [C.c_void_p, c_u32, _u32p, C.POINTER(McArgs), _fp, c_u32, c_u32, _fp]
    L.gnss_b200_mc_correlate_device_async.argtypes = [C.c_void_p, c_u32, _u32p, C.POINTER(McArgs), _fp, c_u32, c_u32,
                                                      C.c_void_p, C.c_uint64]
    L.gnss_b200_mc_sync.argtypes = [C.c_void_p, c_u32, c_u32, _fp]
    L.gnss_b200_mc_chip_indices.argtypes = [C.c_void_p, c_u32, C.POINTER(McArgs), _fp, c_u32, c_u32, C.POINTER(c_i32)]
    L.gnss_b200_mc_stream.argtypes = [C.c_void_p]
    L.gnss_b200_mc_stream.restype = C.c_void_p
    L.gnss_b200_mc_launch_count.argtypes = [C.c_void_p]
    L.gnss_b200_mc_launch_count.restype = C.c_uint64
    L.gnss_b200_debug_wipeoff_phase.argtypes = [c_i32, C.c_double, c_u32, _fp, _u32p]
    L.gnss_b200_debug_fft_plan.argtypes = [c_u32, _u32p, _u32p, _u32p, _u32p]
    L.gnss_b200_debug_transform_length.argtypes = [c_u32, c_u32, _u32p]
    L.gnss_b200_acq_profile.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_int]
    L.gnss_b200_mc_profile.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_int]
    L.gnss_b200_debug_fp32_peak.argtypes = [c_i32, C.POINTER(C.c_double)]
    L.gnss_b200_trk_create.argtypes = [C.POINTER(TrkCfg), C.POINTER(C.c_void_p)]
    L.gnss_b200_trk_destroy.argtypes = [C.c_void_p]
    L.gnss_b200_trk_set_code.argtypes = [C.c_void_p, c_u32, c_u32, _fp]
    L.gnss_b200_trk_set_data_code.argtypes = [C.c_void_p, c_u32, c_u32, _fp]
    L.gnss_b200_trk_start.argtypes = [C.c_void_p, c_u32, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
    L.gnss_b200_trk_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, c_u32, _u32p, c_u32, C.c_void_p]
    L.gnss_b200_trk_stream.argtypes = [C.c_void_p]
    L.gnss_b200_trk_stream.restype = C.c_void_p
    L.gnss_b200_trk_launch_count.argtypes = [C.c_void_p]
    L.gnss_b200_trk_launch_count.restype = C.c_uint64
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise B200Error(f"gnss_b200 status {rc}: {lib().gnss_b200_last_error().decode(errors='replace')}")


def _f32p(a):
    return a.ctypes.data_as(_fp)


def _u32a(x):
    return np.ascontiguousarray(x, dtype=np.uint32)


# --------------------------------------------------------------------------------------------------
# Acquisition geometry: what Acq_Conf::SetFromConfiguration / SetDerivedParams (acq_conf.cc:24-119) and the
# pcps_acquisition constructor (pcps_acquisition.cc:68-92) derive from the configuration keys.
# --------------------------------------------------------------------------------------------------
def acq_geometry(fs, chip_rate, ms_per_code, sampled_ms, bit_transition=False):
    samples_per_ms = np.float32(np.float32(fs) * np.float32(0.001))
    samples_per_chip = int(math.ceil(float(np.float32(fs) / np.float32(chip_rate))))
    samples_per_code = float(np.float32(samples_per_ms * np.float32(ms_per_code)))
    consumed = int(sampled_ms * float(samples_per_ms) * (2.0 if bit_transition else 1.0))
    fft_size = consumed if sampled_ms == ms_per_code else consumed * 2
    return dict(fft_size=fft_size, consumed_samples=consumed, samples_per_chip=samples_per_chip,
                samples_per_code=samples_per_code, effective_size=fft_size // 2 if bit_transition else fft_size)


class PcpsAcquisitionB200:
    """A batch of reference `pcps_acquisition` block instances ("slots") sharing one sample stream.

    Method names follow the reference block / AcquisitionInterface (set_local_code, set_doppler_center,
    reset, ...); `dwell` is one `acquisition_core` pass for a list of slots.
    """

    def __init__(self, fft_size, consumed_samples, doppler_max, doppler_step, fs, samples_per_chip, max_slots,
                 use_cfar=True, bit_transition=False, keep_grid=False, device=0, row_begin=0, row_end=0, step2=None):
        """step2 = (num_doppler_bins_step2, doppler_step2): the narrow second-search handle of make_two_steps."""
        self.L = lib()
        self.cfg = AcqCfg(fft_size, consumed_samples, int(bit_transition), int(doppler_max), int(doppler_step), float(fs),
                          int(samples_per_chip), int(use_cfar), int(keep_grid), int(max_slots), int(device),
                          int(row_begin), int(row_end))
        self.h = C.c_void_p()
        if step2 is None:
            _check(self.L.gnss_b200_acq_create(C.byref(self.cfg), C.byref(self.h)))
        else:
            _check(self.L.gnss_b200_acq_create_step2(C.byref(self.cfg), int(step2[0]), float(step2[1]), C.byref(self.h)))
        self.fft_size = fft_size
        self.consumed = consumed_samples
        self.eff = fft_size // 2 if bit_transition else fft_size
        self.n_doppler = self.L.gnss_b200_acq_num_doppler_bins(self.h)
        self.n_doppler_local = (row_end - row_begin) if row_end > row_begin else self.n_doppler
        self.max_slots = max_slots

    def close(self):
        if getattr(self, "h", None):
            self.L.gnss_b200_acq_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_local_codes(self, slots, codes):
        slots = _u32a(slots)
        codes = np.ascontiguousarray(codes, dtype=np.complex64).reshape(slots.size, -1)
        if codes.shape[1] != self.consumed:
            raise B200Error(f"set_local_codes: each code needs consumed_samples={self.consumed} items, got {codes.shape[1]}")
        _check(self.L.gnss_b200_acq_set_local_codes(self.h, slots.size, slots.ctypes.data_as(_u32p), _f32p(codes.view(np.float32))))

    def set_local_code(self, slot, code):
        self.set_local_codes([slot], np.asarray(code)[None, :])

    def set_doppler_center(self, slot, center_hz, bias_hz=0):
        _check(self.L.gnss_b200_acq_set_doppler_center(self.h, slot, int(center_hz), int(bias_hz)))

    def set_step2_center(self, slot, doppler_center_step_two_hz, input_power_step_one):
        _check(self.L.gnss_b200_acq_set_step2_center(self.h, slot, float(doppler_center_step_two_hz), float(input_power_step_one)))

    def reset(self, slot):
        _check(self.L.gnss_b200_acq_reset(self.h, slot))

    def dwell(self, samples, slots):
        slots = _u32a(slots)
        samples = np.ascontiguousarray(samples)
        if samples.dtype == np.int8:
            is_cshort = 2  # GNSS_B200_SAMPLES_CBYTE
            if samples.size < 2 * self.consumed:
                raise B200Error("dwell: not enough cbyte samples")
        elif samples.dtype == np.int16:
            is_cshort = 1
            if samples.size < 2 * self.consumed:
                raise B200Error("dwell: not enough cshort samples")
        else:
            is_cshort = 0
            samples = samples.astype(np.complex64, copy=False)
            if samples.size < self.consumed:
                raise B200Error("dwell: not enough samples")
        res = np.zeros(slots.size, dtype=ACQ_RESULT_DTYPE)
        _check(self.L.gnss_b200_acq_dwell(self.h, samples.ctypes.data_as(C.c_void_p), is_cshort, slots.size,
                                          slots.ctypes.data_as(_u32p), res.ctypes.data_as(C.c_void_p)))
        return res

    def dwell_device_async(self, d_samples_ptr, slots):
        slots = _u32a(slots)
        _check(self.L.gnss_b200_acq_dwell_device_async(self.h, C.c_void_p(d_samples_ptr), slots.size, slots.ctypes.data_as(_u32p)))

    def sync(self, n_slots):
        res = np.zeros(n_slots, dtype=ACQ_RESULT_DTYPE)
        _check(self.L.gnss_b200_acq_sync(self.h, n_slots, res.ctypes.data_as(C.c_void_p)))
        return res

    def exchange_buffers(self, world):
        """(d_send pointer, d_recv pointer, bytes per shard, rows per shard) of a sharded search (include/gnss_b200.h)."""
        ds, dr, nb, ch = C.c_void_p(), C.c_void_p(), C.c_uint64(), c_u32()
        _check(self.L.gnss_b200_acq_exchange_buffers(self.h, int(world), C.byref(ds), C.byref(dr), C.byref(nb), C.byref(ch)))
        return ds.value, dr.value, int(nb.value), int(ch.value)

    def finalize_gathered_async(self, n_slots):
        _check(self.L.gnss_b200_acq_finalize_gathered_async(self.h, int(n_slots)))

    def peer_setup(self, rank):
        """(64-byte CUDA IPC handle, device pointer) of this shard's peer-exchange region (include/gnss_b200.h)."""
        hd = (C.c_ubyte * 64)()
        ptr = C.c_void_p()
        _check(self.L.gnss_b200_acq_peer_setup(self.h, int(rank), hd, C.byref(ptr)))
        return bytes(hd), ptr.value

    def peer_connect(self, ipc_handles=None, regions=None):
        if regions is not None:
            arr = (C.c_void_p * len(regions))(*regions)
            _check(self.L.gnss_b200_acq_peer_connect(self.h, None, arr))
        else:
            buf = b"".join(ipc_handles)
            _check(self.L.gnss_b200_acq_peer_connect(self.h, buf, None))

    def peer_exchange_async(self, n_slots):
        _check(self.L.gnss_b200_acq_peer_exchange_async(self.h, int(n_slots)))

    def peer_barrier_async(self):
        """device-side rendezvous of all connected shards on the engine's stream (gnss_b200_acq_peer_barrier_async)"""
        _check(self.L.gnss_b200_acq_peer_barrier_async(self.h))

    def capture_begin(self):
        _check(self.L.gnss_b200_acq_capture_begin(self.h))

    def capture_end(self):
        _check(self.L.gnss_b200_acq_capture_end(self.h))

    def replay_async(self):
        _check(self.L.gnss_b200_acq_replay_async(self.h))

    def read_grid(self, slot):
        out = np.zeros((self.n_doppler_local, self.eff), dtype=np.float32)
        _check(self.L.gnss_b200_acq_read_grid(self.h, slot, _f32p(out)))
        return out

    def read_rows(self, slot):
        rows = np.zeros(self.n_doppler, dtype=ACQ_ROW_DTYPE)
        _check(self.L.gnss_b200_acq_read_rows(self.h, slot, rows.ctypes.data_as(C.c_void_p)))
        return rows

    def profile(self, enable, reset=True):
        """(ms_by_kernel_class, launches_by_kernel_class) accumulated since the last reset; then enable/disable."""
        ms = (C.c_double * 8)()
        cnt = (C.c_uint64 * 8)()
        _check(self.L.gnss_b200_acq_profile(self.h, int(enable), ms, cnt, int(reset)))
        return list(ms), list(cnt)

    def stream(self):
        return self.L.gnss_b200_acq_stream(self.h)

    def launch_count(self):
        return int(self.L.gnss_b200_acq_launch_count(self.h))

    @property
    def transform_length(self):
        """fft_size, or the padded length the engine searches on (gnss_b200_acq_transform_length)."""
        return int(self.L.gnss_b200_acq_transform_length(self.h))


class PcpsAcquisitionGroupB200:
    """One search over several GPUs of this process (gnss_b200_acq_group_*): Doppler rows split over `devices`."""

    def __init__(self, fft_size, consumed_samples, doppler_max, doppler_step, fs, samples_per_chip, max_slots, devices, bit_transition=False, keep_grid=False):
        self.L = lib()
        self.cfg = AcqCfg(fft_size, consumed_samples, int(bit_transition), int(doppler_max), int(doppler_step), float(fs),
                          int(samples_per_chip), 1, int(keep_grid), int(max_slots), 0, 0, 0)
        dev = (c_i32 * len(devices))(*[int(d) for d in devices])
        self.h = C.c_void_p()
        _check(self.L.gnss_b200_acq_group_create(C.byref(self.cfg), len(devices), dev, C.byref(self.h)))
        self.consumed = consumed_samples
        self.eff = fft_size // 2 if bit_transition else fft_size
        self.n_doppler = self.L.gnss_b200_acq_group_num_doppler_bins(self.h)
        self.n_shards = self.L.gnss_b200_acq_group_num_shards(self.h)

    def close(self):
        if getattr(self, "h", None):
            self.L.gnss_b200_acq_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_local_codes(self, slots, codes):
        slots = _u32a(slots)
        codes = np.ascontiguousarray(codes, dtype=np.complex64).reshape(slots.size, -1)
        if codes.shape[1] != self.consumed:
            raise B200Error(f"set_local_codes: each code needs consumed_samples={self.consumed} items, got {codes.shape[1]}")
        _check(self.L.gnss_b200_acq_group_set_local_codes(self.h, slots.size, slots.ctypes.data_as(_u32p), _f32p(codes.view(np.float32))))

    def set_doppler_center(self, slot, center_hz, bias_hz=0):
        _check(self.L.gnss_b200_acq_group_set_doppler_center(self.h, slot, int(center_hz), int(bias_hz)))

    def reset(self, slot):
        _check(self.L.gnss_b200_acq_group_reset(self.h, slot))

    def dwell(self, samples, slots):
        slots = _u32a(slots)
        samples = np.ascontiguousarray(samples, dtype=np.complex64)
        if samples.size < self.consumed:
            raise B200Error("dwell: not enough samples")
        res = np.zeros(slots.size, dtype=ACQ_RESULT_DTYPE)
        _check(self.L.gnss_b200_acq_group_dwell(self.h, samples.ctypes.data_as(C.c_void_p), 0, slots.size, slots.ctypes.data_as(_u32p),
                                                res.ctypes.data_as(C.c_void_p)))
        return res

    def read_grid(self, slot):
        out = np.zeros((self.n_doppler, self.eff), dtype=np.float32)
        _check(self.L.gnss_b200_acq_group_read_grid(self.h, slot, _f32p(out)))
        return out

    def launch_count(self):
        return int(self.L.gnss_b200_acq_group_launch_count(self.h))


def finalize_rows(rows, effective_size, doppler_max, doppler_step, doppler_center, dwell_count):
    rows = np.ascontiguousarray(rows, dtype=ACQ_ROW_DTYPE)
    res = np.zeros(1, dtype=ACQ_RESULT_DTYPE)
    _check(lib().gnss_b200_acq_finalize_rows(rows.ctypes.data_as(C.c_void_p), rows.size, effective_size, doppler_max,
                                             doppler_step, doppler_center, dwell_count, res.ctypes.data_as(C.c_void_p)))
    return res[0]


def acq_threshold(pfa, effective_size, n_doppler, max_dwells=1, bit_transition=False):
    return float(lib().gnss_b200_acq_threshold(pfa, effective_size, n_doppler, max_dwells, int(bit_transition)))


def debug_transform_length(fft_size, consumed_samples=None):
    """transform length the engine searches `fft_size` lags on (no device needed)"""
    m = c_u32(0)
    _check(lib().gnss_b200_debug_transform_length(int(fft_size), int(consumed_samples or fft_size), C.byref(m)))
    return int(m.value)


def debug_wipeoff_phase(doppler_hz, fs, n):
    out = np.zeros(n, dtype=np.float32)
    npc = c_u32(0)
    _check(lib().gnss_b200_debug_wipeoff_phase(int(doppler_hz), float(fs), n, _f32p(out), C.byref(npc)))
    return out, npc.value


def fp32_peak_tflops(device=0):
    v = C.c_double(0)
    _check(lib().gnss_b200_debug_fp32_peak(device, C.byref(v)))
    return v.value


def debug_fft_plan(n):
    ns, nr = c_u32(0), c_u32(0)
    rad = np.zeros(32, dtype=np.uint32)
    perm = np.zeros(n, dtype=np.uint32)
    _check(lib().gnss_b200_debug_fft_plan(n, C.byref(ns), rad.ctypes.data_as(_u32p), C.byref(nr), perm.ctypes.data_as(_u32p)))
    return ns.value, rad[:nr.value].tolist(), perm


class MulticorrelatorB200:
    """`max_channels` reference correlator objects (Cpu_Multicorrelator_Real_Codes / Cpu_Multicorrelator) in one engine."""

    def __init__(self, max_channels, max_samples, max_taps, device=0):
        self.L = lib()
        self.h = C.c_void_p()
        _check(self.L.gnss_b200_mc_create(max_channels, max_samples, max_taps, device, C.byref(self.h)))
        self.max_channels, self.max_samples, self.max_taps = max_channels, max_samples, max_taps
        self._keep = []

    def close(self):
        if getattr(self, "h", None):
            self.L.gnss_b200_mc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_local_code(self, ch, code):
        code = np.ascontiguousarray(code)
        if np.iscomplexobj(code):
            code = code.astype(np.complex64, copy=False)
            _check(self.L.gnss_b200_mc_set_code(self.h, ch, code.size, _f32p(code.view(np.float32)), 1))
        else:
            code = code.astype(np.float32, copy=False)
            _check(self.L.gnss_b200_mc_set_code(self.h, ch, code.size, _f32p(code), 0))

    def set_stream(self, samples):
        samples = np.ascontiguousarray(samples, dtype=np.complex64)
        self._keep = [samples]
        _check(self.L.gnss_b200_mc_set_stream(self.h, _f32p(samples.view(np.float32)), samples.size))

    @staticmethod
    def make_args(n_ch, n_samples, rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate, sig_ptrs=None, stream_offsets=None):
        arr = (McArgs * n_ch)()
        bc = lambda v: np.broadcast_to(np.asarray(v), (n_ch,))  # noqa: E731
        ns, rc, cs, cr, rk, ks, kr = map(bc, (n_samples, rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate))
        for i in range(n_ch):
            arr[i].sig_in_c64 = None if sig_ptrs is None else int(sig_ptrs[i])
            arr[i].stream_offset = 0 if stream_offsets is None else int(stream_offsets[i])
            arr[i].n_samples = int(ns[i])
            arr[i].rem_carr_phase_rad = float(rc[i])
            arr[i].carr_step_rad = float(cs[i])
            arr[i].carr_rate_rad = float(cr[i])
            arr[i].rem_code_phase_chips = float(rk[i])
            arr[i].code_step_chips = float(ks[i])
            arr[i].code_rate_chips = float(kr[i])
        return arr

    def correlate_batch(self, channels, args, shifts, flags=0):
        channels = _u32a(channels)
        shifts = np.ascontiguousarray(shifts, dtype=np.float32).reshape(channels.size, -1)
        n_taps = shifts.shape[1]
        out = np.zeros((channels.size, n_taps), dtype=np.complex64)
        _check(self.L.gnss_b200_mc_correlate_batch(self.h, channels.size, channels.ctypes.data_as(_u32p), args, _f32p(shifts),
                                                   n_taps, flags, _f32p(out.view(np.float32))))
        return out

    def correlate_device_async(self, channels, args, shifts, d_sig_ptr, ch_stride_samples, flags=0):
        channels = _u32a(channels)
        shifts = np.ascontiguousarray(shifts, dtype=np.float32).reshape(channels.size, -1)
        _check(self.L.gnss_b200_mc_correlate_device_async(self.h, channels.size, channels.ctypes.data_as(_u32p), args, _f32p(shifts),
                                                          shifts.shape[1], flags, C.c_void_p(d_sig_ptr), ch_stride_samples))

    def sync(self, n_ch, n_taps):
        out = np.zeros((n_ch, n_taps), dtype=np.complex64)
        _check(self.L.gnss_b200_mc_sync(self.h, n_ch, n_taps, _f32p(out.view(np.float32))))
        return out

    def chip_indices(self, ch, args_one, shifts, flags=0):
        shifts = np.ascontiguousarray(shifts, dtype=np.float32)
        n = args_one[0].n_samples
        out = np.zeros((shifts.size, n), dtype=np.int32)
        _check(self.L.gnss_b200_mc_chip_indices(self.h, ch, args_one, _f32p(shifts), shifts.size, flags, out.ctypes.data_as(C.POINTER(c_i32))))
        return out

    def profile(self, enable, reset=True):
        ms = (C.c_double * 8)()
        cnt = (C.c_uint64 * 8)()
        _check(self.L.gnss_b200_mc_profile(self.h, int(enable), ms, cnt, int(reset)))
        return list(ms), list(cnt)

    def stream(self):
        return self.L.gnss_b200_mc_stream(self.h)

    def launch_count(self):
        return int(self.L.gnss_b200_mc_launch_count(self.h))


class TrackingLoopB200:
    """Many `dll_pll_veml_tracking` channels whose DLL/PLL loops are closed on the device.

    start() is start_tracking() + the state-1 alignment; run() processes every code period the sample block holds
    (state 2: correlate -> lock detectors -> discriminators -> loop filters -> NCO update) with no host round trip
    between epochs and returns one record per epoch and channel."""

    def __init__(self, cfg):
        self.L = lib()
        self.cfg = cfg
        self.h = C.c_void_p()
        _check(self.L.gnss_b200_trk_create(C.byref(cfg), C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None):
            self.L.gnss_b200_trk_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_code(self, ch, code):
        code = np.ascontiguousarray(code, dtype=np.float32)
        _check(self.L.gnss_b200_trk_set_code(self.h, ch, code.size, _f32p(code)))

    def set_data_code(self, ch, code):
        code = np.ascontiguousarray(code, dtype=np.float32)
        _check(self.L.gnss_b200_trk_set_data_code(self.h, ch, code.size, _f32p(code)))

    def start(self, ch, acq_delay_samples, acq_doppler_hz, acq_samplestamp, now_sample):
        _check(self.L.gnss_b200_trk_start(self.h, ch, float(acq_delay_samples), float(acq_doppler_hz), int(acq_samplestamp), int(now_sample)))

    def run(self, samples, stream_start, channels, max_epochs):
        samples = np.ascontiguousarray(samples, dtype=np.complex64)
        channels = _u32a(channels)
        out = np.zeros((max_epochs, channels.size), dtype=TRK_EPOCH_DTYPE)
        _check(self.L.gnss_b200_trk_run(self.h, samples.ctypes.data_as(C.c_void_p), 0, int(stream_start), samples.size, channels.size,
                                        channels.ctypes.data_as(_u32p), max_epochs, out.ctypes.data_as(C.c_void_p)))
        return out

    def run_device(self, d_samples_ptr, n_samples, stream_start, channels, max_epochs):
        channels = _u32a(channels)
        out = np.zeros((max_epochs, channels.size), dtype=TRK_EPOCH_DTYPE)
        _check(self.L.gnss_b200_trk_run(self.h, C.c_void_p(d_samples_ptr), 1, int(stream_start), int(n_samples), channels.size,
                                        channels.ctypes.data_as(_u32p), max_epochs, out.ctypes.data_as(C.c_void_p)))
        return out

    def stream(self):
        return self.L.gnss_b200_trk_stream(self.h)

    def launch_count(self):
        return int(self.L.gnss_b200_trk_launch_count(self.h))


def firdes_low_pass(gain, fs, cutoff, transition):
    """gr::filter::firdes::low_pass (Hamming) as the flow graph calls it for the acquisition resampler."""
    taps = np.zeros(4096, dtype=np.float32)
    n = c_u32(0)
    _check(lib().gnss_b200_firdes_low_pass(float(gain), float(fs), float(cutoff), float(transition), taps.size,
                                           taps.ctypes.data_as(C.POINTER(C.c_float)), C.byref(n)))
    return taps[:n.value].copy()


class AcqResamplerB200:
    """The decimating FIR in front of the acquisition blocks (gnss_flowgraph.cc:1063-1118), output resident on the device."""

    def __init__(self, decimation, taps, max_input_items, device=0):
        self.L = lib()
        self.decimation = int(decimation)
        taps = np.ascontiguousarray(taps, dtype=np.float32)
        self.latency = (taps.size - 1) // 2          # Channel::acquisition()->set_resampler_latency (:1111)
        self.h = C.c_void_p()
        _check(self.L.gnss_b200_resampler_create(self.decimation, taps.size, taps.ctypes.data_as(C.POINTER(C.c_float)), int(max_input_items),
                                                 int(device), C.byref(self.h)))

    def reset(self):
        _check(self.L.gnss_b200_resampler_reset(self.h))

    def work(self, x, want_host=True):
        """-> (decimated items on the host or None, device pointer of the same items)"""
        x = np.ascontiguousarray(x, dtype=np.complex64)
        out = np.empty(x.size // self.decimation, dtype=np.complex64) if want_host else None
        dptr = C.c_void_p()
        _check(self.L.gnss_b200_resampler_work(self.h, x.ctypes.data, x.size, out.ctypes.data if want_host else None, C.byref(dptr)))
        return out, dptr.value

    def close(self):
        if self.h:
            self.L.gnss_b200_resampler_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def mc_correlate_noalloc(device, cuda_stream, chans, shifts, d_out_ptr):
    """gnss_b200_mc_correlate_noalloc: `chans` = list of dict(d_sig, d_code, code_len, n_samples, rem_carr, carr_step, rem_code,
    code_step) with DEVICE pointers, `shifts` host [n_ch][n_taps], d_out_ptr DEVICE [n_ch][n_taps] c64. Enqueues only."""
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    arr = (McNoAllocChan * len(chans))()
    for i, c in enumerate(chans):
        arr[i] = McNoAllocChan(c["d_sig"], c["d_code"], int(c["code_len"]), int(c["n_samples"]), float(c["rem_carr"]), float(c["carr_step"]),
                               float(c["rem_code"]), float(c["code_step"]))
    _check(lib().gnss_b200_mc_correlate_noalloc(int(device), cuda_stream, len(chans), arr, _f32p(shifts), shifts.shape[1] if shifts.ndim == 2 else shifts.size,
                                                d_out_ptr))
