"""TEST INFRASTRUCTURE — ctypes front-end to the CPU oracle (oracle/liboracle.so, oracle/_ref/*.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module. The product (gnss-sdr-new_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("GNSS_REF_ROOT", "/root/reference")

c_f = C.c_float
c_i = C.c_int
c_u = C.c_uint32
_fp = C.POINTER(C.c_float)


class AcqCfg(C.Structure):
    _fields_ = [
        ("fft_size", c_u), ("consumed_samples", c_u), ("bit_transition", C.c_int32),
        ("doppler_max", C.c_int32), ("doppler_step", C.c_int32), ("doppler_center", C.c_int32),
        ("doppler_bias", C.c_int32), ("fs", C.c_double), ("samples_per_chip", c_u),
        ("samples_per_code", c_f), ("use_cfar", C.c_int32),
    ]


class AcqResult(C.Structure):
    _fields_ = [
        ("test_statistic", c_f), ("peak", c_f), ("input_power", c_f), ("second_peak", c_f),
        ("code_phase", c_u), ("doppler_index", c_u), ("doppler_hz", C.c_int32), ("dwell_count", c_u),
        ("acq_delay_samples", C.c_double),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class TrkCfg(C.Structure):
    """oracle_trk_cfg == gnss_b200_trk_cfg (include/gnss_b200.h)."""
    _fields_ = [
        ("fs_in", C.c_double), ("code_chip_rate", C.c_double), ("signal_carrier_freq", C.c_double), ("code_period_s", C.c_double),
        ("code_length_chips", c_u), ("code_samples_per_chip", c_u), ("vector_length", c_u), ("veml", C.c_int32),
        ("early_late_space_chips", c_f), ("very_early_late_space_chips", c_f), ("slope", c_f), ("y_intercept", c_f),
        ("pll_bw_hz", c_f), ("dll_bw_hz", c_f), ("fll_bw_hz", c_f),
        ("pll_filter_order", C.c_int32), ("dll_filter_order", C.c_int32),
        ("enable_fll_pull_in", C.c_int32), ("enable_fll_steady_state", C.c_int32), ("carrier_aiding", C.c_int32),
        ("pull_in_time_s", c_u),
        ("cn0_samples", C.c_int32), ("cn0_min", C.c_int32), ("max_code_lock_fail", C.c_int32), ("max_carrier_lock_fail", C.c_int32),
        ("carrier_lock_th", C.c_double),
        ("cn0_smoother_alpha", c_f), ("carrier_lock_test_smoother_alpha", c_f),
        ("cn0_smoother_samples", C.c_int32), ("carrier_lock_test_smoother_samples", C.c_int32),
        ("max_channels", c_u), ("device", C.c_int32),
        # symbol synchronisation, extended integration, pilot tracking, high dynamics (all 0: the state-2 loop only)
        ("symbol_sync", C.c_int32), ("secondary", C.c_int32),
        ("secondary_code_length", c_u), ("symbols_per_bit", c_u), ("data_secondary_code_length", c_u),
        ("extend_correlation_symbols", C.c_int32), ("track_pilot", C.c_int32),
        ("pll_bw_narrow_hz", c_f), ("dll_bw_narrow_hz", c_f), ("early_late_space_narrow_chips", c_f), ("very_early_late_space_narrow_chips", c_f),
        ("high_dyn", C.c_int32), ("smoother_length", c_u), ("bit_synchronization_time_limit_s", c_u),
        ("secondary_code_string", C.c_char * 208), ("data_secondary_code_string", C.c_char * 112),
    ]


TRK_EPOCH_DTYPE = np.dtype([("corr", np.float32, (10,)), ("carrier_doppler_hz", np.float64), ("code_freq_chips", np.float64),
                            ("carr_error_hz", np.float64), ("carr_error_filt_hz", np.float64), ("code_error_chips", np.float64),
                            ("code_error_filt_chips", np.float64), ("rem_code_phase_samples", np.float64),
                            ("acc_carrier_phase_rad", np.float64), ("cn0_db_hz", np.float64), ("carrier_lock_test", np.float64),
                            ("evm", np.float64), ("sample_counter", np.uint64), ("prn_length_samples", np.int32), ("flags", np.uint32),
                            ("accu", np.float32, (10,)), ("prompt_data", np.float32, (2,)), ("state", np.int32), ("reserved", np.int32)], align=True)


def build(ref=None):
    """Compile liboracle.so (always) and _ref/ (when the reference tree is present)."""
    subprocess.run(["make", "-s", "-C", _HERE], check=True)
    if ref is None:
        ref = os.path.isdir(REF_ROOT)
    if ref:
        subprocess.run(["make", "-s", "-C", _HERE, "ref", f"REF={REF_ROOT}"], check=True)


def _ptr(a, t=C.c_float):
    return a.ctypes.data_as(C.POINTER(t))


def _bind_oracle(lib):
    lib.oracle_kernel_flavour.restype = C.c_char_p
    lib.oracle_acq_num_doppler_bins.restype = c_u
    lib.oracle_acq_num_doppler_bins.argtypes = [C.c_int32, C.c_int32]
    lib.oracle_acq_create.restype = C.c_void_p
    lib.oracle_acq_create.argtypes = [C.POINTER(AcqCfg)]
    lib.oracle_acq_destroy.argtypes = [C.c_void_p]
    lib.oracle_acq_n_doppler.restype = c_u
    lib.oracle_acq_n_doppler.argtypes = [C.c_void_p]
    lib.oracle_acq_set_doppler_center.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
    lib.oracle_acq_set_local_code.argtypes = [C.c_void_p, _fp]
    lib.oracle_acq_reset.argtypes = [C.c_void_p]
    lib.oracle_acq_dwell.argtypes = [C.c_void_p, C.c_void_p, c_i, C.POINTER(AcqResult)]
    lib.oracle_acq_dwell_step2.argtypes = [C.c_void_p, C.c_void_p, c_i, c_u, c_f, c_f, C.POINTER(AcqResult)]
    for n in ("oracle_acq_grid", "oracle_acq_fft_codes", "oracle_acq_wipeoffs"):
        getattr(lib, n).restype = _fp
        getattr(lib, n).argtypes = [C.c_void_p]
    lib.oracle_gamma_p_inv.restype = C.c_double
    lib.oracle_gamma_p_inv.argtypes = [C.c_double, C.c_double]
    lib.oracle_acq_threshold.restype = c_f
    lib.oracle_acq_threshold.argtypes = [c_f, c_u, c_u, c_u, c_i]
    lib.oracle_mc_real.argtypes = [_fp, c_i, _fp, c_i, _fp, c_i, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, _fp]
    lib.oracle_mc_cplx.argtypes = [_fp, c_i, _fp, c_i, _fp, c_i, c_f, c_f, c_f, c_f, _fp]
    lib.oracle_mc_chip_indices.argtypes = [c_i, c_i, _fp, c_i, c_f, c_f, c_f, c_i, C.POINTER(C.c_int32)]
    lib.oracle_trk_create.restype = C.c_void_p
    lib.oracle_trk_create.argtypes = [C.POINTER(TrkCfg)]
    lib.oracle_trk_destroy.argtypes = [C.c_void_p]
    lib.oracle_trk_set_code.argtypes = [C.c_void_p, _fp, c_i]
    lib.oracle_trk_set_data_code.argtypes = [C.c_void_p, _fp, c_i]
    lib.oracle_trk_start.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
    lib.oracle_trk_run.restype = c_i
    lib.oracle_trk_run.argtypes = [C.c_void_p, _fp, C.c_uint64, C.c_uint64, c_i, C.c_void_p]
    lib.oracle_trk_run_traced.restype = c_i
    lib.oracle_trk_run_traced.argtypes = [C.c_void_p, _fp, C.c_uint64, C.c_uint64, c_i, C.c_void_p, C.POINTER(C.c_double)]
    lib.oracle_fft_c64.argtypes = [_fp, _fp, c_i, c_i]
    lib.oracle_fft_c128.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), c_i, c_i]
    lib.oracle_samples_per_code.restype = c_i
    lib.oracle_samples_per_code.argtypes = [c_i, c_i, c_i]
    i8p = C.POINTER(C.c_int8)
    lib.oracle_sample_code_lfsr_style.argtypes = [i8p, c_i, c_i, c_i, c_i, c_i, _fp]
    lib.oracle_sample_code_l5.argtypes = [i8p, c_i, c_i, c_i, _fp]
    lib.oracle_sample_code_e1_sinboc11.argtypes = [i8p, c_i, _fp]
    return lib


class Oracle:
    """One flavour of the oracle: 'own' (liboracle.so) or 'ref' (_ref/libref_oracle.so)."""

    def __init__(self, flavour="own"):
        path = os.path.join(_HERE, "liboracle.so" if flavour == "own" else "_ref/libref_oracle.so")
        if not os.path.exists(path):
            build(ref=(flavour == "ref"))
        self.lib = _bind_oracle(C.CDLL(path))
        self.flavour = self.lib.oracle_kernel_flavour().decode()

    # ---- FFT ----
    def fft(self, x, sign=-1):
        x = np.ascontiguousarray(x)
        if x.dtype == np.complex64:
            out = np.empty_like(x)
            self.lib.oracle_fft_c64(_ptr(x.view(np.float32)), _ptr(out.view(np.float32)), x.size, sign)
        else:
            x = x.astype(np.complex128)
            out = np.empty_like(x)
            self.lib.oracle_fft_c128(_ptr(x.view(np.float64), C.c_double), _ptr(out.view(np.float64), C.c_double), x.size, sign)
        return out

    # ---- acquisition ----
    def acq(self, **kw):
        return OracleAcq(self, **kw)

    def threshold(self, pfa, effective_fft_size, n_doppler, max_dwells=1, bit_transition=False):
        return float(self.lib.oracle_acq_threshold(pfa, effective_fft_size, n_doppler, max_dwells, int(bit_transition)))

    def gamma_p_inv(self, a, p):
        return float(self.lib.oracle_gamma_p_inv(a, p))

    # ---- tracking loop (dll_pll_veml_tracking, state 2) ----
    def trk(self, cfg):
        return OracleTrk(self, cfg)

    # ---- multicorrelator ----
    def mc_real(self, sig, code, shifts, rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate,
                high_dyn=False, six_arg=False):
        sig = np.ascontiguousarray(sig, dtype=np.complex64)
        code = np.ascontiguousarray(code, dtype=np.float32)
        shifts = np.ascontiguousarray(shifts, dtype=np.float32)
        out = np.zeros(shifts.size, dtype=np.complex64)
        self.lib.oracle_mc_real(_ptr(sig.view(np.float32)), sig.size, _ptr(code), code.size, _ptr(shifts), shifts.size,
                                rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate,
                                int(high_dyn), int(six_arg), _ptr(out.view(np.float32)))
        return out

    def mc_cplx(self, sig, code, shifts, rem_carr, carr_step, rem_code, code_step):
        sig = np.ascontiguousarray(sig, dtype=np.complex64)
        code = np.ascontiguousarray(code, dtype=np.complex64)
        shifts = np.ascontiguousarray(shifts, dtype=np.float32)
        out = np.zeros(shifts.size, dtype=np.complex64)
        self.lib.oracle_mc_cplx(_ptr(sig.view(np.float32)), sig.size, _ptr(code.view(np.float32)), code.size, _ptr(shifts),
                                shifts.size, rem_carr, carr_step, rem_code, code_step, _ptr(out.view(np.float32)))
        return out

    def mc_chip_indices(self, n, code_len, shifts, rem_code, code_step, code_rate=0.0, high_dyn=False):
        shifts = np.ascontiguousarray(shifts, dtype=np.float32)
        out = np.zeros((shifts.size, n), dtype=np.int32)
        self.lib.oracle_mc_chip_indices(n, code_len, _ptr(shifts), shifts.size, rem_code, code_step, code_rate,
                                        int(high_dyn), _ptr(out, C.c_int32))
        return out

    # ---- code sampling (chips in {+1,-1} int8) ----
    def sample_gps_l1_ca(self, chips, fs):
        n = self.lib.oracle_samples_per_code(fs, 1023000, 1023)
        out = np.zeros(n, dtype=np.complex64)
        chips = np.ascontiguousarray(chips, dtype=np.int8)
        self.lib.oracle_sample_code_lfsr_style(_ptr(chips, C.c_int8), 1023, 1023000, fs, 1, 0, _ptr(out.view(np.float32)))
        return out

    def sample_beidou_b1i(self, chips, fs):
        n = self.lib.oracle_samples_per_code(fs, 2046000, 2046)
        out = np.zeros(n, dtype=np.complex64)
        chips = np.ascontiguousarray(chips, dtype=np.int8)
        self.lib.oracle_sample_code_lfsr_style(_ptr(chips, C.c_int8), 2046, 2046000, fs, 0, 1, _ptr(out.view(np.float32)))
        return out

    def sample_gps_l5i(self, chips, fs):
        n = self.lib.oracle_samples_per_code(fs, 10230000, 10230)
        out = np.zeros(n, dtype=np.complex64)
        chips = np.ascontiguousarray(chips, dtype=np.int8)
        self.lib.oracle_sample_code_l5(_ptr(chips, C.c_int8), 10230, 10230000, fs, _ptr(out.view(np.float32)))
        return out

    def sample_galileo_e1(self, chips, fs):
        n = self.lib.oracle_samples_per_code(fs, 1023000, 4092)
        out = np.zeros(n, dtype=np.complex64)
        chips = np.ascontiguousarray(chips, dtype=np.int8)
        self.lib.oracle_sample_code_e1_sinboc11(_ptr(chips, C.c_int8), fs, _ptr(out.view(np.float32)))
        return out


class OracleAcq:
    """One pcps_acquisition block instance (one PRN / channel) of the oracle."""

    def __init__(self, oracle, fft_size, consumed_samples, doppler_max, doppler_step, fs, samples_per_chip,
                 samples_per_code, use_cfar=True, bit_transition=False, doppler_center=0, doppler_bias=0):
        self.o = oracle
        self.cfg = AcqCfg(fft_size, consumed_samples, int(bit_transition), doppler_max, doppler_step, doppler_center,
                          doppler_bias, float(fs), samples_per_chip, float(samples_per_code), int(use_cfar))
        self.h = oracle.lib.oracle_acq_create(C.byref(self.cfg))
        self.n_doppler = oracle.lib.oracle_acq_n_doppler(self.h)
        self.fft_size = fft_size
        self.consumed = consumed_samples

    def close(self):
        if self.h:
            self.o.lib.oracle_acq_destroy(self.h)
            self.h = None

    __del__ = close

    def set_local_code(self, code):
        code = np.ascontiguousarray(code, dtype=np.complex64)
        need = self.fft_size // 2 if self.cfg.bit_transition else self.consumed
        assert code.size >= need
        self.o.lib.oracle_acq_set_local_code(self.h, _ptr(code.view(np.float32)))

    def set_doppler_center(self, center_hz, bias_hz=0):
        self.o.lib.oracle_acq_set_doppler_center(self.h, center_hz, bias_hz)

    def reset(self):
        self.o.lib.oracle_acq_reset(self.h)

    def dwell(self, samples):
        samples = np.ascontiguousarray(samples)
        r = AcqResult()
        if samples.dtype == np.int16:
            assert samples.size >= 2 * self.consumed
            self.o.lib.oracle_acq_dwell(self.h, samples.ctypes.data_as(C.c_void_p), 1, C.byref(r))
        else:
            samples = samples.astype(np.complex64, copy=False)
            assert samples.size >= self.consumed
            self.o.lib.oracle_acq_dwell(self.h, samples.ctypes.data_as(C.c_void_p), 0, C.byref(r))
        return r.as_dict()

    def dwell_step2(self, samples, num_doppler_bins_step2, doppler_step2, doppler_center_step_two):
        """acquisition_core with d_step_two (make_two_steps) after a first-step dwell() on this object."""
        samples = np.ascontiguousarray(samples)
        r = AcqResult()
        is_cshort = int(samples.dtype == np.int16)
        if not is_cshort:
            samples = samples.astype(np.complex64, copy=False)
        self.o.lib.oracle_acq_dwell_step2(self.h, samples.ctypes.data_as(C.c_void_p), is_cshort, int(num_doppler_bins_step2),
                                          float(doppler_step2), float(doppler_center_step_two), C.byref(r))
        return r.as_dict()

    def grid(self):
        p = self.o.lib.oracle_acq_grid(self.h)
        return np.ctypeslib.as_array(p, shape=(self.n_doppler, self.fft_size)).copy()

    def fft_codes(self):
        p = self.o.lib.oracle_acq_fft_codes(self.h)
        return np.ctypeslib.as_array(p, shape=(self.fft_size * 2,)).copy().view(np.complex64)

    def wipeoffs(self):
        p = self.o.lib.oracle_acq_wipeoffs(self.h)
        return np.ctypeslib.as_array(p, shape=(self.n_doppler, self.fft_size * 2)).copy().view(np.complex64)


class OracleTrk:
    """One channel of the reference block's loop: states 2, 3, 4 (oracle_trk.c)."""

    def __init__(self, oracle, cfg):
        self.lib = oracle.lib
        self.cfg = cfg
        self.h = self.lib.oracle_trk_create(C.byref(cfg))

    def close(self):
        if self.h:
            self.lib.oracle_trk_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def set_code(self, code):
        code = np.ascontiguousarray(code, dtype=np.float32)
        self.lib.oracle_trk_set_code(self.h, _ptr(code), code.size)

    def set_data_code(self, code):
        code = np.ascontiguousarray(code, dtype=np.float32)
        self.lib.oracle_trk_set_data_code(self.h, _ptr(code), code.size)

    def start(self, acq_delay_samples, acq_doppler_hz, acq_samplestamp, now_sample):
        self.lib.oracle_trk_start(self.h, float(acq_delay_samples), float(acq_doppler_hz), int(acq_samplestamp), int(now_sample))

    def run(self, samples, stream_start, max_epochs, trace=False):
        """trace=True: also returns [max_epochs][8] = (rem_carr_phase_rad, carr_step, rem_code, code_step, first sample, ran,
        carr_rate, code_rate) per epoch, the correlator arguments of do_correlation_step."""
        samples = np.ascontiguousarray(samples, dtype=np.complex64)
        out = np.zeros(max_epochs, dtype=TRK_EPOCH_DTYPE)
        if trace:
            tr = np.zeros((max_epochs, 8), dtype=np.float64)
            n = self.lib.oracle_trk_run_traced(self.h, _ptr(samples.view(np.float32)), int(stream_start), samples.size, max_epochs,
                                               out.ctypes.data_as(C.c_void_p), _ptr(tr, C.c_double))
            return out, n, tr
        n = self.lib.oracle_trk_run(self.h, _ptr(samples.view(np.float32)), int(stream_start), samples.size, max_epochs, out.ctypes.data_as(C.c_void_p))
        return out, n


class RefLibs:
    """The reference's own code generators and multicorrelator classes (oracle/_ref). Container-only."""

    def __init__(self):
        d = os.path.join(_HERE, "_ref")
        if not all(os.path.exists(os.path.join(d, f)) for f in ("libref_codes.so", "libref_mc.so", "libref_trk.so")):
            build(ref=True)
        self.codes = C.CDLL(os.path.join(d, "libref_codes.so"))
        self.mc = C.CDLL(os.path.join(d, "libref_mc.so"))
        self.mc.ref_mc_real.argtypes = [_fp, c_i, _fp, c_i, _fp, c_i, c_f, c_f, c_f, c_f, c_f, c_f, c_i, c_i, _fp]
        # the reference's AVX protokernels (timed baseline); only on hosts that have AVX
        self.mc_avx = None
        avx = os.path.join(d, "libref_mc_avx.so")
        try:
            if os.path.exists(avx) and " avx " in open("/proc/cpuinfo").read().replace("\n", " "):
                self.mc_avx = C.CDLL(avx)
                self.mc_avx.ref_mc_real.argtypes = self.mc.ref_mc_real.argtypes
        except OSError:
            self.mc_avx = None
        self.mc.ref_mc_cplx.argtypes = [_fp, c_i, _fp, c_i, _fp, c_i, c_f, c_f, c_f, c_f, _fp]
        t = self.trk = C.CDLL(os.path.join(d, "libref_trk.so"))
        t.ref_trk_loop_filter_run.argtypes = [c_f, c_f, c_i, c_i, c_f, _fp, c_i, _fp]
        t.ref_trk_carrier_filter_run.argtypes = [c_f, c_f, c_i, c_f, _fp, _fp, c_f, c_i, _fp]
        t.ref_trk_smoother_run.argtypes = [c_f, c_f, c_f, c_i, _fp, c_i, _fp]
        t.ref_trk_fll_diff_atan.restype = C.c_double
        t.ref_trk_fll_diff_atan.argtypes = [c_f, c_f, c_f, c_f, C.c_double, C.c_double]
        t.ref_trk_pll_cloop.restype = C.c_double
        t.ref_trk_pll_cloop.argtypes = [c_f, c_f]
        t.ref_trk_dll_e_minus_l.restype = C.c_double
        t.ref_trk_dll_e_minus_l.argtypes = [c_f] * 7
        t.ref_trk_dll_vemlp.restype = C.c_double
        t.ref_trk_dll_vemlp.argtypes = [_fp, _fp, _fp, _fp]
        t.ref_trk_cn0_m2m4.restype = c_f
        t.ref_trk_cn0_m2m4.argtypes = [_fp, c_i, c_f]
        t.ref_trk_carrier_lock.restype = c_f
        t.ref_trk_carrier_lock.argtypes = [_fp, c_i]

    def _i32(self, fn, n, prn):
        a = np.zeros(n, dtype=np.int32)
        fn(_ptr(a, C.c_int32), prn)
        return a.astype(np.int8)

    def gps_l1_ca_chips(self, prn):
        return self._i32(self.codes.ref_gps_l1_ca_chips, 1023, prn)

    def beidou_b1i_chips(self, prn):
        return self._i32(self.codes.ref_beidou_b1i_chips, 2046, prn)

    def gps_l5_chips(self, prn, q=False):
        a = np.zeros(10230, dtype=np.float32)
        (self.codes.ref_gps_l5q_chips if q else self.codes.ref_gps_l5i_chips)(_ptr(a), prn)
        return a.astype(np.int8)

    def galileo_e1_chips(self, prn, pilot=False):
        a = np.zeros(4092, dtype=np.float32)
        self.codes.ref_galileo_e1_chips(_ptr(a), prn, int(pilot))
        return a.astype(np.int8)

    def _sampled(self, fn, prn, fs, n, *extra):
        a = np.zeros(n, dtype=np.complex64)
        fn(_ptr(a.view(np.float32)), prn, fs, n, *extra)
        return a

    def gps_l1_ca_sampled(self, prn, fs):
        return self._sampled(self.codes.ref_gps_l1_ca_sampled, prn, fs, int(fs / (1023000 / 1023)))

    def beidou_b1i_sampled(self, prn, fs):
        return self._sampled(self.codes.ref_beidou_b1i_sampled, prn, fs, int(fs / (2046000 / 2046)))

    def gps_l5i_sampled(self, prn, fs):
        return self._sampled(self.codes.ref_gps_l5i_sampled, prn, fs, int(fs / (10230000 / 10230)))

    def galileo_e1_sampled(self, prn, fs, pilot=False, cboc=False):
        return self._sampled(self.codes.ref_galileo_e1_sampled, prn, fs, int(fs / (1023000 / 4092)), int(pilot), int(cboc))

    def mc_real(self, sig, code, shifts, rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate,
                high_dyn=False, six_arg=False, simd=False):
        sig = np.ascontiguousarray(sig, dtype=np.complex64)
        code = np.ascontiguousarray(code, dtype=np.float32)
        shifts = np.ascontiguousarray(shifts, dtype=np.float32)
        out = np.zeros(shifts.size, dtype=np.complex64)
        (self.mc_avx if simd else self.mc).ref_mc_real(_ptr(sig.view(np.float32)), sig.size, _ptr(code), code.size, _ptr(shifts), shifts.size,
                            rem_carr, carr_step, carr_rate, rem_code, code_step, code_rate,
                            int(high_dyn), int(six_arg), _ptr(out.view(np.float32)))
        return out

    def mc_cplx(self, sig, code, shifts, rem_carr, carr_step, rem_code, code_step):
        sig = np.ascontiguousarray(sig, dtype=np.complex64)
        code = np.ascontiguousarray(code, dtype=np.complex64)
        shifts = np.ascontiguousarray(shifts, dtype=np.float32)
        out = np.zeros(shifts.size, dtype=np.complex64)
        self.mc.ref_mc_cplx(_ptr(sig.view(np.float32)), sig.size, _ptr(code.view(np.float32)), code.size, _ptr(shifts),
                            shifts.size, rem_carr, carr_step, rem_code, code_step, _ptr(out.view(np.float32)))
        return out


def have_reference():
    return os.path.isdir(REF_ROOT)
