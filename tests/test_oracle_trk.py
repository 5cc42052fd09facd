"""The tracking-loop oracle (oracle/oracle_trk.c) — CPU only.

Pins:  * every restated helper (Tracking_loop_filter, Tracking_FLL_PLL_filter, Exponential_Smoother, the discriminators,
         cn0_m2m4_estimator, carrier_lock_detector) bit-for-bit against the reference's own source files compiled
         unmodified into oracle/_ref/libref_trk.so;
       * the closed loop on a synthetic satellite: it pulls in from an acquisition-grade estimate (Doppler off by 200 Hz,
         delay off by 1 sample) and settles on the injected Doppler / code rate, for E-P-L (GPS L1 C/A) and VEML (E1).
The reference's own tracking tests assert nothing numeric about this loop."""
import ctypes as C

import numpy as np
import pytest

import synth


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _own_lib(oracle):
    L = oracle.lib
    L.oracle_loop_filter_apply.restype = C.c_float
    L.oracle_carrier_filter_get_carrier_error.restype = C.c_float
    L.oracle_smoother_smooth.restype = C.c_float
    for n in ("oracle_fll_diff_atan", "oracle_pll_cloop_two_quadrant_atan", "oracle_dll_nc_e_minus_l_normalized", "oracle_dll_nc_vemlp_normalized"):
        getattr(L, n).restype = C.c_double
    L.oracle_cn0_m2m4_estimator.restype = C.c_float
    L.oracle_carrier_lock_detector.restype = C.c_float
    return L


class _LF(C.Structure):
    _fields_ = [("in_coef", C.c_float * 4), ("out_coef", C.c_float * 3), ("n_in", C.c_int), ("n_out", C.c_int), ("idx", C.c_int),
                ("inputs", C.c_float * 4), ("outputs", C.c_float * 4)]


class _CF(C.Structure):
    _fields_ = [("order", C.c_int)] + [(n, C.c_float) for n in ("pll_w", "pll_x", "pll_a2", "pll_a3", "pll_b3", "pll_w0p", "pll_w0p2", "pll_w0p3", "pll_w0f", "pll_w0f2")]


class _SM(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("alpha", "one_minus_alpha", "old_value", "min_value", "offset")] + \
               [(n, C.c_int) for n in ("samples_for_initialization", "init_counter", "initializing")] + [("init_buffer", C.c_float * 256)]


@pytest.mark.parametrize("order,last", [(1, False), (2, False), (3, False), (1, True), (2, True), (3, True)])
def test_loop_filter_bit_exact(oracle, ref_libs, order, last):
    L = _own_lib(oracle)
    rng = np.random.default_rng(order)
    x = (rng.standard_normal(400) * 0.05).astype(np.float32)
    for T, bw in ((0.001, 2.0), (0.004, 0.75), (0.02, 15.0)):
        f = _LF()
        L.oracle_loop_filter_init(C.byref(f), C.c_float(T), C.c_float(bw), order, int(last))
        L.oracle_loop_filter_initialize(C.byref(f), C.c_float(0.3))
        own = np.array([L.oracle_loop_filter_apply(C.byref(f), C.c_float(v)) for v in x], dtype=np.float32)
        ref = np.zeros_like(x)
        ref_libs.trk.ref_trk_loop_filter_run(T, bw, order, int(last), 0.3, _f(x), x.size, _f(ref))
        assert np.array_equal(own.view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("order", [2, 3])
def test_carrier_filter_bit_exact(oracle, ref_libs, order):
    L = _own_lib(oracle)
    rng = np.random.default_rng(10 + order)
    fll = (rng.standard_normal(500) * 3).astype(np.float32)
    pll = (rng.standard_normal(500) * 0.05).astype(np.float32)
    c = _CF()
    L.oracle_carrier_filter_set_params(C.byref(c), C.c_float(35.0), C.c_float(50.0), order)
    L.oracle_carrier_filter_initialize(C.byref(c), C.c_float(-1234.5))
    own = np.array([L.oracle_carrier_filter_get_carrier_error(C.byref(c), C.c_float(a), C.c_float(b), C.c_float(0.001)) for a, b in zip(fll, pll)], dtype=np.float32)
    ref = np.zeros_like(fll)
    ref_libs.trk.ref_trk_carrier_filter_run(35.0, 50.0, order, -1234.5, _f(fll), _f(pll), 0.001, fll.size, _f(ref))
    assert np.array_equal(own.view(np.uint32), ref.view(np.uint32))


def test_smoother_bit_exact(oracle, ref_libs):
    L = _own_lib(oracle)
    rng = np.random.default_rng(3)
    for base, n_init in ((45.0, 200), (30.0, 25), (0.8, 25)):   # the 30 dB-Hz run keeps restarting (min 25 + offset 12)
        x = (base + rng.standard_normal(700)).astype(np.float32)
        mn, off = (25.0, 12.0) if base > 1 else (-1.0, 0.0)
        s = _SM()
        L.oracle_smoother_init(C.byref(s), C.c_float(0.002), C.c_float(mn), C.c_float(off), n_init)
        own = np.array([L.oracle_smoother_smooth(C.byref(s), C.c_float(v)) for v in x], dtype=np.float32)
        ref = np.zeros_like(x)
        ref_libs.trk.ref_trk_smoother_run(0.002, mn, off, n_init, _f(x), x.size, _f(ref))
        assert np.array_equal(own.view(np.uint32), ref.view(np.uint32))


def test_discriminators_and_lock_detectors_bit_exact(oracle, ref_libs):
    L = _own_lib(oracle)
    R = ref_libs.trk
    rng = np.random.default_rng(4)
    cf = C.c_float
    for _ in range(300):
        v = (rng.standard_normal(8) * rng.choice([1.0, 100.0, 1e4])).astype(np.float32)
        a = [cf(float(t)) for t in v]
        assert L.oracle_fll_diff_atan(a[0], a[1], a[2], a[3], C.c_double(0.0), C.c_double(0.001)) == R.ref_trk_fll_diff_atan(float(v[0]), float(v[1]), float(v[2]), float(v[3]), 0.0, 0.001)
        assert L.oracle_pll_cloop_two_quadrant_atan(a[0], a[1]) == R.ref_trk_pll_cloop(float(v[0]), float(v[1]))
        assert L.oracle_dll_nc_e_minus_l_normalized(a[0], a[1], a[2], a[3], cf(0.25), cf(1.0), cf(1.0)) == \
            R.ref_trk_dll_e_minus_l(float(v[0]), float(v[1]), float(v[2]), float(v[3]), 0.25, 1.0, 1.0)
        assert L.oracle_dll_nc_vemlp_normalized(_f(v[0:2].copy()), _f(v[2:4].copy()), _f(v[4:6].copy()), _f(v[6:8].copy())) == \
            R.ref_trk_dll_vemlp(_f(v[0:2].copy()), _f(v[2:4].copy()), _f(v[4:6].copy()), _f(v[6:8].copy()))
    assert L.oracle_pll_cloop_two_quadrant_atan(cf(0.0), cf(1.0)) == 0.0 == R.ref_trk_pll_cloop(0.0, 1.0)
    for n in (1, 20, 64):
        p = (rng.standard_normal(2 * n) * 50 + np.tile([400.0, 0.0], n)).astype(np.float32)
        assert np.float32(L.oracle_cn0_m2m4_estimator(_f(p), n, cf(0.001))).view(np.uint32) == np.float32(R.ref_trk_cn0_m2m4(_f(p), n, 0.001)).view(np.uint32)
        assert np.float32(L.oracle_carrier_lock_detector(_f(p), n)).view(np.uint32) == np.float32(R.ref_trk_carrier_lock(_f(p), n)).view(np.uint32)


def _loop_scene(system, veml, n_ms=600):
    fs = 4e6
    iq, truth, carrier = synth.tracking_stream(system, [7], fs, n_ms, seed=11, cn0=47.0, dopplers=[1830.0], delays=[1234.4])
    return fs, iq, truth[0], carrier


@pytest.mark.parametrize("system,veml", [("gps_l1_ca", False), ("gal_e1b", True)])
def test_closed_loop_converges_on_injected_signal(oracle, oracle_mod, system, veml):
    fs, iq, t, carrier = _loop_scene(system, veml)
    chip_rate, code_len, _ = synth.SYSTEMS[system]
    spc = 2 if system == "gal_e1b" else 1
    cfg = oracle_mod.TrkCfg(fs_in=fs, code_chip_rate=chip_rate, signal_carrier_freq=carrier, code_period_s=code_len / chip_rate,
                            code_length_chips=code_len, code_samples_per_chip=spc, vector_length=int(round(fs * code_len / chip_rate)), veml=int(veml),
                            early_late_space_chips=0.15 if veml else 0.25, very_early_late_space_chips=0.5, slope=1.0, y_intercept=1.0,
                            pll_bw_hz=35.0, dll_bw_hz=2.0, fll_bw_hz=35.0, pll_filter_order=3, dll_filter_order=2, enable_fll_pull_in=1,
                            enable_fll_steady_state=0, carrier_aiding=1, pull_in_time_s=0, cn0_samples=20, cn0_min=25, max_code_lock_fail=50,
                            max_carrier_lock_fail=5000, carrier_lock_th=0.7, cn0_smoother_alpha=0.002, carrier_lock_test_smoother_alpha=0.002,
                            cn0_smoother_samples=200, carrier_lock_test_smoother_samples=25, max_channels=1, device=0)
    trk = oracle.trk(cfg)
    trk.set_code(synth.tracking_code(system, 7))
    trk.start(acq_delay_samples=round(t["delay"]) + 1.0, acq_doppler_hz=t["doppler"] - (200.0 if system == "gps_l1_ca" else 40.0),  # 4 ms epochs alias at 250 Hz
              acq_samplestamp=0, now_sample=3000)
    n_ep = int(600 / (1e3 * code_len / chip_rate)) - 3
    ep, done = trk.run(iq, 0, n_ep)
    assert done >= n_ep - 2
    ok = ep[ep["flags"] & 1 == 1]
    assert not (ok["flags"] & 4).any()                       # no loss of lock
    tail = ok[-100:] if system == "gps_l1_ca" else ok[-40:]
    assert abs(np.mean(tail["carrier_doppler_hz"]) - t["doppler"]) < 2.0
    assert abs(np.mean(tail["code_freq_chips"]) - t["code_rate"]) < 0.5
    assert abs(np.mean(tail["cn0_db_hz"]) - 47.0) < 3.0
    # the prompt carries the energy: |P| well above |E|,|L| tails and the start of each epoch is a code-period boundary
    P = tail["corr"][:, 4] + 1j * tail["corr"][:, 5]
    E = tail["corr"][:, 2] + 1j * tail["corr"][:, 3]
    assert np.mean(np.abs(P)) > 1.1 * np.mean(np.abs(E))
    # the fork's EVM indicator (dll_pll_veml_tracking.cc:1027-1053) over the last cn0_samples prompts, recomputed in float64
    for j in (len(ok) - 1, len(ok) - 37):
        buf = ok["corr"][j - 19:j + 1, 4].astype(np.float64) + 1j * ok["corr"][j - 19:j + 1, 5].astype(np.float64)
        d = np.sqrt(np.mean(buf.real ** 2))
        evm = np.sqrt(np.mean((np.abs(buf.real / d) - 1.0) ** 2 + np.abs(buf.imag / d) ** 2))
        assert ok["evm"][j] == pytest.approx(evm, rel=1e-5) and 0.02 < evm < 0.5
    assert ok["evm"][0] == 0.0                                # nothing before the prompt buffer has filled (:845, :975-979)
    period = fs * code_len / t["code_rate"]
    assert np.all(np.abs(np.mod(tail["sample_counter"].astype(np.float64) - tail["rem_code_phase_samples"] * 0 - t["delay"] + period / 2, period) - period / 2) < 2.5)
    trk.close()
