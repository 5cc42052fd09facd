"""Seeded synthetic inputs for tests and bench: PRN chips from the golden fixture, local codes
sampled the way the reference's acquisition adapters sample them, and IQ blocks with injected satellites + AWGN
following the recipe of the reference's in-tree generator (src/algorithms/signal_generator/gnuradio_blocks/
signal_generator_c.cc:326-420) with fixed seeds.

This module is input generation only: it does not import oracle/ and is not part of the product path.
The numpy float32 arithmetic below restates (and tests check bit-for-bit against) the reference samplers:
  gps_l1_ca_code_gen_complex_sampled   src/algorithms/libs/gps_sdr_signal_replica.cc:136-176
  beidou_b1i_code_gen_complex_sampled  src/algorithms/libs/beidou_b1i_signal_replica.cc:130-170
  galileo_e1_code_gen_complex_sampled  src/algorithms/libs/galileo_e1_signal_replica.cc:143-228 (cboc=false)
  resampler                            src/algorithms/libs/gnss_signal_replica.cc:257-272
"""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_f32 = np.float32
_chips_cache = None

SYSTEMS = {
    # name: (chip_rate, code_len, ms_per_code)
    "gps_l1_ca": (1023000, 1023, 1),
    "bds_b1i": (2046000, 2046, 1),
    "gal_e1b": (1023000, 4092, 4),
    "gal_e1c": (1023000, 4092, 4),
    "gps_l5i": (10230000, 10230, 1),
    "gps_l5q": (10230000, 10230, 1),
}


def load_chips():
    """{system: int8 [n_prn][code_len] of +1/-1} from tests/golden/prn_chips.npz."""
    global _chips_cache
    if _chips_cache is None:
        z = np.load(os.path.join(ROOT, "tests/golden/prn_chips.npz"))
        lens = json.loads(str(z["lengths"]))
        out = {}
        for k, n in lens.items():
            bits = np.unpackbits(z[k], axis=1)[:, :n]
            out[k] = (bits.astype(np.int8) * 2 - 1)
        _chips_cache = out
    return _chips_cache


def chips(system, prn):
    return load_chips()[system][prn - 1]


def _aux_ceil(x_f32):
    return (x_f32 + _f32(1)).astype(np.int64).astype(np.int32)


def samples_per_code(fs, chip_rate, code_len):
    return int(float(fs) / (float(chip_rate) / float(code_len)))


def sample_lfsr_style(ch, chip_rate, fs, imag_code, tc_via_double):
    code_len = ch.size
    n = samples_per_code(fs, chip_rate, code_len)
    tc = _f32(1.0 / float(_f32(chip_rate))) if tc_via_double else _f32(1) / _f32(chip_rate)
    ts = _f32(1) / _f32(fs)
    i = np.arange(n, dtype=np.float32)
    aux = (ts * (i + _f32(1))) / tc
    idx = _aux_ceil(aux) - 1
    idx[n - 1] = code_len - 1
    v = ch[idx].astype(np.float32)
    out = np.zeros(n, dtype=np.complex64)
    if imag_code:
        out.imag = v
    else:
        out.real = v
    return out


def sample_e1_sinboc11(ch, fs):
    code_len, chip_rate, spchip = 4092, 1023000, 2
    boc = np.empty(2 * code_len, dtype=np.float32)
    boc[0::2] = ch
    boc[1::2] = -ch.astype(np.float32)
    n = int(float(fs) / (float(chip_rate) / code_len))
    if fs != spchip * chip_rate:
        fs_in = _f32(spchip * chip_rate)
        t_out = _f32(1) / _f32(fs)
        i = np.arange(n - 1, dtype=np.float32)
        aux = (t_out * (i + _f32(1))) * fs_in
        res = np.empty(n, dtype=np.float32)
        res[: n - 1] = boc[_aux_ceil(aux) - 1]
        res[n - 1] = boc[-1]
    else:
        res = boc
    return res.astype(np.complex64)


def sample_l5(ch, fs):
    code_len, chip_rate = 10230, 10230000
    n = samples_per_code(fs, chip_rate, code_len)
    tc = _f32(1.0 / float(_f32(chip_rate)))
    ts = _f32(1) / _f32(fs)
    i = np.arange(n, dtype=np.float32)
    idx = np.ceil(ts * (i + _f32(1)) / tc).astype(np.int32) - 1
    idx[n - 1] = code_len - 1
    return ch[idx].astype(np.float32).astype(np.complex64)


def local_code(system, prn, fs):
    """One code period sampled at fs, as the adapter's set_local_code() generates it."""
    ch = chips(system, prn)
    if system == "gps_l1_ca":
        return sample_lfsr_style(ch, 1023000, fs, True, False)
    if system == "bds_b1i":
        return sample_lfsr_style(ch, 2046000, fs, False, True)
    if system in ("gal_e1b", "gal_e1c"):
        return sample_e1_sinboc11(ch, fs)
    if system in ("gps_l5i", "gps_l5q"):
        return sample_l5(ch, fs)
    raise KeyError(system)


def adapter_code(system, prn, fs, sampled_ms):
    """The vector the adapter hands to pcps_acquisition::set_local_code: the code period replicated over the
    coherent time (gps_l1_ca_pcps_acquisition.cc:163-167, galileo_e1_pcps_ambiguous_acquisition.cc:190-194)."""
    ms_per_code = SYSTEMS[system][2]
    one = local_code(system, prn, fs)
    reps = max(1, sampled_ms // ms_per_code)
    return np.tile(one, reps)


def make_iq(n_samples, fs, sats, seed, noise_sigma2=1.0):
    """IQ block: sum of satellites + complex AWGN of total variance noise_sigma2.

    sats: list of dict(system, prn, delay (samples, int), doppler (Hz), cn0 (dB-Hz), phase (rad, optional))
    Amplitude follows the reference generator: sqrt(10^(CN0/10) / BW / 2) * sqrt(2) per unit-variance noise,
    i.e. signal power / noise power = CN0 / fs.
    """
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n_samples) + 1j * rng.standard_normal(n_samples)) * np.sqrt(noise_sigma2 / 2.0)
    n = np.arange(n_samples, dtype=np.float64)
    for s in sats:
        code = local_code(s["system"], s["prn"], fs)
        # unit-modulus code value (GPS local code is purely imaginary; its modulus is what carries power)
        c = np.where(np.abs(code.real) > 0, code.real, code.imag).astype(np.float64)
        idx = (np.arange(n_samples) - int(s["delay"])) % code.size
        amp = np.sqrt(noise_sigma2 * 10.0 ** (s["cn0"] / 10.0) / fs)
        ph = s.get("phase", float(rng.uniform(0, 2 * np.pi)))
        x = x + amp * c[idx] * np.exp(1j * (2.0 * np.pi * s["doppler"] * n / fs + ph))
    return x.astype(np.complex64)


def to_cshort(x, scale=64.0):
    """gr_complex -> interleaved int16 (lv_16sc_t) the way a front-end delivers 'cshort' items."""
    y = np.empty(2 * x.size, dtype=np.int16)
    y[0::2] = np.clip(np.rint(x.real * scale), -32768, 32767)
    y[1::2] = np.clip(np.rint(x.imag * scale), -32768, 32767)
    return y


# BASELINE.json configs made concrete
CONFIGS = {
    "C1": dict(system="gps_l1_ca", fs=4000000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=32,
               present=[1, 3, 7, 10, 15, 21, 22, 30], cn0=(38, 50), seed=0),
    "C2": dict(system="gal_e1b", fs=8000000, sampled_ms=4, doppler_max=5000, doppler_step=125, n_prn=50,
               present=[1, 4, 9, 11, 19, 24, 27, 30, 36, 44], cn0=(36, 48), seed=2),
    # weak-signal search: 20 ms coherent (N = 2 * 80 000 zero-padded, pcps_acquisition.cc:85-92) x 10 non-coherent dwells
    "C3": dict(system="gps_l1_ca", fs=4000000, sampled_ms=20, doppler_max=10000, doppler_step=25, n_prn=32,
               present=[4, 11, 18, 29], cn0=(25, 29), seed=3, max_dwells=10),
    "C4_gps": dict(system="gps_l1_ca", fs=25000000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=32,
                   present=[2, 5, 9, 13, 17, 23, 26, 31], cn0=(38, 50), seed=4),
    "C4_gal": dict(system="gal_e1b", fs=25000000, sampled_ms=4, doppler_max=5000, doppler_step=250, n_prn=36,
                   present=[2, 7, 8, 12, 21, 25, 31, 33], cn0=(36, 48), seed=5),
    # sample rates that are multiples of 1.023 MHz: the dwell length carries the prime 31 (direct-DFT pass of the FFT)
    "F4092": dict(system="gps_l1_ca", fs=4092000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=8,
                  present=[1, 3, 7], cn0=(46, 52), seed=7),
    "F16368": dict(system="gps_l1_ca", fs=16368000, sampled_ms=1, doppler_max=2000, doppler_step=250, n_prn=6,
                   present=[2, 5], cn0=(46, 52), seed=8),
    "F65472": dict(system="gal_e1b", fs=16368000, sampled_ms=4, doppler_max=1000, doppler_step=250, n_prn=4,
                   present=[1, 4], cn0=(42, 50), seed=9),
    "F2046": dict(system="gps_l1_ca", fs=2046000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=8,
                  present=[1, 3, 7], cn0=(46, 52), seed=10),
    "F8184": dict(system="gps_l1_ca", fs=8184000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=8,
                  present=[1, 3, 7], cn0=(46, 52), seed=11),
    "F12276": dict(system="gps_l1_ca", fs=12276000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=8,
                   present=[1, 3, 7], cn0=(46, 52), seed=12),
    "F32736": dict(system="gal_e1b", fs=8184000, sampled_ms=4, doppler_max=2000, doppler_step=250, n_prn=4,
                   present=[1, 4], cn0=(42, 50), seed=13),
    "C4_bds": dict(system="bds_b1i", fs=25000000, sampled_ms=1, doppler_max=5000, doppler_step=250, n_prn=63,
                   present=[1, 6, 9, 14, 19, 27, 33, 40, 52, 60], cn0=(38, 50), seed=6),
}


def config_scene(name, n_prn=None, n_dwells=1):
    """(geometry dict, codes [n_prn][consumed] c64, iq [consumed] c64, truth list) for a named config.
    With n_dwells > 1 `iq` is [n_dwells][consumed]: consecutive blocks of ONE continuous stream."""
    c = CONFIGS[name]
    chip_rate, _, ms_per_code = SYSTEMS[c["system"]]
    fs = c["fs"]
    n_prn = n_prn or c["n_prn"]
    consumed = int(c["sampled_ms"] * float(_f32(_f32(fs) * _f32(0.001))))
    fft_size = consumed if c["sampled_ms"] == ms_per_code else 2 * consumed
    rng = np.random.default_rng(1000 + c["seed"])
    one_len = local_code(c["system"], 1, fs).size
    truth = []
    for prn in c["present"]:
        if prn > n_prn:
            continue
        truth.append(dict(system=c["system"], prn=prn, delay=int(rng.integers(0, one_len)),
                          doppler=float(rng.uniform(-0.9, 0.9) * c["doppler_max"]) + 13.7,
                          cn0=float(rng.uniform(*c["cn0"]))))
    iq = make_iq(consumed * n_dwells, fs, truth, c["seed"])
    if n_dwells > 1:
        iq = iq.reshape(n_dwells, consumed)
    codes = np.stack([adapter_code(c["system"], p, fs, c["sampled_ms"])[:consumed] for p in range(1, n_prn + 1)])
    geo = dict(fft_size=fft_size, consumed_samples=consumed, fs=fs, doppler_max=c["doppler_max"],
               doppler_step=c["doppler_step"], samples_per_chip=int(np.ceil(float(_f32(fs) / _f32(chip_rate)))),
               samples_per_code=float(_f32(_f32(_f32(fs) * _f32(0.001)) * _f32(ms_per_code))), n_prn=n_prn,
               system=c["system"], sampled_ms=c["sampled_ms"], max_dwells=c.get("max_dwells", 1))
    return geo, codes.astype(np.complex64), iq, truth


def tracking_stream(system, prns, fs, n_ms, seed, cn0=45.0, dopplers=None, delays=None, code_samples_per_chip=1, data_bits=True):
    """A continuous IQ stream for closed-loop tracking tests: every listed PRN with its own code delay (samples, may be
    fractional), carrier Doppler AND the matching code Doppler (chip rate scaled by 1 + f_d / f_carrier), 50 bps-like
    data-bit flips every 20 code periods, unit-variance complex AWGN. Returns (iq c64, truth list, carrier_hz)."""
    chip_rate, code_len, _ = SYSTEMS[system]
    carrier = {"gps_l1_ca": 1575.42e6, "gal_e1b": 1575.42e6, "bds_b1i": 1561.098e6, "gps_l5i": 1176.45e6}[system]
    rng = np.random.default_rng(seed)
    n = int(round(fs * 1e-3 * n_ms))
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * np.sqrt(0.5)
    t = np.arange(n, dtype=np.float64)
    truth = []
    for k, prn in enumerate(prns):
        fd = float(dopplers[k]) if dopplers is not None else float(rng.uniform(-4000, 4000))
        delay = float(delays[k]) if delays is not None else float(rng.uniform(0, fs / (chip_rate / code_len)))
        ch = chips(system, prn).astype(np.float64)
        rate = chip_rate * (1.0 + fd / carrier)
        phase_chips = (t - delay) * (rate / fs)          # chip phase of every sample
        idx = np.floor(phase_chips).astype(np.int64)
        c = ch[np.mod(idx, code_len)]
        if system == "gal_e1b":                          # sinBOC(1,1): second half of each chip inverted
            c = c * np.where(np.mod(phase_chips, 1.0) < 0.5, 1.0, -1.0)
        if data_bits:
            period = np.floor_divide(idx, code_len)
            bits = rng.integers(0, 2, int(period.max() - period.min()) // 20 + 2) * 2 - 1
            c = c * bits[(period - period.min()) // 20]
        amp = np.sqrt(10.0 ** (cn0 / 10.0) / fs)
        ph0 = float(rng.uniform(0, 2 * np.pi))
        x = x + amp * c * np.exp(1j * (2.0 * np.pi * fd * t / fs + ph0))
        truth.append(dict(system=system, prn=prn, doppler=fd, delay=delay, cn0=cn0, code_rate=rate))
    return x.astype(np.complex64), truth, carrier


SECONDARY = {  # the reference's strings: GPS_L5.h:171-172, Galileo_E1.h:52 ('0' adds, '1' subtracts in save_correlation_results)
    "gps_l5i": "0000110101", "gps_l5q": "00000100110101001110", "gal_e1c": "0011100000001010110110010",
}


def tracking_stream_pilot_data(kind, prn, fs, n_ms, seed, cn0=48.0, doppler=1500.0, delay=100.3, doppler_rate=0.0):
    """One satellite with BOTH components, for pilot tracking with a slave data correlator:
       kind "gps_l5": (L5I * NH10 * data(100 sps) + j * L5Q * NH20) / sqrt(2)            — pilot in quadrature
       kind "gal_e1": (E1B * data(250 sps) - E1C * secondary25) / sqrt(2), sinBOC(1,1)   — pilot in anti-phase
    doppler_rate [Hz/s] adds a carrier (and matching code) acceleration for the high-dynamics mode.
    Returns (iq c64, truth dict with the transmitted symbols and the period index each starts at, carrier_hz)."""
    data_sys, pilot_sys = {"gps_l5": ("gps_l5i", "gps_l5q"), "gal_e1": ("gal_e1b", "gal_e1c")}[kind]
    chip_rate, code_len, _ = SYSTEMS[data_sys]
    carrier = {"gps_l5": 1176.45e6, "gal_e1": 1575.42e6}[kind]
    rng = np.random.default_rng(seed)
    n = int(round(fs * 1e-3 * n_ms))
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * np.sqrt(0.5)
    t = np.arange(n, dtype=np.float64) / fs
    fd = doppler + doppler_rate * t                                   # instantaneous Doppler
    carr_phase = 2.0 * np.pi * (doppler * t + 0.5 * doppler_rate * t * t)
    phase_chips = chip_rate * ((t - delay / fs) + (doppler * t + 0.5 * doppler_rate * t * t) / carrier)
    idx = np.floor(phase_chips).astype(np.int64)
    period = np.floor_divide(idx, code_len)
    p0 = int(period.min())
    d = chips(data_sys, prn).astype(np.float64)[np.mod(idx, code_len)]
    q = chips(pilot_sys, prn).astype(np.float64)[np.mod(idx, code_len)]
    if kind == "gal_e1":
        boc = np.where(np.mod(phase_chips, 1.0) < 0.5, 1.0, -1.0)
        d, q = d * boc, q * boc
    sec_q = np.array([1.0 if ch == "0" else -1.0 for ch in SECONDARY[pilot_sys]])
    sec_offset = int(rng.integers(0, sec_q.size))                      # the receiver has to find it
    q = q * sec_q[np.mod(period - p0 + sec_offset, sec_q.size)]
    if kind == "gps_l5":
        sec_d = np.array([1.0 if ch == "0" else -1.0 for ch in SECONDARY["gps_l5i"]])
        sym_len = 10                                                    # code periods per data symbol, aligned with NH10 and NH20
        d = d * sec_d[np.mod(period - p0 + sec_offset, sec_d.size)]
        sym_index = np.floor_divide(period - p0 + sec_offset, sym_len)
    else:
        sym_len = 1
        sym_index = period - p0
    symbols = rng.integers(0, 2, int(sym_index.max()) + 2) * 2 - 1
    d = d * symbols[sym_index]
    amp = np.sqrt(10.0 ** (cn0 / 10.0) / fs)
    ph0 = float(rng.uniform(0, 2 * np.pi))
    s = (d + 1j * q) if kind == "gps_l5" else (d - q)
    x = x + amp / np.sqrt(2.0) * s * np.exp(1j * (carr_phase + ph0))
    truth = dict(kind=kind, prn=prn, doppler=doppler, doppler_rate=doppler_rate, delay=delay, cn0=cn0, symbols=symbols, sym_len=sym_len,
                 sec_offset=sec_offset, first_period=p0, final_doppler=float(fd[-1]))
    return x.astype(np.complex64), truth, carrier


def tracking_code(system, prn):
    """d_tracking_code as start_tracking() generates it (dll_pll_veml_tracking.cc:663-700): one float per chip, or the
    sinBOC(1,1) replica at 2 samples per chip for Galileo E1 (galileo_e1_code_gen_sinboc11_float)."""
    ch = chips(system, prn).astype(np.float32)
    if system in ("gal_e1b", "gal_e1c"):
        out = np.empty(2 * ch.size, dtype=np.float32)
        out[0::2] = ch
        out[1::2] = -ch
        return out
    return ch
