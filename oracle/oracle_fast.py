"""TEST INFRASTRUCTURE — vectorised numpy/scipy form of the CPU oracle, used as the TIMED CPU baseline.

Same algorithm as oracle_acq.c / oracle_mc.c (reference: pcps_acquisition.cc:615-713,
cpu_multicorrelator_real_codes.cc:103-144) but batched over Doppler bins and run on all host cores, the way the
reference parallelises (one thread per channel / PRN). The FFT is pocketfft (scipy.fft, complex64, `workers`
threads): FFTW3f, which the reference really calls through gr::fft, is not installed and cannot be — every
number produced here carries the footnote "CPU FFT = pocketfft, not FFTW".

tests/test_oracle_fast.py checks this module against the C oracle (decisions exact, statistics to 1e-4).
"""
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import scipy.fft as sfft

_f32 = np.float32


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def wipeoff_table(n_doppler, doppler_max, doppler_step, fs, n, center=0, bias=0, row_begin=0):
    """update_grid_doppler_wipeoffs (:298-305): float32 running-sum phase per row (np.cumsum in float32 is a
    sequential add, i.e. the same rounding sequence as the reference's loop). Rows row_begin .. row_begin + n_doppler - 1."""
    out = np.empty((n_doppler, n), dtype=np.complex64)
    for kk in range(n_doppler):
        k = kk + row_begin
        doppler = -int(doppler_max) + int(center) + int(doppler_step) * k
        step = _f32(_f32(2.0 * np.pi) * _f32(bias + doppler) / _f32(fs))
        inc = np.full(n, -step, dtype=np.float32)
        inc[0] = 0.0
        ph = np.cumsum(inc, dtype=np.float32)
        out[kk].real = np.cos(ph)
        out[kk].imag = np.sin(ph)
    return out


class FastAcq:
    def __init__(self, fft_size, consumed_samples, doppler_max, doppler_step, fs, samples_per_chip, use_cfar=True,
                 bit_transition=False, workers=None, rows=None, center=0):
        """rows = (begin, end): only those Doppler rows of the grid are built (a row-sharded search; `search` then returns
        the per-row records of the shard through `row_records`, the decision needs all rows)."""
        self.N, self.consumed = fft_size, consumed_samples
        self.eff = fft_size // 2 if bit_transition else fft_size
        self.bt = bit_transition
        self.nD = int(np.ceil((2 * doppler_max) / doppler_step))
        self.doppler_max, self.doppler_step, self.fs, self.spc = doppler_max, doppler_step, fs, samples_per_chip
        self.use_cfar = use_cfar
        self.workers = workers or host_cores()
        self.row_begin, self.row_end = rows if rows is not None else (0, self.nD)
        self.center = center
        self.wipe = wipeoff_table(self.row_end - self.row_begin, doppler_max, doppler_step, fs, fft_size, center=center, row_begin=self.row_begin)

    def code_spectrum(self, code):
        buf = np.zeros(self.N, dtype=np.complex64)
        if self.bt:
            buf[self.N // 2:] = code[: self.N // 2]
        elif self.N == self.consumed:
            buf[:] = code[: self.consumed]
        else:
            buf[self.consumed:] = code[: self.consumed]
        return np.conj(sfft.fft(buf)).astype(np.complex64)

    def forward(self, samples):
        x = np.zeros(self.N, dtype=np.complex64)
        x[: self.consumed] = samples[: self.consumed]
        return sfft.fft(x[None, :] * self.wipe, axis=1, workers=self.workers)

    def search(self, X, code_spec, grid=None, dwell=1, workers=None):
        y = sfft.ifft(X * code_spec[None, :], axis=1, workers=workers or self.workers, norm="forward")  # unnormalised backward
        off = self.eff if self.bt else 0
        g = (y.real[:, off:off + self.eff] ** 2 + y.imag[:, off:off + self.eff] ** 2).astype(np.float32)
        if grid is not None:
            g = grid + g if dwell > 1 else g
            grid[...] = g
        idx = g.argmax(axis=1)
        mx = g[np.arange(g.shape[0]), idx]
        self.row_records = dict(max_val=mx.copy(), max_idx=idx.copy(), row_sum=g.sum(axis=1, dtype=np.float32))
        if g.shape[0] != self.nD:
            return None
        k = int(mx.argmax())
        res = dict(peak=float(mx[k]), code_phase=int(idx[k]), doppler_index=k,
                   doppler_hz=-self.doppler_max + self.center + self.doppler_step * k, dwell_count=dwell)
        if self.use_cfar:
            opp = (k + self.nD // 2) % self.nD
            power = _f32(np.float64(_f32(g[opp].sum(dtype=np.float32)) / _f32(self.eff)) / 2.0 / dwell)
            res.update(input_power=float(power), test_statistic=float(_f32(mx[k]) / power))
        else:
            row = g[k].copy()
            ex1, ex2 = res["code_phase"] - self.spc, res["code_phase"] + self.spc
            sel = np.arange(self.eff)
            if ex1 < 0:
                ex1 += self.N
            elif ex2 >= self.N:
                ex2 -= self.N
            mask = ((sel >= ex1) & (sel < ex2)) if ex1 < ex2 else ((sel >= ex1) | (sel < ex2))
            row[mask] = 0.0
            second = float(row.max())
            res.update(second_peak=second, test_statistic=float(_f32(mx[k]) / _f32(second)))
        return res

    def dwell_all(self, samples, code_specs):
        """One dwell for a list of PRNs. The reference recomputes the wiped-off forward FFT per channel; sharing it here
        only makes the CPU baseline faster (conservative for any GPU/CPU ratio)."""
        X = self.forward(samples)
        return [self.search(X, cs) for cs in code_specs]

    def dwell_all_like_reference(self, samples, code_specs):
        """One thread per PRN, each doing its own forward + inverse transforms (the reference's per-channel structure,
        Channels.in_acquisition = all cores)."""
        x = np.zeros(self.N, dtype=np.complex64)
        x[: self.consumed] = samples[: self.consumed]

        def one(cs):
            X = sfft.fft(x[None, :] * self.wipe, axis=1, workers=1)
            return self.search(X, cs, workers=1)

        with ThreadPoolExecutor(max_workers=self.workers) as ex:
            return list(ex.map(one, code_specs))


def mc_real_fast(sig, code, shifts, rem_carr, carr_step, rem_code, code_step):
    """Vectorised Cpu_Multicorrelator_Real_Codes call (plain mode). Chip indices use the exact float32 expression; the
    carrier is evaluated from the effective angle of the rounded per-sample phasor (no recursion), outputs agree with
    the C oracle to ~1e-5."""
    n = sig.size
    fn = np.arange(n, dtype=np.float32)
    a = _f32(code_step) * fn
    inc = np.complex64(np.exp(np.complex64(complex(0.0, -float(_f32(carr_step))))))
    theta = np.arctan2(np.float64(inc.imag), np.float64(inc.real))
    p0 = complex(float(np.cos(_f32(rem_carr))), -float(np.sin(_f32(rem_carr))))
    z = sig.astype(np.complex128) * (p0 * np.exp(1j * theta * np.arange(n)))
    out = np.empty(len(shifts), dtype=np.complex64)
    L = code.size
    for t, s in enumerate(np.asarray(shifts, dtype=np.float32)):
        idx = np.floor((a + s) - _f32(rem_code)).astype(np.int64) % L
        out[t] = (z * code[idx]).sum()
    return out


def mc_batch_threads(sigs, codes, shifts, args, workers=None):
    workers = workers or host_cores()
    def one(c):
        r = args[c]
        return mc_real_fast(sigs[c], codes[c], shifts[c], r[0], r[1], r[3], r[4])
    with ThreadPoolExecutor(max_workers=workers) as ex:
        return np.stack(list(ex.map(one, range(len(sigs)))))
