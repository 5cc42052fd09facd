"""TEST INFRASTRUCTURE ONLY — CPU restatement of the acquisition resampler.

The reference builds it from two GNU Radio pieces (gnss_flowgraph.cc:1063-1118): taps =
gr::filter::firdes::low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 2) and gr::filter::fir_filter_ccf::make(decimation, taps),
with the decimation of Acq_Conf::ConfigureAutomaticResampler (acq_conf.cc:93-111). GNU Radio is a dependency outside the
reference tree (any 3.7.3+ per the reference's CMakeLists; not installed here), so this restates the published algorithm
of gr-filter/lib/firdes.cc and fir_filter.cc. The tap design is pinned on the known_taps vector of GNU Radio's own QA
(qa_firdes.py test_low_pass, quoted in tests/test_resampler.py) and cross-checked against scipy.signal.firwin; the decimating
FIR (fir_filter_ccf history / phase) has no published vector and stays pinned on numpy's convolution only.
"""
import numpy as np


def automatic_resampler(fs_in, opt_freq):
    """Acq_Conf::ConfigureAutomaticResampler (acq_conf.cc:93-111) -> (decimation, resampled_fs)."""
    if fs_in <= opt_freq:
        return 1, int(fs_in)
    decimation = int(fs_in // opt_freq)
    while int(fs_in) % decimation > 0:
        decimation -= 1
    return decimation, int(fs_in) // decimation


def firdes_low_pass(gain, fs, cutoff, transition):
    """firdes::low_pass with the default Hamming window; float32 taps."""
    ntaps = int(53.0 * fs / (22.0 * transition))      # compute_ntaps: max attenuation of the Hamming window = 53 dB
    if ntaps % 2 == 0:
        ntaps += 1
    M = (ntaps - 1) // 2
    n = np.arange(-M, M + 1)
    w = (0.54 - 0.46 * np.cos(2 * np.pi * (n + M) / (ntaps - 1))).astype(np.float32) if ntaps > 1 else np.ones(1, np.float32)
    fwT0 = 2 * np.pi * cutoff / fs
    with np.errstate(invalid="ignore", divide="ignore"):
        t = np.where(n == 0, fwT0 / np.pi, np.sin(n * fwT0) / (n * np.pi))
    taps = (t * w).astype(np.float32)
    fmax = float(taps[M]) + 2.0 * float(np.sum(taps[M + 1:].astype(np.float64)))
    return (taps.astype(np.float64) * (gain / fmax)).astype(np.float32)


class FirDecimator:
    """fir_filter_ccf: y[m] = sum_j taps[j] x[m D - j], ntaps - 1 items of history (zeros at the start of the stream)."""

    def __init__(self, decimation, taps):
        self.D = int(decimation)
        self.taps = np.asarray(taps, dtype=np.float64)
        self.hist = np.zeros(self.taps.size - 1, dtype=np.complex128)

    def work(self, x):
        x = np.asarray(x, dtype=np.complex128)
        assert x.size % self.D == 0
        buf = np.concatenate([self.hist, x])
        full = np.convolve(buf, self.taps)[self.taps.size - 1: self.taps.size - 1 + x.size]   # causal output at every input index
        self.hist = buf[buf.size - (self.taps.size - 1):] if self.taps.size > 1 else self.hist
        return full[::self.D].astype(np.complex64)
