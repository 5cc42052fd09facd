"""The identity behind the engine's padded search (csrc/acq.cu create_impl, stage_kernel<LOAD_CODE> in
csrc/acq_kernels.cuh), checked on the CPU in double precision for the three code layouts of pcps_acquisition::set_local_code
(pcps_acquisition.cc:176-209): the cyclic correlation of length n = fft_size,

    y[m] = sum_i x[i] conj(c[(i - m) mod n]),   x nonzero on [0, consumed),

equals, for every lag 0 <= m < n, the cyclic correlation of ANY length M >= n + consumed - 1 of the samples zero-padded to M
with a code buffer that holds c[0 .. consumed) at its start and c[1 .. n) at its END, scaled by n / M (unnormalised transforms)."""
import numpy as np
import pytest


def reference_lags(x, c):
    n = c.size
    return np.fft.ifft(np.fft.fft(x, n) * np.conj(np.fft.fft(c))) * n


def padded_code_buffer(code, code_off, n, consumed, M):
    """the LOAD_CODE rule of the padded form: position -> index into the code's own period, or nothing"""
    buf = np.zeros(M, dtype=complex)
    for pos in range(M):
        if pos < consumed:
            j = pos
        elif pos + n > M:
            j = pos + n - M
        else:
            continue
        if code_off <= j < code_off + code.size:
            buf[pos] = code[j - code_off] * (n / M)
    return buf


LAYOUTS = {
    # name: (fft_size, consumed, code_len, code_off) for one code period of `p` samples
    "full": lambda p: (p, p, p, 0),
    "zero_padded_double_length": lambda p: (2 * p, p, p, p),        # :200-204
    "bit_transition": lambda p: (2 * p, 2 * p, p, p),               # :188-193 (the search keeps the upper half of the lags)
}


@pytest.mark.parametrize("layout", sorted(LAYOUTS))
@pytest.mark.parametrize("p", [31, 62, 93])
def test_padded_cyclic_correlation_reproduces_every_lag(layout, p):
    rng = np.random.default_rng(p)
    n, consumed, code_len, code_off = LAYOUTS[layout](p)
    x = np.zeros(n, dtype=complex)
    x[:consumed] = rng.standard_normal(consumed) + 1j * rng.standard_normal(consumed)
    code = rng.standard_normal(code_len) + 1j * rng.standard_normal(code_len)
    c = np.zeros(n, dtype=complex)
    c[code_off:code_off + code_len] = code
    want = reference_lags(x, c)
    m_min = n + consumed - 1
    for M in (m_min, m_min + 1, 1 << int(np.ceil(np.log2(m_min))), 3 * n):
        xm = np.zeros(M, dtype=complex)
        xm[:consumed] = x[:consumed]
        got = np.fft.ifft(np.fft.fft(xm) * np.conj(np.fft.fft(padded_code_buffer(code, code_off, n, consumed, M)))) * M
        assert np.abs(got[:n] - want).max() <= 1e-12 * np.abs(want).max(), (layout, p, M)


def test_shorter_than_the_bound_aliases():
    """M = n + consumed - 2 folds one product onto a wrong lag for a code that fills its period: the bound is tight."""
    p = 40
    rng = np.random.default_rng(1)
    x = rng.standard_normal(p) + 1j * rng.standard_normal(p)
    code = rng.standard_normal(p) + 1j * rng.standard_normal(p)
    want = reference_lags(x, code)
    M = 2 * p - 2
    xm = np.zeros(M, dtype=complex)
    xm[:p] = x
    got = np.fft.ifft(np.fft.fft(xm) * np.conj(np.fft.fft(padded_code_buffer(code, 0, p, p, M)))) * M
    assert np.abs(got[:p] - want).max() > 1e-3 * np.abs(want).max()
