"""Host-side checks that need no GPU: the C-ABI library loads and exports every symbol include/gnss_b200.h declares,
the pure-host helpers (wipe-off phase pieces, FFT planner, thresholds, row merge) agree with the oracle, and the product
refuses to run without a CUDA device (there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_gpu


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "gnss_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return sorted(set(re.findall(r"\b(gnss_b200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(b200):
    L = b200.lib()
    names = _header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"{n} is declared in include/gnss_b200.h but not exported by libgnss_b200.so"
    assert sorted(b200.EXPORTS) == names, "binding table and header disagree"


def test_no_torch_types_in_the_boundary():
    text = open(os.path.join(ROOT, "include", "gnss_b200.h")).read()
    text = re.sub(r"//[^\n]*", "", re.sub(r"/\*.*?\*/", "", text, flags=re.S))
    for banned in ("torch", "at::", "c10::", "std::", "template"):
        assert banned not in text, f"{banned!r} in the C header"
    # the header must compile as plain C
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "t.c")
        open(src, "w").write('#include "gnss_b200.h"\nint main(void){gnss_b200_acq_cfg c; (void)c; return 0;}\n')
        subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), src])


@pytest.mark.skipif(has_gpu(), reason="checks the no-device error path")
def test_fails_loudly_without_a_device(b200):
    assert b200.lib().gnss_b200_device_count() == 0
    with pytest.raises(b200.B200Error) as e:
        b200.PcpsAcquisitionB200(4000, 4000, 10000, 500, 4e6, 4, 1)
    assert "device" in str(e.value).lower() or "cuda" in str(e.value).lower()
    with pytest.raises(b200.B200Error):
        b200.MulticorrelatorB200(1, 4000, 3)


def test_missing_library_is_an_error_not_a_fallback(b200, monkeypatch):
    monkeypatch.setattr(b200, "_lib", None)
    monkeypatch.setattr(b200, "LIB_PATH", "/nonexistent/libgnss_b200.so")
    with pytest.raises(b200.B200Error) as e:
        b200.lib()
    assert "no CPU fallback" in str(e.value)


@pytest.mark.parametrize("doppler,fs,n", [(250, 4e6, 4000), (-9750, 4e6, 8000), (4975, 8e6, 32000), (-5000, 8e6, 32000),
                                          (37, 25e6, 100000), (9999, 4e6, 160000), (0, 4e6, 4000)])
def test_wipeoff_pieces_reproduce_the_binary32_running_sum(b200, doppler, fs, n):
    """pcps_acquisition::update_local_carrier (pcps_acquisition.cc:262-279) feeds volk_gnsssdr_s32f_sincos_32fc a
    phase that is a binary32 running sum; the device rebuilds it from a few linear pieces. Must be bit-identical."""
    ph, n_pieces = b200.debug_wipeoff_phase(doppler, fs, n)
    step = np.float32(-(np.float32(6.283185307179586) * np.float32(doppler) / np.float32(fs)))
    want = np.empty(n, dtype=np.float32)
    acc = np.float32(0)
    for i in range(n):
        want[i] = acc
        acc = np.float32(acc + step)
    assert np.array_equal(ph.view(np.uint32), want.view(np.uint32))
    assert 1 <= n_pieces <= 64


def test_wipeoff_table_matches_oracle(b200, oracle):
    acq = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=10000, doppler_step=500, fs=4e6, samples_per_chip=4,
                     samples_per_code=4000)
    tab = acq.wipeoffs()
    for k in (0, 7, 20, 39):
        dop = -10000 + 500 * k
        ph, _ = b200.debug_wipeoff_phase(dop, 4e6, 4000)
        np.testing.assert_allclose(np.cos(ph.astype(np.float64)), tab[k, :4000].real, atol=3e-7)
        np.testing.assert_allclose(np.sin(ph.astype(np.float64)), tab[k, :4000].imag, atol=3e-7)


@pytest.mark.parametrize("n", [4000, 8000, 8192, 16000, 16384, 25000, 32000, 50000, 65536, 100000, 160000, 2000, 1024, 5040,
                               6000, 12000, 7 * 11 * 13 * 4, 4092, 16368, 65472, 31 * 64, 17 * 1000, 127 * 8])
def test_fft_plan_is_a_digit_scramble(b200, n):
    """The forward transform leaves bins in digit-scrambled order; the map must be a permutation of 0..N-1 that keeps DC
    in place, and the radices must multiply to N."""
    stages, radices, perm = b200.debug_fft_plan(n)
    assert stages in (1, 2)
    assert int(np.prod(np.array(radices, dtype=np.int64))) == n
    assert perm[0] == 0
    assert np.array_equal(np.sort(perm), np.arange(n, dtype=np.uint32))
    # digit reversal: bin(pos) = sum_j digit_j * N / (R_j * stride_j)
    stride, want = 1, np.zeros(n, dtype=np.int64)
    pos = np.arange(n, dtype=np.int64)
    for r in radices:
        want += ((pos // stride) % r) * (n // (r * stride))
        stride *= r
    assert np.array_equal(want, perm.astype(np.int64))


@pytest.mark.parametrize("n", [2056, 131 * 64, 257 * 1000])
def test_fft_plan_rejects_uncovered_primes(b200, n):
    with pytest.raises(b200.B200Error) as e:
        b200.debug_fft_plan(n)
    assert "prime" in str(e.value)


@pytest.mark.parametrize("pfa,eff,nd,dw,bt", [(0.01, 4000, 40, 1, False), (1e-3, 32000, 80, 1, False), (1e-5, 160000, 800, 10, False),
                                              (0.1, 8000, 41, 2, True), (1e-7, 25000, 160, 1, False)])
def test_threshold_matches_oracle_and_scipy(b200, oracle, pfa, eff, nd, dw, bt):
    """pcps_acquisition::calculate_threshold (pcps_acquisition.cc:903-908): 2 * gamma_p_inv(2 * max_dwells, (1-pfa)^(1/cells))."""
    from scipy.special import gammaincinv
    got = b200.acq_threshold(pfa, eff, nd, dw, bt)
    assert got == pytest.approx(oracle.threshold(pfa, eff, nd, dw, bt), rel=2e-6)
    cells = float(np.float32(eff * nd))
    a = 2.0 * (1 if bt else dw)
    p = (1.0 - float(np.float32(pfa))) ** (1.0 / cells)
    # p is within 1e-13 of 1 for the big grids: one ulp of p moves the threshold by ~2e-5 relative
    assert got == pytest.approx(2.0 * gammaincinv(a, p), rel=1e-4)


def test_finalize_rows_merges_like_one_search(b200):
    """Sharded search: each rank owns a slice of Doppler rows and the host merges the per-row maxima. The merge must pick
    the lowest row on ties and reproduce the reference's statistic arithmetic."""
    rng = np.random.default_rng(5)
    nd, eff = 40, 4000
    rows = np.zeros(nd, dtype=b200.ACQ_ROW_DTYPE)
    rows["max_val"] = rng.uniform(1, 2, nd).astype(np.float32)
    rows["max_idx"] = rng.integers(0, eff, nd)
    rows["row_sum"] = np.float32(0.37 * eff)
    rows["valid"] = 1
    rows["max_val"][[11, 29]] = np.float32(9.5)   # tie: lowest row wins (volk index_max semantics, pcps_acquisition.cc:451-456)
    rows["max_idx"][11], rows["max_idx"][29] = 123, 456
    res = b200.finalize_rows(rows, eff, 10000, 500, 0, 1)
    assert int(res["doppler_index"]) == 11 and int(res["code_phase"]) == 123
    assert int(res["doppler_hz"]) == -10000 + 500 * 11
    # shuffled arrival order must not matter
    res2 = b200.finalize_rows(rows.copy(), eff, 10000, 500, 0, 1)
    assert res2.tobytes() == res.tobytes()


def test_acq_geometry_matches_reference_configurations(b200):
    """acq_conf.cc:24-119 and the adapters' vector sizing for the BASELINE configurations."""
    g = b200.acq_geometry(4e6, 1.023e6, 1, 1)
    assert (g["fft_size"], g["samples_per_chip"]) == (4000, 4)
    g = b200.acq_geometry(8e6, 1.023e6, 4, 4)
    assert (g["fft_size"], g["samples_per_chip"]) == (32000, 8)
    g = b200.acq_geometry(4e6, 1.023e6, 1, 20)
    assert (g["fft_size"], g["consumed_samples"]) == (160000, 80000)
    g = b200.acq_geometry(4e6, 1.023e6, 1, 1, bit_transition=True)
    assert (g["fft_size"], g["effective_size"]) == (8000, 4000)


def test_transform_length_choice_needs_no_device(b200, monkeypatch):
    """gnss_b200_debug_transform_length = the choice gnss_b200_acq_create makes: codelet lengths as they are,
    lengths with a prime factor >= 31 as a padded correlation on the cheapest 2^a 3^b 5^c length >= fft_size + consumed - 1 (by
    points x cost per point of the plan kind), in every code layout; B200_FFT_PAD forces the choice."""
    monkeypatch.delenv("B200_FFT_PAD", raising=False)
    L = b200.debug_transform_length
    for n in (4000, 8000, 16000, 25000, 32000, 100000, 160000, 7 * 11 * 13 * 4, 17 * 1000):
        assert L(n) == n
    want = {2046: 4096, 4092: 8192, 8184: 16384, 12276: 25000, 16368: 32768, 32736: 65536, 65472: 160000,   # the prime 31
            5003: 10240, 4146: 10000, 2053: 5000}                                                           # primes above 127
    for n, m in want.items():
        assert L(n) == m, n
        assert m >= 2 * n - 1
    # zero-padded double-length dwell (consumed = fft_size / 2): M >= 1.5 fft_size - 1; bit transition: consumed = fft_size
    assert L(16368, 8184) == 25000 and L(8184, 4092) == 12500 and L(8224, 4112) == 12500 and L(4112, 4112) == 10000
    monkeypatch.setenv("B200_FFT_PAD", "0")
    assert L(4092) == 4092 and L(65472) == 65472
    assert L(5003) == 10240                       # no plan of its own: padded whatever the switch says
    monkeypatch.setenv("B200_FFT_PAD", "1")
    assert L(4000) == 8000
    with pytest.raises(b200.B200Error):
        L(4000, 4001)
