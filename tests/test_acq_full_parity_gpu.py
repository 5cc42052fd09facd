"""GPU parity at the FULL sizes of the BASELINE configs that carry the headline claims, against the CPU oracle.

The C oracle (oracle_acq.c) takes minutes for a whole 100 000-point grid, so the checker here is its vectorised form
`oracle_fast.FastAcq` (same float32 running-sum wipe-off, same zero-pad layouts, same statistics; pocketfft for the
transforms), itself pinned to the C oracle by tests/test_oracle_fast.py. Bars (BASELINE.json north_star): statistic and
peak within 1e-4 relative; code-phase index and Doppler bin exact. For a PRN that is absent the "peak" is the largest of
~1e6 noise cells: there the index has to match whenever the oracle's own top-two gap exceeds the float32 error bound of a
cell (GAP), and when it does not, the GPU's cell must be one of the oracle cells inside that bound.
"""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-4
GAP = 2e-5     # relative error bound of one |.|^2 cell between two correct binary32 FFT pipelines (measured: < 4e-6)


def _engine(b200, geo, n_slots, **kw):
    return b200.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                    geo["samples_per_chip"], n_slots, **kw)


def _fast(geo, **kw):
    from b200_import import load_oracle
    load_oracle()                      # puts oracle/ on sys.path (test infrastructure only)
    import oracle_fast
    return oracle_fast.FastAcq(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                               geo["samples_per_chip"], **kw)


def _check_against_grid(got, want, grid, what):
    """Decision record of the engine against the oracle's record and full magnitude grid."""
    assert got["test_statistic"] == pytest.approx(want["test_statistic"], rel=RTOL), what
    assert got["peak"] == pytest.approx(want["peak"], rel=RTOL), what
    k, i = int(got["doppler_index"]), int(got["code_phase"])
    if (k, i) == (want["doppler_index"], want["code_phase"]):
        return True
    # the engine chose another cell: legitimate only if the oracle itself cannot tell the two apart
    top = float(grid.max())
    close = np.argwhere(grid >= top * (1.0 - GAP))
    assert len(close) > 1, (what, "oracle's maximum is unique beyond the float32 bound, the index must match", got, want)
    assert any((k, i) == (int(a), int(b)) for a, b in close), (what, got, want, close.tolist())
    return False


@pytest.mark.parametrize("name", ["C4_gps", "C4_gal", "C4_bds"])
def test_config4_full_length_against_oracle(b200, name):
    """BASELINE configs[3] at 25 MS/s: N = 25 000 = 2^3 5^5 (GPS, BeiDou) and 100 000 = 2^5 5^5 (Galileo), all 40 Doppler rows;
    per constellation two present PRNs and two absent ones, every cell of their grids against the oracle."""
    geo, codes, iq, truth = synth.config_scene(name)
    present = sorted(t["prn"] for t in truth)
    absent = [p for p in range(1, geo["n_prn"] + 1) if p not in present]
    sel = np.array([present[0], absent[0], present[-1], absent[-1]]) - 1
    e = _engine(b200, geo, len(sel), use_cfar=True, keep_grid=True)
    e.set_local_codes(np.arange(len(sel)), codes[sel])
    got = e.dwell(iq, np.arange(len(sel)))
    fa = _fast(geo)
    assert fa.nD == e.n_doppler == 40
    X = fa.forward(iq)
    thr = b200.acq_threshold(0.001, e.eff, e.n_doppler)
    for s, p in enumerate(sel):
        grid = np.zeros((fa.nD, fa.eff), np.float32)
        want = fa.search(X, fa.code_spectrum(codes[p]), grid=grid)
        same = _check_against_grid(got[s], want, grid, (name, int(p) + 1))
        if (p + 1) in present and want["test_statistic"] > thr:
            assert same and got[s]["test_statistic"] > thr
            assert int(got[s]["doppler_hz"]) == want["doppler_hz"]
        g = e.read_grid(s)
        assert np.abs(g - grid).max() <= GAP * grid.max(), (name, int(p) + 1)
    e.close()


def test_config3_full_length_rows_against_oracle(b200):
    """BASELINE configs[2]: N = 160 000 = 2^8 5^4 (20 ms coherent, zero-padded), +-10 kHz / 25 Hz = 800 rows, non-coherent
    accumulation. A 16-row shard (row_begin/row_end, the unit a sharded search hands to one GPU) that contains the weak
    satellite's Doppler, 3 accumulated dwells: the persistent grid after every dwell and the per-row records a shard
    exchanges, cell for cell against the oracle."""
    geo, codes, iq, truth = synth.config_scene("C3", n_prn=4, n_dwells=3)
    t = truth[0]
    assert t["prn"] == 4 and geo["fft_size"] == 160000
    k_true = int(round((t["doppler"] + geo["doppler_max"]) / geo["doppler_step"]))
    lo = max(0, k_true - 8)
    hi = lo + 16
    e = _engine(b200, geo, 1, use_cfar=True, keep_grid=True, row_begin=lo, row_end=hi)
    assert e.n_doppler == 800
    e.set_local_codes([0], codes[3:4])
    fa = _fast(geo, rows=(lo, hi))
    spec = fa.code_spectrum(codes[3])
    grid = np.zeros((16, fa.eff), np.float32)
    for d in range(3):
        e.dwell(iq[d], [0])
        fa.search(fa.forward(iq[d]), spec, grid=grid, dwell=d + 1)
        g = e.read_grid(0)
        assert g.shape == grid.shape
        assert np.abs(g - grid).max() <= GAP * grid.max(), d
        rows = e.read_rows(0)
        assert rows["valid"][lo:hi].all() and not rows["valid"][:lo].any() and not rows["valid"][hi:].any()
        rr = fa.row_records
        assert np.allclose(rows["max_val"][lo:hi], rr["max_val"], rtol=RTOL)
        assert np.allclose(rows["row_sum"][lo:hi], rr["row_sum"], rtol=RTOL)
        for j in range(16):
            if int(rows["max_idx"][lo + j]) != int(rr["max_idx"][j]):
                # a row of noise: two cells the oracle cannot tell apart
                assert grid[j, int(rows["max_idx"][lo + j])] >= rr["max_val"][j] * (1.0 - GAP), (d, j)
    # after 3 dwells the satellite's row holds the shard's maximum at its code phase
    j = int(np.argmax(rows["max_val"][lo:hi]))
    assert abs(lo + j - k_true) <= 1 and abs(int(rows["max_idx"][lo + j]) % 4000 - t["delay"]) <= 1
    e.close()


def test_config2_all_50_prns_against_oracle(b200):
    """BASELINE configs[1] (the bench workload): Galileo E1, 8 MS/s, N = 32 000, 80 rows — every one of the 50 PRNs."""
    geo, codes, iq, truth = synth.config_scene("C2", n_prn=50)
    e = _engine(b200, geo, 50, use_cfar=True)
    e.set_local_codes(np.arange(50), codes)
    got = e.dwell(iq, np.arange(50))
    fa = _fast(geo)
    assert fa.nD == e.n_doppler == 80
    X = fa.forward(iq)
    thr = b200.acq_threshold(0.001, e.eff, e.n_doppler)
    differ = []
    det_g, det_o = set(), set()
    for p in range(50):
        grid = np.zeros((fa.nD, fa.eff), np.float32)
        want = fa.search(X, fa.code_spectrum(codes[p]), grid=grid)
        if not _check_against_grid(got[p], want, grid, ("C2", p + 1)):
            differ.append(p + 1)
        if want["test_statistic"] > thr:
            det_o.add(p + 1)
        if got[p]["test_statistic"] > thr:
            det_g.add(p + 1)
    assert det_g == det_o and det_o >= {t["prn"] for t in truth if t["cn0"] > 40}
    assert not (set(differ) & det_o), differ      # a detected PRN never has an ambiguous cell
    e.close()


def test_config1_index_differences_are_float32_ties(b200, oracle):
    """BASELINE configs[0], all 32 PRNs against the C oracle: wherever the engine's cell differs from the oracle's, the
    oracle's own grid holds both cells within the float32 bound (and the PRN is an undetected, noise-only one)."""
    geo, codes, iq, truth = synth.config_scene("C1")
    e = _engine(b200, geo, 32, use_cfar=True)
    e.set_local_codes(np.arange(32), codes)
    got = e.dwell(iq, np.arange(32))
    thr = oracle.threshold(0.001, 4000, e.n_doppler)
    differ = []
    for p in range(32):
        a = oracle.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"],
                       doppler_step=geo["doppler_step"], fs=geo["fs"], samples_per_chip=geo["samples_per_chip"],
                       samples_per_code=geo["samples_per_code"], use_cfar=True)
        a.set_local_code(codes[p])
        a.reset()
        want = a.dwell(iq)
        grid = a.grid()[:, :geo["fft_size"]]
        if not _check_against_grid(got[p], want, grid, ("C1", p + 1)):
            differ.append((p + 1, want["test_statistic"]))
        a.close()
    for prn, stat in differ:
        assert stat < thr, (prn, stat)
    print("C1 PRNs whose cell is a float32 tie in the oracle grid:", differ)


def test_gsoc2013_four_satellites(b200, oracle):
    """The reference's multi-satellite known-answer case (gps_l1_ca_pcps_acquisition_gsoc2013_test.cc:265-349 config_2,
    :388-392 process_message): PRN 10 / 15 / 21 / 22 at 44 dB-Hz with noise and data bits, fs = 4 MS/s, 1 ms, pfa 0.001,
    +-10 kHz / 250 Hz. The reference asserts detection + correct estimation (delay error < 0.5 chip, Doppler error
    < 2/(3 T) = 666 Hz) on the noise-free config_1 and only PRINTS Pd for the noisy config_2 (44 dB-Hz over 1 ms is ~3 dB
    above the pfa = 0.001 threshold). Here: the noise-free case must be detected and correct; over several noise
    realisations every record equals the oracle's, the detection decisions are the oracle's, and every detection of PRN 10
    is a correct estimation."""
    fs = 4000000
    sats = [(10, 600, 750.0), (15, 100, 1000.0), (21, 200, 2000.0), (22, 300, 3000.0)]
    geo = dict(fft_size=4000, consumed_samples=4000, doppler_max=10000, doppler_step=250, fs=fs, samples_per_chip=4, samples_per_code=4000.0)
    codes = np.stack([synth.adapter_code("gps_l1_ca", p, fs, 1) for p, _, _ in sats] + [synth.adapter_code("gps_l1_ca", 5, fs, 1)])
    e = _engine(b200, geo, 5, use_cfar=True)
    e.set_local_codes(np.arange(5), codes)
    thr = b200.acq_threshold(0.001, 4000, e.n_doppler)
    assert thr == pytest.approx(oracle.threshold(0.001, 4000, 80), rel=1e-6) and e.n_doppler == 80
    correct = detected = 0
    n_real = 8
    for r in range(n_real + 1):
        truth = [dict(system="gps_l1_ca", prn=p, delay=int(round(ch * 4000 / 1023)), doppler=fd, cn0=44.0) for p, ch, fd in sats]
        if r == n_real:                # config_1: PRN 10 alone, noise_flag = false
            iq = synth.make_iq(4000, fs, [dict(truth[0], cn0=90.0)], seed=2013, noise_sigma2=1e-12)
        else:
            iq = synth.make_iq(4000, fs, truth, seed=2013 + r)
        got = e.dwell(iq, np.arange(5))
        for s in range(5):
            a = oracle.acq(use_cfar=True, **geo)
            a.set_local_code(codes[s])
            a.reset()
            want = a.dwell(iq)
            _check_against_grid(got[s], want, a.grid()[:, :4000], ("gsoc2013", r, s))
            assert (got[s]["test_statistic"] > thr) == (want["test_statistic"] > thr)
            a.close()
        if got[0]["test_statistic"] > thr:
            detected += 1
            delay_err_chips = abs(600.0 - float(got[0]["code_phase"]) * 1023.0 / 4000.0)
            if delay_err_chips < 0.5 and abs(750.0 - got[0]["doppler_hz"]) < 2.0 / (3.0 * 1e-3):
                correct += 1
        assert got[4]["test_statistic"] < thr      # PRN 5 is not in the sky
        if r == n_real:
            assert got[0]["test_statistic"] > thr  # config_1: "Expected message: 1=ACQ SUCCESS"
    assert detected >= 2 and correct == detected   # "Incorrect parameters estimation" never happens on a detection
    e.close()
