"""GPU parity of the PCPS acquisition pipeline against the oracle (through the C-ABI).

Bars (BASELINE.json north_star): detected PRN set, code-phase index and Doppler bin identical to the reference
path; test statistics within 1e-4 relative. For PRNs that are absent the "peak" is the maximum of ~1e5..1e6 noise
cells and a 1e-7 relative rounding difference may move it; there the index must match whenever
the oracle's top-two gap exceeds the float32 error bound, and the statistic must always agree to 1e-4.
"""
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-4
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _oracle_search(oracle, geo, codes, iq_blocks, use_cfar=True, bit_transition=False, centers=None, keep=False):
    out, grids = [], []
    for p in range(codes.shape[0]):
        a = oracle.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"],
                       doppler_step=geo["doppler_step"], fs=geo["fs"], samples_per_chip=geo["samples_per_chip"],
                       samples_per_code=geo["samples_per_code"], use_cfar=use_cfar, bit_transition=bit_transition,
                       doppler_center=0 if centers is None else centers[p])
        a.set_local_code(codes[p])
        a.reset()
        r = None
        per_dwell = []
        for blk in iq_blocks:
            r = a.dwell(blk)
            per_dwell.append(r)
        out.append(per_dwell)
        if keep:
            eff = geo["fft_size"] // 2 if bit_transition else geo["fft_size"]
            grids.append(a.grid()[:, :eff])
        a.close()
    return out, grids


def _gpu_engine(b200, geo, n_slots, **kw):
    return b200.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                    geo["samples_per_chip"], n_slots, **kw)


def _compare(got, want, thr=None, strict_idx=True):
    assert got["test_statistic"] == pytest.approx(want["test_statistic"], rel=RTOL)
    assert got["peak"] == pytest.approx(want["peak"], rel=RTOL)
    detected = thr is not None and want["test_statistic"] > thr
    if strict_idx or detected:
        assert int(got["code_phase"]) == want["code_phase"]
        assert int(got["doppler_index"]) == want["doppler_index"]
        assert int(got["doppler_hz"]) == want["doppler_hz"]
    assert int(got["dwell_count"]) == want["dwell_count"]


def test_gps_capture_known_answer(b200, oracle):
    """The reference's own fixture and expected answer (gps_l1_ca_pcps_acquisition_test.cc:290-291,354-361)."""
    d = np.fromfile(os.path.join(GOLD, "GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat"), dtype=np.complex64)
    geo = dict(fft_size=4000, consumed_samples=4000, doppler_max=5000, doppler_step=100, fs=4000000, samples_per_chip=4, samples_per_code=4000.0)
    code = synth.adapter_code("gps_l1_ca", 1, 4000000, 1)[None, :]
    for use_cfar in (False, True):
        e = _gpu_engine(b200, geo, 1, use_cfar=use_cfar)
        e.set_local_codes([0], code)
        r = e.dwell(d[:4000], [0])[0]
        assert r["code_phase"] == 524 and r["doppler_hz"] == 1700 and abs(r["doppler_hz"] - 1680) <= 666
        want, _ = _oracle_search(oracle, geo, code, [d[:4000]], use_cfar=use_cfar)
        _compare(r, want[0][0])
        e.close()


def test_galileo_capture_known_answer(b200, oracle):
    """galileo_e1_pcps_ambiguous_acquisition_test.cc:304-305,359-369: N = 16000 (single-stage plan)."""
    d = np.fromfile(os.path.join(GOLD, "Galileo_E1_ID_1_Fs_4Msps_8ms.dat"), dtype=np.complex64)
    geo = dict(fft_size=16000, consumed_samples=16000, doppler_max=10000, doppler_step=250, fs=4000000, samples_per_chip=4, samples_per_code=16000.0)
    code = synth.adapter_code("gal_e1b", 1, 4000000, 4)[None, :]
    e = _gpu_engine(b200, geo, 1, use_cfar=False)
    e.set_local_codes([0], code)
    r = e.dwell(d[:16000], [0])[0]
    assert r["code_phase"] == 2920 and r["doppler_hz"] == -750
    want, _ = _oracle_search(oracle, geo, code, [d[:16000]], use_cfar=False)
    _compare(r, want[0][0])


def test_config1_all_prns(b200, oracle):
    """BASELINE config 1 (GPS L1 C/A, 4 MS/s, 1 ms, +-5 kHz / 250 Hz = 40 bins, 32 PRNs): every PRN against the oracle."""
    geo, codes, iq, truth = synth.config_scene("C1")
    e = _gpu_engine(b200, geo, 32, use_cfar=True, keep_grid=True)
    e.set_local_codes(np.arange(32), codes)
    got = e.dwell(iq, np.arange(32))
    want, grids = _oracle_search(oracle, geo, codes, [iq], keep=True)
    thr = oracle.threshold(0.001, 4000, e.n_doppler)
    det_g = {p + 1 for p in range(32) if got[p]["test_statistic"] > thr}
    det_o = {p + 1 for p in range(32) if want[p][0]["test_statistic"] > thr}
    assert det_g == det_o and det_o >= {t["prn"] for t in truth if t["cn0"] > 42}
    for p in range(32):
        _compare(got[p], want[p][0], thr=thr, strict_idx=False)
    # the full grid (what the reference's MAT dump holds) for one present and one absent PRN
    for p in (0, 1):
        g = e.read_grid(p)
        assert np.abs(g - grids[p]).max() <= 2e-5 * grids[p].max()
    # exact index agreement wherever it is well defined, i.e. almost everywhere
    same = sum(int(got[p]["code_phase"]) == want[p][0]["code_phase"] and int(got[p]["doppler_index"]) == want[p][0]["doppler_index"] for p in range(32))
    assert same >= 30


# transform length the engine picks for these dwell lengths (gnss_b200_acq_transform_length): the padded form on a length with a
# fast plan where that is cheaper than a direct-DFT pass over the 31
PADDED_LENGTH = {"F2046": 4096, "F4092": 8192, "F8184": 16384, "F12276": 25000, "F16368": 32768, "F32736": 65536, "F65472": 160000}


@pytest.mark.parametrize("form", ["chosen", "direct"])
@pytest.mark.parametrize("name", ["F4092", "F16368", "F65472", "F12276"])
def test_dwell_lengths_with_the_prime_31(b200, oracle, name, form, monkeypatch):
    """Sample rates that are multiples of 1.023 MHz (4.092, 16.368 MS/s front ends): N = 4092 = 4*3*11*31, 16 368 and
    65 472 (two stages). The reference's FFT takes any length; here the 31 runs as a direct-DFT pass ("direct") or, by default,
    the search runs as a padded cyclic correlation on a longer length with a fast plan ("chosen"). Every PRN
    against the oracle in both forms, injected satellites found at their cells."""
    geo, codes, iq, truth = synth.config_scene(name)
    n = geo["n_prn"]
    assert geo["fft_size"] % 31 == 0
    if form == "direct":
        monkeypatch.setenv("B200_FFT_PAD", "0")
    else:
        monkeypatch.delenv("B200_FFT_PAD", raising=False)
    e = _gpu_engine(b200, geo, n, use_cfar=True, keep_grid=True)
    assert e.transform_length == (geo["fft_size"] if form == "direct" else PADDED_LENGTH[name])
    e.set_local_codes(np.arange(n), codes)
    got = e.dwell(iq, np.arange(n))
    want, grids = _oracle_search(oracle, geo, codes, [iq], keep=True)
    thr = oracle.threshold(0.001, geo["fft_size"], e.n_doppler)
    for p in range(n):
        _compare(got[p], want[p][0], thr=thr, strict_idx=False)
        g = e.read_grid(p)
        assert np.abs(g - grids[p]).max() <= RTOL * grids[p].max()
    det_g = {p + 1 for p in range(n) if got[p]["test_statistic"] > thr}
    det_o = {p + 1 for p in range(n) if want[p][0]["test_statistic"] > thr}
    assert det_g == det_o and det_o and det_o <= {t["prn"] for t in truth}
    e.close()


@pytest.mark.parametrize("scene,n_prn", [("C2", 6), ("C1", 8), ("F4092", 4)])
def test_peer_exchange_with_a_single_shard(b200, scene, n_prn):
    """gnss_b200_acq_peer_* with world = 1 on one GPU: the fused tail of a connected shard (per-tile records folded, delivered
    and decided by two kernels instead of the merge / decision / pack launches) gives the records of a plain engine bit for
    bit, eagerly and from a captured graph. The two-GPU form is tests/test_two_devices_gpu.py."""
    import torch
    geo, codes, iq, _ = synth.config_scene(scene, n_prn=n_prn, n_dwells=3)
    N = geo["consumed_samples"]
    iq = np.ascontiguousarray(iq).reshape(3, N)
    slots = np.arange(n_prn, dtype=np.uint32)
    one = _gpu_engine(b200, geo, n_prn, use_cfar=True)
    one.set_local_codes(slots, codes[:n_prn])
    e = _gpu_engine(b200, geo, n_prn, use_cfar=True)
    e.set_local_codes(slots, codes[:n_prn])
    e.exchange_buffers(1)
    _, ptr = e.peer_setup(0)
    e.peer_connect(regions=[ptr])
    buf = torch.empty(2 * N, dtype=torch.float32, device="cuda")
    for k in range(3):
        want = one.dwell(iq[k], slots)
        buf.copy_(torch.from_numpy(iq[k].view(np.float32).copy()))
        torch.cuda.synchronize()
        if k == 1:
            e.capture_begin()
            e.dwell_device_async(buf.data_ptr(), slots)
            e.peer_exchange_async(n_prn)
            e.capture_end()
        e.peer_barrier_async()
        if k >= 1:
            e.replay_async()
        else:
            e.dwell_device_async(buf.data_ptr(), slots)
            with pytest.raises(b200.B200Error):
                e.sync(n_prn)
            e.peer_exchange_async(n_prn)
        got = e.sync(n_prn)
        assert got.tobytes() == want.tobytes(), (scene, k)
        for p in (0, n_prn - 1):
            assert e.read_rows(p).tobytes() == one.read_rows(p).tobytes()
    # a graph that holds the dwell only, the exchange queued eagerly behind every replay
    e.capture_begin()
    e.dwell_device_async(buf.data_ptr(), slots)
    e.capture_end()
    for _ in range(2):
        e.replay_async()
        e.peer_exchange_async(n_prn)
        assert e.sync(n_prn).tobytes() == want.tobytes()
    e.close()
    one.close()


def test_padded_form_with_the_other_engine_features(b200, oracle, monkeypatch):
    """The padded search (4.092 MS/s: 4092 lags on an 8192-point transform) under everything else a handle can be asked for:
    the first-vs-second-peak statistic with and without a kept grid, cshort samples with shifted Doppler centres, a narrow
    second-step grid, and a split into row shards — each against the oracle's 4092-point search."""
    monkeypatch.delenv("B200_FFT_PAD", raising=False)
    geo, codes, iq, truth = synth.config_scene("F4092", n_prn=4)
    slots = np.arange(4)
    present = {t["prn"] for t in truth}
    want, _ = _oracle_search(oracle, geo, codes, [iq], use_cfar=False)
    for keep in (False, True):
        e = _gpu_engine(b200, geo, 4, use_cfar=False, keep_grid=keep)
        assert e.transform_length == 8192
        e.set_local_codes(slots, codes)
        got = e.dwell(iq, slots)
        for p in range(4):
            _compare(got[p], want[p][0], strict_idx=(p + 1) in present)
            assert got[p]["second_peak"] == pytest.approx(want[p][0]["second_peak"], rel=RTOL)
        e.close()
    # peaks in the first / last samples of the code period: the exclusion window wraps at d_fft_size (:580-591), not at 8192
    n = geo["fft_size"]
    edge = [dict(system="gps_l1_ca", prn=1, delay=1, doppler=730.0, cn0=50.0), dict(system="gps_l1_ca", prn=3, delay=n - 2, doppler=-2210.0, cn0=50.0)]
    iq_edge = synth.make_iq(n, geo["fs"], edge, seed=5)
    want_e, _ = _oracle_search(oracle, geo, codes, [iq_edge], use_cfar=False)
    e = _gpu_engine(b200, geo, 4, use_cfar=False)
    e.set_local_codes(slots, codes)
    got = e.dwell(iq_edge, slots)
    assert int(got[0]["code_phase"]) == 1 and int(got[2]["code_phase"]) == n - 2
    for p in (0, 2):
        _compare(got[p], want_e[p][0])
        assert got[p]["second_peak"] == pytest.approx(want_e[p][0]["second_peak"], rel=RTOL)
    e.close()
    # cshort items, Doppler centres
    iqs = synth.to_cshort(iq, scale=64.0)
    centers = [0, 750, -1125, 250]
    e = _gpu_engine(b200, geo, 4, use_cfar=True)
    e.set_local_codes(slots, codes)
    for p, c in enumerate(centers):
        e.set_doppler_center(p, c)
    got = e.dwell(iqs, slots)
    want, _ = _oracle_search(oracle, geo, codes, [iqs], centers=centers)
    for p in range(4):
        _compare(got[p], want[p][0], strict_idx=(p == 0))
    e.close()
    # second step around the coarse Doppler
    e1 = _gpu_engine(b200, geo, 4, use_cfar=True)
    e2 = _gpu_engine(b200, geo, 4, use_cfar=True, step2=(4, 125.0))
    assert e2.transform_length == 8192
    e1.set_local_codes(slots, codes)
    e2.set_local_codes(slots, codes)
    r1 = e1.dwell(iq, slots)
    for p in range(4):
        e2.set_step2_center(p, float(r1[p]["doppler_hz"]), r1[p]["input_power"])
    r2 = e2.dwell(iq, slots)
    for p in range(4):
        a = oracle.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"],
                       doppler_step=geo["doppler_step"], fs=geo["fs"], samples_per_chip=geo["samples_per_chip"],
                       samples_per_code=geo["samples_per_code"], use_cfar=True)
        a.set_local_code(codes[p])
        a.reset()
        w1 = a.dwell(iq)
        w2 = a.dwell_step2(iq, 4, 125.0, float(w1["doppler_hz"]))
        a.close()
        if int(r1[p]["doppler_hz"]) != w1["doppler_hz"]:
            continue
        assert r2[p]["peak"] == pytest.approx(w2["peak"], rel=RTOL) and r2[p]["test_statistic"] == pytest.approx(w2["test_statistic"], rel=RTOL)
        if (p + 1) in present:
            assert int(r2[p]["code_phase"]) == w2["code_phase"] and int(r2[p]["doppler_hz"]) == w2["doppler_hz"]
    e2.close()
    # three row shards merged on the host == the unsharded handle, bit for bit
    nD = e1.n_doppler
    cuts = [0, 13, 27, nD]
    rows = np.zeros((4, nD), dtype=b200.ACQ_ROW_DTYPE)
    for r0, r1_ in zip(cuts[:-1], cuts[1:]):
        sh = _gpu_engine(b200, geo, 4, use_cfar=True, row_begin=r0, row_end=r1_)
        sh.set_local_codes(slots, codes)
        sh.dwell(iq, slots)
        for p in range(4):
            rows[p, r0:r1_] = sh.read_rows(p)[r0:r1_]
        sh.close()
    for p in range(4):
        assert rows[p].tobytes() == e1.read_rows(p).tobytes()
        m = b200.finalize_rows(rows[p], geo["fft_size"], geo["doppler_max"], geo["doppler_step"], 0, 1)
        assert m["test_statistic"] == r1[p]["test_statistic"] and m["code_phase"] == r1[p]["code_phase"] and m["doppler_hz"] == r1[p]["doppler_hz"]
    e1.close()


def test_transform_length_choice(b200, monkeypatch):
    """Codelet lengths are searched as they are; lengths with the prime 31 on the cheapest padded length; a forced choice
    (B200_FFT_PAD) is honoured; zero-padded and bit-transition dwells take the padded form as well."""
    monkeypatch.delenv("B200_FFT_PAD", raising=False)
    def length(n, consumed=None, **kw):
        e = b200.PcpsAcquisitionB200(n, consumed or n, 5000, 250, n * 1000.0, 4, 1, **kw)
        m = e.transform_length
        e.close()
        return m
    for n in (4000, 8000, 16000, 25000, 32000, 100000, 160000, 7 * 11 * 13 * 4):
        assert length(n) == n
    for name, m in PADDED_LENGTH.items():
        assert length(synth.CONFIGS[name]["fs"] // 1000 * (4 if synth.CONFIGS[name]["system"] == "gal_e1b" else 1)) == m
    assert length(8184, consumed=4092) == 12500      # zero-padded double-length dwell: M >= fft_size + consumed - 1
    assert length(4092, bit_transition=True) == 8192
    monkeypatch.setenv("B200_FFT_PAD", "1")
    assert length(4000) == 8000
    monkeypatch.setenv("B200_FFT_PAD", "0")
    assert length(4092) == 4092


@pytest.mark.parametrize("fs,n_dwells", [(5003000, 1), (4146000, 2), (2053000, 1)])
def test_lengths_with_a_large_prime_factor(b200, oracle, fs, n_dwells):
    """d_fft_size = fs * 1 ms with a prime factor the mixed-radix kernels do not take (5003 is prime, 4146 = 2 * 3 * 691,
    2053 is prime): the reference's FFTW plan takes any length (pcps_acquisition.cc:85-92). The engine computes the same cyclic
    correlation on a padded length (acq.cu, create_impl) — results against the oracle's any-length DFT: statistic, peak, grid,
    exact cell."""
    n = fs // 1000
    prns = [3, 9, 17, 25]
    sats = [dict(system="gps_l1_ca", prn=3, delay=1234, doppler=1830.0, cn0=47.0), dict(system="gps_l1_ca", prn=17, delay=n - 7, doppler=-3320.0, cn0=45.0)]
    iq = synth.make_iq(n * n_dwells, fs, sats, seed=fs % 997)
    codes = np.stack([synth.local_code("gps_l1_ca", p, fs)[:n] for p in prns])
    geo = dict(fft_size=n, consumed_samples=n, doppler_max=5000, doppler_step=250, fs=fs, samples_per_chip=int(np.ceil(fs / 1.023e6)), samples_per_code=float(n))
    e = _gpu_engine(b200, geo, len(prns), use_cfar=True, keep_grid=True)
    e.set_local_codes(np.arange(len(prns)), codes)
    blocks = [iq[k * n:(k + 1) * n] for k in range(n_dwells)]
    for blk in blocks:
        got = e.dwell(blk, np.arange(len(prns)))
    want, grids = _oracle_search(oracle, geo, codes, blocks, keep=True)
    thr = oracle.threshold(0.001, n, e.n_doppler)
    for p in range(len(prns)):
        _compare(got[p], want[p][-1], thr=thr, strict_idx=False)
        g = e.read_grid(p)
        assert g.shape == grids[p].shape and np.abs(g - grids[p]).max() <= RTOL * grids[p].max()
    for k, p in ((0, 3), (2, 17)):
        assert got[k]["test_statistic"] > thr and int(got[k]["code_phase"]) == want[k][-1]["code_phase"] and int(got[k]["doppler_hz"]) == want[k][-1]["doppler_hz"]
    assert got[1]["test_statistic"] < thr and got[3]["test_statistic"] < thr
    e.close()


def test_config2_two_stage_plan(b200, oracle):
    """BASELINE config 2 geometry (Galileo E1, 8 MS/s, 4 ms, N = 32000 -> two-stage FFT), 6 of the 50 PRNs against the oracle."""
    geo, codes, iq, truth = synth.config_scene("C2", n_prn=50)
    sel = np.array([0, 3, 8, 10, 20, 49])
    e = _gpu_engine(b200, geo, 50, use_cfar=True)
    e.set_local_codes(np.arange(50), codes)
    got = e.dwell(iq, np.arange(50))
    want, _ = _oracle_search(oracle, geo, codes[sel], [iq])
    thr = oracle.threshold(0.001, 32000, e.n_doppler)
    for i, p in enumerate(sel):
        _compare(got[p], want[i][0], thr=thr, strict_idx=False)
    det = {p + 1 for p in range(50) if got[p]["test_statistic"] > thr}
    assert det >= {t["prn"] for t in truth if t["cn0"] > 40}
    assert not (det - {t["prn"] for t in truth})
    for t in truth:
        if t["prn"] in det:
            r = got[t["prn"] - 1]
            assert abs(int(r["code_phase"]) % 32000 - t["delay"]) <= 1
            assert abs(r["doppler_hz"] - t["doppler"]) <= geo["doppler_step"]


def test_noncoherent_dwells_and_first_vs_second(b200, oracle):
    """max_dwells > 1 (:672-680): persistent grid accumulation, statistic after every dwell; both statistics."""
    geo, codes, _, truth = synth.config_scene("C1", n_prn=4)
    blocks = [synth.make_iq(4000, geo["fs"], [dict(t, cn0=36.0, phase=0.1 * k) for t in truth], seed=70 + k) for k in range(4)]
    for use_cfar in (True, False):
        e = _gpu_engine(b200, geo, 4, use_cfar=use_cfar, keep_grid=True)
        e.set_local_codes(np.arange(4), codes)
        want, grids = _oracle_search(oracle, geo, codes, blocks, use_cfar=use_cfar, keep=True)
        for k, blk in enumerate(blocks):
            got = e.dwell(blk, np.arange(4))
            for p in range(4):
                _compare(got[p], want[p][k], strict_idx=(p in (0, 2)))
        g = e.read_grid(0)
        assert np.abs(g - grids[0]).max() <= 2e-5 * grids[0].max()
        e.reset(0)
        got = e.dwell(blocks[0], [0])
        _compare(got[0], want[0][0])
        e.close()


def test_first_vs_second_without_grid(b200, oracle):
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=8)
    e = _gpu_engine(b200, geo, 8, use_cfar=False, keep_grid=False)
    e.set_local_codes(np.arange(8), codes)
    got = e.dwell(iq, np.arange(8))
    want, _ = _oracle_search(oracle, geo, codes, [iq], use_cfar=False)
    for p in range(8):
        strong = want[p][0]["test_statistic"] > 3.0
        _compare(got[p], want[p][0], strict_idx=strong)
        assert got[p]["second_peak"] == pytest.approx(want[p][0]["second_peak"], rel=RTOL)


@pytest.mark.parametrize("fs,form,lengths", [(4000000, "own", (16000, 8000)), (4092000, "chosen", (25000, 16384)), (4092000, "direct", (16368, 8184)),
                                            (2056000, "chosen", (12500, 10000))])
def test_zero_padded_long_coherent_and_bit_transition(b200, oracle, monkeypatch, fs, form, lengths):
    """sampled_ms != ms_per_code doubles the FFT and zero-pads (:85-92,:200-204); bit_transition keeps the upper half
    (:188-193,:671). At 4 MS/s on the lengths' own plans; at 4.092 MS/s (prime 31) both as a padded correlation on a fast-plan
    length and with the direct-DFT pass; at 2.056 MS/s (4112 = 2^4 * 257: no plan of its own) padded only."""
    if form == "direct":
        monkeypatch.setenv("B200_FFT_PAD", "0")
    else:
        monkeypatch.delenv("B200_FFT_PAD", raising=False)
    n = fs // 1000
    spc = int(np.ceil(fs / 1.023e6))
    truth = [dict(system="gps_l1_ca", prn=5, delay=1234, doppler=1130.0, cn0=42.0)]
    # 2 ms coherent: consumed 2n, N = 4n
    iq = synth.make_iq(2 * n, fs, truth, seed=91)
    geo = dict(fft_size=4 * n, consumed_samples=2 * n, doppler_max=5000, doppler_step=250, fs=fs, samples_per_chip=spc, samples_per_code=float(n))
    codes = np.stack([synth.adapter_code("gps_l1_ca", p, fs, 2) for p in (5, 6)])
    e = _gpu_engine(b200, geo, 2, use_cfar=True, keep_grid=True)
    assert e.transform_length == lengths[0]
    e.set_local_codes([0, 1], codes)
    got = e.dwell(iq, [0, 1])
    want, grids = _oracle_search(oracle, geo, codes, [iq], keep=True)
    _compare(got[0], want[0][0])
    _compare(got[1], want[1][0], strict_idx=False)
    assert int(got[0]["code_phase"]) % n == 1234
    assert np.abs(e.read_grid(0) - grids[0]).max() <= 2e-5 * grids[0].max()
    e.close()
    # bit transition: consumed = N = 2n, effective n; both statistics
    geo = dict(fft_size=2 * n, consumed_samples=2 * n, doppler_max=5000, doppler_step=250, fs=fs, samples_per_chip=spc, samples_per_code=float(n))
    for use_cfar in (True, False):
        e = _gpu_engine(b200, geo, 2, use_cfar=use_cfar, bit_transition=True, keep_grid=True)
        assert e.transform_length == lengths[1]
        e.set_local_codes([0, 1], codes)
        got = e.dwell(iq, [0, 1])
        want, grids = _oracle_search(oracle, geo, codes, [iq], use_cfar=use_cfar, bit_transition=True, keep=True)
        _compare(got[0], want[0][0])
        assert np.abs(e.read_grid(0) - grids[0]).max() <= 2e-5 * grids[0].max()
        e.close()


def test_cshort_input_and_doppler_center(b200, oracle):
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=3)
    iqs = synth.to_cshort(iq, scale=64.0)
    e = _gpu_engine(b200, geo, 3, use_cfar=True)
    e.set_local_codes(np.arange(3), codes)
    centers = [0, 1000, -2375]
    for p, c in enumerate(centers):
        e.set_doppler_center(p, c)
    got = e.dwell(iqs, np.arange(3))
    want, _ = _oracle_search(oracle, geo, codes, [iqs], centers=centers)
    for p in range(3):
        _compare(got[p], want[p][0], strict_idx=(p == 0))


def test_doppler_group_cache_never_aliases_within_a_call(b200):
    """The engine caches the wipe-off piece tables of up to 64 Doppler centres. A call whose centres overflow the cache must
    not hand the table slot of one of ITS OWN groups to another: fill the cache to 63 entries, then search three
    centres of which the first is the oldest cached one; every slot must equal a fresh engine's answer bit for bit."""
    geo, codes, iq, _ = synth.config_scene("C1", n_prn=3)
    e = _gpu_engine(b200, geo, 3, use_cfar=True)
    e.set_local_codes(np.arange(3), codes)
    first = None
    for i in range(21):                                  # 21 calls x 3 new centres = 63 cached groups
        cs = [7 * (3 * i + s) - 400 for s in range(3)]
        first = first or cs
        for s, c in enumerate(cs):
            e.set_doppler_center(s, c)
        e.dwell(iq, np.arange(3))
    probe = [first[0], 1234, -2222]                      # cached at ordinal 0, new, new (the second one overflows)
    for s, c in enumerate(probe):
        e.set_doppler_center(s, c)
    got = e.dwell(iq, np.arange(3))
    fresh = _gpu_engine(b200, geo, 3, use_cfar=True)
    fresh.set_local_codes(np.arange(3), codes)
    for s, c in enumerate(probe):
        fresh.set_doppler_center(s, c)
    want = fresh.dwell(iq, np.arange(3))
    assert got.tobytes() == want.tobytes()
    e.close()
    fresh.close()


def test_cbyte_input(b200, oracle):
    """interleaved int8 items (the adapters' `cbyte` item type, gps_l1_ca_pcps_acquisition.cc:76-80,191-198)
    converted on the device: identical to handing the block the converted gr_complex stream."""
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=4)
    q = np.empty(2 * iq.size, dtype=np.int8)
    q[0::2] = np.clip(np.rint(iq.real * 24.0), -128, 127)
    q[1::2] = np.clip(np.rint(iq.imag * 24.0), -128, 127)
    as_complex = (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)
    e = _gpu_engine(b200, geo, 4, use_cfar=True)
    e.set_local_codes(np.arange(4), codes)
    got = e.dwell(q, np.arange(4))
    same = e.dwell(as_complex, np.arange(4))
    assert got.tobytes() == same.tobytes()
    want, _ = _oracle_search(oracle, geo, codes, [as_complex])
    for p in range(4):
        _compare(got[p], want[p][0], strict_idx=(p in (0, 2)))
    with pytest.raises(b200.B200Error):
        b200._check(e.L.gnss_b200_acq_dwell(e.h, q.ctypes.data, 3, 1, np.zeros(1, np.uint32).ctypes.data_as(b200._u32p), got.ctypes.data))


def test_row_sharding_merge(b200, oracle):
    """Doppler rows of one PRN split over two handles (two GPUs in production), merged with finalize_rows."""
    geo, codes, iq, _ = synth.config_scene("C1", n_prn=2)
    full = _gpu_engine(b200, geo, 2, use_cfar=True)
    full.set_local_codes([0, 1], codes)
    ref = full.dwell(iq, [0, 1])
    nD = full.n_doppler
    halves = []
    for lo, hi in ((0, nD // 2), (nD // 2, nD)):
        e = _gpu_engine(b200, geo, 2, use_cfar=True, row_begin=lo, row_end=hi)
        e.set_local_codes([0, 1], codes)
        e.dwell(iq, [0, 1])
        halves.append([e.read_rows(0), e.read_rows(1)])
    for p in range(2):
        rows = halves[0][p].copy()
        take = halves[1][p]["valid"] == 1
        rows[take] = halves[1][p][take]
        assert rows["valid"].all()
        m = b200.finalize_rows(rows, 4000, geo["doppler_max"], geo["doppler_step"], 0, 1)
        for k in ("test_statistic", "peak", "code_phase", "doppler_index", "doppler_hz", "input_power"):
            assert m[k] == ref[p][k]


def test_a_captured_search_replays_bit_identically(b200):
    """gnss_b200_acq_capture_begin / _capture_end / _replay_async: the launches of one dwell recorded into a CUDA graph (what a
    small grid such as C1 is bound by is launch latency). Replays on new samples in the same device buffer equal eager dwells bit
    for bit; a persistent grid or a running profiler refuse capture."""
    import torch
    geo, codes, iq, truth = synth.config_scene("C1", n_dwells=3)
    n, N = geo["n_prn"], geo["consumed_samples"]
    e = _gpu_engine(b200, geo, n, use_cfar=True)
    slots = np.arange(n, dtype=np.uint32)
    e.set_local_codes(slots, codes)
    iq = np.ascontiguousarray(iq).reshape(3, N)
    d_iq = torch.from_numpy(iq[0].view(np.float32).copy()).cuda()
    e.dwell_device_async(d_iq.data_ptr(), slots)
    e.sync(n)
    l0 = e.launch_count()
    e.capture_begin()
    e.dwell_device_async(d_iq.data_ptr(), slots)
    e.capture_end()
    per_dwell = e.launch_count() - l0
    for k in range(3):
        blk = iq[k]
        want = e.dwell(blk, slots)
        d_iq.copy_(torch.from_numpy(blk.view(np.float32).copy()))
        torch.cuda.synchronize()
        e.replay_async()
        got = e.sync(n)
        for f in ("test_statistic", "peak", "code_phase", "doppler_index", "doppler_hz", "input_power"):
            assert np.array_equal(got[f], want[f]), f
    assert 3 <= per_dwell <= 6
    g = _gpu_engine(b200, geo, n, use_cfar=True, keep_grid=True)
    with pytest.raises(b200.B200Error):
        g.capture_begin()
    g.close()
    e.close()


def test_unsupported_and_state_errors(b200):
    for kw in (dict(), dict(bit_transition=True)):                                               # 257 | 2056: padded correlation, every layout
        b200.PcpsAcquisitionB200(2 * 2056, 2056, 5000, 250, 2.056e6, 3, 1, **kw).close()
    b200.PcpsAcquisitionB200(2056, 2056, 5000, 250, 2.056e6, 3, 1).close()
    e = b200.PcpsAcquisitionB200(4000, 4000, 5000, 250, 4e6, 4, 2)
    with pytest.raises(b200.B200Error):
        e.dwell(np.zeros(4000, np.complex64), [0])  # no local code
    with pytest.raises(b200.B200Error):
        e.read_grid(0)  # keep_grid = 0
    with pytest.raises(b200.B200Error):
        e.set_local_codes([7], np.zeros((1, 4000), np.complex64))  # slot out of range
    e.close()
    # second-step handles (make_two_steps)
    with pytest.raises(b200.B200Error):
        b200.PcpsAcquisitionB200(4000, 4000, 5000, 250, 4e6, 4, 2, step2=(0, 125.0))      # second_nbins = 0
    with pytest.raises(b200.B200Error):
        b200.PcpsAcquisitionB200(4000, 4000, 5000, 250, 4e6, 4, 2, step2=(4, 0.0))        # second_doppler_step = 0
    with pytest.raises(b200.B200Error):
        b200.PcpsAcquisitionB200(4000, 4000, 5000, 250, 4e6, 4, 2, step2=(4, 125.0), row_begin=0, row_end=2)   # not row-sharded
    e2 = b200.PcpsAcquisitionB200(4000, 4000, 5000, 250, 4e6, 4, 2, step2=(4, 125.0))
    with pytest.raises(b200.B200Error):
        e2.set_step2_center(2, 0.0, 1.0)          # slot out of range
    with pytest.raises(b200.B200Error):
        e2.dwell(np.zeros(4000, np.complex64), [0])  # no local code
    e2.close()


@pytest.mark.parametrize("name,min_cn0", [("C4_gps", 47.0), ("C4_gal", 41.0), ("C4_bds", 47.0)])
def test_full_size_config4_properties(b200, name, min_cn0):
    """BASELINE config 4 at full size (25 MS/s; N = 25 000 / 100 000; 32 / 36 / 63 PRNs x 40 bins), where the oracle would take
    minutes: size-independent properties instead. (1) every strong injected satellite is found at its injected cell and nothing
    else crosses the threshold; (2) homogeneity: samples x 2 gives peak x 4 exactly (power-of-two scaling commutes with every
    binary32 operation of the pipeline), identical cells and an identical statistic."""
    geo, codes, iq, truth = synth.config_scene(name)
    n = geo["n_prn"]
    e = _gpu_engine(b200, geo, n, use_cfar=True)
    e.set_local_codes(np.arange(n), codes)
    r1 = e.dwell(iq, np.arange(n))
    thr = b200.acq_threshold(0.001, e.eff, e.n_doppler)
    one = int(round(geo["samples_per_code"]))
    present = {t["prn"] for t in truth}
    for t in truth:
        if t["cn0"] >= min_cn0:
            r = r1[t["prn"] - 1]
            assert r["test_statistic"] > thr
            assert abs(int(r["code_phase"]) % one - t["delay"]) <= 1 and abs(r["doppler_hz"] - t["doppler"]) <= geo["doppler_step"]
    assert not ({p + 1 for p in range(n) if r1[p]["test_statistic"] > thr} - present)
    r2 = e.dwell((iq * np.float32(2.0)).astype(np.complex64), np.arange(n))
    assert np.array_equal(r2["code_phase"], r1["code_phase"]) and np.array_equal(r2["doppler_index"], r1["doppler_index"])
    assert np.array_equal(r2["peak"], r1["peak"] * np.float32(4.0))
    assert np.array_equal(r2["test_statistic"], r1["test_statistic"])
    e.close()


def test_full_size_config3_properties(b200):
    """BASELINE config 3 geometry at full row size (N = 160 000, 800 bins, persistent grid), 2 PRNs: (1) searching the SAME block
    twice doubles every grid cell exactly (x + x in binary32), so the peak doubles, the cell stays, and the CFAR statistic
    doubles too (the noise row is divided by the dwell count, pcps_acquisition.cc:533-536); (2) reset() returns to the one-dwell result bit for bit; (3) the weak satellite comes out after the 10 dwells."""
    geo, codes, iq, truth = synth.config_scene("C3", n_prn=4, n_dwells=10)
    e = _gpu_engine(b200, geo, 2, use_cfar=True, keep_grid=True)
    sel = np.array([3, 0])                       # PRN 4 (present, weak) and PRN 1 (absent)
    e.set_local_codes([0, 1], codes[sel])
    a = e.dwell(iq[0], [0, 1])
    b2 = e.dwell(iq[0], [0, 1])
    assert np.array_equal(b2["peak"], a["peak"] * np.float32(2.0))
    assert np.array_equal(b2["code_phase"], a["code_phase"]) and np.array_equal(b2["doppler_index"], a["doppler_index"])
    assert np.allclose(b2["test_statistic"], 2.0 * a["test_statistic"], rtol=1e-6) and list(b2["dwell_count"]) == [2, 2]
    assert np.array_equal(b2["input_power"], a["input_power"])
    for s in (0, 1):
        e.reset(s)
    c = e.dwell(iq[0], [0, 1])
    assert c.tobytes() == a.tobytes()
    for k in range(1, 10):
        c = e.dwell(iq[k], [0, 1])
    thr = b200.acq_threshold(0.001, e.eff, e.n_doppler, max_dwells=10)
    t = truth[0]
    # (the absent PRN is not required to stay below the pfa threshold: with the zero-padded 2N transform the noise variance of a
    # code phase grows with its overlap, while the reference normalises by the row AVERAGE (:533-536), so noise peaks read ~2x high)
    assert t["prn"] == 4 and c[0]["test_statistic"] > thr and c[0]["test_statistic"] > 2.0 * c[1]["test_statistic"]
    assert abs(int(c[0]["code_phase"]) % 4000 - t["delay"]) <= 1 and abs(c[0]["doppler_hz"] - t["doppler"]) <= 25
    g = e.read_grid(0)
    assert g.shape == (800, 160000) and float(g.max()) == float(c[0]["peak"])
    e.close()


@pytest.mark.parametrize("use_cfar", [True, False])
def test_two_step_acquisition(b200, oracle, use_cfar):
    """SURVEY 8.6 N3, make_two_steps (pcps_acquisition.cc:308-315,715-775): after the coarse search a narrow grid of
    second_nbins bins spaced second_doppler_step around each PRN's own coarse Doppler; float Doppler arithmetic, the
    Doppler reported by :539, and a CFAR statistic that keeps the first step's input power."""
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=8)
    iq2 = synth.make_iq(4000, geo["fs"], [dict(t, phase=0.7) for t in truth], seed=123)   # the block the second step runs on
    e1 = _gpu_engine(b200, geo, 8, use_cfar=use_cfar)
    e2 = _gpu_engine(b200, geo, 8, use_cfar=use_cfar, step2=(4, 125.0))
    assert e2.n_doppler == 4
    e1.set_local_codes(np.arange(8), codes)
    e2.set_local_codes(np.arange(8), codes)
    r1 = e1.dwell(iq, np.arange(8))
    for p in range(8):
        e2.set_step2_center(p, float(r1[p]["doppler_hz"]), r1[p]["input_power"])
    r2 = e2.dwell(iq2, np.arange(8))
    present = {t["prn"]: t for t in truth}
    for p in range(8):
        a = oracle.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"],
                       doppler_step=geo["doppler_step"], fs=geo["fs"], samples_per_chip=geo["samples_per_chip"],
                       samples_per_code=geo["samples_per_code"], use_cfar=use_cfar)
        a.set_local_code(codes[p])
        a.reset()
        w1 = a.dwell(iq)
        w2 = a.dwell_step2(iq2, 4, 125.0, float(w1["doppler_hz"]))
        strong = (p + 1) in present
        _compare(r1[p], w1, strict_idx=strong)
        if int(r1[p]["doppler_hz"]) != w1["doppler_hz"]:
            continue        # a noise-only PRN whose coarse peak moved: the second grids differ by construction
        assert r2[p]["peak"] == pytest.approx(w2["peak"], rel=RTOL)
        assert r2[p]["test_statistic"] == pytest.approx(w2["test_statistic"], rel=RTOL)
        if strong:
            assert int(r2[p]["code_phase"]) == w2["code_phase"] and int(r2[p]["doppler_index"]) == w2["doppler_index"]
            assert int(r2[p]["doppler_hz"]) == w2["doppler_hz"]
            t = present[p + 1]
            # a 1 ms dwell has a 1 kHz main lobe: the fine bin is noise-limited, it only has to stay inside the narrow grid
            assert abs(int(r2[p]["doppler_hz"]) - int(r1[p]["doppler_hz"])) <= 2 * 125 and abs(r2[p]["doppler_hz"] - t["doppler"]) < 500
        a.close()
    with pytest.raises(b200.B200Error):
        e1.set_step2_center(0, 0.0, 1.0)          # not a second-step handle
    e1.close()
    e2.close()


def test_runtime_geometry_kernels_agree_with_fixed_ones():
    """The common dwell lengths have inverse-stage instantiations with compile-time geometry (acq_fast_lower.cu /
    acq_fast_upper.cu: launch_fixed); every other fast-plan length runs the runtime-geometry instantiations of the same
    kernels. B200_NO_FIXED=1 sends the whole suite's lengths through the latter: same parity tests, in a child process
    because the switch is read once per process."""
    import subprocess
    import sys
    env = dict(os.environ, B200_NO_FIXED="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k",
                        "known_answer or config1 or config2 or noncoherent or zero_padded or cshort or cbyte or two_step"], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


_FIXED_CHILD = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tools")
from b200_import import load_package
b = load_package()
out = {}
rng = np.random.default_rng(77)
for N in (4000, 8000, 16000, 25000, 32000, 50000, 100000, 160000):
    for keep in (False, True):
        e = b.PcpsAcquisitionB200(N, N, 500, 250, 4e6, 4, 2, use_cfar=True, keep_grid=keep)
        codes = np.sign(rng.standard_normal((2, N))).astype(np.complex64)
        x = (rng.standard_normal(N) + 1j * rng.standard_normal(N)).astype(np.complex64)
        x += 0.2 * np.roll(codes[0], 1234 % N) * np.exp(2j * np.pi * 250.0 / 4e6 * np.arange(N))
        e.set_local_codes([0, 1], codes)
        r = e.dwell(x, [0, 1])
        if keep:
            r = e.dwell(x[::-1].copy(), [0, 1])          # second dwell: accumulation path
            out["grid%d" % N] = e.read_grid(0)
        out["res%d_%d" % (N, keep)] = r.view(np.uint8).copy()
        e.close()
np.savez(sys.argv[2], **out)
"""


def test_fixed_geometry_kernels_are_bit_identical_to_runtime_ones(tmp_path):
    """Every dwell length with compile-time-geometry instantiations (forward, both inverse stages, records-only and
    persistent-grid modes): result records and accumulated grids must equal those of the runtime-geometry kernels BIT FOR
    BIT — same arithmetic in the same order, only the address computations differ."""
    import subprocess
    import sys
    outs = []
    # B200_ACQ_CLUSTER=0: N = 32000 on the two-kernel path in both runs (its row-resident kernel has no runtime-geometry twin)
    for tag, extra in (("fixed", {"B200_ACQ_CLUSTER": "0"}), ("runtime", {"B200_NO_FIXED": "1", "B200_ACQ_CLUSTER": "0"})):
        path = str(tmp_path / (tag + ".npz"))
        r = subprocess.run([sys.executable, "-c", _FIXED_CHILD, os.path.dirname(os.path.dirname(os.path.abspath(__file__))), path],
                           env=dict(os.environ, **extra), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        outs.append(np.load(path))
    a, c = outs
    assert sorted(a.files) == sorted(c.files) and len(a.files) == 24
    for k in a.files:
        assert np.array_equal(a[k], c[k]), k


@pytest.mark.parametrize("n_shards", [2, 3, 8])
def test_group_of_shards_equals_one_engine(b200, n_shards):
    """gnss_b200_acq_group_*: one search split Doppler-major over several engines (several GPUs in production; here the same
    device listed n times, which exercises the same pack -> device-to-device copy -> decide-on-the-root path). Every field of
    every record must equal the unsharded engine's bit for bit — same rows, same tie rule, same input-power arithmetic —
    including Doppler centres, several non-coherent dwells with a persistent grid, reset, and the gathered grid."""
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=6)
    blocks = [iq] + [synth.make_iq(4000, geo["fs"], [dict(t, phase=0.3 * k) for t in truth], seed=50 + k) for k in range(2)]
    one = _gpu_engine(b200, geo, 6, use_cfar=True, keep_grid=True)
    grp = b200.PcpsAcquisitionGroupB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                        geo["samples_per_chip"], 6, [0] * n_shards, keep_grid=True)
    assert grp.n_doppler == one.n_doppler == 40 and grp.n_shards == n_shards
    for e in (one, grp):
        e.set_local_codes(np.arange(6), codes)
        e.set_doppler_center(1, 750)
        e.set_doppler_center(4, -1125)
    sel = np.array([0, 1, 2, 3, 4, 5])
    for k, blk in enumerate(blocks):
        a = one.dwell(blk, sel)
        g = grp.dwell(blk, sel)
        assert g.tobytes() == a.tobytes(), k
    assert np.array_equal(grp.read_grid(2), one.read_grid(2))
    for e in (one, grp):
        e.reset(3)
    a = one.dwell(blocks[0], [3, 0])
    g = grp.dwell(blocks[0], [3, 0])                      # a slot subset in another order
    assert g.tobytes() == a.tobytes() and list(a["dwell_count"]) == [1, 4]
    with pytest.raises(b200.B200Error):
        b200.PcpsAcquisitionGroupB200(4000, 4000, 5000, 250, 4e6, 4, 2, [7, 99])      # no such devices
    one.close()
    grp.close()
