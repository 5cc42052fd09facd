"""The host C++ mirror (gnss-sdr-new_b200/host): pcps_acquisition_b200 and the B200 correlator classes, driven by
tests/host/host_mirror_test.cc the way the reference's own tests drive pcps_acquisition
(gps_l1_ca_pcps_acquisition_test.cc:290-361: Channel call order, set_state(1), feed the capture, check
Gnss_Synchro against the expected delay / Doppler) and the way dll_pll_veml_tracking drives its correlator object.
Expected values come from the oracle; the C++ driver computes none."""
import importlib.util
import os
import struct
import subprocess

import numpy as np
import pytest

import synth
from conftest import ROOT, has_gpu

GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def driver():
    spec = importlib.util.spec_from_file_location("b200_build", os.path.join(ROOT, "gnss-sdr-new_b200", "build.py"))
    bld = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bld)
    return bld.build_host()[1]


@pytest.mark.skipif(has_gpu(), reason="checks the no-device behaviour")
def test_mirror_refuses_to_run_without_a_device(driver):
    r = subprocess.run([driver, "nodevice"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "mc.init=false acq_threw=true" in r.stdout
    assert "no CPU path" in r.stdout


def test_mirror_keeps_the_reference_method_set():
    """Same member names and argument lists as the reference classes (cpu_multicorrelator_real_codes.h:40-50,
    cpu_multicorrelator.h:40-47, pcps_acquisition.h:99-209)."""
    def text(name):
        return open(os.path.join(ROOT, "gnss-sdr-new_b200", "host", name)).read()
    h = text("b200_multicorrelator_real_codes.h")
    for sig in ("void set_high_dynamics_resampler(bool use_high_dynamics_resampler);",
                "bool init(int max_signal_length_samples, int n_correlators);",
                "bool set_local_code_and_taps(int code_length_chips, const float *local_code_in, float *shifts_chips);",
                "bool set_input_output_vectors(std::complex<float> *corr_out, const std::complex<float> *sig_in);",
                "float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples);",
                "bool free();"):
        assert sig in h, sig
    h = text("b200_multicorrelator.h")
    assert "bool set_local_code_and_taps(int code_length_chips, const std::complex<float> *local_code_in, float *shifts_chips);" in h
    assert "bool Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips, int signal_length_samples);" in h
    h = text("pcps_acquisition_b200.h")
    for name in ("void init();", "void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro);", "void set_local_code(std::complex<float>* code);",
                 "void set_state(int32_t state);", "void set_resampler_latency(uint32_t latency_samples);", "uint32_t mag() const",
                 "void set_active(bool active);", "void set_channel(uint32_t channel)", "void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm)",
                 "void set_threshold(float threshold);", "void set_doppler_max(uint32_t doppler_max);", "void set_doppler_step(uint32_t doppler_step);",
                 "void set_doppler_center(int32_t doppler_center);", "int general_work(int noutput_items,"):
        assert name in h, name


def _write_acq_input(path, fs, sampled_ms, ms_per_code, spc, doppler_max, doppler_step, pfa, threshold, max_dwells, bit_transition,
                     use_cfar, cshort, blocking, codes, stream, chunk, use_fsm, coalesce_us, two_steps=None, dump=None, repeat_events=None,
                     resampler=None):
    with open(path, "wb") as f:
        f.write(struct.pack("<q", int(fs)))
        f.write(struct.pack("<IIIiI", sampled_ms, ms_per_code, spc, doppler_max, doppler_step))
        f.write(struct.pack("<ff", pfa, threshold))
        n_items = stream.size // 2 if cshort else stream.size
        f.write(struct.pack("<IIIIIIIIII", max_dwells, int(bit_transition), int(use_cfar), int(cshort), int(blocking), len(codes), n_items,
                            chunk, int(use_fsm), coalesce_us))
        for c in codes:
            f.write(np.ascontiguousarray(c, dtype=np.complex64).tobytes())
        f.write(np.ascontiguousarray(stream, dtype=np.int16 if cshort else np.complex64).tobytes())
        repeat_flag = repeat_events is not None
        if resampler is not None and repeat_events is None:
            repeat_events = 1
        if two_steps is not None or dump is not None or repeat_events is not None:
            ts = two_steps or (0, 125.0, 0.0)             # (second_nbins, second_doppler_step, pfa2)
            f.write(struct.pack("<IIff", int(two_steps is not None), int(ts[0]), float(ts[1]), float(ts[2])))
            name = (dump[1] if dump else "").encode()     # dump = (dump_channel, dump_filename)
            f.write(struct.pack("<III", int(dump is not None), int(dump[0]) if dump else 0, len(name)))
            f.write(name)
            if repeat_events is not None:             # make_repeat_steps: wait for this many decisions per block
                f.write(struct.pack("<II", int(repeat_flag), int(repeat_events)))
            if resampler is not None:                 # (resampler_ratio, resampler_latency_samples, undecimated fs_in); `fs` above = resampled_fs
                f.write(struct.pack("<IfIq", 1, float(resampler[0]), int(resampler[1]), int(resampler[2])))


def _read_acq_output(path):
    raw = open(path, "rb").read()
    n, = struct.unpack_from("<I", raw, 0)
    out, off = [], 4
    for _ in range(n):
        ev, delay, dop, stamp, stat, thr, nd, nfft, dstep, st2, npos, nneg = struct.unpack_from("<iddQffIIIIII", raw, off)
        off += struct.calcsize("<iddQffIIIIII")
        out.append(dict(event=ev, delay=delay, doppler=dop, stamp=stamp, stat=stat, threshold=thr, n_doppler=nd, fft_size=nfft,
                        doppler_step=dstep, step_two=st2, n_positive=npos, n_negative=nneg))
    return out


def _run(driver, mode, inp, outp):
    r = subprocess.run([driver, mode, str(inp), str(outp)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.gpu
def test_gps_capture_through_the_block_interface(driver, oracle, tmp_path):
    """gps_l1_ca_pcps_acquisition_test.cc ValidationOfResults: PRN 1, 4 Msps, +-5000/100 Hz... here step 250, threshold
    0.001 with the first-vs-second-peak test (pfa = 0): positive acquisition, delay 524 samples, Doppler within 666 Hz of 1680."""
    iq = np.fromfile(os.path.join(GOLD, "GPS_L1_CA_ID_1_Fs_4Msps_2ms.dat"), dtype=np.complex64)
    code = synth.adapter_code("gps_l1_ca", 1, 4e6, 1)[:4000]
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, 4e6, 1, 1, 4, 5000, 250, 0.0, 0.001, 1, False, False, False, True, [code], iq, 1000, True, 0)
    _run(driver, "acq", inp, outp)
    got = _read_acq_output(outp)[0]
    acq = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=5000, doppler_step=250, fs=4e6, samples_per_chip=4,
                     samples_per_code=4000, use_cfar=False)
    acq.set_local_code(code)
    # set_state(1) skips the restart state: the dwell buffer fills from the first item on
    want = acq.dwell(iq[:4000])
    assert got["event"] == 1
    assert got["n_doppler"] == 40 and got["fft_size"] == 4000
    assert got["delay"] == float(np.fmod(np.float32(want["code_phase"]), np.float32(4000)))
    assert got["doppler"] == float(want["doppler_hz"])
    assert got["stat"] == pytest.approx(want["test_statistic"], rel=1e-4)
    assert got["stamp"] == 4000
    # the reference test's own acceptance window
    delay_chips = abs(got["delay"] - 524) * 1023.0 / 4000.0
    assert delay_chips <= 0.5 and abs(got["doppler"] - 1680) <= 666


@pytest.mark.gpu
def test_block_behind_the_acquisition_resampler(driver, oracle, tmp_path):
    """GNSS-SDR.use_acquisition_resampler: the block receives the DECIMATED stream (resampled_fs = 5 MS/s) while
    Acq_Conf::fs_in stays the undecimated 25 MS/s (acq_conf.cc:93-111). The carrier wipe-off must be generated at
    resampled_fs (update_local_carrier, pcps_acquisition.cc:236-243), and the answer goes back to the input rate:
    Acq_delay_samples = delay * ratio - latency, Acq_samplestamp_samples = rint(stamp * ratio) (:699-705)."""
    fs_in, acq_fs, ratio, latency = 25000000, 5000000, 5.0, 37
    N = 5000
    truth = [dict(system="gps_l1_ca", prn=9, delay=1403, doppler=-1830.0, cn0=50.0)]
    iq = synth.make_iq(2 * N, acq_fs, truth, seed=18)          # what the decimating FIR hands to the block
    code = synth.adapter_code("gps_l1_ca", 9, acq_fs, 1)[:N]
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, acq_fs, 1, 1, 5, 5000, 250, 0.001, 0.0, 1, False, True, False, True, [code], iq, 1000, True, 0,
                     resampler=(ratio, latency, fs_in))
    _run(driver, "acq", inp, outp)
    got = _read_acq_output(outp)[0]
    a = oracle.acq(fft_size=N, consumed_samples=N, doppler_max=5000, doppler_step=250, fs=acq_fs, samples_per_chip=5, samples_per_code=float(N))
    a.set_local_code(code)
    want = a.dwell(iq[:N])
    assert got["event"] == 1 and got["fft_size"] == N
    assert got["doppler"] == float(want["doppler_hz"]) and abs(got["doppler"] - truth[0]["doppler"]) <= 250
    assert got["stat"] == pytest.approx(want["test_statistic"], rel=1e-4)
    assert got["delay"] == float(np.fmod(np.float32(want["code_phase"]), np.float32(N))) * ratio - latency
    assert got["stamp"] == int(np.rint(N * ratio))
    a.close()


@pytest.mark.gpu
def test_block_interface_at_a_sample_rate_with_the_prime_31(driver, oracle, tmp_path):
    """4.092 MS/s front end: the block derives d_fft_size = 4092 = 4*3*11*31 from fs as the adapter does; decision,
    delay, Doppler and statistic equal the oracle's."""
    geo, codes, iq, truth = synth.config_scene("F4092", n_prn=1)
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, geo["fs"], 1, 1, geo["samples_per_chip"], geo["doppler_max"], geo["doppler_step"], 0.001, 0.0, 1, False, True, False,
                     True, [codes[0]], np.concatenate([iq, iq]), 1023, True, 0)
    _run(driver, "acq", inp, outp)
    got = _read_acq_output(outp)[0]
    acq = oracle.acq(fft_size=4092, consumed_samples=4092, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"], fs=geo["fs"],
                     samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"], use_cfar=True)
    acq.set_local_code(codes[0])
    want = acq.dwell(iq)
    assert got["fft_size"] == 4092 and got["n_doppler"] == 40
    assert got["event"] == 1 and truth[0]["prn"] == 1
    assert got["delay"] == float(np.fmod(np.float32(want["code_phase"]), np.float32(geo["samples_per_code"])))
    assert got["doppler"] == float(want["doppler_hz"])
    assert got["stat"] == pytest.approx(want["test_statistic"], rel=1e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("cshort,blocking,coalesce_us", [(False, True, 0), (False, True, 20000), (True, False, 0)])
def test_many_blocks_share_one_engine(driver, oracle, tmp_path, cshort, blocking, coalesce_us):
    """Six channels (PRN 1..6 of config 1's scene, three of them present) run concurrently, one scheduler thread each;
    CFAR statistic with pfa-derived threshold; decisions and Gnss_Synchro fields must equal the oracle's per channel."""
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=6)
    stream = np.concatenate([iq, iq])
    scale = 1500.0
    s_in = synth.to_cshort(stream, scale) if cshort else stream
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, geo["fs"], 1, 1, geo["samples_per_chip"], geo["doppler_max"], geo["doppler_step"], 0.01, 0.0, 1, False, True,
                     cshort, blocking, list(codes), s_in, 4000, False, coalesce_us)
    _run(driver, "acq", inp, outp)
    got = _read_acq_output(outp)
    thr = oracle.threshold(0.01, 4000, 40, 1, False)
    present = {t["prn"] for t in truth}
    n_pos = 0
    for p in range(6):
        acq = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"],
                         fs=geo["fs"], samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"])
        acq.set_local_code(codes[p])
        block = s_in.reshape(-1, 2)[:4000].reshape(-1) if cshort else stream[:4000]
        want = acq.dwell(block)
        g = got[p]
        assert g["threshold"] == pytest.approx(thr, rel=2e-6)
        assert g["stat"] == pytest.approx(want["test_statistic"], rel=1e-4)
        assert g["event"] == (1 if want["test_statistic"] > thr else 2)
        assert g["delay"] == float(want["code_phase"]) and g["doppler"] == float(want["doppler_hz"]) and g["stamp"] == 4000
        n_pos += g["event"] == 1
        if (p + 1) in present:
            assert g["event"] == 1
    assert n_pos >= len([t for t in truth if t["prn"] <= 6])


def _write_mc_input(path, mode, high_dyn, coalesce_us, codes, shifts, sigs, args):
    n_ch, n_epochs, n_samples = sigs.shape
    with open(path, "wb") as f:
        code_len = codes.shape[1] // (2 if mode >= 2 else 1)
        f.write(struct.pack("<IIIIIIII", n_ch, shifts.shape[1], n_samples, code_len, n_epochs, mode, int(high_dyn), coalesce_us))
        for c in range(n_ch):
            f.write(codes[c].astype(np.float32).tobytes())
            f.write(shifts[c].astype(np.float32).tobytes())
            f.write(sigs[c].astype(np.complex64).tobytes())
            f.write(args[c].astype(np.float32).tobytes())


@pytest.mark.gpu
@pytest.mark.parametrize("mode,high_dyn,coalesce_us", [(0, False, 0), (0, True, 0), (1, True, 0), (2, False, 0), (0, False, 20000)])
def test_correlator_objects_match_oracle(driver, oracle, tmp_path, mode, high_dyn, coalesce_us):
    """One correlator object per channel, each called for 3 code periods from its own thread (as tracking blocks do)."""
    rng = np.random.default_rng(11 + mode)
    n_ch, n_ep, n, L, T = 5, 3, 4000, 1023, 3
    chips = np.stack([synth.chips("gps_l1_ca", p + 1) for p in range(n_ch)]).astype(np.float32)
    if mode == 2:
        cc = (chips + 1j * np.roll(chips, 7, axis=1)).astype(np.complex64)
        codes = cc.view(np.float32).reshape(n_ch, 2 * L)
    else:
        codes = chips
    shifts = np.tile(np.array([-0.5, 0.0, 0.5], np.float32), (n_ch, 1)) + rng.uniform(-0.01, 0.01, (n_ch, 1)).astype(np.float32)
    sigs = (rng.standard_normal((n_ch, n_ep, n)) + 1j * rng.standard_normal((n_ch, n_ep, n))).astype(np.complex64)
    args = np.zeros((n_ch, n_ep, 6), dtype=np.float32)
    args[..., 0] = rng.uniform(0, 6.28, (n_ch, n_ep))
    args[..., 1] = rng.uniform(-0.01, 0.01, (n_ch, n_ep))
    args[..., 2] = rng.uniform(-1e-9, 1e-9, (n_ch, n_ep)) if high_dyn else 0
    args[..., 3] = rng.uniform(0, 1, (n_ch, n_ep))
    args[..., 4] = 1.023e6 / 4e6 * (1 + rng.uniform(-1e-5, 1e-5, (n_ch, n_ep)))
    args[..., 5] = rng.uniform(-1e-10, 1e-10, (n_ch, n_ep)) if high_dyn else 0
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_mc_input(inp, mode, high_dyn, coalesce_us, codes, shifts, sigs, args)
    _run(driver, "mc", inp, outp)
    raw = open(outp, "rb").read()
    failures, batches, calls = struct.unpack_from("<IQQ", raw, 0)
    got = np.frombuffer(raw, dtype=np.complex64, offset=20).reshape(n_ch, n_ep, T)
    assert failures == 0 and calls == n_ch * n_ep
    if coalesce_us:
        assert batches < calls   # concurrent calls were combined into fewer engine launches
    for c in range(n_ch):
        for e in range(n_ep):
            a = args[c, e]
            if mode == 2:
                want = oracle.mc_cplx(sigs[c, e], cc[c], shifts[c], a[0], a[1], a[3], a[4])
            else:
                want = oracle.mc_real(sigs[c, e], chips[c], shifts[c], a[0], a[1], a[2], a[3], a[4], a[5], high_dyn=high_dyn,
                                      six_arg=(mode == 1))
            scale = np.abs(want).max() + np.sqrt(n)
            assert np.abs(got[c, e] - want).max() <= 1e-4 * scale


@pytest.mark.gpu
def test_real_and_complex_correlators_share_the_pool(driver, oracle, tmp_path):
    """B200_Multicorrelator_Real_Codes and B200_Multicorrelator objects with EQUAL tap counts and flags call at the same
    time through the one pool (e.g. GPS L1 dll_pll_veml next to a complex-code tracker). The engine refuses a batch that
    mixes code types, so the pool must not coalesce them into one: every call succeeds and matches the oracle."""
    rng = np.random.default_rng(5)
    n_ch, n_ep, n, L, T = 6, 3, 4000, 1023, 3
    chips = np.stack([synth.chips("gps_l1_ca", p + 1) for p in range(n_ch)]).astype(np.float32)
    cc = (chips + 1j * np.roll(chips, 7, axis=1)).astype(np.complex64)
    codes = np.empty((n_ch, 2 * L), np.float32)
    for c in range(n_ch):
        codes[c] = cc[c].view(np.float32) if c & 1 else np.concatenate([chips[c], np.zeros(L, np.float32)])
    shifts = np.tile(np.array([-0.5, 0.0, 0.5], np.float32), (n_ch, 1))
    sigs = (rng.standard_normal((n_ch, n_ep, n)) + 1j * rng.standard_normal((n_ch, n_ep, n))).astype(np.complex64)
    args = np.zeros((n_ch, n_ep, 6), dtype=np.float32)
    args[..., 0] = rng.uniform(0, 6.28, (n_ch, n_ep))
    args[..., 1] = rng.uniform(-0.01, 0.01, (n_ch, n_ep))
    args[..., 3] = rng.uniform(0, 1, (n_ch, n_ep))
    args[..., 4] = 1.023e6 / 4e6 * (1 + rng.uniform(-1e-5, 1e-5, (n_ch, n_ep)))
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_mc_input(inp, 3, False, 20000, codes, shifts, sigs, args)
    _run(driver, "mc", inp, outp)
    raw = open(outp, "rb").read()
    failures, batches, calls = struct.unpack_from("<IQQ", raw, 0)
    got = np.frombuffer(raw, dtype=np.complex64, offset=20).reshape(n_ch, n_ep, T)
    assert failures == 0 and calls == n_ch * n_ep and batches < calls
    for c in range(n_ch):
        for e in range(n_ep):
            a = args[c, e]
            want = (oracle.mc_cplx(sigs[c, e], cc[c], shifts[c], a[0], a[1], a[3], a[4]) if c & 1 else
                    oracle.mc_real(sigs[c, e], chips[c], shifts[c], a[0], a[1], 0.0, a[3], a[4], 0.0))
            assert np.abs(got[c, e] - want).max() <= 1e-4 * (np.abs(want).max() + np.sqrt(n))


@pytest.mark.gpu
@pytest.mark.parametrize("blocking,coalesce_us", [(True, 0), (True, 20000)])
def test_two_step_acquisition_through_the_block_interface(driver, oracle, tmp_path, blocking, coalesce_us):
    """make_two_steps (pcps_acquisition.cc:786-803, 823-828, 937-943): a positive coarse search re-arms the block around
    Acq_doppler_hz, the second search runs num_doppler_bins_step2 bins of doppler_step2 on a LATER block of the stream
    and only its decision is reported. Six channels share one bank (and its second-step engine). Only the blocking mode
    (the reference's default) is deterministic: with a worker thread, a general_work call that wins the lock between the
    thread's creation and its first statement re-reads Acq_doppler_hz — zeroed by the restart state — as the narrow
    grid's centre (:937-943 run whenever d_worker_active). The mirror keeps that order; the test stays away from it."""
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=6, n_dwells=6)
    stream = iq.reshape(-1)
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, geo["fs"], 1, 1, geo["samples_per_chip"], geo["doppler_max"], geo["doppler_step"], 0.01, 0.0, 1, False, True,
                     False, blocking, list(codes), stream, 4000, False, coalesce_us, two_steps=(4, 125.0, 0.02))
    _run(driver, "acq", inp, outp)
    got = _read_acq_output(outp)
    thr1 = oracle.threshold(0.01, 4000, 40, 1, False)
    thr2 = oracle.threshold(0.02, 4000, 4, 1, False)
    n_two = 0
    for p in range(6):
        acq = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"],
                         fs=geo["fs"], samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"])
        acq.set_local_code(codes[p])
        w1 = acq.dwell(stream[:4000])
        g = got[p]
        assert g["step_two"] == 0 and g["threshold"] == pytest.approx(thr1, rel=1e-6)   # both exits restore the first threshold
        if w1["test_statistic"] > thr1:
            # call sequence: fill, switch, search (stamp 4000), idle call consumes a block and re-arms, restart state
            # consumes another, fill takes stream[12000:16000], switch, search (stamp 16000)
            w2 = acq.dwell_step2(stream[12000:16000], 4, 125.0, float(w1["doppler_hz"]))
            n_two += 1
            assert g["stamp"] == 16000
            assert g["stat"] == pytest.approx(w2["test_statistic"], rel=1e-4)
            assert g["doppler"] == float(w2["doppler_hz"]) and g["delay"] == float(w2["code_phase"] % 4000)
            assert g["doppler_step"] == 125
            assert g["event"] == (1 if w2["test_statistic"] > thr2 else 2)
        else:
            assert g["event"] == 2 and g["stamp"] == 4000
            assert g["stat"] == pytest.approx(w1["test_statistic"], rel=1e-4)
        acq.close()
    assert n_two >= 1


def test_mat_writer_is_read_by_scipy(driver, tmp_path):
    """b200_mat_writer: Level-5 MAT container with the classes the acquisition dump uses (pcps_acquisition.cc:425-500)."""
    from scipy.io import loadmat
    out = tmp_path / "t.mat"
    r = subprocess.run([driver, "mat", str(out)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    m = loadmat(str(out))
    assert m["acq_grid"].dtype == np.float32 and m["acq_grid"].shape == (7, 3)
    assert np.array_equal(m["acq_grid"], (0.5 * np.arange(21, dtype=np.float32)).reshape(3, 7).T)
    assert m["doppler_max"].dtype == np.int32 and int(m["doppler_max"][0, 0]) == -5000
    assert m["PRN"].dtype == np.uint32 and int(m["PRN"][0, 0]) == 4000000000
    assert m["sample_counter"].dtype == np.uint64 and int(m["sample_counter"][0, 0]) == 0x123456789ABCDEF
    assert m["threshold"].dtype == np.float32 and float(m["threshold"][0, 0]) == 2.5
    assert m["x"].dtype == np.float64 and float(m["x"][0, 0]) == 1.0 / 3.0
    assert np.array_equal(m["odd_length_name_"], np.array([[1, -2, 3, -4, 5]], dtype=np.int32))
    # the same file through the HDF5-free parser behind tools/acq_dump_reader.py (matio's Mat_Open / Mat_VarRead walk, restated)
    import acq_dump_reader
    v = acq_dump_reader.mat_open(str(out))
    assert v["acq_grid"]["rank"] == 2 and tuple(v["acq_grid"]["dims"]) == (7, 3) and v["acq_grid"]["data_type"] == "f4"
    assert np.array_equal(v["acq_grid"]["data"], 0.5 * np.arange(21, dtype=np.float32))          # column-major, as Mat_VarRead hands it over
    assert v["sample_counter"]["data_type"] == "u8" and int(v["sample_counter"]["data"][0]) == 0x123456789ABCDEF
    assert v["x"]["data_type"] == "f8" and v["PRN"]["data_type"] == "u4" and v["doppler_max"]["data_type"] == "i4"
    with pytest.raises(acq_dump_reader.MatError):
        acq_dump_reader.read_binary_acq(str(out), 7, 4)


@pytest.mark.gpu
@pytest.mark.parametrize("two_steps", [False, True])
def test_acquisition_dump_file(driver, oracle, tmp_path, two_steps):
    """dump = true (pcps_acquisition.cc:135-165, 408-508, 875-878): channel `dump_channel` writes
    <dump_filename>_<System>_<Signal>_ch_<channel>_<n>_sat_<PRN>.mat with the magnitude grid (effective size x Doppler bins,
    column per bin) and the scalars of the decision; with make_two_steps also the narrow grid and its geometry."""
    from scipy.io import loadmat
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=4, n_dwells=6)
    stream = iq.reshape(-1)
    def first_stat(p):
        a = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"],
                       fs=geo["fs"], samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"])
        a.set_local_code(codes[p])
        v = a.dwell(stream[:4000])["test_statistic"]
        a.close()
        return v
    strong = int(np.argmax([first_stat(p) for p in range(4)]))      # the channel that certainly reaches the second step
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, geo["fs"], 1, 1, geo["samples_per_chip"], geo["doppler_max"], geo["doppler_step"], 0.01, 0.0, 1, False, True,
                     False, True, list(codes), stream, 4000, False, 0, two_steps=(4, 125.0, 0.02) if two_steps else None,
                     dump=(strong, str(tmp_path / "dumps" / "acq_dump.dat")))
    _run(driver, "acq", inp, outp)
    got = _read_acq_output(outp)[strong]
    files = sorted(os.listdir(tmp_path / "dumps"))
    assert files == ["acq_dump_G_1C_ch_%d_1_sat_%d.mat" % (strong, strong + 1)]     # only the dump channel writes, extension dropped
    m = loadmat(str(tmp_path / "dumps" / files[0]))
    acq = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"],
                     fs=geo["fs"], samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"])
    acq.set_local_code(codes[strong])
    w1 = acq.dwell(stream[:4000])
    assert m["acq_grid"].dtype == np.float32 and m["acq_grid"].shape == (4000, 40)
    np.testing.assert_allclose(m["acq_grid"].T, acq.grid()[:, :4000], rtol=2e-4, atol=2e-4 * w1["peak"])
    assert int(m["doppler_max"][0, 0]) == geo["doppler_max"] and int(m["doppler_step"][0, 0]) == geo["doppler_step"]
    assert int(m["PRN"][0, 0]) == strong + 1 and int(m["d_positive_acq"][0, 0]) == int(got["event"] == 1) and int(m["num_dwells"][0, 0]) == 1
    assert w1["test_statistic"] > got["threshold"]
    assert float(m["input_power"][0, 0]) == pytest.approx(w1["input_power"], rel=1e-4)
    assert float(m["threshold"][0, 0]) == pytest.approx(got["threshold"], rel=1e-6)
    assert float(m["acq_doppler_hz"][0, 0]) == got["doppler"] and float(m["acq_delay_samples"][0, 0]) == got["delay"]
    assert float(m["test_statistic"][0, 0]) == pytest.approx(got["stat"], rel=1e-6)
    assert m["sample_counter"].dtype == np.uint64 and int(m["sample_counter"][0, 0]) == got["stamp"]
    # ... and what the reference's own reader takes from the file (Acquisition_Dump_Reader::read_binary_acq, restated in
    # tools/acq_dump_reader.py over an HDF5-free Level-5 parser: its rank / dims / single-precision checks and its pointer casts)
    import acq_dump_reader
    rd = acq_dump_reader.read_binary_acq(str(tmp_path / "dumps" / files[0]), 4000, 40)
    assert rd["doppler_max"] == geo["doppler_max"] and rd["doppler_step"] == geo["doppler_step"] and rd["PRN"] == strong + 1
    assert rd["positive_acq"] == int(got["event"] == 1) and rd["num_dwells"] == 1 and rd["sample_counter"] == got["stamp"]
    assert rd["acq_doppler_hz"] == got["doppler"] and rd["acq_delay_samples"] == got["delay"]
    assert rd["test_statistic"] == pytest.approx(got["stat"], rel=1e-6) and rd["threshold"] == pytest.approx(got["threshold"], rel=1e-6)
    k_dop, k_smp = np.unravel_index(np.argmax(rd["mag"]), rd["mag"].shape)
    assert (k_dop, k_smp) == (w1["doppler_index"], w1["code_phase"])                 # mag[doppler][sample], as the reader's plots index it
    with pytest.raises(acq_dump_reader.MatError):
        acq_dump_reader.read_binary_acq(str(tmp_path / "dumps" / files[0]), 4000, 41)  # "dimension matrix error"
    if two_steps:
        w2 = acq.dwell_step2(stream[12000:16000], 4, 125.0, float(w1["doppler_hz"]))
        assert m["acq_grid_narrow"].shape == (4000, 4)
        assert float(m["acq_grid_narrow"].max()) == pytest.approx(w2["peak"], rel=2e-4)
        assert np.unravel_index(np.argmax(m["acq_grid_narrow"].T), (4, 4000)) == (w2["doppler_index"], w2["code_phase"])
        assert float(m["doppler_step_narrow"][0, 0]) == 125.0
        assert float(m["doppler_grid_narrow_min"][0, 0]) == float(w1["doppler_hz"]) - 2 * 125.0
    else:
        assert "acq_grid_narrow" not in m


# one record of dll_pll_veml_tracking::log_data() (dll_pll_veml_tracking.cc:1445-1502), in file order
TRK_DUMP_DTYPE = np.dtype([("abs_VE", "<f4"), ("abs_E", "<f4"), ("abs_P", "<f4"), ("abs_L", "<f4"), ("abs_VL", "<f4"), ("Prompt_I", "<f4"),
                           ("Prompt_Q", "<f4"), ("PRN_start_sample_count", "<u8"), ("acc_carrier_phase_rad", "<f4"),
                           ("carrier_doppler_hz", "<f4"), ("carrier_doppler_rate_hz", "<f4"), ("code_freq_chips", "<f4"),
                           ("code_freq_rate_chips", "<f4"), ("carr_error_hz", "<f4"), ("carr_error_filt_hz", "<f4"),
                           ("code_error_chips", "<f4"), ("code_error_filt_chips", "<f4"), ("CN0_SNV_dB_Hz", "<f4"),
                           ("carrier_lock_test", "<f4"), ("aux1", "<f4"), ("aux2", "<f8"), ("PRN", "<u4"),
                           ("acq_code_phase_samples", "<f4"), ("acq_carrier_doppler_hz", "<f4"), ("EVM", "<f4")])


@pytest.mark.parametrize("veml", [False, True])
def test_tracking_dump_files(driver, oracle, oracle_mod, tmp_path, veml):
    """B200_Tracking_Dump: epoch records of the tracking loop (here the oracle's, same layout as gnss_b200_trk_epoch) become
    the reference's <dump_filename><channel>.dat (108-byte records, log_data :1403-1509) and .mat (save_matfile :1511-1735:
    one 1 x num_epoch variable per field, same names and classes)."""
    from scipy.io import loadmat
    assert TRK_DUMP_DTYPE.itemsize == 108
    system = "gal_e1b" if veml else "gps_l1_ca"
    chip_rate, code_len, _ = synth.SYSTEMS[system]
    fs = 4e6
    iq, truth, carrier = synth.tracking_stream(system, [7], fs, 120, seed=5, cn0=47.0, dopplers=[1500.0], delays=[321.6],
                                               code_samples_per_chip=2 if veml else 1)
    cfg = oracle_mod.TrkCfg(fs_in=fs, code_chip_rate=chip_rate, signal_carrier_freq=carrier, code_period_s=code_len / chip_rate,
                            code_length_chips=code_len, code_samples_per_chip=2 if veml else 1, vector_length=int(round(fs * code_len / chip_rate)),
                            veml=int(veml), early_late_space_chips=0.15 if veml else 0.25, very_early_late_space_chips=0.5, slope=1.0, y_intercept=1.0,
                            pll_bw_hz=35.0, dll_bw_hz=2.0, fll_bw_hz=35.0, pll_filter_order=3, dll_filter_order=2, enable_fll_pull_in=1,
                            enable_fll_steady_state=0, carrier_aiding=1, pull_in_time_s=0, cn0_samples=20, cn0_min=25, max_code_lock_fail=50,
                            max_carrier_lock_fail=5000, carrier_lock_th=0.7, cn0_smoother_alpha=0.002, carrier_lock_test_smoother_alpha=0.002,
                            cn0_smoother_samples=200, carrier_lock_test_smoother_samples=25, max_channels=1, device=0)
    trk = oracle.trk(cfg)
    trk.set_code(synth.tracking_code(system, 7))
    trk.start(acq_delay_samples=322.0, acq_doppler_hz=1460.0, acq_samplestamp=0, now_sample=2000)
    ep, done = trk.run(iq, 0, 40)
    trk.close()
    ep = ep[:done + 1]                                            # the record after the last one is not valid: it must be skipped
    assert (ep["flags"][:done] & 1).all() and done >= 20
    inp = tmp_path / "ep.bin"
    with open(inp, "wb") as f:
        f.write(struct.pack("<IIIIdd", len(ep), int(veml), 3, 7, 322.0, 1460.0))
        f.write(ep.tobytes())
    r = subprocess.run([driver, "trkdump", str(inp), str(tmp_path / "trk" / "track_ch_.bin")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    assert sorted(os.listdir(tmp_path / "trk")) == ["track_ch_3.dat", "track_ch_3.mat"]
    rec = np.fromfile(tmp_path / "trk" / "track_ch_3.dat", dtype=TRK_DUMP_DTYPE)
    ok = ep[:done]
    assert len(rec) == done
    z = ok["corr"].reshape(-1, 5, 2)
    mag = np.abs(z[..., 0].astype(np.complex64) + 1j * z[..., 1].astype(np.complex64)).astype(np.float32)
    for k, name in enumerate(["abs_VE", "abs_E", "abs_P", "abs_L", "abs_VL"]):
        want = mag[:, k] if (veml or k in (1, 2, 3)) else np.zeros(done, np.float32)
        np.testing.assert_allclose(rec[name], want, rtol=1e-6)
    assert np.array_equal(rec["Prompt_I"], ok["corr"][:, 4]) and np.array_equal(rec["Prompt_Q"], ok["corr"][:, 5])
    stamp = ok["sample_counter"] + ok["prn_length_samples"].astype(np.uint64)
    assert np.array_equal(rec["PRN_start_sample_count"], stamp) and np.array_equal(rec["aux2"], stamp.astype(np.float64))
    for name, src in [("acc_carrier_phase_rad", "acc_carrier_phase_rad"), ("carrier_doppler_hz", "carrier_doppler_hz"),
                      ("code_freq_chips", "code_freq_chips"), ("carr_error_hz", "carr_error_hz"), ("carr_error_filt_hz", "carr_error_filt_hz"),
                      ("code_error_chips", "code_error_chips"), ("code_error_filt_chips", "code_error_filt_chips"), ("CN0_SNV_dB_Hz", "cn0_db_hz"),
                      ("carrier_lock_test", "carrier_lock_test"), ("aux1", "rem_code_phase_samples"), ("EVM", "evm")]:
        assert np.array_equal(rec[name], ok[src].astype(np.float32)), name
    assert (rec["carrier_doppler_rate_hz"] == 0).all() and (rec["code_freq_rate_chips"] == 0).all()
    assert (rec["PRN"] == 7).all() and (rec["acq_code_phase_samples"] == 322.0).all() and (rec["acq_carrier_doppler_hz"] == 1460.0).all()
    m = loadmat(str(tmp_path / "trk" / "track_ch_3.mat"))
    for name in TRK_DUMP_DTYPE.names:
        assert m[name].shape == (1, done) and m[name].dtype == TRK_DUMP_DTYPE[name].newbyteorder("="), name
        assert np.array_equal(m[name][0], rec[name]), name


@pytest.mark.gpu
def test_repeat_mode_keeps_reporting(driver, oracle, tmp_path):
    """The fork's make_repeat_steps (pcps_acquisition.cc:324, 360-368, 944-947, 956): after the first positive acquisition
    the block re-arms itself and searches again and again without clearing Gnss_Synchro in between; only the first
    detection writes a dump. Every search is one block of the stream; blocks in between are consumed by the idle and
    restart calls, so the k-th search sees stream[12000 k : 12000 k + 4000]."""
    from scipy.io import loadmat
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=4, n_dwells=10)
    stream = iq.reshape(-1)

    def search(p, k):
        a = oracle.acq(fft_size=4000, consumed_samples=4000, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"],
                       fs=geo["fs"], samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"])
        a.set_local_code(codes[p])
        r = a.dwell(stream[12000 * k:12000 * k + 4000])
        a.close()
        return r
    thr = oracle.threshold(0.01, 4000, 40, 1, False)
    strong = int(np.argmax([search(p, 0)["test_statistic"] for p in range(4)]))
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    _write_acq_input(inp, geo["fs"], 1, 1, geo["samples_per_chip"], geo["doppler_max"], geo["doppler_step"], 0.01, 0.0, 1, False, True,
                     False, True, [codes[strong]], stream, 4000, False, 0, dump=(0, str(tmp_path / "rep" / "acq")), repeat_events=3)
    _run(driver, "acq", inp, outp)
    g = _read_acq_output(outp)[0]
    want = [search(strong, k) for k in range(3)]
    assert want[0]["test_statistic"] > thr
    assert g["n_positive"] + g["n_negative"] == 3
    assert g["n_positive"] == sum(w["test_statistic"] > thr for w in want)
    assert g["stamp"] == 4000 + 12000 * 2 and g["stat"] == pytest.approx(want[2]["test_statistic"], rel=1e-4)
    assert g["doppler"] == float(want[2]["doppler_hz"]) and g["delay"] == float(want[2]["code_phase"] % 4000)
    # the dump counter moves with every finished search, but d_positive_acq is only raised by the first detection
    files = sorted(os.listdir(tmp_path / "rep"))
    assert files[0] == "acq_G_1C_ch_0_1_sat_1.mat" and int(loadmat(str(tmp_path / "rep" / files[0]))["d_positive_acq"][0, 0]) == 1
    for f in files[1:]:
        assert int(loadmat(str(tmp_path / "rep" / f))["d_positive_acq"][0, 0]) == 0


def _symsync_expected(prompts, flags, secondary, code, spb, data_code):
    """dll_pll_veml_tracking.cc:1890-1946 (collect + acquire_secondary), :1290-1400 and :2041-2066 (accumulate + emit), restated
    in Python for the prompt only; binary32 sums in arrival order."""
    out, buf, locked, cur_d, acc, flag180 = [], [], False, 0, np.complex64(0), False
    for i, (p, fl) in enumerate(zip(prompts, flags)):
        if not fl & 1:
            continue
        if fl & 4:
            buf, locked, cur_d, acc, flag180 = [], False, 0, np.complex64(0), False
            continue
        if not locked:
            if fl & 2:
                continue
            if secondary or spb > 1:
                buf = (buf + [p])[-len(code):]
                if len(buf) == len(code):
                    corr = sum((1 if c == "0" else -1) if b.real < 0 else (-1 if c == "0" else 1) for b, c in zip(buf, code))
                    if abs(corr) == len(code):
                        locked, flag180, buf, cur_d, acc = True, corr < 0, [], 0, np.complex64(0)
            else:
                locked = True
            continue
        if spb > 1:
            if data_code:
                acc = np.complex64(acc + p) if data_code[cur_d] == "0" else np.complex64(acc - p)
                cur_d = (cur_d + 1) % len(data_code)
            else:
                acc = np.complex64(acc + p)
                cur_d = (cur_d + 1) % spb
        else:
            acc = p
        if cur_d == 0:
            out.append((i, float(acc.real), float(acc.imag), flag180))
            acc = np.complex64(0)
    return out, locked


GPS_CA_PREAMBLE_SYMBOLS_STR = "".join(c * 20 for c in "10001011")   # GPS_L1_CA.h:73
BEIDOU_B1I_SECONDARY_CODE_STR = "00000100110101001110"              # Beidou_B1I.h


@pytest.mark.parametrize("case", ["gps_l1_ca", "gps_l1_ca_inverted", "bds_b1i_nh", "one_symbol_per_period"])
def test_symbol_synchronisation_on_epoch_records(driver, b200, tmp_path, case):
    """B200_Symbol_Sync: bit / secondary-code lock and per-bit prompt accumulation from the device loop's records."""
    rng = np.random.default_rng(11)
    if case.startswith("gps_l1_ca"):
        secondary, code, spb, data_code = False, GPS_CA_PREAMBLE_SYMBOLS_STR, 20, ""
        bits = list(rng.integers(0, 2, 7)) + [1, 0, 0, 0, 1, 0, 1, 1] + list(rng.integers(0, 2, 12))
        signs = np.repeat(np.array(bits) * 2 - 1, 20)[7:]             # the stream starts in the middle of a bit
        if case.endswith("inverted"):
            signs = -signs                                             # Costas loop locked half a cycle off
    elif case == "bds_b1i_nh":
        secondary, code, spb, data_code = True, BEIDOU_B1I_SECONDARY_CODE_STR, 20, BEIDOU_B1I_SECONDARY_CODE_STR
        nh = np.array([1 if c == "1" else -1 for c in code])
        bits = rng.integers(0, 2, 14) * 2 - 1
        signs = (np.repeat(bits, 20) * np.tile(nh, 14))[13:]
    else:
        secondary, code, spb, data_code = False, "", 1, ""
        signs = rng.integers(0, 2, 60) * 2 - 1
    n = signs.size
    ep = np.zeros(n, dtype=b200.TRK_EPOCH_DTYPE)
    ep["corr"][:, 4] = (1000.0 * signs + rng.standard_normal(n) * 40).astype(np.float32)
    ep["corr"][:, 5] = (rng.standard_normal(n) * 40).astype(np.float32)
    ep["sample_counter"] = 5000 + 4000 * np.arange(n)
    ep["prn_length_samples"] = 4000
    ep["flags"] = 1
    ep["flags"][:9] |= 2                               # pull-in transitory: nothing collected
    ep["flags"][n - 30] |= 4                           # a loss of lock near the end: back to the search
    ep["flags"][n - 12] = 0                            # an unused record
    prompts = (ep["corr"][:, 4] + 1j * ep["corr"][:, 5]).astype(np.complex64)
    want, want_locked = _symsync_expected(prompts, ep["flags"], secondary, code, spb, data_code)
    inp, outp = tmp_path / "s.bin", tmp_path / "s.out"
    with open(inp, "wb") as f:
        f.write(struct.pack("<IIIII", n, int(secondary), len(code), spb, len(data_code)))
        f.write(code.encode() + data_code.encode())
        f.write(ep.tobytes())
    r = subprocess.run([driver, "symsync", str(inp), str(outp)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = open(outp, "rb").read()
    cnt, = struct.unpack_from("<I", raw, 0)
    got = [struct.unpack_from("<IddQI", raw, 4 + k * struct.calcsize("<IddQI")) for k in range(cnt)]
    locked, _ = struct.unpack_from("<II", raw, 4 + cnt * struct.calcsize("<IddQI"))
    assert cnt == len(want) and cnt >= (3 if spb > 1 else 40)
    for g, w in zip(got, want):
        assert g[0] == w[0] and g[1] == w[1] and g[2] == w[2] and bool(g[4]) == w[3]
        assert g[3] == 5000 + 4000 * g[0]
    assert bool(locked) == want_locked
    if case.startswith("gps_l1_ca"):
        # the first symbols after the preamble are whole bits: 20 prompts of one sign (inverted stream: flag set, sign kept)
        assert all(abs(abs(g[1]) - 20000.0) < 1500.0 for g in got[:3]) and bool(got[0][4]) == case.endswith("inverted")


# ------------------------------------------------------------------------------------------------------------------------
# integration/: the adapters, the gr::block shell and the factory arms a maintainer adds to the GNSS-SDR tree
# ------------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def adapter_driver():
    spec = importlib.util.spec_from_file_location("b200_build", os.path.join(ROOT, "gnss-sdr-new_b200", "build.py"))
    bld = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bld)
    return bld.build_integration()


def _write_adapter_input(path, fs, doppler_max, doppler_step, pfa, coherent_ms, which, prn, rf_channel, code, stream):
    with open(path, "wb") as f:
        f.write(struct.pack("<qiIfIIII", int(fs), int(doppler_max), int(doppler_step), float(pfa), int(coherent_ms), int(which), int(prn), int(rf_channel)))
        f.write(struct.pack("<I", code.size))
        f.write(np.ascontiguousarray(code, dtype=np.complex64).tobytes())
        f.write(struct.pack("<I", stream.size))
        f.write(np.ascontiguousarray(stream, dtype=np.complex64).tobytes())


def _read_adapter_output(path):
    ev, delay, dop, stamp, consumed, gen_calls, mag = struct.unpack("<iddQQii", open(path, "rb").read())
    return dict(event=ev, delay=delay, doppler=dop, stamp=stamp, consumed=consumed, generator_calls=gen_calls, mag=mag)


def test_integration_sources_compile_against_the_interface_stubs(adapter_driver):
    """integration/src/** (gr::block shell, adapter base + GPS L1 C/A / Galileo E1 / BeiDou B1I adapters, factory arms) build with
    -Wall -Wextra -Werror against stand-ins of the reference's interfaces (integration/stubs/) — runs on the CPU box."""
    assert os.path.exists(adapter_driver)
    patch = open(os.path.join(ROOT, "integration/src/core/receiver/gnss_block_factory.patch")).read()
    for name in ("GPS_L1_CA_PCPS_Acquisition_B200", "Galileo_E1_PCPS_Ambiguous_Acquisition_B200", "BEIDOU_B1I_PCPS_Acquisition_B200"):
        assert name in open(os.path.join(ROOT, "integration/src/core/receiver/gnss_block_factory_b200.cc")).read()
    assert "IsB200AcquisitionImplementation(implementation)" in patch and "GetAcqBlock" in patch


@pytest.mark.skipif(has_gpu(), reason="checks the no-device behaviour")
def test_adapter_refuses_to_run_without_a_device(adapter_driver, tmp_path):
    geo, codes, iq, _ = synth.config_scene("C1", n_prn=1)
    inp, outp = tmp_path / "a.bin", tmp_path / "a.out"
    _write_adapter_input(inp, geo["fs"], 5000, 250, 0.001, 1, 0, 1, 0, codes[0], iq)
    r = subprocess.run([adapter_driver, str(inp), str(outp)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 7 and "cannot create the B200 engine" in r.stderr, r.stdout + r.stderr      # init() throws: no CPU path


@pytest.mark.gpu
@pytest.mark.parametrize("which,scene,system", [(0, "C1", "gps_l1_ca"), (1, "C2", "gal_e1b"), (2, "C4_bds", "bds_b1i")])
def test_adapter_through_the_channel_call_order(adapter_driver, oracle, tmp_path, which, scene, system):
    """GetAcqBlock -> Channel's constructor order (set_channel, set_channel_fsm, set_gnss_synchro, set_doppler_step,
    set_threshold, init) -> set_signal's set_local_code -> connect / start -> reset -> the scheduler's general_work calls,
    for each of the three B200 adapters: the ChannelFsm gets Event_valid_acquisition and Gnss_Synchro holds what the oracle's
    block instance finds on the dwell the state machine really searched."""
    geo, codes, iq1, truth = synth.config_scene(scene, n_prn=max(t["prn"] for t in synth.config_scene(scene)[3][:1]))
    t = truth[0]
    n = geo["consumed_samples"]
    ms = geo["sampled_ms"]
    iq = synth.make_iq(3 * n, geo["fs"], [dict(t, cn0=max(t["cn0"], 47.0))], seed=77)
    one = synth.local_code(system, t["prn"], geo["fs"])
    inp, outp = tmp_path / "a.bin", tmp_path / "a.out"
    _write_adapter_input(inp, geo["fs"], geo["doppler_max"], geo["doppler_step"], 0.001, ms, which, t["prn"], 1, one, iq)
    r = subprocess.run([adapter_driver, str(inp), str(outp)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = _read_adapter_output(outp)
    assert got["event"] == 1 and got["generator_calls"] == 1
    # which samples did the block search? state 0 drops the first call's items (1000), the buffer then fills from there
    start = got["stamp"] - n
    assert start >= 1000 and start % 1000 == 0
    a = oracle.acq(fft_size=geo["fft_size"], consumed_samples=n, doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"], fs=geo["fs"],
                   samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"], use_cfar=True)
    a.set_local_code(np.tile(one, max(1, ms // synth.SYSTEMS[system][2]))[:n])
    want = a.dwell(iq[start:start + n])
    assert got["doppler"] == float(want["doppler_hz"]) and abs(got["doppler"] - t["doppler"]) <= geo["doppler_step"]
    assert got["delay"] == float(np.fmod(np.float32(want["code_phase"]), np.float32(geo["samples_per_code"])))
    one_len = one.size
    assert abs(((got["delay"] + start) - t["delay"] + one_len / 2) % one_len - one_len / 2) <= 1
    a.close()
