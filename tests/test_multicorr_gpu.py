"""GPU parity of the tracking multicorrelator kernel against the oracle (through the C-ABI).

Bars (BASELINE.json north_star): chip indices bit-exact; correlator outputs within 1e-4 relative
(relative to the largest tap magnitude of the channel, since far taps are noise-like and near zero).
"""
import ctypes as C

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-4


def _rel_err(got, want):
    scale = np.abs(want).max(axis=-1, keepdims=True)
    return (np.abs(got - want) / scale).max()


def _scene(rng, n_ch, n, code_len, ntaps, fs=30e6, chip_rate=10.23e6, cn0=45.0):
    chips = synth.load_chips()["gps_l5i"] if code_len == 10230 else None
    codes, sigs, args = [], [], []
    shifts = np.tile(np.linspace(-1, 1, ntaps).astype(np.float32) if ntaps > 1 else np.zeros(1, np.float32), (n_ch, 1))
    for c in range(n_ch):
        code = (chips[c % 32].astype(np.float32) if chips is not None else rng.choice([-1.0, 1.0], code_len).astype(np.float32))
        fd = rng.uniform(-5000, 5000)
        code_step = np.float32(chip_rate * (1 + fd / 1176.45e6) / fs)
        rem_code = np.float32(rng.uniform(0, 1))
        carr_step = np.float32(2 * np.pi * fd / fs)
        rem_carr = np.float32(rng.uniform(0, 2 * np.pi))
        t = np.arange(n)
        idx = np.floor(code_step * t - rem_code).astype(np.int64) % code_len
        amp = np.sqrt(10 ** (cn0 / 10) / fs)
        sig = amp * code[idx] * np.exp(1j * (rem_carr + carr_step * t)) + (rng.standard_normal(n) + 1j * rng.standard_normal(n)) / np.sqrt(2)
        codes.append(code)
        sigs.append(sig.astype(np.complex64))
        args.append((rem_carr, carr_step, np.float32(rng.uniform(-2e-10, 2e-10)), rem_code, code_step, np.float32(rng.uniform(-1e-10, 1e-10))))
    return codes, sigs, np.array(args, dtype=np.float32), shifts


def _run_gpu(b200, codes, sigs, args, shifts, flags=0):
    n_ch, n = len(sigs), sigs[0].size
    mc = b200.MulticorrelatorB200(n_ch, n, shifts.shape[1])
    for c in range(n_ch):
        mc.set_local_code(c, codes[c])
    host = np.ascontiguousarray(np.stack(sigs))
    ptrs = [host[c].ctypes.data for c in range(n_ch)]
    a = mc.make_args(n_ch, n, args[:, 0], args[:, 1], args[:, 2], args[:, 3], args[:, 4], args[:, 5], sig_ptrs=ptrs)
    out = mc.correlate_batch(np.arange(n_ch), a, shifts, flags)
    return mc, a, out


def _run_oracle(oracle, codes, sigs, args, shifts, high_dyn=False, six=False):
    out = []
    for c in range(len(sigs)):
        r = args[c]
        out.append(oracle.mc_real(sigs[c], codes[c], shifts[c], float(r[0]), float(r[1]), float(r[2]), float(r[3]), float(r[4]), float(r[5]),
                                  high_dyn=high_dyn, six_arg=six))
    return np.stack(out)


@pytest.mark.parametrize("n,code_len,ntaps,n_ch", [(2048, 1023, 3, 4), (4096, 1023, 3, 2), (8192, 1023, 5, 3), (30000, 10230, 11, 6), (8111, 4092, 1, 2),
                                                  (30000, 10230, 20, 2)])
def test_outputs_match_oracle(b200, oracle, n, code_len, ntaps, n_ch):
    rng = np.random.default_rng(100 + n + ntaps)
    codes, sigs, args, shifts = _scene(rng, n_ch, n, code_len, ntaps)
    _, _, got = _run_gpu(b200, codes, sigs, args, shifts)
    want = _run_oracle(oracle, codes, sigs, args, shifts)
    assert _rel_err(got, want) < RTOL


def test_chip_indices_bit_exact(b200, oracle):
    rng = np.random.default_rng(5)
    for n, L, T in ((30000, 10230, 11), (8192, 1023, 3), (16000, 8184, 5)):
        codes, sigs, args, shifts = _scene(rng, 1, n, L if L != 8184 else 1023, T)
        codes = [np.resize(codes[0], L)]
        mc, a, _ = _run_gpu(b200, codes, sigs, args, shifts)
        got = mc.chip_indices(0, a, shifts[0])
        want = oracle.mc_chip_indices(n, L, shifts[0], float(args[0, 3]), float(args[0, 4]))
        assert np.array_equal(got, want)
        # negative and large shifts, rem > 1
        sh = np.array([-3.5, -1.0, 0.0, 2.25, 700.5], np.float32)
        a2 = mc.make_args(1, n, 0.1, 0.001, 0.0, 1.75, float(args[0, 4]), 0.0)
        mc2 = b200.MulticorrelatorB200(1, n, 5)
        mc2.set_local_code(0, codes[0])
        assert np.array_equal(mc2.chip_indices(0, a2, sh), oracle.mc_chip_indices(n, L, sh, 1.75, float(args[0, 4])))


@pytest.mark.parametrize("case", ["random", "dyadic", "misaligned", "small_phase"])
def test_short_chip_phase_form_is_exact(b200, monkeypatch, case):
    """The window kernel replaces floor(fl(fl(a + shift) - rem)) by floor(a + k) wherever a, a + shift and a + shift - rem share a
    binade (multicorr.cu, sample_short). Both forms walk the samples in the same order with the same arithmetic, so if every chip
    index agrees the correlator outputs are BIT-identical to a run with the short form switched off (B200_MC_EXACT_ONLY=1) —
    one differing index anywhere in 6 x 30000 x 11 lookups would show. Cases: random shifts/rem; dyadic shifts and rem (the tie
    candidates: multiples of the binade spacing and exact halves of it); a stream that starts on an 8-byte, not 16-byte, boundary
    (the peeled first sample); chip phases below 64 (short form mostly unavailable) and a whole code period."""
    rng = np.random.default_rng({"random": 1, "dyadic": 2, "misaligned": 3, "small_phase": 4}[case])
    n_ch, n, ntaps = 6, 30000, 11
    if case == "small_phase":
        n = 9000
    codes, sigs, args, shifts = _scene(rng, n_ch, n, 10230, ntaps)
    if case == "random":
        shifts = np.sort(rng.uniform(-1.5, 1.5, (n_ch, ntaps)).astype(np.float32), axis=1)
    if case == "dyadic":
        base = np.array([-1, -0.5, -0.25, -2.0 ** -9, -2.0 ** -11, 0, 2.0 ** -12, 2.0 ** -10, 0.25, 0.5, 1], np.float32)
        shifts = np.tile(base, (n_ch, 1))
        shifts[1] += np.float32(2.0 ** -13)
        args[:, 3] = np.array([0.5, 0.25 + 2.0 ** -11, 2.0 ** -12, 0.0, 0.75 + 2.0 ** -13, 1 - 2.0 ** -10], np.float32)    # rem_code_phase_chips
        args[2, 4] = np.float32(0.3125)                                                                                  # a dyadic code step too
    results = []
    for exact_only in ("0", "1"):
        monkeypatch.setenv("B200_MC_EXACT_ONLY", exact_only)
        mc = b200.MulticorrelatorB200(n_ch, n, ntaps)
        for c in range(n_ch):
            mc.set_local_code(c, codes[c])
        if case == "misaligned":
            stream = np.concatenate([np.zeros(1, np.complex64)] + sigs)          # every channel window starts at an odd sample
            mc.set_stream(stream)
            offs = 1 + n * np.arange(n_ch)
            a = mc.make_args(n_ch, n, args[:, 0], args[:, 1], args[:, 2], args[:, 3], args[:, 4], args[:, 5], stream_offsets=offs)
        else:
            host = np.ascontiguousarray(np.stack(sigs))
            a = mc.make_args(n_ch, n, args[:, 0], args[:, 1], args[:, 2], args[:, 3], args[:, 4], args[:, 5], sig_ptrs=[host[c].ctypes.data for c in range(n_ch)])
        results.append(np.array(mc.correlate_batch(np.arange(n_ch), a, shifts), copy=True))
        mc.close()
    assert np.array_equal(results[0].view(np.uint32), results[1].view(np.uint32))
    assert np.abs(results[0]).max() > 1.0


@pytest.mark.parametrize("six", [False, True])
def test_high_dynamics_mode(b200, oracle, six):
    rng = np.random.default_rng(17)
    codes, sigs, args, shifts = _scene(rng, 3, 8192, 1023, 5, fs=4e6, chip_rate=1.023e6)
    args[:, 2] = rng.uniform(-1e-7, 1e-7, 3)   # carrier phase rate [rad/sample^2]
    args[:, 5] = rng.uniform(-1e-9, 1e-9, 3)   # code phase rate
    flags = b200.MC_HIGH_DYN | (b200.MC_SIX_ARG if six else 0)
    mc, a, got = _run_gpu(b200, codes, sigs, args, shifts, flags)
    want = _run_oracle(oracle, codes, sigs, args, shifts, high_dyn=True, six=six)
    assert _rel_err(got, want) < RTOL
    idx = mc.chip_indices(1, mc.make_args(1, 8192, *[float(v) for v in args[1]]), shifts[1], flags)
    assert np.array_equal(idx, oracle.mc_chip_indices(8192, 1023, shifts[1], float(args[1, 3]), float(args[1, 4]), float(args[1, 5]), high_dyn=True))


def test_complex_codes(b200, oracle):
    rng = np.random.default_rng(23)
    n, L, T = 4000, 511, 3
    sig = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    code = (rng.choice([-1.0, 1.0], L) + 1j * rng.choice([-1.0, 1.0], L)).astype(np.complex64)
    shifts = np.array([[-0.5, 0.0, 0.5]], np.float32)
    mc = b200.MulticorrelatorB200(1, n, T)
    mc.set_local_code(0, code)
    a = mc.make_args(1, n, 0.25, 0.1, 0.0, 0.3, 0.255, 0.0, sig_ptrs=[sig.ctypes.data])
    got = mc.correlate_batch([0], a, shifts)
    want = oracle.mc_cplx(sig, code, shifts[0], 0.25, 0.1, 0.3, 0.255)
    assert _rel_err(got, want[None, :]) < RTOL


def test_shared_stream_windows_and_linearity(b200, oracle):
    """Channels windowing one uploaded stream give the same result as private copies; outputs are linear in the input."""
    rng = np.random.default_rng(31)
    n, L, T, n_ch = 4000, 1023, 3, 5
    stream = (rng.standard_normal(2 * n) + 1j * rng.standard_normal(2 * n)).astype(np.complex64)
    offs = rng.integers(0, n, n_ch)
    code = rng.choice([-1.0, 1.0], L).astype(np.float32)
    shifts = np.tile(np.array([-0.5, 0.0, 0.5], np.float32), (n_ch, 1))
    mc = b200.MulticorrelatorB200(n_ch, n, T)
    for c in range(n_ch):
        mc.set_local_code(c, code)
    mc.set_stream(stream)
    a = mc.make_args(n_ch, n, 0.7, 0.01, 0.0, 0.2, 0.25575, 0.0, stream_offsets=offs)
    got = mc.correlate_batch(np.arange(n_ch), a, shifts)
    for c in range(n_ch):
        want = oracle.mc_real(stream[offs[c]:offs[c] + n], code, shifts[c], 0.7, 0.01, 0.0, 0.2, 0.25575, 0.0)
        assert _rel_err(got[c][None], want[None]) < RTOL
    mc.set_stream(stream * np.complex64(2.0))
    got2 = mc.correlate_batch(np.arange(n_ch), a, shifts)
    assert np.allclose(got2, 2 * got, rtol=1e-6, atol=0)


def test_ragged_batch_and_tiny_blocks(b200, oracle):
    """One batch whose channels differ in length (down to one sample: fewer samples than threads, partial copy groups,
    a split boundary that is not a multiple of anything), code length and tap spread (shifts of several chips, negative
    and beyond a code period); every channel against the oracle, chip indices of the extremes bit-exact."""
    rng = np.random.default_rng(404)
    lens = [1, 2, 31, 255, 256, 257, 1023, 4097, 12289, 29999]
    n_ch, n_max, T = len(lens), max(lens), 5
    code_lens = [1023, 1023, 511, 4092, 1023, 10230, 1023, 2046, 10230, 10230]
    shifts = np.stack([np.linspace(-w, w, T).astype(np.float32) for w in (0.5, 1.0, 2.5, 0.1, 7.0, 0.5, 1100.0, 0.25, 3.0, 0.5)])
    mc = b200.MulticorrelatorB200(n_ch, n_max, T)
    sigs, codes = [], []
    for c in range(n_ch):
        codes.append(rng.choice([-1.0, 1.0], code_lens[c]).astype(np.float32))
        sigs.append((rng.standard_normal(lens[c]) + 1j * rng.standard_normal(lens[c])).astype(np.complex64))
        mc.set_local_code(c, codes[c])
    rem_carr = rng.uniform(0, 2 * np.pi, n_ch).astype(np.float32)
    carr_step = rng.uniform(-0.01, 0.01, n_ch).astype(np.float32)
    rem_code = rng.uniform(0, 1, n_ch).astype(np.float32)
    code_step = (np.float32(0.341) * (1 + rng.uniform(-1e-5, 1e-5, n_ch))).astype(np.float32)
    a = mc.make_args(n_ch, lens, rem_carr, carr_step, 0.0, rem_code, code_step, 0.0, sig_ptrs=[x.ctypes.data for x in sigs])
    got = mc.correlate_batch(np.arange(n_ch), a, shifts)
    for c in range(n_ch):
        want = oracle.mc_real(sigs[c], codes[c], shifts[c], float(rem_carr[c]), float(carr_step[c]), 0.0, float(rem_code[c]), float(code_step[c]), 0.0)
        scale = max(np.abs(want).max(), np.sqrt(lens[c]))  # a few samples: the sum itself can be near zero
        assert (np.abs(got[c] - want) / scale).max() < RTOL, (c, lens[c])
    for c in (0, 6, 9):
        one = mc.make_args(1, lens[c], float(rem_carr[c]), float(carr_step[c]), 0.0, float(rem_code[c]), float(code_step[c]), 0.0)
        assert np.array_equal(mc.chip_indices(c, one, shifts[c]), oracle.mc_chip_indices(lens[c], code_lens[c], shifts[c], float(rem_code[c]), float(code_step[c]), 0.0))


def test_full_size_config5_properties(b200, oracle):
    """BASELINE config 5 at full per-channel size (30 000 samples, 11 taps, L5I), 64 channels: the prompt tap of every
    channel carries the injected signal, and EVERY channel is checked against the oracle."""
    rng = np.random.default_rng(55)
    codes, sigs, args, shifts = _scene(rng, 64, 30000, 10230, 11)
    _, _, got = _run_gpu(b200, codes, sigs, args, shifts)
    prompt = np.abs(got[:, 5])
    far = 0.5 * (np.abs(got[:, 0]) + np.abs(got[:, 10]))  # +-1 chip taps: noise only (~sqrt(N) sigma)
    assert (prompt > 2 * far).all() and prompt.mean() > 4 * far.mean()
    want = _run_oracle(oracle, codes, sigs, args, shifts)
    assert want.shape == got.shape == (64, 11)
    assert _rel_err(got, want) < RTOL


def test_error_behaviour(b200):
    mc = b200.MulticorrelatorB200(2, 1000, 3)
    a = mc.make_args(1, 1000, 0, 0, 0, 0, 0.25, 0)
    with pytest.raises(b200.B200Error):
        mc.correlate_batch([0], a, np.zeros((1, 3), np.float32))   # no local code yet
    mc.set_local_code(0, np.ones(100, np.float32))
    with pytest.raises(b200.B200Error):
        mc.correlate_batch([5], a, np.zeros((1, 3), np.float32))   # channel out of range
    with pytest.raises(b200.B200Error):
        mc.correlate_batch([0], mc.make_args(1, 5000, 0, 0, 0, 0, 0.25, 0), np.zeros((1, 3), np.float32))  # too long
