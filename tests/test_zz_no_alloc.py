"""gnss_b200_mc_correlate_noalloc (include/gnss_b200.h): the multicorrelator entry point that allocates nothing — its scratch is
a fixed device array inside the library. No counterpart in the reference.
Checked against the existing entry point (gnss_b200_mc_correlate_batch) through the C-ABI on two calls in a row, in the
tolerance of tests/test_multicorr_gpu.py. This file is named to be collected LAST."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu

RTOL = 1e-4   # relative to the largest tap of a channel, as in tests/test_multicorr_gpu.py


def _rel_err(got, want):
    scale = np.abs(want).max(axis=-1, keepdims=True)
    return (np.abs(got - want) / scale).max()


def _scene(rng, n_ch, n, ntaps, fs=30e6, chip_rate=10.23e6):
    chips = synth.load_chips()["gps_l5i"]
    shifts = np.tile(np.linspace(-1, 1, ntaps).astype(np.float32) if ntaps > 1 else np.zeros(1, np.float32), (n_ch, 1))
    codes, sigs, args = [], [], []
    for c in range(n_ch):
        code = chips[(c * 5 + 3) % 32].astype(np.float32)
        fd = rng.uniform(-5000, 5000)
        code_step = np.float32(chip_rate * (1 + fd / 1176.45e6) / fs)
        rem_code = np.float32(rng.uniform(0, 1))
        carr_step = np.float32(2 * np.pi * fd / fs)
        rem_carr = np.float32(rng.uniform(0, 2 * np.pi))
        t = np.arange(n)
        idx = np.floor(code_step * t - rem_code).astype(np.int64) % code.size
        sig = 0.05 * code[idx] * np.exp(1j * (rem_carr + carr_step * t)) + (rng.standard_normal(n) + 1j * rng.standard_normal(n)) / np.sqrt(2)
        codes.append(code)
        sigs.append(sig.astype(np.complex64))
        args.append((rem_carr, carr_step, rem_code, code_step))
    return codes, sigs, np.array(args, dtype=np.float32), shifts


def _existing(b200, codes, sigs, args, shifts):
    n_ch, n = len(sigs), sigs[0].size
    mc = b200.MulticorrelatorB200(n_ch, n, shifts.shape[1])
    for c in range(n_ch):
        mc.set_local_code(c, codes[c])
    host = np.ascontiguousarray(np.stack(sigs))
    z = np.zeros(n_ch, np.float32)
    a = mc.make_args(n_ch, n, args[:, 0], args[:, 1], z, args[:, 2], args[:, 3], z, sig_ptrs=[host[c].ctypes.data for c in range(n_ch)])
    out = mc.correlate_batch(np.arange(n_ch), a, shifts, 0)
    mc.close()
    return out


def _noalloc(b200, torch, codes, sigs, args, shifts, d_out):
    # the CALLER owns every device buffer (torch here); the entry point itself allocates nothing
    d_sig = [torch.from_numpy(s.view(np.float32).copy()).cuda() for s in sigs]
    d_code = [torch.from_numpy(c.copy()).cuda() for c in codes]
    chans = [dict(d_sig=d_sig[c].data_ptr(), d_code=d_code[c].data_ptr(), code_len=codes[c].size, n_samples=sigs[c].size, rem_carr=args[c, 0],
                  carr_step=args[c, 1], rem_code=args[c, 2], code_step=args[c, 3]) for c in range(len(sigs))]
    stream = torch.cuda.current_stream()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    b200.mc_correlate_noalloc(0, stream.cuda_stream, chans, shifts, d_out.data_ptr())
    stream.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    n_ch, n_taps = shifts.shape
    return d_out[:2 * n_ch * n_taps].cpu().numpy().view(np.complex64).reshape(n_ch, n_taps), free_before - free_after


def test_noalloc_entry_point_matches_the_existing_one_twice_in_a_row(b200):
    import torch
    assert 0 < b200.lib().gnss_b200_mc_noalloc_scratch_bytes() < 160 * 1024
    rng = np.random.default_rng(2024)
    d_out = torch.full((2 * 64 * 16,), float("nan"), dtype=torch.float32, device="cuda")
    # warm-up call: the first call on a device resolves the scratch address and loads the kernels (module load may take memory)
    w = _scene(rng, 1, 4096, 1)
    _noalloc(b200, torch, *w, d_out)
    # call 1: few channels, split over many CTAs (partials + tickets in the resident scratch); call 2 right after it with other
    # shapes, so that whatever call 1 left in the scratch would show
    for n_ch, n, ntaps in ((3, 30000, 5), (40, 12000, 11)):
        codes, sigs, args, shifts = _scene(rng, n_ch, n, ntaps)
        d_out.fill_(float("nan"))
        got, _ = _noalloc(b200, torch, codes, sigs, args, shifts, d_out)
        want = _existing(b200, codes, sigs, args, shifts)
        assert np.isfinite(got.view(np.float32)).all()
        assert _rel_err(got, want) <= RTOL
        # the same call again on the same inputs: bit-identical (deterministic split order, tickets cleared), and — the kernels
        # of this tap count being loaded by now — the device's free memory does not move across the call
        again, grew = _noalloc(b200, torch, codes, sigs, args, shifts, d_out)
        assert np.array_equal(again, got)
        assert grew <= 0, "device memory in use grew by %d bytes across the call" % grew


def test_noalloc_entry_point_against_the_oracle_and_under_graph_capture(b200, oracle):
    """(1) directly against the oracle (not only against the handle-based entry point); (2) recorded under CUDA stream capture
    and replayed twice on NEW sample data in the caller's buffers: the call issues kernels only (parameters travel as by-value
    kernel arguments), so the graph holds no pointer into this process's host memory."""
    import torch
    rng = np.random.default_rng(77)
    n_ch, n, ntaps = 20, 8000, 5           # 20 channels: two by-value parameter launches (16 + 4)
    codes, sigs, args, shifts = _scene(rng, n_ch, n, ntaps)
    d_out = torch.full((2 * n_ch * ntaps,), float("nan"), dtype=torch.float32, device="cuda")
    got, _ = _noalloc(b200, torch, codes, sigs, args, shifts, d_out)
    for c in range(n_ch):
        want = oracle.mc_real(sigs[c], codes[c], shifts[c], float(args[c, 0]), float(args[c, 1]), 0.0, float(args[c, 2]), float(args[c, 3]), 0.0)
        assert np.abs(got[c] - want).max() <= RTOL * np.abs(want).max(), c
    # capture: the caller's device buffers are fixed, their CONTENT changes between replays
    d_sig = [torch.from_numpy(s.view(np.float32).copy()).cuda() for s in sigs]
    d_code = [torch.from_numpy(c.copy()).cuda() for c in codes]
    chans = [dict(d_sig=d_sig[c].data_ptr(), d_code=d_code[c].data_ptr(), code_len=codes[c].size, n_samples=n, rem_carr=args[c, 0],
                  carr_step=args[c, 1], rem_code=args[c, 2], code_step=args[c, 3]) for c in range(n_ch)]
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(graph, stream=side):
        b200.mc_correlate_noalloc(0, side.cuda_stream, chans, shifts, d_out.data_ptr())
    for rep in range(2):
        new = [(s * np.exp(1j * 0.3 * (rep + 1)) + np.float32(0.01 * rep)).astype(np.complex64) for s in sigs]
        for c in range(n_ch):
            d_sig[c].copy_(torch.from_numpy(new[c].view(np.float32).copy()))
        d_out.fill_(float("nan"))
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        out = d_out.cpu().numpy().view(np.complex64).reshape(n_ch, ntaps)
        for c in (0, 7, 19):
            want = oracle.mc_real(new[c], codes[c], shifts[c], float(args[c, 0]), float(args[c, 1]), 0.0, float(args[c, 2]), float(args[c, 3]), 0.0)
            assert np.abs(out[c] - want).max() <= RTOL * np.abs(want).max(), (rep, c)
    assert torch.cuda.current_device() == 0


def test_noalloc_entry_point_rejects_sizes_beyond_the_scratch(b200):
    import torch
    d = torch.zeros(64, dtype=torch.float32, device="cuda")
    one = dict(d_sig=d.data_ptr(), d_code=d.data_ptr(), code_len=8, n_samples=8, rem_carr=0.0, carr_step=0.0, rem_code=0.0, code_step=0.5)
    with pytest.raises(b200.B200Error, match="at most 64"):
        b200.mc_correlate_noalloc(0, None, [one] * 65, np.zeros((65, 1), np.float32), d.data_ptr())
    with pytest.raises(b200.B200Error, match="at most 16"):
        b200.mc_correlate_noalloc(0, None, [one], np.zeros((1, 17), np.float32), d.data_ptr())
    with pytest.raises(b200.B200Error, match="samples"):
        b200.mc_correlate_noalloc(0, None, [dict(one, n_samples=(1 << 22) + 1)], np.zeros((1, 1), np.float32), d.data_ptr())
    with pytest.raises(b200.B200Error):
        b200.mc_correlate_noalloc(0, None, [dict(one, d_code=None)], np.zeros((1, 1), np.float32), d.data_ptr())
