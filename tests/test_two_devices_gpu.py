"""One host process driving two GPUs through the C-ABI (the reference receiver is a single process: its blocks would
carry different `device` ordinals). Kernel attributes (dynamic shared memory, carve-out) are per device, so the second
device must be configured on its own; results are bit-identical to the first device's.

Skipped on a one-GPU box.
"""
import numpy as np
import pytest

import synth
from test_acq_gpu import _gpu_engine
from test_multicorr_gpu import _scene

pytestmark = pytest.mark.gpu


@pytest.fixture()
def two_devices(b200):
    if b200.lib().gnss_b200_device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    return b200


def _correlate_on(b200, device, codes, sigs, args, shifts):
    n_ch, n = len(sigs), sigs[0].size
    mc = b200.MulticorrelatorB200(n_ch, n, shifts.shape[1], device=device)
    for c in range(n_ch):
        mc.set_local_code(c, codes[c])
    host = np.ascontiguousarray(np.stack(sigs))
    a = mc.make_args(n_ch, n, args[:, 0], args[:, 1], args[:, 2], args[:, 3], args[:, 4], args[:, 5], sig_ptrs=[host[c].ctypes.data for c in range(n_ch)])
    out = mc.correlate_batch(np.arange(n_ch), a, shifts, 0).copy()
    mc.close()
    return out


@pytest.mark.parametrize("n,code_len,ntaps,n_ch", [(30000, 10230, 11, 6), (30000, 10230, 20, 2), (8192, 1023, 3, 3)])
def test_multicorrelator_on_the_second_device(two_devices, n, code_len, ntaps, n_ch):
    b200 = two_devices
    rng = np.random.default_rng(77 + ntaps)
    codes, sigs, args, shifts = _scene(rng, n_ch, n, code_len, ntaps)
    first = _correlate_on(b200, 0, codes, sigs, args, shifts)
    second = _correlate_on(b200, 1, codes, sigs, args, shifts)
    again = _correlate_on(b200, 0, codes, sigs, args, shifts)
    assert np.array_equal(first.view(np.uint32), second.view(np.uint32))
    assert np.array_equal(first.view(np.uint32), again.view(np.uint32))


@pytest.mark.parametrize("scene", ["C1", "C2"])
def test_acquisition_on_the_second_device(two_devices, scene):
    """C1: single-stage plan (tile > 48 KB of dynamic shared memory); C2: the two-stage fast kernels."""
    b200 = two_devices
    n_prn = 8
    geo, codes, iq, _ = synth.config_scene(scene, n_prn=n_prn)
    records = []
    for device in (0, 1, 0):
        e = _gpu_engine(b200, geo, n_prn, use_cfar=True, device=device)
        e.set_local_codes(np.arange(n_prn), codes[:n_prn])
        got = e.dwell(iq, np.arange(n_prn))
        records.append([(int(r["code_phase"]), int(r["doppler_index"]), np.float32(r["test_statistic"]).view(np.uint32),
                         np.float32(r["peak"]).view(np.uint32)) for r in got])
        e.close()
    assert records[0] == records[1] == records[2]


@pytest.mark.parametrize("scene,n_prn", [("C1", 8), ("C2", 8), ("C4_gal", 4)])
def test_one_search_over_two_devices(two_devices, scene, n_prn):
    """gnss_b200_acq_group_*: the Doppler rows of one search split over the two GPUs, sample block broadcast host -> GPU 0 ->
    GPU 1, row records sent GPU 1 -> GPU 0 (cudaMemcpyPeerAsync), decision on GPU 0. Records bit-identical to one engine."""
    b200 = two_devices
    geo, codes, iq, _ = synth.config_scene(scene, n_prn=n_prn)
    one = _gpu_engine(b200, geo, n_prn, use_cfar=True)
    one.set_local_codes(np.arange(n_prn), codes[:n_prn])
    want = one.dwell(iq, np.arange(n_prn))
    for devices in ([0, 1], [1, 0]):
        grp = b200.PcpsAcquisitionGroupB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                            geo["samples_per_chip"], n_prn, devices)
        grp.set_local_codes(np.arange(n_prn), codes[:n_prn])
        for _ in range(3):
            got = grp.dwell(iq, np.arange(n_prn))
            assert got.tobytes() == want.tobytes(), devices
        grp.close()
    one.close()


@pytest.mark.parametrize("scene,n_prn", [("C2", 8), ("C1", 6)])
def test_record_exchange_over_peer_memory(two_devices, scene, n_prn):
    """gnss_b200_acq_peer_*: two shards of one search on the two GPUs store their row records straight into each other's
    region over NVLink and raise a flag; each decision kernel waits for both flags. No collective, no host hop — and the whole
    step of a shard (search + exchange + decision) replays from one CUDA graph. Records on BOTH shards bit-identical to one
    engine's, over several searches (the regions are double-buffered by exchange parity)."""
    import torch
    b200 = two_devices
    geo, codes, iq, _ = synth.config_scene(scene, n_prn=n_prn, n_dwells=4)
    N = geo["consumed_samples"]
    iq = np.ascontiguousarray(iq).reshape(4, N)
    slots = np.arange(n_prn, dtype=np.uint32)
    one = _gpu_engine(b200, geo, n_prn, use_cfar=True)
    one.set_local_codes(slots, codes[:n_prn])
    nD = one.n_doppler
    chunk = (nD + 1) // 2
    shards, regions, bufs = [], [], []
    for r in range(2):
        e = _gpu_engine(b200, geo, n_prn, use_cfar=True, device=r, row_begin=r * chunk, row_end=min(nD, (r + 1) * chunk))
        e.set_local_codes(slots, codes[:n_prn])
        e.exchange_buffers(2)
        _, ptr = e.peer_setup(r)
        shards.append(e)
        regions.append(ptr)
        bufs.append(torch.empty(2 * N, dtype=torch.float32, device=f"cuda:{r}"))
    for e in shards:
        e.peer_connect(regions=regions)
    for k in range(4):
        want = one.dwell(iq[k], slots)
        for r in range(2):
            bufs[r].copy_(torch.from_numpy(iq[k].view(np.float32).copy()))
        torch.cuda.synchronize(0)
        torch.cuda.synchronize(1)
        if k == 2:     # from the third search on: one graph per shard
            for r, e in enumerate(shards):
                e.capture_begin()
                e.dwell_device_async(bufs[r].data_ptr(), slots)
                e.peer_exchange_async(n_prn)
                e.capture_end()
        for e in shards:
            e.peer_barrier_async()           # device-side rendezvous of the two shards in front of the search
        for r, e in enumerate(shards):
            if k >= 2:
                e.replay_async()
            else:
                e.dwell_device_async(bufs[r].data_ptr(), slots)
                if k == 0:       # a connected shard has no results of its own before the exchange
                    with pytest.raises(b200.B200Error):
                        e.sync(n_prn)
                e.peer_exchange_async(n_prn)
        for e in shards:
            got = e.sync(n_prn)
            assert got.tobytes() == want.tobytes(), (k, scene)
    for e in shards:
        e.close()
    one.close()
