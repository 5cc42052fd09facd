"""N>1 host logic on CPU: world_size-2 (and 3) gloo groups run the product's partition / gather / merge code
(gnss-sdr-new_b200/sharding.py + gnss_b200_acq_finalize_rows) on per-row records derived from the oracle's grids.
The merged decision must equal the oracle's single-process decision."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import synth
from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run_world(world, work, extra_env=None):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(_free_port()), WORLD_SIZE=str(world),
               GLOO_SOCKET_IFNAME="lo", OMP_NUM_THREADS="1")
    env.update(extra_env or {})
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "_gloo_worker.py"), str(work)],
                              env=dict(env, RANK=str(r), LOCAL_RANK=str(r)), stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
             for r in range(world)]
    outs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(o.decode(errors="replace"))
    for r, p in enumerate(procs):
        assert p.returncode == 0, f"rank {r} failed:\n{outs[r][-3000:]}"
    return [np.load(os.path.join(work, f"out{r}.npz")) for r in range(world)]


@pytest.fixture(scope="module")
def scene(b200, oracle, tmp_path_factory):
    """Three GPS PRNs on config 1's grid; oracle grids -> the per-row records a device would emit."""
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=3)
    nd = int(np.ceil(2.0 * geo["doppler_max"] / geo["doppler_step"]))
    rows = np.zeros((3, nd), dtype=b200.ACQ_ROW_DTYPE)
    want = []
    for s in range(3):
        acq = oracle.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"],
                         doppler_step=geo["doppler_step"], fs=geo["fs"], samples_per_chip=geo["samples_per_chip"],
                         samples_per_code=geo["samples_per_code"])
        acq.set_local_code(codes[s])
        want.append(acq.dwell(iq))
        g = acq.grid()[:, :geo["fft_size"]]
        rows[s]["max_val"] = g.max(axis=1)
        rows[s]["max_idx"] = g.argmax(axis=1)
        for k in range(nd):
            acc = np.float32(0)
            for v in g[k].reshape(-1, 100).sum(axis=1, dtype=np.float32):  # blocked float32 sum, like a device reduction
                acc = np.float32(acc + v)
            rows[s]["row_sum"][k] = acc
        rows[s]["valid"] = 1
        acq.close()
    work = tmp_path_factory.mktemp("gloo")
    rng = np.random.default_rng(3)
    n_ch = 7
    ch_out = (rng.standard_normal((n_ch, 5)) + 1j * rng.standard_normal((n_ch, 5))).astype(np.complex64)
    np.savez(os.path.join(work, "in.npz"), rows=rows.view(np.uint8), n_slots=3, n_doppler=nd, eff=geo["fft_size"],
             doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"], samples=iq, n_ch=n_ch, ch_out=ch_out)
    return dict(work=work, want=want, iq=iq, ch_out=ch_out, geo=geo, nd=nd)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_search_equals_single_process(b200, scene, world):
    outs = _run_world(world, scene["work"])
    for r, o in enumerate(outs):
        assert str(o["err"]) == ""
        res = o["res"].view(b200.ACQ_RESULT_DTYPE)
        for s, w in enumerate(scene["want"]):
            assert int(res[s]["code_phase"]) == w["code_phase"]
            assert int(res[s]["doppler_index"]) == w["doppler_index"]
            assert int(res[s]["doppler_hz"]) == w["doppler_hz"]
            assert float(res[s]["peak"]) == w["peak"]
            assert float(res[s]["test_statistic"]) == pytest.approx(w["test_statistic"], rel=1e-5)
        assert np.array_equal(o["samples"], scene["iq"])           # broadcast of the sample block
        assert np.array_equal(o["ch_out"], scene["ch_out"])        # tracking channel gather
    # every rank reached the same answer bit for bit
    for o in outs[1:]:
        assert o["res"].tobytes() == outs[0]["res"].tobytes()
    # the Doppler-major split of the same search (what the engines exchange on the device): the same decision on every rank
    for o in outs:
        assert o["res_dm"].tobytes() == outs[0]["res"].tobytes()
    chunk = -(-scene["nd"] // world)
    assert [int(o["rows_dm"][0]) for o in outs] == [max(0, min(scene["nd"], (r + 1) * chunk) - min(scene["nd"], r * chunk)) for r in range(world)]
    if world == 8:  # 3 slots on 8 ranks: rows of a slot are split over 2 ranks, 2 ranks idle
        assert sorted(int(o["owned"][0]) for o in outs) == [0, 0, 1, 1, 1, 1, 1, 1]


def test_merge_detects_double_claims_and_holes(b200, scene):
    outs = _run_world(2, scene["work"], {"B200_TEST_DOUBLE_CLAIM": "1"})
    assert all("more than one rank" in str(o["err"]) for o in outs)
    outs = _run_world(2, scene["work"], {"B200_TEST_DROP_ROW": "1"})
    assert all("no rank" in str(o["err"]) for o in outs)


@pytest.mark.parametrize("n_slots,nd,world", [(32, 40, 1), (32, 40, 8), (50, 80, 8), (3, 800, 8), (1, 40, 8), (1, 3, 8), (0, 40, 4),
                                               (131, 40, 8), (5, 7, 3)])
def test_partition_covers_every_unit_once(b200, n_slots, nd, world):
    plan = b200.sharding.partition_search(n_slots, nd, world)
    assert len(plan) == world
    seen = np.zeros((n_slots, nd), dtype=np.int32)
    for r in plan:
        for slot, r0, r1 in r:
            assert 0 <= r0 < r1 <= nd
            seen[slot, r0:r1] += 1
    assert (seen == 1).all()
    loads = [sum(r1 - r0 for _, r0, r1 in r) for r in plan]
    if n_slots >= world:
        assert max(loads) - min(loads) <= nd   # PRN-major: at most one slot of imbalance
    parts = b200.sharding.partition_channels(2048, world)
    assert parts[0][0] == 0 and parts[-1][1] == 2048 and all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
