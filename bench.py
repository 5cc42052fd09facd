#!/usr/bin/env python3
"""bench.py — acquisition cells/s (headline, BASELINE config 2) and tracking channel-ms/s (config 5) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2|C1]

One "step" = one coherent dwell of the whole PRN x Doppler x code-phase grid of the workload on one batch of
synthetic IQ (ours: through libgnss_b200.so; reference: the CPU path on the host cores).
  value  : cells/s with the sample block already resident in HBM (device-timed with CUDA events on the engine stream)
  e2e    : cells/s through the C-ABI call a GNSS-SDR adapter makes, HOST sample buffer in, host results out
  N > 1  : one process per GPU (torchrun); the Doppler rows of ONE search are sharded over the ranks (strong scaling: the
           workload is the same 128 M-cell grid at every N), the per-row records are all-gathered device-to-device with NCCL
           and every rank decides on the device — all inside the CUDA-event bracket.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import synth  # noqa: E402

METRIC = "acq_cells_per_s"
UNIT = "cells/s"


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class RawCuda:
    """A device buffer of the engine (raw pointer from the C-ABI) as something torch.as_tensor understands."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def ncu_dram_bytes(path, tag):
    """dram__bytes_read + dram__bytes_write of one captured launch from a committed ncu summary under profiles/ (None if absent)."""
    try:
        txt = open(os.path.join(ROOT, "profiles", path)).read()
        blk = txt[txt.index("== " + tag + ":"):]
        nxt = blk.find("\n== ", 4)
        blk = blk if nxt < 0 else blk[:nxt]
        tot = 0.0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            line = [ln for ln in blk.splitlines() if key + " " in ln + " "][0].split()
            tot += float(line[-2]) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[line[-1]]
        return tot
    except Exception:
        return None


def ncu_dram_bytes_multicorr(name):
    """dram bytes of one 2048-channel multicorrelator launch from a committed ncu summary under profiles/ (None if absent)."""
    try:
        tot = 0.0
        for ln in open(os.path.join(ROOT, "profiles", name)):
            f = ln.replace(",", " ").split()
            for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                if key in f:
                    i = f.index(key)
                    nums = [x for x in f[i + 1:] if x.replace(".", "", 1).isdigit()]
                    unit = [x for x in f[i + 1:] if x in ("byte", "Kbyte", "Mbyte", "Gbyte")]
                    if nums and unit:
                        tot += float(nums[0]) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit[0]]
        return tot if tot > 0 else None
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def acq_flops_per_cell(n):
    return 10.0 * math.log2(n) + 17.0


# ----------------------------------------------------------------------------------------------
# reference arm: the CPU implementation of the same path on the host cores
# ----------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from b200_import import load_oracle
    load_oracle()
    import oracle_fast
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        time.sleep(8.0)  # the other ranks of the launcher exit without work: let their interpreter start-up drain off the host cores
    geo, codes, iq, _ = synth.config_scene(args.workload)
    cores = oracle_fast.host_cores()
    n_prn_sample = min(geo["n_prn"], max(cores, 8))  # bounded sample: one PRN per host thread, at least 8
    fa = oracle_fast.FastAcq(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                             geo["samples_per_chip"], use_cfar=True, workers=cores)
    specs = [fa.code_spectrum(c) for c in codes[:n_prn_sample]]
    cells = n_prn_sample * fa.nD * fa.eff
    for _ in range(args.warmup):
        fa.dwell_all_like_reference(iq, specs)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fa.dwell_all_like_reference(iq, specs)
    dt = (time.perf_counter() - t0) / args.steps
    value = cells / dt
    sample = (f"{n_prn_sample} of {geo['n_prn']} PRNs x {fa.nD} Doppler bins x {fa.eff} code phases per step, one thread per PRN "
              f"(reference structure: per-channel forward+inverse FFT); numpy/scipy pocketfft complex64 port of "
              f"pcps_acquisition::acquisition_core — CPU FFT = pocketfft, not FFTW (FFTW/GNU Radio/VOLK not installable here)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, geo, fa.nD),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def connect_peers(eng, dist, world, rank):
    """CUDA IPC handles of every rank's exchange region, gathered on the host once; False when this box refuses peer mappings."""
    ok = True
    try:
        hd, _ = eng.peer_setup(rank)
    except Exception as ex:   # noqa: BLE001
        hd, ok = b"", False
        sys.stderr.write(f"bench.py: peer exchange unavailable on rank {rank}: {ex}\n")
    handles = [None] * world
    dist.all_gather_object(handles, hd)
    if ok and all(len(h) == 64 for h in handles):
        try:
            eng.peer_connect(ipc_handles=handles)
        except Exception as ex:   # noqa: BLE001
            ok = False
            sys.stderr.write(f"bench.py: peer exchange unavailable on rank {rank}: {ex}\n")
    else:
        ok = False
    flags = [None] * world
    dist.all_gather_object(flags, ok)
    return all(flags)


def workload_config(name, geo, n_doppler, world=1, exchange="nccl"):
    desc = {"C2": "BASELINE configs[1]: Galileo E1b PCPS acquisition, 4 ms code, 8 MS/s, +-5 kHz / 125 Hz, 50 PRNs",
            "C1": "BASELINE configs[0]: GPS L1 C/A PCPS acquisition, 1 ms, 4 MS/s, +-5 kHz / 250 Hz, 32 PRNs"}.get(name, name)
    return {"workload": desc, "fft_size": geo["fft_size"], "n_doppler_bins": n_doppler, "n_prn": geo["n_prn"],
            "cells_per_step": geo["n_prn"] * n_doppler * geo["fft_size"],
            "l2_policy": "L2 flushed (256 MiB write) between timed iterations",
            "parallelism": "one GPU; the search's five launches replayed as ONE CUDA graph (gnss_b200_acq_capture_*), per-kernel times taken eagerly afterwards" if world == 1 else
                           f"ONE search, its {n_doppler} Doppler rows sharded over {world} GPUs ({-(-n_doppler // world)} rows x all PRNs each, no forward FFT computed twice); "
                           + ("row records stored into every rank's region over peer memory (NVLink stores + system-scope flags, no collective), decision on "
                              "every device; a rank's whole step (search + exchange + decision) replayed as ONE CUDA graph" if exchange == "peer" else
                              "row records all-gathered device-to-device (NCCL), decision on every device; a shard's dwell replayed as one CUDA graph")}


# ----------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from b200_import import load_package
    b = load_package()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or b.lib().gnss_b200_device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device; libgnss_b200 has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    hbm_peak, peak_src = measured_peaks()
    cs0 = ClockSampler(local)
    cs0.start()
    fp32_peak = max(b.fp32_peak_tflops(local) for _ in range(3))
    fp32_clock = cs0.stop()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def flush_l2():
        flush.fill_(rank + 1)

    # ------------------------------ acquisition ------------------------------
    geo, codes, iq, truth = synth.config_scene(args.workload)
    n_prn = geo["n_prn"]
    nD = int(math.ceil(2.0 * geo["doppler_max"] / geo["doppler_step"]))
    # N > 1: ONE search, its Doppler rows sharded over the ranks (rank r owns rows [r * chunk, (r + 1) * chunk))
    chunk = -(-nD // world)
    rb, re_ = (rank * chunk, min(nD, (rank + 1) * chunk)) if world > 1 else (0, 0)
    if world > 1 and rb >= re_:
        raise RuntimeError(f"bench.py: {world} ranks but only {nD} Doppler rows")
    eng = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                geo["samples_per_chip"], n_prn, use_cfar=True, device=local, row_begin=rb, row_end=re_)
    slots = np.arange(n_prn, dtype=np.uint32)
    eng.set_local_codes(slots, codes)
    eff = eng.eff
    assert eng.n_doppler == nD
    cells = n_prn * nD * eff                 # the whole search (all ranks together)
    cells_local = n_prn * eng.n_doppler_local * eff
    est = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local))
    d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
    h_iq = torch.from_numpy(iq.view(np.float32).copy()).pin_memory()
    h_iq_np = h_iq.numpy().view(np.complex64)
    send_t = recv_t = None
    if world > 1:
        ds, dr, _, ch = eng.exchange_buffers(world)
        assert ch == chunk
        send_t = torch.as_tensor(RawCuda(ds, n_prn * chunk * 16), device=torch.device("cuda", local))
        recv_t = torch.as_tensor(RawCuda(dr, world * n_prn * chunk * 16), device=torch.device("cuda", local))
    peer = world > 1 and args.exchange == "peer"
    if peer:
        peer = connect_peers(eng, dist, world, rank)
    align_t = recv_t[:4].view(torch.float32) if world > 1 else None   # engine-owned memory (nothing torch-owned lives on the engine's stream)

    def exchange_and_decide():
        """rows of all shards -> every rank, decision on the device: peer-memory stores + flags (no collective), or an NCCL all-gather
        between the engines' device buffers"""
        if peer:
            eng.peer_exchange_async(n_prn)
            return
        with torch.cuda.stream(est):
            dist.all_gather_into_tensor(recv_t, send_t)
        eng.finalize_gathered_async(n_prn)

    def step_device():
        eng.dwell_device_async(d_iq.data_ptr(), slots)
        if world > 1:
            exchange_and_decide()
        return eng.sync(n_prn)

    def step_e2e():
        """host sample block in, host records out: the call a receiver makes"""
        if world == 1:
            return eng.dwell(h_iq_np, slots)
        cur = torch.cuda.current_stream()
        if rank == 0:
            d_iq.copy_(h_iq, non_blocking=True)
        dist.broadcast(d_iq, src=0)
        est.wait_stream(cur)   # (torch-owned tensors stay on torch's stream: the engine's stream is not torch's to record on)
        eng.dwell_device_async(d_iq.data_ptr(), slots)
        exchange_and_decide()
        return eng.sync(n_prn)

    torch.cuda.synchronize()
    for _ in range(args.warmup):
        res = step_device()
    # sanity: the injected satellites are found (the bench refuses to time a wrong search)
    thr = b.acq_threshold(0.001, eff, nD)
    found = {int(p) + 1 for p in range(n_prn) if res[p]["test_statistic"] > thr}
    want = {t["prn"] for t in truth if t["cn0"] > 41.0}
    if not want <= found:
        raise RuntimeError(f"bench.py: search is wrong, strong PRNs {sorted(want - found)} not detected")
    use_graph = True   # the search's launches go out as one CUDA graph (gnss_b200_acq_capture_*); per-kernel times are taken afterwards, eagerly
    if use_graph:
        eng.capture_begin()
        eng.dwell_device_async(d_iq.data_ptr(), slots)
        if peer:
            eng.peer_exchange_async(n_prn)     # the exchange and the decision are kernels too: the rank's whole step is ONE graph
        eng.capture_end()
        barrier()
        eng.replay_async()
        if world > 1 and not peer:
            exchange_and_decide()
        if eng.sync(n_prn).tobytes() != res.tobytes():
            raise RuntimeError("bench.py: the replayed dwell differs from the recorded one")

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if not use_graph:
        eng.profile(True)
    l0 = eng.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    wall0 = time.perf_counter()
    for i in range(args.steps):
        flush_l2()
        barrier()
        if world > 1 and peer:
            # the host barrier releases the ranks tens of microseconds apart, and so does a collective; the engine's own rendezvous
            # kernel (flags in every rank's region) right before the start event aligns the ranks ON THE DEVICE to within a store
            # latency, so the timed region holds the search and its exchange, not that skew
            eng.peer_barrier_async()
        elif world > 1:
            with torch.cuda.stream(est):
                dist.all_reduce(align_t)
        ev[i][0].record(est)
        if use_graph:
            eng.replay_async()
        else:
            eng.dwell_device_async(d_iq.data_ptr(), slots)
        if world > 1 and not (use_graph and peer):
            exchange_and_decide()
        ev[i][1].record(est)
        res = eng.sync(n_prn)
    barrier()
    wall = time.perf_counter() - wall0
    dev_ms = sum(a.elapsed_time(c) for a, c in ev) / args.steps
    dev_ms = max_over_ranks(dev_ms)
    launches = eng.launch_count() - l0
    if use_graph:   # per-kernel times of this rank's shard, from a few un-graphed dwells after the timed region
        eng.profile(True)
        for _ in range(5):
            flush_l2()
            torch.cuda.synchronize()   # the flush runs on torch's stream: keep it out of the engine stream's first kernel
            step_device()
        prof_steps = 5
    else:
        prof_steps = args.steps
    prof_ms, prof_cnt = eng.profile(False)
    value = cells / (dev_ms * 1e-3)

    # e2e through the host-buffer path
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
    e2e = cells / e2e_s
    clocks = sampler.stop() if rank == 0 else None

    # dominant acquisition kernel and rooflines (this rank's shard)
    names = ["fwd_fft_stages", "inv_first_stage", "inv_last_stage_mag_reduce", "code_spectra", "decision"]
    dom = int(np.argmax(prof_ms[:5]))
    dom_ms = prof_ms[dom] / max(1, prof_cnt[dom])
    cells_per_launch = cells_local * prof_steps / max(1, prof_cnt[dom])
    alg_bytes_per_cell = 8.0 / nD  # sample block + code spectra + row records per dwell
    achieved_gbs = cells_per_launch * alg_bytes_per_cell / (dom_ms * 1e-3) / 1e9
    fp = acq_flops_per_cell(geo["fft_size"])
    # executed arithmetic: the inverse transform of every (PRN, bin) row, the forward transform once per bin (shared by all PRNs)
    lg = math.log2(geo["fft_size"])
    fp_exec = (5.0 * lg + 11.0) + (5.0 * lg + 6.0) / n_prn
    row_resident = prof_cnt[1] == 0 and geo["fft_size"] == 32000   # the row-resident inverse kernel ran (no "first stage -> scratch" launches)
    traffic, traffic_src = None, "no ncu capture of this kernel variant under profiles/"
    if row_resident and dom == 2:
        per_dwell = None
        for tag in ("v7", "v6", "v4"):                                    # the newest capture of this kernel; one launch = 4000 rows
            per_dwell = ncu_dram_bytes("r02_cluster_ncu_summary.txt", tag)
            if per_dwell is not None:
                break
        if per_dwell is not None:
            traffic = per_dwell * (n_prn * eng.n_doppler_local) / 4000.0
            traffic_src = f"ncu --set full, profiles/r02_cluster_ncu_summary.txt ({tag}: dram read + write of one 4000-row launch, scaled to the rows of this launch)"
    roofline = {"kernel": names[dom] + (" (row-resident: cluster + DSMEM)" if row_resident and dom == 2 else ""), "bound": "hbm", "achieved": achieved_gbs,
                "peak": hbm_peak, "unit": "GB/s", "frac": achieved_gbs / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src,
                "algorithmic_bytes_per_cell": alg_bytes_per_cell, "cells_per_launch": cells_per_launch, "avg_launch_ms": dom_ms,
                "note": "fused search: algorithmic HBM traffic is ~8/nDoppler B per cell, so HBM is not the binding roofline; see roofline_fp32"}
    step_kernel_ms = sum(prof_ms[:5]) / prof_steps
    roofline_fp32 = {"bound": "fp32", "achieved": cells * fp / (dev_ms * 1e-3) / 1e12, "peak": fp32_peak * world, "unit": "TFLOP/s",
                     "frac": cells * fp / (dev_ms * 1e-3) / 1e12 / (fp32_peak * world), "flops_per_cell": fp,
                     "executed_flops_per_cell": fp_exec, "frac_executed": cells * fp_exec / (dev_ms * 1e-3) / 1e12 / (fp32_peak * world),
                     "peak_source": f"measured FFMA loop (gnss_b200_debug_fp32_peak) at {fp32_clock.get('sm_mhz')} MHz; nominal 148 x 128 x 2 x 1.965 GHz = 74.5",
                     "peak_clocks": fp32_clock,
                     "kernel_ms_per_step_this_rank": {names[i]: prof_ms[i] / prof_steps for i in range(5)}, "kernel_ms_sum": step_kernel_ms,
                     "note": "frac = algorithmic flops (SURVEY 8.4: 10 log2 N + 17 per cell, a forward FFT per row as the reference does) / ms_per_step / peak; "
                             "frac_executed counts what the engine really computes (forward FFTs shared by all PRNs)"}

    # ------------------------------ tracking (config 5, per-GPU share) ------------------------------
    trk = bench_tracking(b, torch, est.device, args, world, rank, local, flush_l2, barrier, max_over_ranks, hbm_peak, peak_src)

    trk_loop = bench_tracking_loop(b, torch, args, world, rank, local, barrier, max_over_ranks)

    # ------------------------------ multi-GNSS cold start, all GPUs on ONE search (strong scaling) ------------------------------
    cold = bench_cold_start(b, torch, dist, args, world, rank, local, barrier, max_over_ranks)

    # ------------------------------ weak-signal search (config 3), Doppler rows sharded, grids stay sharded ------------------------------
    weak = bench_weak_signal(b, torch, dist, args, world, rank, local, barrier, max_over_ranks, hbm_peak) if not args.no_weak else None

    # ------------------------------ small grid (config 1): launch-bound, eager launches vs one CUDA-graph replay ------------------------------
    small = bench_small_grid(b, torch, args, local) if (rank == 0 and world == 1 and args.workload != "C1") else None

    # ------------------------------ CPU baseline on rank 0 ------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.workload, geo, codes, iq)
        trk["cpu_baseline"] = cpu_baseline_tracking()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.workload, geo, nD, world, "peer" if peer else "nccl"),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(iq.nbytes), "d2h_bytes_per_step": int(n_prn * 32), "ms_per_step": e2e_s * 1e3,
                    "path": "gnss_b200_acq_dwell (host block in, records out)" if world == 1 else
                            "host block -> rank 0 (H2D) -> NCCL broadcast -> sharded dwell -> row records to every rank (" + ("peer memory" if peer else "NCCL all-gather") +
                            ") -> decision on the device -> records to the host"},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_fp32": roofline_fp32,
            "wall_ms_per_step_incl_flush": wall / args.steps * 1e3, "tracking": trk, "tracking_loop": trk_loop, "cold_start": cold,
            "weak_signal": weak, "small_grid": small,
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
        sys.stdout.flush()
    # orderly teardown: the tensors that alias the engine's exchange buffers, then the engine, then the process group
    send_t = recv_t = align_t = None
    eng.close()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bench_tracking(b, torch, device, args, world, rank, local, flush_l2, barrier, max_over_ranks, hbm_peak, peak_src):
    """BASELINE config 5: 2048 channels x 11 taps, GPS L5I at 30 MS/s, 1 ms (30 000 samples), the 2048 channels sharded over
    the GPUs (256 per GPU at N = 8): strong scaling, no exchange between the shards (channels are independent)."""
    n_ch, n, T = 2048 // world, 30000, 11
    rng = np.random.default_rng(5 + rank)
    chips = synth.load_chips()["gps_l5i"]
    mc = b.MulticorrelatorB200(n_ch, n, T, device=local)
    for c in range(n_ch):
        mc.set_local_code(c, chips[c % 32].astype(np.float32))
    shifts = np.tile(np.array([-1, -0.5, -0.25, -0.1, -0.05, 0, 0.05, 0.1, 0.25, 0.5, 1], np.float32), (n_ch, 1))
    fd = rng.uniform(-5000, 5000, n_ch)
    args_arr = mc.make_args(n_ch, n, rng.uniform(0, 2 * np.pi, n_ch), 2 * np.pi * fd / 30e6, 0.0, rng.uniform(0, 1, n_ch),
                            10.23e6 * (1 + fd / 1176.45e6) / 30e6, 0.0)
    g = torch.Generator(device="cuda")
    g.manual_seed(7 + rank)
    d_sig = torch.randn((n_ch, n, 2), generator=g, device="cuda", dtype=torch.float32)
    chans = np.arange(n_ch, dtype=np.uint32)
    est = torch.cuda.ExternalStream(mc.stream(), device=torch.device("cuda", local))
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        mc.correlate_device_async(chans, args_arr, shifts, d_sig.data_ptr(), n)
        mc.sync(n_ch, T)
    mc.profile(True)
    steps = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    for i in range(steps):
        if n_ch * n * 8 < (200 << 20):
            flush_l2()
            torch.cuda.synchronize()
        ev[i][0].record(est)
        mc.correlate_device_async(chans, args_arr, shifts, d_sig.data_ptr(), n)
        ev[i][1].record(est)
        mc.sync(n_ch, T)
    barrier()
    dev_ms = max_over_ranks(sum(a.elapsed_time(c) for a, c in ev) / steps)
    pm, pc = mc.profile(False)
    k_ms = pm[0] / max(1, pc[0])
    # e2e: host sample buffers (pinned, contiguous) -> H2D inside the call
    e2e_ch = min(n_ch, 512)
    h_sig = torch.randn((e2e_ch, n, 2), dtype=torch.float32).pin_memory()
    base = h_sig.data_ptr()
    a2 = mc.make_args(e2e_ch, n, rng.uniform(0, 2 * np.pi, e2e_ch), 2 * np.pi * fd[:e2e_ch] / 30e6, 0.0, rng.uniform(0, 1, e2e_ch),
                      10.23e6 * (1 + fd[:e2e_ch] / 1176.45e6) / 30e6, 0.0, sig_ptrs=[base + c * n * 8 for c in range(e2e_ch)])
    mc.correlate_batch(chans[:e2e_ch], a2, shifts[:e2e_ch])
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, steps // 4)
    for _ in range(e2e_steps):
        mc.correlate_batch(chans[:e2e_ch], a2, shifts[:e2e_ch])
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
    # e2e in the receiver's own input mode: every tracking channel reads the SAME signal-conditioner stream
    # (dll_pll_veml_tracking::general_work consumes one gr_complex stream per channel, all fed by one source), so an epoch
    # uploads two code periods of the antenna stream once and every channel correlates its own window of it
    h_stream = torch.randn((2 * n, 2), dtype=torch.float32).pin_memory()
    h_stream_np = h_stream.numpy().view(np.complex64).reshape(-1)
    offs = rng.integers(0, n, n_ch)
    a3 = mc.make_args(n_ch, n, rng.uniform(0, 2 * np.pi, n_ch), 2 * np.pi * fd / 30e6, 0.0, rng.uniform(0, 1, n_ch),
                      10.23e6 * (1 + fd / 1176.45e6) / 30e6, 0.0, stream_offsets=offs)
    for _ in range(2):
        mc.set_stream(h_stream_np)
        mc.correlate_batch(chans, a3, shifts)
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        mc.set_stream(h_stream_np)
        mc.correlate_batch(chans, a3, shifts)
    barrier()
    sh_s = max_over_ranks((time.perf_counter() - t0) / steps)
    alg_bytes = n_ch * (8.0 * n + 8.0 * T)
    mc_traffic, mc_traffic_file = None, None
    for cand in ("r02_multicorr_ncu_full_summary.txt", "r01_multicorr_ncu_full_summary.txt"):   # the newest capture of this kernel
        mc_traffic = ncu_dram_bytes_multicorr(cand)
        if mc_traffic:
            mc_traffic_file = cand
            break
    mc_traffic = mc_traffic * n_ch / 2048.0 if mc_traffic else None
    return {
        "metric": "tracking_channel_ms_per_s", "value": world * n_ch / (dev_ms * 1e-3), "unit": "channel-ms/s", "ms_per_step": dev_ms,
        "scaling": "strong", "n_gpus": world,
        "config": {"workload": "BASELINE configs[4]: multicorrelator, GPS L5I, 30 MS/s, 1 ms, 11 taps, 2048 channels in total", "channels_per_gpu": n_ch,
                   "samples": n, "taps": T, "parallelism": "one GPU" if world == 1 else f"2048 channels sharded over {world} GPUs, no exchange",
                   "input_mode": f"independent per-channel sample streams resident in HBM ({n_ch * n * 8 / 1e6:.0f} MB per GPU; L2 flushed between steps when that fits L2)"},
        "e2e": {"value": world * n_ch / sh_s, "unit": "channel-ms/s", "channels": n_ch, "h2d_bytes_per_step": int(2 * n * 8),
                "d2h_bytes_per_step": int(n_ch * T * 8), "ms_per_step": sh_s * 1e3,
                "input_mode": "one shared antenna stream per epoch (H2D once), each channel correlates its own window — the receiver's case"},
        "e2e_independent_streams": {"value": world * e2e_ch / e2e_s, "unit": "channel-ms/s", "channels": e2e_ch, "h2d_bytes_per_step": int(e2e_ch * n * 8),
                                    "d2h_bytes_per_step": int(e2e_ch * T * 8), "ms_per_step": e2e_s * 1e3,
                                    "input_mode": "a separate host sample buffer per channel (PCIe-bound: 240 kB per channel-ms)"},
        "roofline": {"kernel": "multicorr_window_kernel", "bound": "hbm", "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak, "traffic": mc_traffic,
                     "traffic_source": f"ncu --set full, profiles/{mc_traffic_file} (dram read + write of one 2048-channel launch, scaled to this launch's channels)"
                                       if mc_traffic else "no ncu capture under profiles/",
                     "peak_source": peak_src, "avg_launch_ms": k_ms,
                     "algorithmic_bytes_per_channel_ms": 8.0 * n + 8.0 * T},
        "gpu_launches": int(pc[0]),
    }


def bench_cold_start(b, torch, dist, args, world, rank, local, barrier, max_over_ranks):
    """BASELINE config 4 / north_star target: ONE multi-GNSS cold start — GPS L1 C/A (32 PRNs, N = 25 000) + Galileo E1b
    (36 PRNs, N = 100 000) + BeiDou B1I (63 PRNs, N = 25 000) at 25 MS/s, 40 Doppler bins each = 239 M cells — searched by
    all N GPUs together (STRONG scaling: the total work is fixed). Doppler-major partition: every rank takes 40 / N rows of
    ALL 131 PRNs, so no forward transform is computed twice and every rank carries the same share of every constellation.
    Inside the device-timed region: H2D of the 800 kB sample block on rank 0, NCCL broadcast, the three constellations' dwells
    (one CUDA graph each, concurrently on their engines' streams), NCCL all-gather of the row records straight between the
    engines' device buffers, the decision on the device, the result copy."""
    names = ["C4_gps", "C4_gal", "C4_bds"]
    scenes = {k: synth.config_scene(k) for k in names}
    truth = [t for k in names for t in scenes[k][3]]
    n_big = max(scenes[k][0]["consumed_samples"] for k in names)
    fs = scenes["C4_gal"][0]["fs"]
    iq = synth.make_iq(n_big, fs, truth, seed=44)      # one antenna stream holding all three constellations
    dev = torch.device("cuda", local)
    cur = torch.cuda.current_stream()
    h_iq = torch.from_numpy(iq.view(np.float32).copy()).pin_memory()
    d_iq = torch.empty(2 * n_big, dtype=torch.float32, device="cuda")
    engines, total_cells = [], 0
    for k in names:
        geo, codes = scenes[k][0], scenes[k][1]
        nD = int(math.ceil(2.0 * geo["doppler_max"] / geo["doppler_step"]))
        chunk = -(-nD // world)
        rb, re_ = (rank * chunk, min(nD, (rank + 1) * chunk)) if world > 1 else (0, 0)
        n = geo["n_prn"]
        e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                  geo["samples_per_chip"], n, use_cfar=True, device=local, row_begin=rb, row_end=re_)
        slots = np.arange(n, dtype=np.uint32)
        e.set_local_codes(slots, codes)
        send_t = recv_t = None
        if world > 1:
            ds, dr, _, ch = e.exchange_buffers(world)
            send_t = torch.as_tensor(RawCuda(ds, n * ch * 16), device=dev)
            recv_t = torch.as_tensor(RawCuda(dr, world * n * ch * 16), device=dev)
        pe = world > 1 and args.exchange == "peer" and connect_peers(e, dist, world, rank)
        engines.append(dict(name=k, geo=geo, e=e, n=n, slots=slots, est=torch.cuda.ExternalStream(e.stream(), device=dev), send=send_t, recv=recv_t, graph=False, peer=pe))
        total_cells += n * nD * e.eff

    def search():
        d_iq.copy_(h_iq, non_blocking=True) if rank == 0 else None     # the antenna block enters on rank 0 ...
        if world > 1:
            dist.broadcast(d_iq, src=0)                                # ... and goes to every GPU over NVLink
        for g in engines:
            g["est"].wait_stream(cur)
            if g["graph"]:
                g["e"].replay_async()          # with the peer exchange: search + record exchange + decision in this one graph
            else:
                g["e"].dwell_device_async(d_iq.data_ptr(), g["slots"])
                if g["peer"]:
                    g["e"].peer_exchange_async(g["n"])
        if world > 1:
            for g in engines:
                if g["peer"]:
                    continue
                with torch.cuda.stream(g["est"]):
                    dist.all_gather_into_tensor(g["recv"], g["send"])
                g["e"].finalize_gathered_async(g["n"])
        for g in engines:
            cur.wait_stream(g["est"])

    def fetch():
        return [g["e"].sync(g["n"]) for g in engines]

    for _ in range(max(3, args.warmup)):
        search()
        first = fetch()
    for g in engines:   # every engine's dwell as one graph launch
        g["e"].capture_begin()
        g["e"].dwell_device_async(d_iq.data_ptr(), g["slots"])
        if g["peer"]:
            g["e"].peer_exchange_async(g["n"])
        g["e"].capture_end()
        g["graph"] = True
    search()
    res = fetch()
    if any(x.tobytes() != y.tobytes() for x, y in zip(first, res)):
        raise RuntimeError("bench.py cold start: the replayed dwells differ from the recorded ones")
    steps = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    l0 = sum(g["e"].launch_count() for g in engines)
    barrier()
    wall = []
    for i in range(steps):
        barrier()
        t0 = time.perf_counter()
        ev[i][0].record(cur)
        search()
        ev[i][1].record(cur)
        res = fetch()
        wall.append(time.perf_counter() - t0)
    barrier()
    dev_ms = max_over_ranks(sum(a.elapsed_time(c) for a, c in ev) / steps)
    wall_ms = max_over_ranks(float(np.mean(wall)) * 1e3)
    launches = sum(g["e"].launch_count() for g in engines) - l0
    # every rank holds the complete decision: every strong satellite is found where it was injected
    found, want = 0, 0
    for g, r in zip(engines, res):
        geo, e = g["geo"], g["e"]
        thr = b.acq_threshold(0.001, e.eff, e.n_doppler)
        one = int(round(geo["samples_per_code"]))
        for t in scenes[g["name"]][3]:
            if t["cn0"] < (47.0 if geo["sampled_ms"] == 1 else 41.0):
                continue
            want += 1
            x = r[t["prn"] - 1]
            if x["test_statistic"] > thr and abs(int(x["code_phase"]) % one - t["delay"]) <= 1 and abs(x["doppler_hz"] - t["doppler"]) <= geo["doppler_step"]:
                found += 1
    if found != want:
        raise RuntimeError(f"bench.py cold start: {found} of {want} strong satellites found at their injected cell")
    for g in engines:
        g["e"].close()
    return {"metric": "cold_start_search_ms", "value": dev_ms, "unit": "ms", "higher_is_better": False, "scaling": "strong", "n_gpus": world,
            "cells": int(total_cells), "cells_per_s": total_cells / (dev_ms * 1e-3), "wall_ms_host_clock": wall_ms,
            "satellites_checked": want, "gpu_launches_per_search_this_rank": launches / steps,
            "config": {"workload": "BASELINE configs[3]: GPS L1 C/A (32) + Galileo E1b (36) + BeiDou B1I (63) at 25 MS/s, +-5 kHz / 250 Hz, 1 dwell",
                       "parallelism": f"Doppler-major over {world} GPU(s): {-(-40 // world)} of 40 rows x all 131 PRNs per GPU; sample block H2D on rank 0 + NCCL broadcast, "
                                      + ("row records stored into every rank over peer memory (flags, no collective) and decided on the device inside ONE CUDA graph per constellation"
                                         if all(g["peer"] for g in engines) and world > 1 else
                                         "row records all-gathered device-to-device (NCCL), decision on the device, one CUDA graph per constellation") + " — all inside the timed region",
                       "h2d_bytes": int(iq.nbytes), "note": "north_star target: < 1 ms on 8 GPUs"}}


def bench_weak_signal(b, torch, dist, args, world, rank, local, barrier, max_over_ranks, hbm_peak):
    """BASELINE config 3: weak-signal GPS L1 C/A search — 20 ms coherent (N = 2 x 80 000 zero-padded), +-10 kHz / 25 Hz = 800
    Doppler rows, 32 PRNs, 10 non-coherent dwells. The persistent magnitude grids (16.4 GB) stay sharded by Doppler row over
    the GPUs for the whole search; after every dwell only the 16-byte row records are exchanged and every rank decides, as the
    reference tests its statistic after every dwell (pcps_acquisition.cc:781-866). A step = all 10 dwells. Strong scaling."""
    geo, codes, iq, truth = synth.config_scene("C3", n_dwells=10)
    n = geo["n_prn"]
    nD = int(math.ceil(2.0 * geo["doppler_max"] / geo["doppler_step"]))
    chunk = -(-nD // world)
    rb, re_ = (rank * chunk, min(nD, (rank + 1) * chunk)) if world > 1 else (0, 0)
    dev = torch.device("cuda", local)
    e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                              geo["samples_per_chip"], n, use_cfar=True, keep_grid=True, device=local, row_begin=rb, row_end=re_)
    slots = np.arange(n, dtype=np.uint32)
    e.set_local_codes(slots, codes)
    est = torch.cuda.ExternalStream(e.stream(), device=dev)
    d_iq = torch.from_numpy(np.ascontiguousarray(iq).view(np.float32).copy()).cuda()      # [10][80 000] c64: the whole 200 ms resident
    send_t = recv_t = None
    if world > 1:
        ds, dr, _, ch = e.exchange_buffers(world)
        send_t = torch.as_tensor(RawCuda(ds, n * ch * 16), device=dev)
        recv_t = torch.as_tensor(RawCuda(dr, world * n * ch * 16), device=dev)
    row_bytes = geo["consumed_samples"] * 8

    def search():
        res = None
        for d in range(10):
            e.dwell_device_async(d_iq.data_ptr() + d * row_bytes, slots)
            if world > 1:
                with torch.cuda.stream(est):
                    dist.all_gather_into_tensor(recv_t, send_t)
                e.finalize_gathered_async(n)
        return res

    def rearm():
        for s in range(n):
            e.reset(s)
        torch.cuda.synchronize()

    steps = max(2, min(3, args.steps // 8))
    rearm()
    search()
    res = e.sync(n)
    e.profile(True)
    times = []
    for i in range(steps):
        rearm()
        barrier()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(est)
        search()
        c.record(est)
        res = e.sync(n)
        times.append(a.elapsed_time(c))
    barrier()
    dev_ms = max_over_ranks(float(np.mean(times)))
    pm, pc = e.profile(False)
    thr = b.acq_threshold(0.001, e.eff, nD, max_dwells=10)
    found = sum(1 for t in truth if res[t["prn"] - 1]["test_statistic"] > thr and abs(int(res[t["prn"] - 1]["code_phase"]) % 4000 - t["delay"]) <= 1
                and abs(res[t["prn"] - 1]["doppler_hz"] - t["doppler"]) <= 25)
    if found != len(truth):
        raise RuntimeError(f"bench.py weak-signal search: {found} of {len(truth)} weak satellites found after 10 dwells")
    cells = n * nD * e.eff * 10
    cells_local = n * e.n_doppler_local * e.eff
    last_ms = pm[2] / (10.0 * steps)    # the last inverse stage per dwell (all its chunk launches): |.|^2 + grid read-modify-write (8 B per cell on dwells 2..10, 4 B on the first)
    e.close()
    return {"metric": "acq_cells_per_s", "value": cells / (dev_ms * 1e-3), "unit": "cells/s", "higher_is_better": True, "scaling": "strong", "n_gpus": world,
            "ms_per_step": dev_ms, "ms_per_dwell": dev_ms / 10.0, "cells_per_step": int(cells), "steps": steps, "satellites_found": found,
            "grid_bytes_per_gpu": int(4 * cells_local),
            "roofline": {"kernel": "inv_last_stage_mag_reduce (persistent grid)", "bound": "hbm", "achieved": cells_local * 7.6 / (last_ms * 1e-3) / 1e9, "peak": hbm_peak,
                         "unit": "GB/s", "frac": cells_local * 7.6 / (last_ms * 1e-3) / 1e9 / hbm_peak, "ms_per_dwell_this_kernel": last_ms,
                         "kernel_ms_per_dwell_this_rank": {"fwd_fft_stages": pm[0] / (10.0 * steps), "inv_first_stage": pm[1] / (10.0 * steps), "inv_last_stage_mag_reduce": last_ms},
                         "algorithmic_bytes_per_cell": 7.6, "note": "grid traffic only: 4 B/cell on the first dwell, 8 B/cell on the nine later ones"},
            "config": {"workload": "BASELINE configs[2]: GPS L1 C/A, 4 MS/s, 20 ms coherent x 10 non-coherent dwells, +-10 kHz / 25 Hz (800 rows), 32 PRNs, N = 160 000",
                       "parallelism": "one GPU" if world == 1 else f"800 Doppler rows sharded over {world} GPUs ({chunk} rows x 32 PRNs each); grids stay sharded, "
                                      "row records all-gathered (NCCL) and decided on the device after every dwell"}}


def bench_small_grid(b, torch, args, local):
    """BASELINE config 1 (GPS L1 C/A, 4 MS/s, 1 ms, 40 bins x 32 PRNs = 5.12 M cells): five short kernels per dwell, so the search is
    bound by launch latency — timed as eager launches and as one CUDA-graph replay (gnss_b200_acq_capture_*), L2 flushed between."""
    geo, codes, iq, truth = synth.config_scene("C1")
    n = geo["n_prn"]
    e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], n,
                              use_cfar=True, device=local)
    slots = np.arange(n, dtype=np.uint32)
    e.set_local_codes(slots, codes)
    d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
    est = torch.cuda.ExternalStream(e.stream())
    flush = torch.empty(192 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        e.dwell_device_async(d_iq.data_ptr(), slots)
        ref = e.sync(n)
    l0 = e.launch_count()
    e.capture_begin()
    e.dwell_device_async(d_iq.data_ptr(), slots)
    e.capture_end()
    per_dwell = e.launch_count() - l0

    def timed(fn):
        ts = []
        for _ in range(max(10, args.steps)):
            flush.fill_(1)
            torch.cuda.synchronize()
            a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(est)
            fn()
            z.record(est)
            z.synchronize()
            ts.append(a.elapsed_time(z))
        return float(np.mean(ts))

    eager = timed(lambda: e.dwell_device_async(d_iq.data_ptr(), slots))
    graph = timed(e.replay_async)
    got = e.sync(n)
    same = bool(np.array_equal(got["test_statistic"], ref["test_statistic"]) and np.array_equal(got["code_phase"], ref["code_phase"]))
    cells = n * e.n_doppler * geo["fft_size"]
    out = {"workload": "C1: GPS L1 C/A, 4 MS/s, 1 ms, 40 bins x 32 PRNs", "cells": cells, "kernels_per_dwell": int(per_dwell),
           "eager_ms_per_dwell": eager, "graph_ms_per_dwell": graph, "cells_per_s_eager": cells / (eager * 1e-3), "cells_per_s_graph": cells / (graph * 1e-3),
           "replay_equals_eager": same, "l2": "flushed (192 MiB write) before every dwell"}
    del flush
    e.close()
    return out


def bench_tracking_loop(b, torch, args, world, rank, local, barrier, max_over_ranks):
    """the dll_pll_veml loop closed on the device. 2048 GPS L5I channels per GPU at 30 MS/s track the
    satellites of one shared antenna stream; ONE C-ABI call takes a 20 ms block of HOST samples and returns the per-epoch
    records of every channel (correlate -> lock detectors -> discriminators -> loop filters -> NCO update per code period,
    two launches per epoch for all channels, no host round trip in between). e2e by construction: H2D of the block and
    D2H of the records are inside the timed call."""
    fs, n_ms, n_ch, n_sat = 30e6, 20, 2048, 16
    prns = list(range(1, n_sat + 1))
    iq, truth, carrier = synth.tracking_stream("gps_l5i", prns, fs, n_ms, seed=300 + rank, cn0=48.0, data_bits=False)
    cfg = b.trk_config(fs, 10.23e6, 10230, carrier, n_ch, veml=False, enable_fll_pull_in=1, pull_in_time_s=0, device=local)
    trk = b.TrackingLoopB200(cfg)
    for c in range(n_ch):
        trk.set_code(c, synth.tracking_code("gps_l5i", prns[c % n_sat]))
    h_iq = torch.from_numpy(iq.view(np.float32).copy()).pin_memory()
    h_np = h_iq.numpy().view(np.complex64)
    chans = np.arange(n_ch, dtype=np.uint32)
    n_ep = n_ms - 2

    def start_all():
        for c in range(n_ch):
            t = truth[c % n_sat]
            trk.start(c, round(t["delay"]), t["doppler"] + 20.0, 0, 0)

    steps = max(3, args.steps // 4)
    start_all()
    rec = trk.run(h_np, 0, chans, n_ep)
    valid = int((rec["flags"] & 1).sum())
    last = np.array([np.flatnonzero(rec[:, c]["flags"] & 1)[-1] for c in range(n_ch)])   # last epoch each channel ran
    dop = rec["carrier_doppler_hz"][last, np.arange(n_ch)]
    want = np.array([truth[c % n_sat]["doppler"] for c in range(n_ch)])
    locked = int((np.abs(dop - want) < 25.0).sum())
    if locked < 0.98 * n_ch:
        raise RuntimeError(f"bench.py tracking loop: only {locked} of {n_ch} channels hold their satellite")
    times = []
    l0 = trk.launch_count()
    for _ in range(steps):
        start_all()
        barrier()
        t0 = time.perf_counter()
        rec = trk.run(h_np, 0, chans, n_ep)
        times.append(time.perf_counter() - t0)
    barrier()
    dt = max_over_ranks(float(np.mean(times)))
    launches = (trk.launch_count() - l0) / steps
    trk.close()
    return {"metric": "tracking_channel_ms_per_s", "value": world * valid / dt, "unit": "channel-ms/s", "ms_per_call": dt * 1e3,
            "epochs_per_call": n_ep, "channels_per_gpu": n_ch, "channel_epochs_per_call": valid, "channels_locked": locked,
            "h2d_bytes_per_call": int(iq.nbytes), "d2h_bytes_per_call": int(rec.nbytes), "gpu_launches_per_call": launches,
            "config": {"workload": "dll_pll_veml loop on device: GPS L5I, 30 MS/s, 1 ms epochs, E/P/L taps, 3rd-order PLL + 2nd-order DLL, "
                                   "2048 channels per GPU on one shared 20 ms antenna block (16 satellites)",
                       "timing": "host clock around ONE gnss_b200_trk_run call (host samples in, per-epoch records out)"}}


def cpu_baseline_tracking(n=30000, T=11):
    """The reference multicorrelator on the host cores: one thread per channel (the reference runs one tracking block per
    thread), each a Carrier_wipeoff_multicorrelator_resampler call through the reference's own classes compiled from
    /root/reference (oracle/_ref): over its AVX protokernels of the resampler and the rotator/dot-product — the variant
    volk_gnsssdr_profile selects on an AVX host, and the TIMED baseline — and over its generic kernels (the parity oracle's
    arithmetic), both reported. Falls back to the oracle port when oracle/_ref did not travel with the repo."""
    from concurrent.futures import ThreadPoolExecutor
    from b200_import import load_oracle
    om = load_oracle()
    import oracle_fast
    cores = oracle_fast.host_cores()
    variants = {}
    try:
        if os.path.exists(os.path.join(ROOT, "oracle/_ref/libref_mc.so")):
            rl = om.RefLibs()
            variants["reference classes, generic volk_gnsssdr kernels"] = ("reference", lambda *a: rl.mc_real(*a))
            if rl.mc_avx is not None:
                variants["reference classes, AVX volk_gnsssdr protokernels (resampler _u_avx, rotator_dot_prod _u_avx)"] = ("reference", lambda *a: rl.mc_real(*a, simd=True))
    except Exception:
        variants = {}
    if not variants:
        orc = om.Oracle("own")
        variants["oracle port of Cpu_Multicorrelator_Real_Codes (generic kernels)"] = ("port", lambda *a: orc.mc_real(*a))
    rng = np.random.default_rng(9)
    n_ch = max(cores, 8)
    sig = (rng.standard_normal((n_ch, n)) + 1j * rng.standard_normal((n_ch, n))).astype(np.complex64)
    code = synth.load_chips()["gps_l5i"][0].astype(np.float32)
    shifts = np.array([-1, -0.5, -0.25, -0.1, -0.05, 0, 0.05, 0.1, 0.25, 0.5, 1], np.float32)[:T]
    out = {}
    for name, (kind, fn) in variants.items():
        def one(c, fn=fn):
            return fn(sig[c], code, shifts, 0.3, 1e-3, 0.0, 0.4, 0.341, 0.0)

        with ThreadPoolExecutor(max_workers=cores) as ex:
            list(ex.map(one, range(n_ch)))
            reps, t0 = 0, time.perf_counter()
            while reps < 3 or time.perf_counter() - t0 < 4.0:
                list(ex.map(one, range(n_ch)))
                reps += 1
                if time.perf_counter() - t0 > 10.0:
                    break
            dt = (time.perf_counter() - t0) / reps
        out[name] = dict(value=n_ch / dt, kind=kind, reps=reps)
    best = max(out, key=lambda k: out[k]["value"])
    return {"value": out[best]["value"], "unit": "channel-ms/s", "cores": cores, "kind": out[best]["kind"],
            "sample": f"{n_ch} channels x {T} taps x {n} samples per pass, {out[best]['reps']} passes, one thread per channel; {best}",
            "variants": {k: v["value"] for k, v in out.items()}}


def cpu_baseline(workload, geo, codes, iq):
    """acquisition_core on the host cores, one thread per PRN like the reference's channels. Two FFT substitutes for the absent
    FFTW3f, both reported: pocketfft (scipy, complex64) — the faster one, the value of the line and of
    `--impl reference` — and the oracle's own C mixed-radix FFT with libm sine/cosine (oracle_acq.c), every thread a whole
    reference block instance."""
    from concurrent.futures import ThreadPoolExecutor
    from b200_import import load_oracle
    om = load_oracle()
    import oracle_fast
    cores = oracle_fast.host_cores()
    n_s = min(geo["n_prn"], max(cores, 8))
    fa = oracle_fast.FastAcq(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                             geo["samples_per_chip"], use_cfar=True, workers=cores)
    specs = [fa.code_spectrum(c) for c in codes[:n_s]]
    fa.dwell_all_like_reference(iq, specs)
    reps, t0 = 0, time.perf_counter()
    while reps < 3 or time.perf_counter() - t0 < 8.0:
        fa.dwell_all_like_reference(iq, specs)
        reps += 1
        if time.perf_counter() - t0 > 20.0:
            break
    dt = (time.perf_counter() - t0) / reps
    # the C leg: one oracle block instance per thread
    orc = om.Oracle("own")
    n_c = min(n_s, cores)
    blocks = []
    for p in range(n_c):
        a = orc.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"], doppler_step=geo["doppler_step"],
                    fs=geo["fs"], samples_per_chip=geo["samples_per_chip"], samples_per_code=geo["samples_per_code"], use_cfar=True)
        a.set_local_code(codes[p])
        blocks.append(a)

    def one(a):
        a.reset()
        return a.dwell(iq)

    with ThreadPoolExecutor(max_workers=cores) as ex:
        t1 = time.perf_counter()
        list(ex.map(one, blocks))
        dt_c = time.perf_counter() - t1
    for a in blocks:
        a.close()
    return {"value": n_s * fa.nD * fa.eff / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n_s} of {geo['n_prn']} PRNs x {fa.nD} bins x {fa.eff} code phases, {reps} repetitions, one thread per PRN; "
                      "numpy/scipy pocketfft complex64 port of acquisition_core (CPU FFT = pocketfft, not FFTW)",
            "variants": {"pocketfft (scipy.fft complex64), numpy float32 wipe-off": n_s * fa.nD * fa.eff / dt,
                         "C oracle: own mixed-radix FFT + libm sincosf, one block instance per thread": n_c * fa.nD * fa.eff / dt_c}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-weak", action="store_true", help="skip the config-3 (weak-signal, persistent grid) leg")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: row records over peer memory (stores + flags fused behind the search, gnss_b200_acq_peer_*) or an NCCL all-gather")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: file descriptor 1 is pointed at stderr for everything native code prints there
    # (NCCL's version banner at any NCCL_DEBUG level), Python's own sys.stdout keeps the real one
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
