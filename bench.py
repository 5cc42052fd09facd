#!/usr/bin/env python3
"""bench.py — acquisition cells/s (headline, BASELINE config 2) and tracking channel-ms/s (config 5) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C2|C1]

One "step" = one coherent dwell of the whole PRN x Doppler x code-phase grid of the workload on one batch of
synthetic IQ (ours: through libgnss_b200.so; reference: the CPU path on the host cores).
  value  : cells/s with the sample block already resident in HBM (device-timed with CUDA events on the engine stream)
  e2e    : cells/s through the C-ABI call a GNSS-SDR adapter makes, HOST sample buffer in, host results out
  N > 1  : one process per GPU (torchrun); every rank searches the full workload on its own sample block (weak scaling),
           per-PRN results are all-gathered with NCCL inside the step.
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
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, geo, fa.nD),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(name, geo, n_doppler):
    desc = {"C2": "BASELINE configs[1]: Galileo E1b PCPS acquisition, 4 ms code, 8 MS/s, +-5 kHz / 125 Hz, 50 PRNs",
            "C1": "BASELINE configs[0]: GPS L1 C/A PCPS acquisition, 1 ms, 4 MS/s, +-5 kHz / 250 Hz, 32 PRNs"}.get(name, name)
    return {"workload": desc, "fft_size": geo["fft_size"], "n_doppler_bins": n_doppler, "n_prn": geo["n_prn"],
            "cells_per_step_per_gpu": geo["n_prn"] * n_doppler * geo["fft_size"],
            "l2_policy": "L2 flushed (256 MiB write) between timed iterations", "parallelism": "one process per GPU, PRN x Doppler grid per GPU"}


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
    fp32_peak = b.fp32_peak_tflops(local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def flush_l2():
        flush.fill_(rank + 1)

    # ------------------------------ acquisition ------------------------------
    geo, codes, iq0, truth = synth.config_scene(args.workload)
    # each rank searches the whole grid on its own sample block (weak scaling): same satellites, rank-specific noise
    iq = iq0 if rank == 0 else synth.make_iq(geo["consumed_samples"], geo["fs"], truth, seed=1000 + rank)
    n_prn = geo["n_prn"]
    eng = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                geo["samples_per_chip"], n_prn, use_cfar=True, device=local)
    slots = np.arange(n_prn, dtype=np.uint32)
    eng.set_local_codes(slots, codes)
    nD, eff = eng.n_doppler, eng.eff
    cells = n_prn * nD * eff
    est = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local))
    d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
    h_iq = torch.from_numpy(iq.view(np.float32).copy()).pin_memory()
    h_iq_np = h_iq.numpy().view(np.complex64)
    gathered = torch.empty(world * n_prn * 32, dtype=torch.uint8, device="cuda") if world > 1 else None

    def gather(res):
        if world > 1:
            t = torch.from_numpy(res.view(np.uint8)).cuda(non_blocking=False)
            dist.all_gather_into_tensor(gathered, t)

    def step_device():
        eng.dwell_device_async(d_iq.data_ptr(), slots)
        res = eng.sync(n_prn)
        gather(res)
        return res

    def step_e2e():
        res = eng.dwell(h_iq_np, slots)
        gather(res)
        return res

    torch.cuda.synchronize()
    for _ in range(args.warmup):
        res = step_device()
    # sanity: the injected satellites are found (the bench refuses to time a wrong search)
    thr = b.acq_threshold(0.001, eff, nD)
    found = {int(p) + 1 for p in range(n_prn) if res[p]["test_statistic"] > thr}
    want = {t["prn"] for t in truth if t["cn0"] > 41.0}
    if not want <= found:
        raise RuntimeError(f"bench.py: search is wrong, strong PRNs {sorted(want - found)} not detected")

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    eng.profile(True)
    l0 = eng.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    wall0 = time.perf_counter()
    for i in range(args.steps):
        flush_l2()
        torch.cuda.synchronize()
        ev[i][0].record(est)
        eng.dwell_device_async(d_iq.data_ptr(), slots)
        ev[i][1].record(est)
        res = eng.sync(n_prn)
        gather(res)
    barrier()
    wall = time.perf_counter() - wall0
    dev_ms = sum(a.elapsed_time(c) for a, c in ev) / args.steps
    dev_ms = max_over_ranks(dev_ms)
    launches = eng.launch_count() - l0
    prof_ms, prof_cnt = eng.profile(False)
    value = world * cells / (dev_ms * 1e-3)

    # e2e through the host-buffer C-ABI call
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
    e2e = world * cells / e2e_s
    clocks = sampler.stop() if rank == 0 else None

    # dominant acquisition kernel and rooflines
    names = ["fwd_fft_stages", "inv_first_stage", "inv_last_stage_mag_reduce", "code_spectra", "decision"]
    dom = int(np.argmax(prof_ms[:5]))
    dom_ms = prof_ms[dom] / max(1, prof_cnt[dom])
    cells_per_launch = cells * args.steps / max(1, prof_cnt[dom])
    alg_bytes_per_cell = 8.0 / nD  # sample block + code spectra + row records per dwell
    achieved_gbs = cells_per_launch * alg_bytes_per_cell / (dom_ms * 1e-3) / 1e9
    step_kernel_ms = sum(prof_ms[:5]) / args.steps
    fp = acq_flops_per_cell(geo["fft_size"])
    # DRAM bytes per (PRN, Doppler) row of the two inverse kernels from the ncu --set full capture under profiles/
    # (r01_acq_ncu_full_summary.txt: dram__bytes_read + dram__bytes_write of an 800-row launch / 800): the inter-stage
    # intermediate (8 B/cell written by the first stage, read by the last) is what exceeds the algorithmic bytes
    ncu_bytes_per_row = {"inv_first_stage": (23.196160e6 + 150.798080e6) / 800.0, "inv_last_stage_mag_reduce": (216.051712e6 + 4.115968e6) / 800.0}
    traffic = ncu_bytes_per_row.get(names[dom])
    traffic = traffic * (geo["fft_size"] / 32000.0) * n_prn * nD * args.steps / max(1, prof_cnt[dom]) if traffic else None
    roofline = {"kernel": names[dom], "bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved_gbs / hbm_peak, "traffic": traffic, "traffic_source": "ncu --set full, profiles/r01_acq_ncu_full_summary.txt (bytes per launch)",
                "peak_source": peak_src,
                "algorithmic_bytes_per_cell": alg_bytes_per_cell, "cells_per_launch": cells_per_launch, "avg_launch_ms": dom_ms,
                "note": "fused search: algorithmic HBM traffic is ~8/nDoppler B per cell, so HBM is not the binding roofline; see roofline_fp32"}
    roofline_fp32 = {"bound": "fp32", "achieved": cells * fp / (step_kernel_ms * 1e-3) / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": cells * fp / (step_kernel_ms * 1e-3) / 1e12 / fp32_peak, "flops_per_cell": fp,
                     "peak_source": "measured FFMA loop (gnss_b200_debug_fp32_peak)",
                     "kernel_ms_per_step": {names[i]: prof_ms[i] / args.steps for i in range(5)},
                     "note": "algorithmic flops = SURVEY §8.4 (10 log2 N + 17 per cell, reference formulation with a forward FFT per row); "
                             "the engine shares forward FFTs across PRNs, so executed flops are about half"}

    # ------------------------------ tracking (config 5, per-GPU share) ------------------------------
    trk = bench_tracking(b, torch, est.device, args, world, rank, local, flush_l2, barrier, max_over_ranks, hbm_peak, peak_src)

    trk_loop = bench_tracking_loop(b, torch, args, world, rank, local, barrier, max_over_ranks)

    # ------------------------------ multi-GNSS cold start, all GPUs on ONE search (strong scaling) ------------------------------
    cold = bench_cold_start(b, torch, dist, args, world, rank, local, barrier, max_over_ranks)

    # ------------------------------ CPU baseline on rank 0 ------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.workload, geo, codes, iq)
        trk["cpu_baseline"] = cpu_baseline_tracking()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.workload, geo, nD),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(iq.nbytes), "d2h_bytes_per_step": int(n_prn * 32), "ms_per_step": e2e_s * 1e3},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_fp32": roofline_fp32,
            "wall_ms_per_step_incl_flush": wall / args.steps * 1e3, "tracking": trk, "tracking_loop": trk_loop, "cold_start": cold,
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def bench_tracking(b, torch, device, args, world, rank, local, flush_l2, barrier, max_over_ranks, hbm_peak, peak_src):
    """BASELINE config 5: 2048 channels x 11 taps, GPS L5I at 30 MS/s, 1 ms (30 000 samples); channels are sharded over the
    GPUs; every GPU takes a full 2048-channel epoch at every N so that per-GPU work is fixed — weak scaling)."""
    n_ch, n, T = 2048, 30000, 11
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
    return {
        "metric": "tracking_channel_ms_per_s", "value": world * n_ch / (dev_ms * 1e-3), "unit": "channel-ms/s", "ms_per_step": dev_ms,
        "config": {"workload": "BASELINE configs[4]: multicorrelator, GPS L5I, 30 MS/s, 1 ms, 11 taps", "channels_per_gpu": n_ch, "samples": n, "taps": T,
                   "input_mode": "independent per-channel sample streams resident in HBM (491 MB > L2 at 2048 channels)"},
        "e2e": {"value": world * n_ch / sh_s, "unit": "channel-ms/s", "channels": n_ch, "h2d_bytes_per_step": int(2 * n * 8),
                "d2h_bytes_per_step": int(n_ch * T * 8), "ms_per_step": sh_s * 1e3,
                "input_mode": "one shared antenna stream per epoch (H2D once), each channel correlates its own window — the receiver's case"},
        "e2e_independent_streams": {"value": world * e2e_ch / e2e_s, "unit": "channel-ms/s", "channels": e2e_ch, "h2d_bytes_per_step": int(e2e_ch * n * 8),
                                    "d2h_bytes_per_step": int(e2e_ch * T * 8), "ms_per_step": e2e_s * 1e3,
                                    "input_mode": "a separate host sample buffer per channel (PCIe-bound: 240 kB per channel-ms)"},
        "roofline": {"kernel": "multicorr_kernel", "bound": "hbm", "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak, "traffic": (575.693056e6 + 4.326144e6) * n_ch / 2048.0,
                     "traffic_source": "ncu --set full, profiles/r01_multicorr_ncu_full_summary.txt (bytes per launch)",
                     "peak_source": peak_src, "avg_launch_ms": k_ms,
                     "algorithmic_bytes_per_channel_ms": 8.0 * n + 8.0 * T},
        "gpu_launches": int(pc[0]),
    }


def bench_cold_start(b, torch, dist, args, world, rank, local, barrier, max_over_ranks):
    """BASELINE config 4 / north_star target: ONE multi-GNSS cold start — GPS L1 C/A (32 PRNs, N = 25 000) + Galileo E1b
    (36 PRNs, N = 100 000) + BeiDou B1I (63 PRNs, N = 25 000) at 25 MS/s, 40 Doppler bins each = 239 M cells — searched by
    all N GPUs together (STRONG scaling: the total work is fixed). PRN-major partition (gnss-sdr-new_b200/sharding.py);
    per search the 800 kB sample block is broadcast from rank 0 and the 32-byte per-PRN records are all-gathered, both over
    NCCL inside the timed region; the three constellations of a rank run concurrently on their engines' streams.
    Device-timed: CUDA events on torch's current stream bracket broadcast -> dwells -> gather."""
    names = ["C4_gps", "C4_gal", "C4_bds"]
    scenes = {k: synth.config_scene(k) for k in names}
    truth = [t for k in names for t in scenes[k][3]]
    n_big = max(scenes[k][0]["consumed_samples"] for k in names)
    fs = scenes["C4_gal"][0]["fs"]
    iq = synth.make_iq(n_big, fs, truth, seed=44)      # one antenna stream holding all three constellations
    cur = torch.cuda.current_stream()
    d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda() if rank == 0 else torch.empty(2 * n_big, dtype=torch.float32, device="cuda")
    engines, total_cells, max_local = [], 0, 0
    for k in names:
        geo, codes = scenes[k][0], scenes[k][1]
        mine = np.arange(rank, geo["n_prn"], world)
        max_local = max(max_local, -(-geo["n_prn"] // world))
        e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                                  geo["samples_per_chip"], max(1, mine.size), use_cfar=True, device=local)
        slots = np.arange(mine.size, dtype=np.uint32)
        if mine.size:
            e.set_local_codes(slots, codes[mine])
        engines.append((k, geo, e, mine, slots, torch.cuda.ExternalStream(e.stream(), device=torch.device("cuda", local))))
        total_cells += geo["n_prn"] * e.n_doppler * e.eff
    rec = 32
    local_buf = torch.zeros(3 * max_local * rec, dtype=torch.uint8, device="cuda")
    gathered = torch.zeros(world * local_buf.numel(), dtype=torch.uint8, device="cuda")
    h_local = torch.zeros(3 * max_local * rec, dtype=torch.uint8).pin_memory()

    def search():
        if world > 1:
            dist.broadcast(d_iq, src=0)
        for _, _, e, mine, slots, est in engines:
            if mine.size:
                est.wait_stream(cur)
                e.dwell_device_async(d_iq.data_ptr(), slots)
        res = []
        for i, (_, _, e, mine, slots, est) in enumerate(engines):
            r = e.sync(mine.size) if mine.size else None
            res.append(r)
            if r is not None:
                h_local.numpy()[i * max_local * rec: i * max_local * rec + r.nbytes] = r.view(np.uint8)
        if world > 1:
            local_buf.copy_(h_local, non_blocking=True)
            dist.all_gather_into_tensor(gathered, local_buf)
        return res

    for _ in range(max(3, args.warmup)):
        search()
    steps = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    l0 = sum(e.launch_count() for _, _, e, _, _, _ in engines)
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        ev[i][0].record(cur)
        search()
        ev[i][1].record(cur)
        ev[i][1].synchronize()
    barrier()
    wall_ms = (time.perf_counter() - t0) / steps * 1e3
    # the events bracket the device work of this rank; the host sync inside search() is part of a real search
    dev_ms = max_over_ranks(sum(a.elapsed_time(c) for a, c in ev) / steps)
    wall_ms = max_over_ranks(wall_ms)
    launches = sum(e.launch_count() for _, _, e, _, _, _ in engines) - l0
    # merge + check on rank 0: every strong satellite is found where it was injected
    if world > 1:
        allr = gathered.cpu().numpy().reshape(world, 3, max_local * rec)
    else:
        allr = h_local.numpy().reshape(1, 3, max_local * rec)
    found, want = 0, 0
    for i, (k, geo, e, _, _, _) in enumerate(engines):
        thr = b.acq_threshold(0.001, e.eff, e.n_doppler)
        one = int(round(geo["samples_per_code"]))
        for t in scenes[k][3]:
            if t["cn0"] < (47.0 if geo["sampled_ms"] == 1 else 41.0):
                continue
            want += 1
            owner, idx = (t["prn"] - 1) % world, (t["prn"] - 1) // world
            r = allr[owner, i, idx * rec:(idx + 1) * rec].view(b.ACQ_RESULT_DTYPE)[0]
            if r["test_statistic"] > thr and abs(int(r["code_phase"]) % one - t["delay"]) <= 1 and abs(r["doppler_hz"] - t["doppler"]) <= geo["doppler_step"]:
                found += 1
    if found != want:
        raise RuntimeError(f"bench.py cold start: {found} of {want} strong satellites found at their injected cell")
    for _, _, e, _, _, _ in engines:
        e.close()
    return {"metric": "cold_start_search_ms", "value": dev_ms, "unit": "ms", "higher_is_better": False, "scaling": "strong", "n_gpus": world,
            "cells": int(total_cells), "cells_per_s": total_cells / (dev_ms * 1e-3), "wall_ms_host_clock": wall_ms,
            "satellites_checked": want, "gpu_launches_per_search_this_rank": launches / steps,
            "config": {"workload": "BASELINE configs[3]: GPS L1 C/A (32) + Galileo E1b (36) + BeiDou B1I (63) at 25 MS/s, +-5 kHz / 250 Hz, 1 dwell",
                       "parallelism": f"PRN-major over {world} GPU(s); sample block broadcast + per-PRN records all-gathered (NCCL) inside the timed region",
                       "h2d_bytes": 0, "note": "north_star target: < 1 ms on 8 GPUs"}}


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
    thread), each a Carrier_wipeoff_multicorrelator_resampler call — through the reference's own classes and generic
    volk_gnsssdr kernels (oracle/_ref/libref_mc.so) when that library travelled with the repo, else the oracle port."""
    from concurrent.futures import ThreadPoolExecutor
    from b200_import import load_oracle
    om = load_oracle()
    import oracle_fast
    cores = oracle_fast.host_cores()
    kind, fn = "port", None
    try:
        if os.path.exists(os.path.join(ROOT, "oracle/_ref/libref_mc.so")):
            fn = om.RefLibs().mc_real
            kind = "reference"
    except Exception:
        fn = None
    if fn is None:
        fn = om.Oracle("own").mc_real
    rng = np.random.default_rng(9)
    n_ch = max(cores, 8)
    sig = (rng.standard_normal((n_ch, n)) + 1j * rng.standard_normal((n_ch, n))).astype(np.complex64)
    code = synth.load_chips()["gps_l5i"][0].astype(np.float32)
    shifts = np.array([-1, -0.5, -0.25, -0.1, -0.05, 0, 0.05, 0.1, 0.25, 0.5, 1], np.float32)[:T]

    def one(c):
        return fn(sig[c], code, shifts, 0.3, 1e-3, 0.0, 0.4, 0.341, 0.0)

    with ThreadPoolExecutor(max_workers=cores) as ex:
        list(ex.map(one, range(n_ch)))
        reps, t0 = 0, time.perf_counter()
        while reps < 3 or time.perf_counter() - t0 < 5.0:
            list(ex.map(one, range(n_ch)))
            reps += 1
            if time.perf_counter() - t0 > 15.0:
                break
        dt = (time.perf_counter() - t0) / reps
    return {"value": n_ch / dt, "unit": "channel-ms/s", "cores": cores, "kind": kind,
            "sample": f"{n_ch} channels x {T} taps x {n} samples per pass, {reps} passes, one thread per channel; "
                      + ("the reference's Cpu_Multicorrelator_Real_Codes over its generic volk_gnsssdr kernels (oracle/_ref)" if kind == "reference"
                         else "oracle port of Cpu_Multicorrelator_Real_Codes (generic kernels)")}


def cpu_baseline(workload, geo, codes, iq):
    from b200_import import load_oracle
    load_oracle()
    import oracle_fast
    cores = oracle_fast.host_cores()
    n_s = min(geo["n_prn"], max(cores, 8))
    fa = oracle_fast.FastAcq(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                             geo["samples_per_chip"], use_cfar=True, workers=cores)
    specs = [fa.code_spectrum(c) for c in codes[:n_s]]
    fa.dwell_all_like_reference(iq, specs)
    reps, t0 = 0, time.perf_counter()
    while reps < 3 or time.perf_counter() - t0 < 10.0:
        fa.dwell_all_like_reference(iq, specs)
        reps += 1
        if time.perf_counter() - t0 > 30.0:
            break
    dt = (time.perf_counter() - t0) / reps
    return {"value": n_s * fa.nD * fa.eff / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n_s} of {geo['n_prn']} PRNs x {fa.nD} bins x {fa.eff} code phases, {reps} repetitions, one thread per PRN; "
                      "numpy/scipy pocketfft complex64 port of acquisition_core (CPU FFT = pocketfft, not FFTW)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
