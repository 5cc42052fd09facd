#!/usr/bin/env python3
"""Every BASELINE.json config on one B200, device-timed (CUDA events on the engine stream, L2 flushed between steps).

  C1  GPS L1 C/A, 4 MS/s, 1 ms, 40 bins, 32 PRNs                         (N = 4000)
  C2  Galileo E1b, 8 MS/s, 4 ms, 80 bins, 50 PRNs                        (N = 32000)
  C3  weak-signal GPS L1 C/A, 20 ms x 10 dwells, 800 bins, persistent grid (N = 160000)  [n_prn from argv, default 8]
  C4  cold start at 25 MS/s: GPS (N = 25000, 32 PRNs) + Galileo (N = 100000, 36 PRNs) + BeiDou (N = 25000, 63 PRNs)
  C5  multicorrelator, 2048 channels x {3, 5, 11} taps, GPS L5I at 30 MS/s, 1 ms

Prints one JSON object; every acquisition entry also checks that the injected satellites come out at the injected
code phase / Doppler bin (a wrong search is not timed). Usage: python tools/config_sweep.py [c3_prns] [steps]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth  # noqa: E402
from b200_import import load_package  # noqa: E402
import torch  # noqa: E402

b = load_package()
c3_prns = int(sys.argv[1]) if len(sys.argv) > 1 else 8
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
fp32_peak = b.fp32_peak_tflops(0)
out = {"fp32_peak_tflops": fp32_peak, "configs": {}}


def check_truth(res, truth, geo, thr, min_cn0):
    one = int(round(geo["samples_per_code"]))
    for t in truth:
        if t["cn0"] < min_cn0:
            continue
        r = res[t["prn"] - 1]
        assert r["test_statistic"] > thr, (t, r)
        assert abs(int(r["code_phase"]) % one - t["delay"]) <= 1, (t, r)
        assert abs(r["doppler_hz"] - t["doppler"]) <= geo["doppler_step"], (t, r)


def acq_config(name, n_prn=None, min_cn0=41.0):
    geo, codes, iq, truth = synth.config_scene(name, n_prn=n_prn)
    n_prn = geo["n_prn"]
    e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                              geo["samples_per_chip"], n_prn, use_cfar=True)
    slots = np.arange(n_prn, dtype=np.uint32)
    e.set_local_codes(slots, codes)
    d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
    est = torch.cuda.ExternalStream(e.stream())
    for _ in range(3):
        e.dwell_device_async(d_iq.data_ptr(), slots)
        res = e.sync(n_prn)
    thr = b.acq_threshold(0.001, e.eff, e.n_doppler)
    check_truth(res, truth, geo, thr, min_cn0)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        ev[i][0].record(est)
        e.dwell_device_async(d_iq.data_ptr(), slots)
        ev[i][1].record(est)
        e.sync(n_prn)
    ms = sum(a.elapsed_time(c) for a, c in ev) / steps
    cells = n_prn * e.n_doppler * e.eff
    flops = 10.0 * np.log2(geo["fft_size"]) + 17.0
    e.close()
    return {"fft_size": geo["fft_size"], "n_doppler": e.n_doppler, "n_prn": n_prn, "cells": cells, "ms_per_dwell": ms,
            "cells_per_s": cells / (ms * 1e-3), "fp32_frac": cells * flops / (ms * 1e-3) / 1e12 / fp32_peak}


def c3_config(n_prn):
    n_dw = 10
    geo, codes, iq, truth = synth.config_scene("C3", n_prn=n_prn, n_dwells=n_dw)
    e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                              geo["samples_per_chip"], n_prn, use_cfar=True, keep_grid=True)
    slots = np.arange(n_prn, dtype=np.uint32)
    e.set_local_codes(slots, codes)
    d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
    est = torch.cuda.ExternalStream(e.stream())
    times, stats = [], None
    for rep in range(2):  # the first search warms up, the second is timed
        for s in slots:
            e.reset(int(s))
        times = []
        for k in range(n_dw):
            a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record(est)
            e.dwell_device_async(d_iq[k].data_ptr(), slots)
            c.record(est)
            res = e.sync(n_prn)
            times.append(a.elapsed_time(c))
        stats = res
    thr = b.acq_threshold(0.001, e.eff, e.n_doppler, max_dwells=n_dw)
    found = []
    for t in truth:
        r = stats[t["prn"] - 1]
        ok = bool(r["test_statistic"] > thr and abs(int(r["code_phase"]) % 4000 - t["delay"]) <= 1 and abs(r["doppler_hz"] - t["doppler"]) <= 25)
        found.append({"prn": t["prn"], "cn0": round(t["cn0"], 1), "stat": float(r["test_statistic"]), "found_at_truth": ok})
    assert all(f["found_at_truth"] for f in found), found
    assert int(stats[0]["dwell_count"]) == n_dw
    cells = n_prn * e.n_doppler * e.eff
    flops = 10.0 * np.log2(geo["fft_size"]) + 17.0
    first, later = times[0], float(np.mean(times[1:]))
    total = sum(times)
    e.close()
    return {"fft_size": geo["fft_size"], "n_doppler": e.n_doppler, "n_prn": n_prn, "dwells": n_dw, "cells_per_dwell": cells, "threshold": thr,
            "ms_first_dwell": first, "ms_later_dwell": later, "ms_search": total, "cells_per_s": cells * n_dw / (total * 1e-3),
            "fp32_frac": cells * n_dw * flops / (total * 1e-3) / 1e12 / fp32_peak,
            "grid_hbm_gbs_later_dwells": cells * 8.0 / (later * 1e-3) / 1e9, "detections": found,
            "note": "persistent grid: 4 B/cell written on dwell 1, 8 B/cell read-modify-write afterwards (SURVEY 8.4)"}


def mc_config(T, n_ch=2048, n=30000):
    rng = np.random.default_rng(0)
    chips = synth.load_chips()["gps_l5i"]
    mc = b.MulticorrelatorB200(n_ch, n, T)
    for c in range(n_ch):
        mc.set_local_code(c, chips[c % 32].astype(np.float32))
    base = np.array([-1, -0.5, -0.25, -0.1, -0.05, 0, 0.05, 0.1, 0.25, 0.5, 1], np.float32)
    sel = {3: [3, 5, 7], 5: [1, 3, 5, 7, 9], 11: list(range(11))}[T]
    shifts = np.tile(base[sel], (n_ch, 1))
    fd = rng.uniform(-5000, 5000, n_ch)
    a = mc.make_args(n_ch, n, rng.uniform(0, 6.28, n_ch), 2 * np.pi * fd / 30e6, 0.0, rng.uniform(0, 1, n_ch), 10.23e6 * (1 + fd / 1176.45e6) / 30e6, 0.0)
    d_sig = torch.randn((n_ch, n, 2), device="cuda", dtype=torch.float32)
    chans = np.arange(n_ch, dtype=np.uint32)
    for _ in range(3):
        mc.correlate_device_async(chans, a, shifts, d_sig.data_ptr(), n)
        mc.sync(n_ch, T)
    mc.profile(True)
    for _ in range(steps):
        mc.correlate_device_async(chans, a, shifts, d_sig.data_ptr(), n)
        mc.sync(n_ch, T)
    pm, pc = mc.profile(False)
    k = pm[0] / max(1, pc[0])
    gbs = n_ch * (8.0 * n + 8 * T) / (k * 1e-3) / 1e9
    return {"channels": n_ch, "taps": T, "samples": n, "kernel_ms": k, "channel_ms_per_s": n_ch / (k * 1e-3), "hbm_gbs": gbs}


out["configs"]["C1"] = acq_config("C1")
out["configs"]["C2"] = acq_config("C2")
out["configs"]["C4_gps"] = acq_config("C4_gps", min_cn0=47.0)  # 1 ms at 25 MS/s: 1e6 cells push the pfa threshold to ~48
out["configs"]["C4_gal"] = acq_config("C4_gal")
out["configs"]["C4_bds"] = acq_config("C4_bds", min_cn0=47.0)
c4 = [out["configs"][k] for k in ("C4_gps", "C4_gal", "C4_bds")]
out["configs"]["C4_total"] = {"cells": sum(c["cells"] for c in c4), "ms_one_gpu": sum(c["ms_per_dwell"] for c in c4),
                              "cells_per_s": sum(c["cells"] for c in c4) / (sum(c["ms_per_dwell"] for c in c4) * 1e-3),
                              "note": "full GPS + Galileo + BeiDou cold-start grid searched back to back on ONE B200"}
out["configs"]["C3"] = c3_config(c3_prns)
for T in (3, 5, 11):
    out["configs"][f"C5_T{T}"] = mc_config(T)
print(json.dumps(out))
