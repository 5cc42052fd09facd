#!/usr/bin/env python3
"""Small fixed workloads for ncu captures (short: ncu replays every kernel ~40 times).
   python tools/prof_run.py acq [workload] [n_prn]   |   python tools/prof_run.py mc [n_ch] [n_taps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth  # noqa: E402
from b200_import import load_package  # noqa: E402

b = load_package()
what = sys.argv[1] if len(sys.argv) > 1 else "acq"
if what == "acq":
    wl = sys.argv[2] if len(sys.argv) > 2 else "C2"
    n_prn = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    geo, codes, iq, truth = synth.config_scene(wl, n_prn=n_prn)
    e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                              geo["samples_per_chip"], n_prn, use_cfar=True)
    e.set_local_codes(np.arange(n_prn), codes)
    for _ in range(2):
        r = e.dwell(iq, np.arange(n_prn))
    print("acq", wl, n_prn, "launches", e.launch_count(), "stat[0]", float(r[0]["test_statistic"]))
else:
    n_ch = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    T = int(sys.argv[3]) if len(sys.argv) > 3 else 11
    n = 30000
    rng = np.random.default_rng(0)
    chips = synth.load_chips()["gps_l5i"]
    mc = b.MulticorrelatorB200(n_ch, n, T)
    for c in range(n_ch):
        mc.set_local_code(c, chips[c % 32].astype(np.float32))
    sig = (rng.standard_normal((n_ch, n, 2))).astype(np.float32)
    shifts = np.tile(np.linspace(-1, 1, T).astype(np.float32), (n_ch, 1))
    fd = rng.uniform(-5000, 5000, n_ch)
    a = mc.make_args(n_ch, n, rng.uniform(0, 6.28, n_ch), 2 * np.pi * fd / 30e6, 0.0, rng.uniform(0, 1, n_ch), 10.23e6 * (1 + fd / 1176.45e6) / 30e6, 0.0,
                     sig_ptrs=[sig[c].ctypes.data for c in range(n_ch)])
    for _ in range(2):
        out = mc.correlate_batch(np.arange(n_ch), a, shifts)
    print("mc", n_ch, T, "launches", mc.launch_count(), abs(out[0, T // 2]))
