#!/usr/bin/env python3
"""Host-side cost of one batched multicorrelator call (time spent inside the C-ABI call before it returns) vs the device time."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth  # noqa: E402
from b200_import import load_package  # noqa: E402
import torch  # noqa: E402

b = load_package()
n, T = 30000, 11
for n_ch in (256, 2048):
    rng = np.random.default_rng(0)
    chips = synth.load_chips()["gps_l5i"]
    mc = b.MulticorrelatorB200(n_ch, n, T)
    for c in range(n_ch):
        mc.set_local_code(c, chips[c % 32].astype(np.float32))
    shifts = np.tile(np.linspace(-1, 1, T).astype(np.float32), (n_ch, 1))
    fd = rng.uniform(-5000, 5000, n_ch)
    a = mc.make_args(n_ch, n, rng.uniform(0, 6.28, n_ch), 2 * np.pi * fd / 30e6, 0.0, rng.uniform(0, 1, n_ch), 10.23e6 * (1 + fd / 1176.45e6) / 30e6, 0.0)
    d_sig = torch.randn((n_ch, n, 2), device="cuda", dtype=torch.float32)
    chans = np.arange(n_ch, dtype=np.uint32)
    for _ in range(3):
        mc.correlate_device_async(chans, a, shifts, d_sig.data_ptr(), n)
        mc.sync(n_ch, T)
    th, tt = [], []
    for _ in range(20):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mc.correlate_device_async(chans, a, shifts, d_sig.data_ptr(), n)
        t1 = time.perf_counter()
        mc.sync(n_ch, T)
        t2 = time.perf_counter()
        th.append(t1 - t0)
        tt.append(t2 - t0)
    print(f"n_ch={n_ch}: host call {1e6*np.median(th):.1f} us, call+sync {1e6*np.median(tt):.1f} us")
