#!/usr/bin/env python3
"""Kernel time of the config-5 multicorrelator batch (2048 ch x 11 taps x 30000 samples, device-resident) — tuning aid.
   Honours B200_MC_SPLITS. Prints avg kernel ms from the engine's own CUDA-event profiler."""
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
n_ch = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
n, T = 30000, int(sys.argv[2]) if len(sys.argv) > 2 else 11
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
mc.profile(True)
for _ in range(20):
    mc.correlate_device_async(chans, a, shifts, d_sig.data_ptr(), n)
    mc.sync(n_ch, T)
pm, pc = mc.profile(False)
k = pm[0] / max(1, pc[0])
print(f"splits={os.environ.get('B200_MC_SPLITS', 'auto')} n_ch={n_ch} T={T}: kernel {k*1e3:.1f} us  -> {n_ch*(8.0*n+8*T)/k/1e6:.0f} GB/s")
