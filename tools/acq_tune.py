#!/usr/bin/env python3
"""Per-kernel-class time of one C2 dwell (50 PRNs x 80 bins x 32000) — tuning aid. Honours B200_FFT_PLAN."""
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
wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
n_prn = int(sys.argv[2]) if len(sys.argv) > 2 else 50
geo, codes, iq, truth = synth.config_scene(wl, n_prn=n_prn)
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                          geo["samples_per_chip"], n_prn, use_cfar=True)
slots = np.arange(n_prn, dtype=np.uint32)
e.set_local_codes(slots, codes)
d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    e.dwell_device_async(d_iq.data_ptr(), slots)
    r = e.sync(n_prn)
steps = 10
est = torch.cuda.ExternalStream(e.stream())
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
for i in range(steps):   # wall time of a step without the per-kernel profiler
    flush.fill_(1)
    torch.cuda.synchronize()
    ev[i][0].record(est)
    e.dwell_device_async(d_iq.data_ptr(), slots)
    ev[i][1].record(est)
    r = e.sync(n_prn)
wall = sum(a.elapsed_time(c) for a, c in ev) / steps
e.profile(True)
for _ in range(steps):
    flush.fill_(1)
    torch.cuda.synchronize()
    e.dwell_device_async(d_iq.data_ptr(), slots)
    r = e.sync(n_prn)
ms, cnt = e.profile(False)
names = ["fwd", "inv_first", "inv_last", "code", "decide"]
tot = sum(ms[:5]) / steps
print(f"plan={os.environ.get('B200_FFT_PLAN', 'auto')} {wl} n_prn={n_prn}: " + ", ".join(f"{n} {m/steps:.3f}" for n, m in zip(names, ms)) +
      f" | sum {tot:.3f} | WALL {wall:.3f} ms/step, {n_prn * e.n_doppler * e.eff / wall / 1e6:.1f} Gcells/s; stat0 {float(r[0]['test_statistic']):.3f} cp {int(r[0]['code_phase'])}")
