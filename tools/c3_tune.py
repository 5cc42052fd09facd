#!/usr/bin/env python3
"""Per-dwell time of a reduced config-3 search (N = 160000, 800 bins, persistent grid) — tuning aid.
Honours B200_FFT_PLAN / B200_FFT_COLS / B200_FFT_NB / B200_ACQ_SCRATCH_MB. Usage: c3_tune.py [n_prn] [n_dwells]"""
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
n_prn = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n_dw = int(sys.argv[2]) if len(sys.argv) > 2 else 4
geo, codes, iq, truth = synth.config_scene("C3", n_prn=n_prn, n_dwells=n_dw)
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                          geo["samples_per_chip"], n_prn, use_cfar=True, keep_grid=True)
slots = np.arange(n_prn, dtype=np.uint32)
e.set_local_codes(slots, codes)
d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
est = torch.cuda.ExternalStream(e.stream())
for rep in range(2):
    for s in slots:
        e.reset(int(s))
    if rep == 1:
        e.profile(True)
    times = []
    for k in range(n_dw):
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(est)
        e.dwell_device_async(d_iq[k].data_ptr(), slots)
        c.record(est)
        res = e.sync(n_prn)
        times.append(a.elapsed_time(c))
ms, cnt = e.profile(False)
cells = n_prn * e.n_doppler * e.eff
env = {k: os.environ[k] for k in ("B200_FFT_PLAN", "B200_FFT_COLS", "B200_FFT_NB", "B200_ACQ_SCRATCH_MB") if k in os.environ}
print(f"{env} C3 n_prn={n_prn}: first {times[0]:.2f} ms ({cells/times[0]/1e6:.1f} Gcells/s), later {np.mean(times[1:]):.2f} ms "
      f"({cells/np.mean(times[1:])/1e6:.1f} Gcells/s, grid {cells*8/np.mean(times[1:])/1e6:.0f} GB/s) | per-class ms/dwell: "
      + ", ".join(f"{n} {m/n_dw:.2f}" for n, m in zip(["fwd", "inv_first", "inv_last", "code", "decide"], ms)) + f" | stat0 {float(res[0]['test_statistic']):.2f}")
