"""C1 (5.12 M cells) per dwell: eager launches vs one CUDA-graph replay, and the per-kernel-class profile."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth, torch
from b200_import import load_package
b = load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "C1"
geo, codes, iq, truth = synth.config_scene(name)
n = geo["n_prn"]
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], n, use_cfar=True)
slots = np.arange(n, dtype=np.uint32)
e.set_local_codes(slots, codes)
d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
est = torch.cuda.ExternalStream(e.stream())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    e.dwell_device_async(d_iq.data_ptr(), slots); ref = e.sync(n)
def timeit(fn, steps=20):
    ts = []
    for _ in range(steps):
        flush.fill_(1); torch.cuda.synchronize()
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(est); fn(); z.record(est); z.synchronize(); ts.append(a.elapsed_time(z))
    return float(np.median(ts)), float(np.min(ts))
print(name, "eager  ms median/min", timeit(lambda: e.dwell_device_async(d_iq.data_ptr(), slots)), "launches/dwell", end=" ")
l0 = e.launch_count(); e.dwell_device_async(d_iq.data_ptr(), slots); e.sync(n); print(e.launch_count() - l0)
e.capture_begin(); e.dwell_device_async(d_iq.data_ptr(), slots); e.capture_end()
e.replay_async(); got = e.sync(n)
assert all(got[k]["test_statistic"] == ref[k]["test_statistic"] and got[k]["code_phase"] == ref[k]["code_phase"] for k in range(n))
print(name, "graph  ms median/min", timeit(lambda: e.replay_async()))
# no flush: warm L2
ts = []
for _ in range(20):
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(est); e.replay_async(); z.record(est); z.synchronize(); ts.append(a.elapsed_time(z))
print(name, "graph warm-L2 ms median", float(np.median(ts)))
e.profile(True)
for _ in range(10):
    e.dwell_device_async(d_iq.data_ptr(), slots); e.sync(n)
ms, cnt = e.profile(False)
print("profile ms/dwell by class", [round(m / 10, 4) for m in ms], "launches", [c // 10 for c in cnt])
