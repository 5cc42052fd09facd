#!/usr/bin/env python3
"""Row-resident (cluster) inverse against the two-kernel path: records and grids bit for bit, and the time of a dwell.
Runs each variant in a child process because B200_ACQ_CLUSTER is read once per process.  usage: cluster_check.py [WL] [n_prn]"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tools")
import synth, torch
from b200_import import load_package
b = load_package()
wl, n_prn, out = sys.argv[2], int(sys.argv[3]), sys.argv[4]
geo, codes, iq, truth = synth.config_scene(wl, n_prn=n_prn)
res = {}
slots = np.arange(n_prn, dtype=np.uint32)
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], n_prn, use_cfar=True)
e.set_local_codes(slots, codes)
d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    e.dwell_device_async(d_iq.data_ptr(), slots); r = e.sync(n_prn)
res["rec"] = r.view(np.uint8).copy()
est = torch.cuda.ExternalStream(e.stream())
steps = 10
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
for i in range(steps):
    flush.fill_(1); torch.cuda.synchronize()
    ev[i][0].record(est); e.dwell_device_async(d_iq.data_ptr(), slots); ev[i][1].record(est)
    r = e.sync(n_prn)
wall = sum(a.elapsed_time(c) for a, c in ev) / steps
e.profile(True)
for _ in range(steps):
    flush.fill_(1); torch.cuda.synchronize()
    e.dwell_device_async(d_iq.data_ptr(), slots); r = e.sync(n_prn)
ms, cnt = e.profile(False)
print("variant", os.environ.get("B200_ACQ_CLUSTER", "default"), wl, "n_prn", n_prn, "kernels(ms):", [round(m / steps, 4) for m in ms[:5]],
      "WALL %.4f ms  %.1f Gcells/s  launches/dwell %d" % (wall, n_prn * e.n_doppler * e.eff / wall / 1e6, sum(cnt) // steps), flush=True)
e.close()
# persistent grid (mode 1), two dwells, 4 PRNs
g = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], 4, use_cfar=True, keep_grid=True)
g.set_local_codes(np.arange(4), codes[:4])
g.dwell(iq, np.arange(4))
r2 = g.dwell(iq[::-1].copy(), np.arange(4))
res["rec_grid"] = r2.view(np.uint8).copy()
res["grid0"] = g.read_grid(0)
res["grid3"] = g.read_grid(3)
g.close()
np.savez(out, **res)
"""


VARIANTS = {
    "ref": dict(B200_ACQ_CLUSTER="0"),                                   # two-kernel path, its own plan
    "ref_sameplan": dict(B200_ACQ_CLUSTER="0", B200_FFT_PLAN="8,8,25,20"),  # two-kernel path on the row-resident kernel's plan
    "cl_tma": dict(B200_ACQ_CLUSTER="1"),
    "cl_halfwarp": dict(B200_ACQ_CLUSTER="2"),                           # registers, one upper-stage column per half-warp
    "cl": dict(),                                                        # default: registers, grouped upper stage
}
REC = np.dtype([("test_statistic", np.float32), ("peak", np.float32), ("input_power", np.float32), ("second_peak", np.float32),
                ("code_phase", np.uint32), ("doppler_index", np.uint32), ("doppler_hz", np.int32), ("dwell_count", np.uint32)])


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
    n_prn = sys.argv[2] if len(sys.argv) > 2 else "50"
    variants = sys.argv[3].split(",") if len(sys.argv) > 3 else list(VARIANTS)
    base_name = sys.argv[4] if len(sys.argv) > 4 else "ref_sameplan"
    outs = {}
    for v in variants:
        path = f"/tmp/cluster_check_{v}.npz"
        r = subprocess.run([sys.executable, "-c", CHILD, ROOT, wl, n_prn, path], env=dict(os.environ, **VARIANTS[v]), capture_output=True, text=True, timeout=600)
        print(v, r.stdout.strip(), r.stderr.strip()[-300:] if "co-resident" in r.stderr else "")
        if r.returncode != 0:
            print("variant", v, "FAILED:", r.stderr[-1500:])
            continue
        outs[v] = np.load(path)
    base = outs.get(base_name)
    for v, o in outs.items():
        if v == base_name or base is None or not v.startswith("cl"):
            continue
        for k in base.files:
            if k.startswith("grid"):
                same = np.array_equal(base[k], o[k])
                print(f"{v} vs {base_name}: {k}: " + ("bit-identical" if same else "DIFFERENT max rel diff %.3g" % (np.abs(base[k] - o[k]).max() / base[k].max())))
            else:
                a, c = base[k].view(REC), o[k].view(REC)
                exact = all(np.array_equal(a[f], c[f]) for f in ("peak", "code_phase", "doppler_index", "doppler_hz", "dwell_count"))
                rel = float(np.abs(a["test_statistic"] / c["test_statistic"] - 1).max())
                print(f"{v} vs {base_name}: {k}: peak / cell / Doppler {'identical' if exact else 'DIFFERENT'}, statistic rel diff {rel:.2e}")


if __name__ == "__main__":
    main()
