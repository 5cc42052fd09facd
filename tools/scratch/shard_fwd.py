"""Per-kernel-class times of ONE shard of a C2 search (rows [0, n_rows) of 80) — what bounds the sharded search at N = 8."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth, torch
from b200_import import load_package
b = load_package()
n_rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10
geo, codes, iq, truth = synth.config_scene("C2")
n = geo["n_prn"]
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], n, use_cfar=True,
                          row_begin=0, row_end=n_rows)
slots = np.arange(n, dtype=np.uint32)
e.set_local_codes(slots, codes)
e.exchange_buffers(80 // n_rows)
d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    e.dwell_device_async(d_iq.data_ptr(), slots); e.sync(n)
e.profile(True)
for _ in range(10):
    flush.fill_(1); torch.cuda.synchronize()
    e.dwell_device_async(d_iq.data_ptr(), slots); e.sync(n)
ms, cnt = e.profile(False)
print("rows", n_rows, "env", {k: v for k, v in os.environ.items() if k.startswith("B200_")}, "ms/dwell by class", [round(m / 10, 4) for m in ms[:5]], "launches", [c // 10 for c in cnt[:5]])
