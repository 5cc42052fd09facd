"""Device time of ONE shard's captured dwell (rows [0, n_rows) of a C2 search) on one GPU: the compute part of a sharded step."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import synth, torch
from b200_import import load_package
b = load_package()
geo, codes, iq, truth = synth.config_scene("C2")
n = geo["n_prn"]
slots = np.arange(n, dtype=np.uint32)
d_iq = torch.from_numpy(iq.view(np.float32).copy()).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for n_rows in (80, 40, 20, 10):
    e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], n, use_cfar=True,
                              row_begin=0, row_end=n_rows)
    e.set_local_codes(slots, codes)
    e.exchange_buffers(80 // n_rows)
    for _ in range(3):
        e.dwell_device_async(d_iq.data_ptr(), slots); e.sync(n)
    e.capture_begin(); e.dwell_device_async(d_iq.data_ptr(), slots); e.capture_end()
    est = torch.cuda.ExternalStream(e.stream())
    ts = []
    for _ in range(12):
        flush.fill_(1); torch.cuda.synchronize()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(est); e.replay_async(); c.record(est); e.sync(n)
        ts.append(a.elapsed_time(c))
    print("rows", n_rows, "graph replay ms: median %.4f min %.4f" % (float(np.median(ts[2:])), min(ts[2:])))
    e.close()
