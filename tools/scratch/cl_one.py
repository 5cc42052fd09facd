"""One C2 dwell sequence for ncu: 3 dwells of 50 PRNs (variant from B200_ACQ_CLUSTER)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + "/tools")
import synth
from b200_import import load_package
b = load_package()
n_prn = int(sys.argv[1]) if len(sys.argv) > 1 else 50
wl = sys.argv[2] if len(sys.argv) > 2 else "C2"
geo, codes, iq, truth = synth.config_scene(wl, n_prn=n_prn)
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], n_prn, use_cfar=True)
e.set_local_codes(np.arange(n_prn), codes)
for _ in range(3):
    r = e.dwell(iq, np.arange(n_prn))
print("ok", float(r[0]["test_statistic"]))
