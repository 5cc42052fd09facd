"""The row-resident inverse kernel (csrc/acq_cluster.cuh: thread-block cluster, DSMEM stage transpose) against the two-kernel
path on the same FFT plan: magnitude grids bit for bit, decisions identical, statistics to 1e-6; and against the oracle through
the ordinary parity tests (tests/test_acq_full_parity_gpu.py runs on whatever path is the default). Child processes, because the
path switch (B200_ACQ_CLUSTER) is read once per process."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_CHILD = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tools")
import synth
from b200_import import load_package
b = load_package()
geo, codes, iq, truth = synth.config_scene("C2", n_prn=12)
out = {}
# records only (CFAR, no grid kept): MODE 2
e = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], 12, use_cfar=True)
e.set_local_codes(np.arange(12), codes)
e.profile(True)
out["rec"] = e.dwell(iq, np.arange(12))
out["rec_subset"] = e.dwell(iq, np.array([7, 2, 11]))          # fewer rows than co-resident clusters x ... and a slot permutation
out["first_stage_launches"] = np.array([e.profile(False)[1][1]])   # "first inverse stage -> scratch" launches: none on the row-resident path
e.close()
# persistent grid, accumulation over two dwells: MODE 1
g = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], 3, use_cfar=True, keep_grid=True)
g.set_local_codes(np.arange(3), codes[:3])
g.dwell(iq, np.arange(3))
out["rec_grid"] = g.dwell(iq[::-1].copy(), np.arange(3))
out["grid0"] = g.read_grid(0)
out["grid2"] = g.read_grid(2)
g.close()
# first-vs-second peak without a grid: the winning row is recomputed by the two-kernel path into a row buffer
f = b.PcpsAcquisitionB200(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"], geo["samples_per_chip"], 4, use_cfar=False)
f.set_local_codes(np.arange(4), codes[:4])
out["rec_2nd"] = f.dwell(iq, np.arange(4))
f.close()
# 50 PRNs x 400 Doppler rows (25 Hz step) = 20 000 rows: more than one launch of the row-resident kernel (74 clusters x 256 rows each)
geo50, codes50, iq50, _ = synth.config_scene("C2", n_prn=50)
m = b.PcpsAcquisitionB200(geo50["fft_size"], geo50["consumed_samples"], geo50["doppler_max"], 25, geo50["fs"], geo50["samples_per_chip"], 50, use_cfar=True)
assert m.n_doppler == 400
m.set_local_codes(np.arange(50), codes50)
out["rec_many"] = m.dwell(iq50, np.arange(50))
m.close()
np.savez(sys.argv[2], **out)
"""


def _run(tmp_path, tag, env):
    path = str(tmp_path / (tag + ".npz"))
    r = subprocess.run([sys.executable, "-c", _CHILD, ROOT, path], env=dict(os.environ, **env), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return np.load(path)


@pytest.mark.parametrize("variant", ["default", "tma", "halfwarp"])
def test_row_resident_inverse_equals_two_kernel_path(tmp_path, variant):
    ref = _run(tmp_path, "ref", dict(B200_ACQ_CLUSTER="0", B200_FFT_PLAN="8,8,25,20"))
    got = _run(tmp_path, variant, {"default": {}, "tma": dict(B200_ACQ_CLUSTER="1"), "halfwarp": dict(B200_ACQ_CLUSTER="2")}[variant])
    assert int(ref["first_stage_launches"][0]) == 2 and int(got["first_stage_launches"][0]) == 0   # the path under test really ran
    for k in ("grid0", "grid2"):
        assert np.array_equal(got[k], ref[k]), k                        # same arithmetic in the same order
    for k in ("rec", "rec_subset", "rec_grid", "rec_2nd", "rec_many"):
        a, c = ref[k], got[k]
        for f in ("peak", "code_phase", "doppler_index", "doppler_hz", "dwell_count", "second_peak"):
            assert np.array_equal(a[f], c[f]), (k, f)
        # the row sum is added up in another order (per-warp records): input power and statistic move in the last bits
        assert np.allclose(a["test_statistic"], c["test_statistic"], rtol=1e-6, atol=0), k
        assert np.allclose(a["input_power"], c["input_power"], rtol=1e-6, atol=0), k
