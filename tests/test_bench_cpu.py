"""bench.py without a GPU: the reference arm's JSON line (the driver runs `bench.py --impl reference` beside the GPU arm and
divides the two), the committed ncu summaries the live line quotes its DRAM traffic from, and the algorithmic work per cell."""
import json
import math
import os
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_line_has_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "acq_cells_per_s" and line["unit"] == "cells/s"
    assert line["higher_is_better"] is True and line["n_gpus"] == 1 and line["steps"] == 1 and line["gpu_launches"] == 0
    assert line["config"]["fft_size"] == 32000 and line["config"]["n_doppler_bins"] == 80 and line["config"]["n_prn"] == 50
    cb = line["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == line["value"] > 0 and "sample" in cb
    assert line["e2e"] == {"value": line["value"], "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    # value = cells of the bounded sample / time of the step
    assert line["value"] > 1e6 and line["ms_per_step"] > 0


def test_committed_ncu_summaries_parse():
    sys.path.insert(0, ROOT)
    import bench
    newest = None
    for tag in ("v7", "v6", "v4"):
        newest = bench.ncu_dram_bytes("r02_cluster_ncu_summary.txt", tag)
        if newest is not None:
            break
    # the row-resident inverse reads both operands once and writes records only: ~34 MB per 4000-row dwell, far from the 2.05 GB
    # of the two-kernel path
    assert newest is not None and 30e6 < newest < 60e6
    assert bench.ncu_dram_bytes("r02_cluster_ncu_summary.txt", "no such tag") is None
    mc = bench.ncu_dram_bytes_multicorr("r02_multicorr_ncu_full_summary.txt")
    assert mc is not None and 4.9e8 < mc < 7e8           # 2048 channels x 240 kB algorithmic = 492 MB
    assert bench.ncu_dram_bytes_multicorr("missing.txt") is None


def test_algorithmic_flops_per_cell():
    sys.path.insert(0, ROOT)
    import bench
    # 10 log2 N + 17
    for n, want in ((4000, 137), (25000, 163), (32000, 167), (100000, 183)):
        assert round(bench.acq_flops_per_cell(n)) == want
        assert abs(bench.acq_flops_per_cell(n) - (10 * math.log2(n) + 17)) < 1e-9
