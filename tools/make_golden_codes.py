#!/usr/bin/env python3
"""Write tests/golden/prn_chips.npz from the REFERENCE's own PRN code generators (oracle/_ref).

Container-only (needs /root/reference). The chips are ICD data / LFSR outputs, stored bit-packed
(1 = chip value +1, 0 = chip value -1), so the GPU box can build local codes without the reference tree.
Also stores sha256 digests of the reference's *sampled* codes at the BASELINE sampling rates, which pin
the oracle's own samplers (oracle/oracle_codes.c) where the reference is not available.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from oracle import RefLibs  # noqa: E402


def pack(rows):
    a = np.stack(rows)
    assert set(np.unique(a)) <= {-1, 1}
    return np.packbits((a > 0).astype(np.uint8), axis=1)


def main():
    R = RefLibs()
    out = {
        "gps_l1_ca": pack([R.gps_l1_ca_chips(p) for p in range(1, 33)]),
        "bds_b1i": pack([R.beidou_b1i_chips(p) for p in range(1, 64)]),
        "gal_e1b": pack([R.galileo_e1_chips(p, pilot=False) for p in range(1, 51)]),
        "gal_e1c": pack([R.galileo_e1_chips(p, pilot=True) for p in range(1, 51)]),
        "gps_l5i": pack([R.gps_l5_chips(p, q=False) for p in range(1, 33)]),
        "gps_l5q": pack([R.gps_l5_chips(p, q=True) for p in range(1, 33)]),
    }
    lens = {"gps_l1_ca": 1023, "bds_b1i": 2046, "gal_e1b": 4092, "gal_e1c": 4092, "gps_l5i": 10230, "gps_l5q": 10230}
    np.savez_compressed(os.path.join(ROOT, "tests/golden/prn_chips.npz"), **out,
                        lengths=np.array(json.dumps(lens)))
    dig = {}
    for fs in (4000000, 8000000, 25000000, 6625000):
        for prn in (1, 7, 32):
            dig[f"gps_l1_ca/{prn}/{fs}"] = hashlib.sha256(R.gps_l1_ca_sampled(prn, fs).tobytes()).hexdigest()
            dig[f"gal_e1b/{prn}/{fs}"] = hashlib.sha256(R.galileo_e1_sampled(prn, fs).tobytes()).hexdigest()
            dig[f"bds_b1i/{prn}/{fs}"] = hashlib.sha256(R.beidou_b1i_sampled(prn, fs).tobytes()).hexdigest()
    for fs in (30000000, 25000000):
        for prn in (1, 7, 32):
            dig[f"gps_l5i/{prn}/{fs}"] = hashlib.sha256(R.gps_l5i_sampled(prn, fs).tobytes()).hexdigest()
    with open(os.path.join(ROOT, "tests/golden/sampled_code_sha256.json"), "w") as f:
        json.dump(dig, f, indent=1, sort_keys=True)
    print("wrote", {k: v.shape for k, v in out.items()}, len(dig), "digests")


if __name__ == "__main__":
    main()
