"""Worker of tests/test_sharding_gloo.py: one rank of a world_size-N gloo group on CPU.

Reads the full per-row record tables the parent derived from the oracle's grids, keeps only the rows its partition
owns, runs the product's gather + merge (gnss-sdr-new_b200/sharding.py + gnss_b200_acq_finalize_rows through the
C-ABI) and writes what it got. No GPU work happens here."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch.distributed as dist
    from b200_import import load_package
    b = load_package()
    sh = b.sharding
    work = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z = np.load(os.path.join(work, "in.npz"))
        full = z["rows"].view(b.ACQ_ROW_DTYPE).reshape(int(z["n_slots"]), int(z["n_doppler"]))
        n_slots, n_doppler = full.shape
        plan = sh.partition_search(n_slots, n_doppler, world)
        local = np.zeros_like(full)
        for slot, r0, r1 in plan[rank]:
            local[slot, r0:r1] = full[slot, r0:r1]
        if os.environ.get("B200_TEST_DOUBLE_CLAIM") and rank == 1:
            local[0, 0] = full[0, 0]
        if os.environ.get("B200_TEST_DROP_ROW") and rank == 0:
            s, r0, _ = plan[0][0]
            local[s, r0]["valid"] = 0
        err = ""
        try:
            merged = sh.gather_rows(local)
        except RuntimeError as e:
            err, merged = str(e), None
        res = np.zeros(n_slots, dtype=b.ACQ_RESULT_DTYPE)
        if merged is not None:
            for s in range(n_slots):
                res[s] = b.finalize_rows(merged[s], int(z["eff"]), int(z["doppler_max"]), int(z["doppler_step"]), 0, 1)
        # the same search split Doppler-major (the engines' own split: gnss_b200_acq_exchange_buffers / _group / _peer)
        chunk, parts = sh.partition_doppler_major(n_doppler, world)
        mine = np.zeros_like(full)
        mine[:, parts[rank][0]:parts[rank][1]] = full[:, parts[rank][0]:parts[rank][1]]
        merged_dm = sh.gather_rows_doppler_major(mine, n_doppler)
        res_dm = np.zeros(n_slots, dtype=b.ACQ_RESULT_DTYPE)
        for s in range(n_slots):
            res_dm[s] = b.finalize_rows(merged_dm[s], int(z["eff"]), int(z["doppler_max"]), int(z["doppler_step"]), 0, 1)
        # sample block broadcast and channel gather
        blk = z["samples"] if rank == 0 else np.zeros_like(z["samples"])
        blk = sh.broadcast_samples(blk)
        n_ch = int(z["n_ch"])
        cb, ce = sh.partition_channels(n_ch, world)[rank]
        ch_out = sh.gather_channels(z["ch_out"][cb:ce], n_ch)
        np.savez(os.path.join(work, f"out{rank}.npz"), res=res.view(np.uint8), err=np.array(err), samples=blk, ch_out=ch_out,
                 owned=np.array([len(plan[rank])]), res_dm=res_dm.view(np.uint8), rows_dm=np.array([parts[rank][1] - parts[rank][0]]))
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
