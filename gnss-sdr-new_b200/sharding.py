"""Multi-GPU partition of one acquisition search and of a tracking epoch.

One process per GPU (torchrun). The units are independent. The engines themselves split ONE search Doppler-major (partition_doppler_major,
pack_rows, unpack_gathered below restate that layout for the gloo tests; on the device it is gnss_b200_acq_exchange_buffers with the
peer-memory or NCCL exchange). The PRN-major helpers are the general (slot, row) form:

* acquisition: a unit is a (PRN slot, Doppler row). Partition PRN-major (rank g owns slots g, g+G, ...); when there
  are fewer slots than ranks, the Doppler rows of a slot are split over `G // n_slots` ranks (config 3: 800 rows per
  PRN). The only exchange is an all-gather of the 16-byte per-row records {max, argmax, row sum, valid}
  (gnss_b200_acq_row); the CFAR noise row `(k* + nD/2) % nD` (pcps_acquisition.cc:533-534) may live on another rank,
  which is why the row sums travel with the maxima. The merge itself is gnss_b200_acq_finalize_rows.
* tracking: a unit is a channel; static block partition, results gathered as [channels][taps] complex64.

`torch.distributed` is used only as plumbing (NCCL on the GPU box, gloo in the CPU tests); there is no reduction
heavier than a gather, so no fused compute+collective kernel exists on this path.
"""
from __future__ import annotations

import numpy as np


def partition_search(n_slots: int, n_doppler: int, world: int):
    """-> list over ranks of [(slot, row_begin, row_end), ...]. Every (slot, row) is owned by exactly one rank."""
    if world < 1 or n_slots < 0 or n_doppler < 1:
        raise ValueError("partition_search: bad arguments")
    plan = [[] for _ in range(world)]
    if n_slots == 0:
        return plan
    if n_slots >= world:
        for s in range(n_slots):
            plan[s % world].append((s, 0, n_doppler))
        return plan
    per = min(world // n_slots, n_doppler)  # ranks sharing one slot; left-over ranks idle
    for s in range(n_slots):
        for j in range(per):
            r0 = (n_doppler * j) // per
            r1 = (n_doppler * (j + 1)) // per
            plan[s * per + j].append((s, r0, r1))
    return plan


def partition_doppler_major(n_doppler: int, world: int):
    """The split the engines themselves use for ONE search over `world` shards (gnss_b200_acq_exchange_buffers,
    gnss_b200_acq_group_*): shard r owns rows [r * chunk, min(n_doppler, (r + 1) * chunk)) of EVERY slot, chunk = ceil(n_doppler /
    world) — no forward transform is computed twice and every shard carries an equal share of every PRN.
    -> (chunk, [(row_begin, row_end) per rank]); trailing ranks may own nothing."""
    if world < 1 or n_doppler < 1:
        raise ValueError("partition_doppler_major: bad arguments")
    chunk = -(-n_doppler // world)
    return chunk, [(min(n_doppler, r * chunk), min(n_doppler, (r + 1) * chunk)) for r in range(world)]


def pack_rows(rows: np.ndarray, row_begin: int, row_end: int, chunk: int) -> np.ndarray:
    """What pack_rows_kernel leaves in d_send: this shard's rows of every slot as [n_slots][chunk], padded with invalid records."""
    out = np.zeros((rows.shape[0], chunk), dtype=rows.dtype)
    out[:, :row_end - row_begin] = rows[:, row_begin:row_end]
    return out


def unpack_gathered(gathered: np.ndarray, n_doppler: int) -> np.ndarray:
    """[world][n_slots][chunk] (what d_recv / the peer regions hold) -> [n_slots][n_doppler], the table finalize_gathered_kernel decides on."""
    world, n_slots, chunk = gathered.shape
    return np.ascontiguousarray(np.transpose(gathered, (1, 0, 2)).reshape(n_slots, world * chunk)[:, :n_doppler])


def gather_rows_doppler_major(local_rows: np.ndarray, n_doppler: int, group=None) -> np.ndarray:
    """All-gather of the packed records of a Doppler-major split (the host form of the device exchange): every rank passes its
    full-size [n_slots][n_doppler] table with only its own rows valid and gets the complete table back."""
    import torch
    dist = _dist()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    chunk, parts = partition_doppler_major(n_doppler, world)
    mine = pack_rows(np.ascontiguousarray(local_rows), parts[rank][0], parts[rank][1], chunk)
    send = torch.from_numpy(mine.view(np.uint8).reshape(-1).copy())
    recv = torch.empty(world * send.numel(), dtype=torch.uint8)
    dist.all_gather_into_tensor(recv, send, group=group)
    gathered = recv.numpy().view(local_rows.dtype).reshape(world, local_rows.shape[0], chunk)
    merged = unpack_gathered(gathered, n_doppler)
    if not merged["valid"].all():
        raise RuntimeError("gather_rows_doppler_major: no rank delivered some row")
    return merged


def partition_channels(n_channels: int, world: int):
    """Static block partition of tracking channels -> list over ranks of (begin, end)."""
    if world < 1 or n_channels < 0:
        raise ValueError("partition_channels: bad arguments")
    return [((n_channels * r) // world, (n_channels * (r + 1)) // world) for r in range(world)]


def _dist():
    import torch.distributed as dist
    return dist


def gather_rows(local_rows: np.ndarray, device=None, group=None) -> np.ndarray:
    """All-gather per-row records and merge them.

    local_rows: structured array [n_slots][n_doppler] of ACQ_ROW_DTYPE with valid=1 only on the rows this rank
    searched. Returns the merged [n_slots][n_doppler] table on every rank; raises if two ranks claim the same row or
    a row is missing."""
    import torch
    dist = _dist()
    rows = np.ascontiguousarray(local_rows)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        merged = rows.copy()
    else:
        world = dist.get_world_size(group)
        t = torch.from_numpy(rows.view(np.uint8).reshape(-1).copy())
        if device is not None:
            t = t.to(device, non_blocking=False)
        out = torch.empty(world * t.numel(), dtype=torch.uint8, device=t.device)
        dist.all_gather_into_tensor(out, t, group=group)
        allr = out.cpu().numpy().view(rows.dtype).reshape((world,) + rows.shape)
        owners = allr["valid"].astype(np.int64).sum(axis=0)
        if (owners > 1).any():
            raise RuntimeError("gather_rows: a (slot, row) was searched by more than one rank")
        pick = allr["valid"].argmax(axis=0)
        merged = np.take_along_axis(allr, pick[None, ...], axis=0)[0]
    if (merged["valid"] == 0).any():
        missing = np.argwhere(merged["valid"] == 0)[0]
        raise RuntimeError(f"gather_rows: row {tuple(int(x) for x in missing)} was searched by no rank")
    return merged


def gather_channels(local_out: np.ndarray, n_channels: int, device=None, group=None) -> np.ndarray:
    """All-gather [local_channels][taps] complex64 correlator outputs of a block partition into [n_channels][taps]."""
    import torch
    dist = _dist()
    local_out = np.ascontiguousarray(local_out, dtype=np.complex64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local_out.copy()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    parts = partition_channels(n_channels, world)
    taps = local_out.shape[1]
    width = max(e - b for b, e in parts)
    pad = np.zeros((width, taps), dtype=np.complex64)
    b, e = parts[rank]
    if local_out.shape[0] != e - b:
        raise ValueError("gather_channels: local block does not match the partition")
    pad[: e - b] = local_out
    t = torch.from_numpy(pad.view(np.float32).reshape(-1))
    if device is not None:
        t = t.to(device)
    out = torch.empty(world * t.numel(), dtype=torch.float32, device=t.device)
    dist.all_gather_into_tensor(out, t, group=group)
    allp = out.cpu().numpy().view(np.complex64).reshape(world, width, taps)
    return np.concatenate([allp[r, : pe - pb] for r, (pb, pe) in enumerate(parts)], axis=0)


def broadcast_samples(samples: np.ndarray, device=None, src: int = 0, group=None) -> np.ndarray:
    """The sample block of a dwell is captured on one rank and searched by all (8*N bytes <= 1.28 MB)."""
    import torch
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return samples
    arr = np.ascontiguousarray(samples)
    t = torch.from_numpy(arr.view(np.uint8).reshape(-1).copy())
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, src=src, group=group)
    return t.cpu().numpy().view(arr.dtype).reshape(arr.shape)
