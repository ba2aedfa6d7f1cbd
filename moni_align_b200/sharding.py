"""Read sharding across the GPUs of one box: the index is replicated, reads are split into
contiguous ranges of (nearly) equal base count, there is NO collective on the data path, and results come back
in input order.  torch.distributed is used only for the barrier and the max-over-ranks of device times
(backend nccl on GPUs, gloo in the CPU tests)."""
from __future__ import annotations

import numpy as np


def shard_bounds(offsets: np.ndarray, parts: int) -> np.ndarray:
    """cut[0..parts]: rank g owns reads [cut[g], cut[g+1]); cuts fall where the cumulative base count crosses
    g/parts of the total (same rule as split_by_bases() in csrc/moni_gpu.cu)."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    n = offsets.shape[0] - 1
    total = int(offsets[n] - offsets[0])
    cut = np.empty(parts + 1, dtype=np.int64)
    cut[0], cut[parts] = 0, n
    for g in range(1, parts):
        want = int(offsets[0]) + total // parts * g
        cut[g] = min(max(int(np.searchsorted(offsets, np.uint64(want), side="left")), int(cut[g - 1])), n)
    return cut


def shard_of(seqs: np.ndarray, offsets: np.ndarray, rank: int, world: int):
    """(seqs, offsets) of this rank's shard, offsets rebased to 0, plus the global read range."""
    cut = shard_bounds(offsets, world)
    a, b = int(cut[rank]), int(cut[rank + 1])
    o = offsets[a:b + 1]
    return seqs[int(o[0]):int(o[-1])], (o - o[0]).astype(np.uint64), (a, b)


def max_over_ranks(value: float, device=None) -> float:
    """MAX all-reduce of a scalar; identity when torch.distributed is not initialised."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, device=None) -> float:
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def gather_in_order(local_offsets: np.ndarray, local_values: np.ndarray, world: int, rank: int):
    """Host-side gather of per-rank results into input order (rank order == read order because shards are
    contiguous).  Returns (offsets, values) on every rank."""
    import torch.distributed as dist
    parts = [None] * world
    dist.all_gather_object(parts, (np.asarray(local_offsets), np.asarray(local_values)))
    offs = [np.zeros(1, dtype=np.uint64)]
    vals = []
    base = np.uint64(0)
    for o, v in parts:
        offs.append(o[1:].astype(np.uint64) + base)
        vals.append(v)
        base = base + np.uint64(o[-1])
    return np.concatenate(offs), np.concatenate(vals) if vals else np.zeros(0, dtype=np.uint64)
