"""Host logic of the multi-GPU map phase: which input splits a GPU map slot takes, and how
the per-slot partials are combined.

The reference parallelises by input split, one map task per split, any slot
(mapred/JobQueueTaskScheduler.java:334-387), GPU slot i bound to device i
(mapred/TaskTrackerStatus.java:536-551).  Inside one box this project gives every GPU a
contiguous run of splits (static) and replaces the shuffle of per-task partial records by one
sum-allreduce of [k*d sums][k counts].
"""
from __future__ import annotations

import numpy as np


def split_ranges(total_bytes: int, split_bytes: int, row_bytes: int) -> list[tuple[int, int]]:
    """FileInputFormat-style splits of a file of fixed-size rows: (offset, length) pairs that
    never cut a row (a split boundary is moved down to a row boundary)."""
    if split_bytes < row_bytes:
        raise ValueError("split smaller than a row")
    per = (split_bytes // row_bytes) * row_bytes
    out, off = [], 0
    total = (total_bytes // row_bytes) * row_bytes
    while off < total:
        out.append((off, min(per, total - off)))
        off += per
    return out


def splits_for_rank(n_splits: int, world: int, rank: int) -> range:
    """Contiguous, balanced share of splits 0..n_splits-1 for `rank` of `world` slots: the
    first n_splits % world ranks take one extra."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(n_splits, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def allreduce_partials(sums: np.ndarray, counts: np.ndarray):
    """Sum partials over all ranks of the default torch.distributed group (gloo on CPU hosts;
    on GPUs the library's own hgm_allreduce_partials does this over NCCL).  Counts stay exact
    (int64); sums are FP64 and order-dependent across ranks within the stated 1e-5 tolerance."""
    import torch
    import torch.distributed as dist

    s = torch.from_numpy(np.ascontiguousarray(sums, dtype=np.float64).copy())
    c = torch.from_numpy(np.ascontiguousarray(counts, dtype=np.int64).copy())
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    dist.all_reduce(c, op=dist.ReduceOp.SUM)
    return s.numpy(), c.numpy()


def update_centroids(sums: np.ndarray, counts: np.ndarray, old: np.ndarray) -> np.ndarray:
    """The reducer's centroid update (oracle/kmeans_pipes_cpu.cc KMeansReduce): FP32(sum/count);
    an empty cluster keeps its centroid (no record reaches the reducer for it)."""
    new = np.array(old, dtype=np.float32, copy=True)
    live = counts > 0
    new[live] = (sums[live] / counts[live, None]).astype(np.float32)
    return new
