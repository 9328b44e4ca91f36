"""The K-means job loop around the GPU map slot: what the JobClient + reducer do per
iteration in the reference (one MapReduce job per iteration; the reduce side is the
centroid update, pipes/PipesReducer.java:59-124 driving the program's Reducer), with the
input kept in the slot's HBM split cache from the second iteration on.

    iteration 1:   every split is staged from the host into the cache (hgm_cache_put)
    every iteration: the cached splits are mapped from HBM (hgm_cache_map_all: neighbours in the
                   cache's arena share a launch); what did not fit is streamed from the host
    each iteration: partials -> (allreduce across slots) -> centroid update -> hgm_set_centroids
"""
from __future__ import annotations

import numpy as np

from . import shard
from .capi import FMT_F32LE, GpuMapSlot, HgmError


def run_kmeans(slot: GpuMapSlot, splits: list, centroids: np.ndarray, iterations: int, fmt: int = FMT_F32LE,
               allreduce=None, on_iteration=None):
    """Lloyd iterations over `splits` (bytes-like or arrays of rows).  `allreduce(sums, counts)`
    combines the partials of several slots (shard.allreduce_partials, or None for one slot).
    Returns (centroids, history) where history[i] = (sums, counts) of iteration i."""
    cent = np.ascontiguousarray(centroids, dtype=np.float32)
    history = []
    slot.set_centroids(cent)
    slot.cache_drop()  # the cache holds this job's splits only
    streamed = []
    for sid, split in enumerate(splits):
        try:
            slot.cache_put(sid, split, fmt)
        except HgmError as err:
            if err.code != -12:  # anything but "does not fit": a real failure
                raise
            streamed.append(split)
    for it in range(iterations):
        slot.set_centroids(cent)
        slot.reset_partials()
        slot.cache_map_all()  # adds into the slot's device partials
        sums, counts = slot.read_partials()
        # the cache is only a cache: what did not fit is streamed from the host every iteration
        # (hgm_map_split starts from zeroed partials, hence after the read above)
        for split in streamed:
            s, c, _ = slot.map_split(split, fmt=fmt)
            sums = sums + s
            counts = counts + c
        if allreduce is not None:
            sums, counts = allreduce(sums, counts)
        history.append((sums, counts))
        cent = shard.update_centroids(sums, counts, cent)
        if on_iteration is not None:
            on_iteration(it, cent, sums, counts)
    return cent, history
