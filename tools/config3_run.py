#!/usr/bin/env python
"""BASELINE config 3 on ONE B200, measured: 20 GB of points (d=2, k=128) in 64 MiB splits,
10 Lloyd iterations; iteration 1 stages every split from pinned host memory into the HBM
split cache (hgm_cache_put), iterations 2..10 map the splits from HBM (hgm_cache_map_all: one launch per 2 GiB arena block).

The full configuration (8 GPU slots + CPU slots under the hybrid scheduler) is simulated in
tools/config3_sim.py from per-task times; this run measures those per-task times and shows
that the whole 20 GB stay resident on one device.

    python tools/config3_run.py [--gb 20] [--iterations 10] [--out profiles/rNN_config3_run.json]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hadoop-1.0.3-gpu_b200"))
from hgpumap import capi, shard, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=20.0)
    ap.add_argument("--iterations", type=int, default=10)
    ap.add_argument("--k", type=int, default=128)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    d, k = 2, args.k
    split_bytes = 64 << 20
    pts_per_split = split_bytes // (4 * d)
    n_splits = max(1, int(args.gb * 1e9) // split_bytes)
    seed = 20261106
    centres = synth.centres(seed, k, d)
    cent = synth.perturb(seed + 1, centres)

    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        d_scratch = slot.device_alloc(split_bytes)
        h_ptr = slot.host_alloc(split_bytes)
        host = np.ctypeslib.as_array(C.cast(h_ptr, C.POINTER(C.c_float)), shape=(pts_per_split, d))
        # iteration 1: generate each split (on the device, copied to the host so that the staged
        # bytes really cross PCIe), stage it into the cache, map it
        stage_s = map1_s = 0.0
        slot.reset_partials()
        t_all = time.perf_counter()
        for sid in range(n_splits):
            slot.synth_points_device(seed, sid * pts_per_split, pts_per_split, centres, d_scratch)
            slot.memcpy_d2h(host, d_scratch)
            t0 = time.perf_counter()
            slot.cache_put(sid, host)          # H2D from pinned memory, synchronous
            t1 = time.perf_counter()
            slot.cache_map(sid)
            stage_s += t1 - t0
        slot.sync()
        sums, counts = slot.read_partials()
        iter1_wall = time.perf_counter() - t_all
        n_points = n_splits * pts_per_split
        assert counts.sum() == n_points
        iters = [{"iteration": 1, "staging_s": stage_s, "wall_s_incl_generation": iter1_wall,
                  "stage_gbs": n_splits * split_bytes / stage_s / 1e9}]
        cent = shard.update_centroids(sums, counts, cent)
        costs = []
        for it in range(2, args.iterations + 1):
            t0 = time.perf_counter()
            slot.set_centroids(cent)
            slot.reset_partials()
            _, groups = slot.cache_map_all()
            sums, counts = slot.read_partials()   # synchronises
            dt = time.perf_counter() - t0
            assert counts.sum() == n_points
            new = shard.update_centroids(sums, counts, cent)
            costs.append(float(np.abs(new - cent).max()))
            cent = new
            iters.append({"iteration": it, "wall_s": dt, "points_per_s": n_points / dt,
                          "hbm_gbs": n_points * 8 / dt / 1e9, "launch_groups": groups,
                          "max_centroid_move": costs[-1]})
        n_cached, cached_bytes = slot.cache_info()
        slot.device_free(d_scratch)
        slot.host_free(h_ptr)
    later = [x["wall_s"] for x in iters[1:]]
    summary = {
        "workload": "config3 on one GPU: %.1f GB = %d splits x 64 MiB, d=2, k=%d, %d iterations" % (
            n_splits * split_bytes / 1e9, n_splits, k, args.iterations),
        "points": n_points, "cached_splits": n_cached, "cached_bytes": cached_bytes,
        "iteration1_staging_s": stage_s, "iteration1_stage_gbs": iters[0]["stage_gbs"],
        "iterations_2plus_mean_s": float(np.mean(later)) if later else None,
        "iterations_2plus_points_per_s": n_points / float(np.mean(later)) if later else None,
        "job_s_excluding_generation": stage_s + sum(later),
        "centroid_moves": costs,
        "iterations": iters,
    }
    line = json.dumps(summary)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(json.dumps(summary, indent=1) + "\n")


if __name__ == "__main__":
    main()
