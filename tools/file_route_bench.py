#!/usr/bin/env python
"""Throughput of the file route (hgm_map_file / hgm_cache_put_file: what the worker's MAP_FILE does)
on a page-cache-resident file of binary rows, next to the host-memory routes.

    python tools/file_route_bench.py [--mib 1024]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hadoop-1.0.3-gpu_b200")]

import numpy as np  # noqa: E402

from hgpumap import capi, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    k, d = 128, 2
    n = args.mib * (1 << 20) // 8
    centres = synth.centres(synth.SEED_POINTS, k, d)
    cent = synth.perturb(synth.SEED_CENTROIDS, centres)
    tmp = os.path.join(ROOT, "gpurun_out", "file_tmp")
    os.makedirs(tmp, exist_ok=True)
    path = os.path.join(tmp, "points.bin")
    res = {}
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        pts = capi.synth_points_host(synth.SEED_POINTS, 0, n, centres)
        pts.tofile(path)
        nbytes = pts.nbytes

        def timed(name, fn, reps=3):
            best = 1e9
            for _ in range(reps):
                t0 = time.perf_counter()
                fn()
                best = min(best, time.perf_counter() - t0)
            res[name] = {"s": best, "gbs": nbytes / best / 1e9}
            print(name, json.dumps(res[name]))

        timed("map_file (page cache -> bounce -> H2D -> map)", lambda: slot.map_file(path))
        timed("cache_put_file", lambda: slot.cache_put_file(1, path))
        timed("map_split from pageable memory", lambda: slot.map_split(pts))
        h = slot.host_alloc(nbytes)
        import ctypes as C
        host = np.ctypeslib.as_array(C.cast(h, C.POINTER(C.c_float)), shape=(n, d))
        host[:] = pts
        sums = np.zeros((k, d)); counts = np.zeros(k, np.int64)
        timed("map_split from pinned memory", lambda: slot.map_split_raw(h, nbytes, sums, counts))
        slot.host_free(h)
    os.remove(path)
    if args.out:
        json.dump({"bytes": nbytes, "routes": res}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
