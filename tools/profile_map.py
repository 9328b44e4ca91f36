#!/usr/bin/env python
"""A handful of hgm_map_device launches on HBM-resident synthetic points -- the short command to wrap in ncu
(bench.py runs thousands of launches for its clock sampling).

  python tools/profile_map.py [--points N] [--k K] [--dims D] [--kernel F] [--launches L] [--data blobs|uniform|skew]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "hadoop-1.0.3-gpu_b200"), os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=200_000_000)
    ap.add_argument("--k", type=int, default=128)
    ap.add_argument("--dims", type=int, default=2)
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--launches", type=int, default=6)
    ap.add_argument("--data", default="blobs")
    ap.add_argument("--lib", default=None, help="another build of libhgpumap.so (same-box A/B runs)")
    ap.add_argument("--tensor-form", type=int, default=0, help="HGM_TENSOR_FORM_* (0 auto, 3 pair with fused sums, 4 pair with K4)")
    args = ap.parse_args()
    import torch

    from hgpumap import capi, synth

    if args.lib:
        capi.load_library(args.lib)
    n, k, d = args.points, args.k, args.dims
    dev = torch.device("cuda", 0)
    centres = synth.centres(synth.SEED_POINTS, k, d)
    slot = capi.GpuMapSlot(0)
    if args.kernel:
        slot.set_kernel(args.kernel)
    slot.set_tensor_form(args.tensor_form)
    slot.set_centroids(synth.perturb(synth.SEED_CENTROIDS, centres))
    points = torch.empty(n * d, dtype=torch.float32, device=dev)
    if args.data == "blobs":
        slot.synth_points_device(synth.SEED_POINTS, 0, n, centres, points.data_ptr())
    else:
        import bench

        points.copy_(bench.other_points(torch, dev, args.data, n, centres).reshape(-1))
    torch.cuda.synchronize()
    times = []
    for _ in range(args.launches):
        slot.reset_partials()
        slot.map_device(points.data_ptr(), n)
        slot.sync()
        times.append(slot.last_map_ms()[0])
    _, counts = slot.read_partials()
    assert int(counts.sum()) == n
    print(f"kernel family {slot.last_kernel()}, {n} points, d={d}, k={k}: map ms per launch {np.round(times, 4).tolist()}")
    if d > 2:
        print(f"  last launch: K2 / K3 / sums ms {np.round(slot.last_tensor_phases(), 3).tolist()}, {slot.last_rechecks()} rows re-checked")
    slot.close()


if __name__ == "__main__":
    main()
