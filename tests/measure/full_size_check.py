#!/usr/bin/env python
"""full_size_check.py -- size-independent parity properties at BASELINE.json's FULL per-GPU sizes.

The oracle finishes 2^20 points in seconds, not 125,000,000 x 128 floats, so at the sizes of
configs[3] (d=64, k=1024, 500 M points / 8 GPUs = 62.5 M per GPU, 16 GB) and configs[4] (d=128,
k=4096, 1 G points / 8 GPUs = 125 M per GPU, 64 GB) the map task is checked through properties
that do not need a second full pass on the CPU:

  whole vs chunks   one hgm_map_device over the whole resident shard against the same shard mapped
                    in CHUNKS pieces that accumulate into the same partials: counts must be
                    bit-identical, sums within the 1e-5 bar (conftest.check_partials' formula);
  count closure     sum(counts) == n;
  assignments       the whole run also writes its n assignments; windows of it -- head, tail and
                    the window straddling float index 2^31 (where a 32-bit offset would wrap) --
                    must equal, bit for bit, what the exact FP32 scan kernel (HGM_KERNEL_EXACT,
                    the arithmetic the oracle pins at small sizes) assigns to the same rows, and
                    the window's counts must equal the histogram of those assignments.

  python tests/measure/full_size_check.py --d 128 --k 4096 --points 125000000 --out profiles/x.json

Prints one JSON line (and writes it to --out).  Needs a GPU; not collected by pytest."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "hadoop-1.0.3-gpu_b200")]

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--d", type=int, required=True)
    ap.add_argument("--k", type=int, required=True)
    ap.add_argument("--points", type=int, required=True)
    ap.add_argument("--chunks", type=int, default=16)
    ap.add_argument("--window", type=int, default=1 << 20, help="points per assignment window")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    import torch

    from hgpumap import capi, synth

    n, d, k = args.points, args.d, args.k
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    centres = synth.centres(synth.SEED_POINTS, k, d)
    centroids = synth.perturb(synth.SEED_CENTROIDS, centres)
    slot = capi.GpuMapSlot(0)
    slot.set_centroids(centroids)
    points = torch.empty(n * d, dtype=torch.float32, device=dev)
    slot.synth_points_device(synth.SEED_POINTS, 0, n, centres, points.data_ptr())
    assign = torch.empty(n, dtype=torch.int32, device=dev)
    slot.sync()

    # ---- whole shard in one call, assignments written ------------------------------------
    slot.reset_partials()
    t0 = time.perf_counter()
    slot.map_device(points.data_ptr(), n, assign.data_ptr())
    slot.sync()
    whole_s = time.perf_counter() - t0
    sums_w, counts_w = slot.read_partials()
    family = slot.last_kernel()
    rechecks = slot.last_rechecks()
    assert int(counts_w.sum()) == n, f"count closure: {int(counts_w.sum())} != {n}"
    hist = torch.bincount(assign.to(torch.int64), minlength=k).cpu().numpy()
    assert np.array_equal(hist, counts_w), "counts differ from the histogram of the assignments"

    # ---- the same shard in chunks accumulating into the same partials ----------------------
    slot.reset_partials()
    per = (n + args.chunks - 1) // args.chunks
    for first in range(0, n, per):
        m = min(per, n - first)
        slot.map_device(points.data_ptr() + first * d * 4, m)
    slot.sync()
    sums_c, counts_c = slot.read_partials()
    assert np.array_equal(counts_c, counts_w), "whole vs chunks: counts differ"
    # sum|x| per cluster is bounded by count * (|centre| + a few sigma); use |sum| + count as a safe floor
    ref_abs = np.abs(sums_w) + counts_w[:, None]
    tol = 1e-5 * np.maximum(np.abs(sums_w), 1e-5 * ref_abs)
    err = np.abs(sums_c - sums_w)
    worst = float(np.max(err / np.maximum(tol, 1e-300)))
    assert worst <= 1.0, f"whole vs chunks: sums out of tolerance, worst err/tol = {worst}"

    # ---- assignment windows against the exact scan ---------------------------------------------
    w = min(args.window, n)
    wrap = (1 << 31) // d  # first point whose first float has index >= 2^31
    starts = {"head": 0, "tail": n - w}
    if w // 2 < wrap < n - w // 2:
        starts["float_index_2^31"] = wrap - w // 2
    slot.set_kernel(capi.KERNEL_EXACT)
    exact = torch.empty(w, dtype=torch.int32, device=dev)
    windows = {}
    for name, first in starts.items():
        slot.reset_partials()
        slot.map_device(points.data_ptr() + first * d * 4, w, exact.data_ptr())
        slot.sync()
        assert slot.last_kernel() == capi.KERNEL_EXACT
        _, counts_e = slot.read_partials()
        same = bool(torch.equal(exact, assign[first:first + w]))
        assert same, f"window {name}: assignments of the full-size run differ from the exact scan"
        h = torch.bincount(exact.to(torch.int64), minlength=k).cpu().numpy()
        assert np.array_equal(h, counts_e), f"window {name}: exact-scan counts differ from its assignments"
        windows[name] = {"first_point": first, "points": w, "assignments_bit_exact": same}
    slot.set_kernel(capi.KERNEL_AUTO)
    slot.close()

    line = {"check": "full_size", "d": d, "k": k, "points": n, "bytes": n * d * 4, "kernel_family": family,
            "whole_map_seconds": whole_s, "points_per_sec_with_assignments": n / whole_s,
            "rechecked_points": rechecks, "chunks": args.chunks, "counts_whole_eq_chunks": True,
            "sums_worst_err_over_tol": worst, "count_closure": True, "windows": windows}
    text = json.dumps(line)
    print(text, flush=True)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
