#!/usr/bin/env python
"""One small parity case per kernel family and form, for `compute-sanitizer --tool memcheck|racecheck|synccheck`
(SURVEY.md section 5).  Every case goes through the C ABI and is checked against the oracle, so a sanitizer run
that changes results is caught too.  n <= 50 k: the tools slow kernels down by one to two orders of magnitude.

  compute-sanitizer --tool racecheck python tools/sanitizer_subset.py [case ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hadoop-1.0.3-gpu_b200"), os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402

import cases  # noqa: E402
from conftest import ORACLE_DIR, Oracle, check_partials  # noqa: E402
from hgpumap import capi  # noqa: E402


def run(slot, oracle, pts, cent, kernel, form=capi.TENSOR_FORM_AUTO, want_assign=True):
    slot.set_tensor_form(form)
    slot.set_kernel(kernel)
    slot.set_centroids(cent)
    sums, counts, assign = slot.map_split(pts, want_assign=want_assign)
    r_sums, r_counts, r_abs, r_assign = oracle.map_combine(pts, cent)
    if want_assign:
        np.testing.assert_array_equal(assign, r_assign)
    check_partials(sums, counts, r_sums, r_counts, r_abs)
    slot.set_kernel(capi.KERNEL_AUTO)
    slot.set_tensor_form(capi.TENSOR_FORM_AUTO)


def tensor_case(d, k, n, seed):
    pts, cent = cases.general_d(n, d, k, seed)
    cent = cent.copy()
    cent[1] = cent[0]  # a duplicated centroid: its cluster's points take the exact re-check (K3)
    return pts, cent


def skewed_rows(d, k, n, seed):
    rng = np.random.default_rng(seed)
    _, cent = cases.general_d(16, d, k, seed)
    which = np.where(rng.random(n) < 0.8, rng.integers(0, 3, n), rng.integers(0, k, n))
    return (cent[which] + rng.normal(0, 0.5, size=(n, d))).astype(np.float32), cent


CASES = {
    "grid2d_fx": lambda s, o: run(s, o, *cases.blobs(50001, 2, 128, seed=3), capi.KERNEL_GRID2D),
    "grid2d_fx_no_assign": lambda s, o: run(s, o, *cases.blobs(50001, 2, 128, seed=4), capi.KERNEL_GRID2D, want_assign=False),
    "grid2d_fx_narrow": lambda s, o: run(s, o, *cases.blobs(30011, 2, 200, seed=5), capi.KERNEL_GRID2D),
    "grid2d_fx_edges": lambda s, o: run(s, o, *cases.adversarial_2d(), capi.KERNEL_GRID2D),
    "grid2d_fx_tiny": lambda s, o: run(s, o, *cases.tiny_values_2d(), capi.KERNEL_GRID2D),
    "grid2d_sort": lambda s, o: run(s, o, *cases.blobs(50001, 2, 128, seed=6), capi.KERNEL_GRID2D_SORT),
    "grid2d_sort_k1000": lambda s, o: run(s, o, *cases.blobs(30011, 2, 1000, seed=7), capi.KERNEL_GRID2D_SORT),
    "grid2d_sort_skew": lambda s, o: run(s, o, *cases.skewed_2d(50001, share=0.9), capi.KERNEL_GRID2D_SORT),
    "exact_d5": lambda s, o: run(s, o, *cases.general_d(5003, 5, 17, seed=8), capi.KERNEL_EXACT),
    "exact_d300": lambda s, o: run(s, o, *cases.general_d(777, 300, 9, seed=9), capi.KERNEL_EXACT),
    "tensor_single_d32": lambda s, o: run(s, o, *tensor_case(32, 128, 5000, 10), capi.KERNEL_TENSOR, capi.TENSOR_FORM_SINGLE),
    "tensor_single_d64": lambda s, o: run(s, o, *tensor_case(64, 256, 4097, 11), capi.KERNEL_TENSOR, capi.TENSOR_FORM_SINGLE),
    "tensor_pair_d64": lambda s, o: run(s, o, *tensor_case(64, 1024, 6001, 12), capi.KERNEL_TENSOR, capi.TENSOR_FORM_PAIR),
    "tensor_pair_d128": lambda s, o: run(s, o, *tensor_case(128, 512, 3001, 13), capi.KERNEL_TENSOR, capi.TENSOR_FORM_PAIR),
    # the per-cluster sums made by the pair kernel's decider warps (three point-tile buffers at d = 64), forced at a small k,
    # and on rows that mostly share three clusters (the in-register merge before the reductions)
    "tensor_pair_fused_d64": lambda s, o: run(s, o, *tensor_case(64, 300, 20011, 14), capi.KERNEL_TENSOR, capi.TENSOR_FORM_PAIR_FUSED),
    "tensor_pair_fused_d100_skew": lambda s, o: run(s, o, *skewed_rows(100, 300, 9001, 15), capi.KERNEL_TENSOR, capi.TENSOR_FORM_PAIR_FUSED),
    "tensor_pair_segmented_d64": lambda s, o: run(s, o, *tensor_case(64, 1024, 6001, 16), capi.KERNEL_TENSOR, capi.TENSOR_FORM_PAIR_SEGMENTED),
}


def main():
    import argparse
    import ctypes as C

    ap = argparse.ArgumentParser()
    ap.add_argument("cases", nargs="*")
    ap.add_argument("--lib", help="another build of the library, e.g. hadoop-1.0.3-gpu_b200/libhgpumap_checked.so (make CHECKED build)")
    ap.add_argument("--repeat", type=int, default=1, help="run every case this many times (a race shows as a parity miss on some run)")
    args = ap.parse_args()
    names = args.cases or list(CASES)
    lib = capi.load_library(args.lib) if args.lib else capi.load_library()
    oracle = Oracle(os.path.join(ORACLE_DIR, "libkmeans_oracle.so"))
    with capi.GpuMapSlot(0) as slot:
        for name in names:
            for _ in range(args.repeat):
                CASES[name](slot, oracle)
            print(f"case {name}: parity ok x{args.repeat}", flush=True)
        count, line = C.c_int32(-1), C.c_int32(0)
        rc = lib.hgm_checked_failures(C.byref(count), C.byref(line))
    print(f"{len(names)} cases x {args.repeat}, all match the oracle")
    if rc == 0:
        print(f"checked build: {count.value} index bound misses" + (f", first at grid2d source line {line.value}" if count.value else ""))
        if count.value:
            sys.exit(1)
    else:
        print("product build (no bounds accounting compiled in)")


if __name__ == "__main__":
    main()
