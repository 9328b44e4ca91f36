#!/usr/bin/env python
"""lsu_model.py -- Monte-Carlo bank-conflict model of the d = 2 map kernel's shared-memory accesses.

Shared memory on sm_100 serves one wavefront (32 banks x 4 bytes) per clock per SM; a warp-wide access
costs as many wavefronts as its busiest bank has distinct addresses (atomics: lanes), 64-bit accesses
are served per half-warp.  The kernel's addresses are decided by the point's cluster or cell, so the
conflict degree is a balls-in-bins maximum.  This script reproduces the per-instruction wavefront
counts ncu measured (profiles/r01_v11_grid2d_summary.md) from that model alone and prices the layouts
DESIGN.md section 9 lists, so that the next kernel change can be chosen on paper first.

  python tools/microbench/lsu_model.py [--k 128] [--trials 20000] [--multi 0.06]

No GPU, no oracle; prints a markdown table."""
import argparse

import numpy as np


def max_multiplicity(bins, nbins):
    """busiest bin per row of `bins` (trials x lanes)"""
    out = np.zeros(bins.shape[0], dtype=np.int64)
    for b in range(nbins):
        np.maximum(out, (bins == b).sum(axis=1), out=out)
    return out


def distinct_per_bank(addr_words, nbanks=32):
    """wavefronts of a 32-bit-or-narrower load/store: per bank, DISTINCT 4-byte words (same word = broadcast)"""
    trials, lanes = addr_words.shape
    waves = np.zeros(trials, dtype=np.int64)
    bank = addr_words % nbanks
    for b in range(nbanks):
        words = np.where(bank == b, addr_words, -1)
        words.sort(axis=1)
        distinct = (np.diff(words, axis=1) != 0).sum(axis=1) + 1 - (words[:, 0] == -1)
        np.maximum(waves, distinct, out=waves)
    return waves


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=128)
    ap.add_argument("--trials", type=int, default=20000)
    ap.add_argument("--multi", type=float, default=0.06, help="fraction of points in multi-candidate cells")
    ap.add_argument("--G", type=int, default=256)
    args = ap.parse_args()
    rng = np.random.default_rng(20261107)
    k, T = args.k, args.trials
    rs = (k // 16) * 16 + 8  # row stride of the striped buffer at one owner lane per cluster
    if rs < k:
        rs += 16

    c = rng.integers(0, k, size=(T, 32))                 # cluster of each lane's point
    queued = rng.random((T, 32)) < args.multi            # multi-candidate cell -> queue
    rank = rng.integers(0, 16, size=(T, 32))             # row of the striped buffer (any value: only parity matters)
    rows = []

    # table lookup: LDS.U8 at a random cell of a G x G byte table
    cell = rng.integers(0, args.G * args.G, size=(T, 32))
    rows.append(("LDS.U8 cell table (shipped)", distinct_per_bank(cell // 4).mean(), 3.89))

    # rank atomic: one counter per cluster, one shared queue counter; atomics serialise per lane of a bank
    ctr = np.where(queued, k, c)
    rows.append(("ATOMS.ADD rank (shipped: hist[cluster])", max_multiplicity(ctr % 32, 32).mean(), 3.81))

    # scatter: STS.64 to slot rank*rs + cluster, served per half-warp over 16 eight-byte bank pairs
    slot = rank * rs + c
    qslot = rng.integers(0, 192, size=(T, 32))           # queue slots are consecutive per arrival: model as random
    slot = np.where(queued, 16 * rs + qslot, slot)
    w = max_multiplicity(slot[:, :16] % 16, 16) + max_multiplicity(slot[:, 16:] % 16, 16)
    rows.append(("STS.64 scatter (shipped: slot = rank*rs + cluster)", w.mean(), 6.18))

    # ---- alternatives -------------------------------------------------------------------------
    lane = np.arange(32)[None, :]
    for classes in (2, 4, 8, 16):
        q = lane % classes
        # counter per (cluster, lane class), laid out so that classes sit in different banks
        ctr_bank = ((c % (32 // classes)) * classes + q) % 32
        a = max_multiplicity(np.where(queued, 0, ctr_bank), 32).mean()
        # slot bank pair = class-major: pair = q * (16/classes) + cluster % (16/classes)
        per = max(16 // classes, 1)
        pair = (q % 16) * per % 16 + (c % per)
        pair = np.where(queued, qslot % 16, pair % 16)
        w = max_multiplicity(pair[:, :16], 16) + max_multiplicity(pair[:, 16:], 16)
        mean_rows = 1024 / k / classes
        rows.append((f"sub-columns by lane & {classes - 1}: ATOMS / STS.64 (mean {mean_rows:.1f} points per sub-column per 1024-point tile)",
                     (a, w.mean()), None))

    print("| access | model: wavefronts per warp instruction | ncu |")
    print("|---|---|---|")
    for name, val, meas in rows:
        if isinstance(val, tuple):
            print(f"| {name} | {val[0]:.2f} / {val[1]:.2f} | - |")
        else:
            print(f"| {name} | {val:.2f} | {meas:.2f} |")


if __name__ == "__main__":
    main()
