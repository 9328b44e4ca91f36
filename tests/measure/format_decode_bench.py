#!/usr/bin/env python
"""Decode rates of the on-device input formats (K6 text, K7 SequenceFile) next to the CPU restatement
(sscanf per field / record walk, one core): a 2-D split of --points points as
  text          "%.6f %.6f\\n" lines                      -> hgm_cache_put(HGM_FMT_TEXT)
  sequencefile  <LongWritable, BytesWritable> records     -> hgm_cache_put_file(HGM_FMT_SEQ_F32BE)
  binary        rows of LE float32 (the baseline: copy only)
Times include the host->device copy from pageable memory (the bytes come from a file or a Python
buffer) and the decode kernels; the map itself is separate (hgm_cache_map).

    python tests/measure/format_decode_bench.py [--points 4000000] [--out profiles/rNN_format_decode.json]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "hadoop-1.0.3-gpu_b200"), os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402

import seqfile  # noqa: E402
from conftest import Oracle  # noqa: E402
from hgpumap import capi, synth  # noqa: E402


def best(fn, reps=3):
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append(time.perf_counter() - t0)
    return min(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=4_000_000)
    ap.add_argument("--rows-per-record", type=int, default=64)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    n, d, k = args.points, 2, 128
    centres = synth.centres(synth.SEED_POINTS, k, d)
    cent = synth.perturb(synth.SEED_CENTROIDS, centres)
    pts = synth.points(synth.SEED_POINTS, 0, n, centres)
    tmp = os.path.join(ROOT, "gpurun_out", "format_tmp")
    os.makedirs(tmp, exist_ok=True)
    lines = np.char.add(np.char.add(np.char.mod("%.6f", pts[:, 0]), " "), np.char.mod("%.6f", pts[:, 1]))
    text = ("\n".join(lines.tolist()) + "\n").encode()
    seq = seqfile.write(pts, args.rows_per_record, big_endian=True)
    seq_path = os.path.join(tmp, "points.seq")
    open(seq_path, "wb").write(seq)
    raw = pts.astype("<f4").tobytes()
    oracle = Oracle(os.path.join(ROOT, "oracle", "libkmeans_oracle.so"))
    out = {"points": n, "d": d, "rows_per_sequencefile_record": args.rows_per_record, "formats": {}}
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        slot.cache_put(0, raw)  # warm up: context, bounce buffers, arena
        slot.cache_put(1, text[: 1 << 20], fmt=capi.FMT_TEXT)

        def record(name, nbytes, gpu_s, cpu_s, rows, want=pts):
            np.testing.assert_array_equal(rows.view(np.uint32), want.view(np.uint32))
            out["formats"][name] = {"input_bytes": nbytes, "gpu_s": gpu_s, "gpu_input_gbs": nbytes / gpu_s / 1e9,
                                    "gpu_points_per_s": n / gpu_s, "cpu_1core_s": cpu_s,
                                    "cpu_points_per_s": (n / cpu_s) if cpu_s else None}
            print(name, json.dumps(out["formats"][name]))

        g = best(lambda: slot.cache_put(10, raw))
        record("binary", len(raw), g, None, slot.cache_read(10))
        g = best(lambda: slot.cache_put(11, text, fmt=capi.FMT_TEXT))
        c = best(lambda: oracle.text_split_decode(text, 0, len(text), d), reps=1)
        record("text", len(text), g, c, slot.cache_read(11), want=oracle.text_split_decode(text, 0, len(text), d)[0])
        g = best(lambda: slot.cache_put_file(12, seq_path, fmt=capi.FMT_SEQ_F32BE))
        c = best(lambda: oracle.seq_split_decode(seq, 0, len(seq), d, True), reps=1)
        record("sequencefile", len(seq), g, c, slot.cache_read(12))
        slot.reset_partials()
        t0 = time.perf_counter()
        slot.cache_map(11)
        slot.sync()
        out["map_of_the_decoded_rows_s"] = time.perf_counter() - t0
    if args.out:
        with open(args.out, "w") as f:
            f.write(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
