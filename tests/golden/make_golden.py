#!/usr/bin/env python
"""Generate tests/golden/*.  Run in the build container (needs /root/reference through
oracle/_ref); the outputs are committed so the GPU box, which has no reference tree, can
check both the oracle and the CUDA path against what the REFERENCE RUNTIME produced here.

  python tests/golden/make_golden.py

Files:
  wordcount_part.uplink / wordcount_simple.uplink
      raw uplink bytes of the reference's own wordcount-part (2 reduces) and
      wordcount-simple (0 reduces) replaying TestPipes.java:138-147 -- these pin the wire
      codec (hgpumap/wire.py) and the harness to TestPipes.java:101-136.
  kmeans_blobs_d2_k128.npz, kmeans_text_d5_k7.npz
      counts/sums parsed from the uplink of oracle/kmeans_pipes_cpu.cc running INSIDE the
      reference runtime (record pump, CombineRunner, vint framing = reference code), plus
      per-point assignments from the C restatement (oracle/kmeans_oracle.c).
  kmeans_adversarial_2d.npz
      assignments/counts/sums of the C restatement on the edge-case set (tests/cases.py).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "hadoop-1.0.3-gpu_b200"), os.path.join(ROOT, "tests")]

import cases  # noqa: E402
import pipes_harness as ph  # noqa: E402
from conftest import Oracle, REF_DIR, ORACLE_DIR  # noqa: E402
from hgpumap import wire  # noqa: E402


def main():
    oracle = Oracle(os.path.join(ORACLE_DIR, "libkmeans_oracle.so"))
    tmp_cent = {}
    jobs = [(os.path.join(REF_DIR, "wordcount-part"), ph.wordcount_downlink(2), None),
            (os.path.join(REF_DIR, "wordcount-simple"), ph.wordcount_downlink(0), None)]

    pts_a, cent_a = cases.blobs(20000, 2, 128)
    pts_b, cent_b = cases.general_d(3000, 5, 7, seed=5)
    for name, cent in (("a", cent_a), ("b", cent_b)):
        path = os.path.join("/tmp", f"golden_cent_{name}_{os.getpid()}.bin")
        cent.astype("<f4").tofile(path)
        tmp_cent[name] = path
    jobs.append((os.path.join(REF_DIR, "kmeans-pipes-cpu"),
                 ph.kmeans_downlink(pts_a, tmp_cent["a"], 128, 2, "binary", num_reduces=1), None))
    jobs.append((os.path.join(REF_DIR, "kmeans-pipes-cpu"),
                 ph.kmeans_downlink(pts_b, tmp_cent["b"], 7, 5, "text", num_reduces=3), None))
    results = ph.run_children(jobs)
    for (proc, data, _), job in zip(results, jobs):
        assert proc.returncode == 1, (job[0], proc.returncode, proc.stderr)  # bool true -> exit 1
        assert wire.parse_uplink(data).done

    open(os.path.join(HERE, "wordcount_part.uplink"), "wb").write(results[0][1])
    open(os.path.join(HERE, "wordcount_simple.uplink"), "wb").write(results[1][1])

    for fname, pts, cent, res in (("kmeans_blobs_d2_k128.npz", pts_a, cent_a, results[2]),
                                  ("kmeans_text_d5_k7.npz", pts_b, cent_b, results[3])):
        up = wire.parse_uplink(res[1])
        k, d = cent.shape
        counts, sums = wire.kmeans_partials([(key, v) for _, key, v in up.partitioned], k, d)
        parts = np.array([p for p, _, _ in up.partitioned], np.int32)
        keys = np.array([int(key) for _, key, _ in up.partitioned], np.int32)
        o_sums, o_counts, o_abs, o_assign = oracle.map_combine(pts, cent)
        assert np.array_equal(counts, o_counts), "C restatement disagrees with the Pipes program"
        assert np.allclose(sums, o_sums, rtol=1e-12, atol=1e-9)
        np.savez_compressed(os.path.join(HERE, fname), points=pts, centroids=cent, counts=counts, sums=sums,
                            abs_sums=o_abs, assign=o_assign.astype(np.int16), emitted_keys=keys,
                            emitted_partitions=parts, points_counter=up.counter("KMEANS", "POINTS"))

    pts, cent = cases.adversarial_2d()
    o_sums, o_counts, o_abs, o_assign = oracle.map_combine(pts, cent)
    np.savez_compressed(os.path.join(HERE, "kmeans_adversarial_2d.npz"), points=pts, centroids=cent,
                        counts=o_counts, sums=o_sums, abs_sums=o_abs, assign=o_assign.astype(np.int16))
    for p in tmp_cent.values():
        os.remove(p)
    print("golden files written to", HERE)


if __name__ == "__main__":
    main()
