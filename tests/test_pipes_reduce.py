"""CPU: the reduce side of a K-means iteration -- the centroid update -- through the Pipes protocol.

In the fork a reduce task always starts the child with runOnGPU = false
(mapred/org/apache/hadoop/mapred/pipes/PipesReducer.java:59-93), i.e. normally the -cpubin; the -gpubin
carries the same Reducer so that either binary can be handed to `-program`.  Its reduce path never opens
the device, so all four builds run here: CPU program and GPU program, each on the reference's Pipes
runtime and on this project's.  RUN_REDUCE / REDUCE_KEY / REDUCE_VALUE framing:
pipes/BinaryProtocol.java:326-347 == c++/pipes/impl/HadoopPipes.cc:526-543."""
import os

import numpy as np
import pytest

import cases
import pipes_harness as ph
from conftest import REF_DIR, ROOT
from hgpumap import wire

BINARIES = {
    "cpu program, reference runtime": os.path.join(REF_DIR, "kmeans-pipes-cpu"),
    "cpu program, this runtime": os.path.join(REF_DIR, "kmeans-pipes-cpu-hgmrt"),
    "gpu program, this runtime": os.path.join(ROOT, "hadoop-1.0.3-gpu_b200", "kmeans-pipes-gpu"),
    "gpu program, reference runtime": os.path.join(REF_DIR, "kmeans-pipes-gpu-refrt"),
}


def reduce_downlink(groups, k, d):
    dl = (wire.Downlink().authenticate().start()
          .set_job_conf({"kmeans.k": str(k), "kmeans.d": str(d), "mapred.reduce.tasks": "1"})
          .set_input_types("org.apache.hadoop.io.Text", "org.apache.hadoop.io.Text").run_reduce(0, True))
    for key, values in groups:
        dl.reduce_key(key)
        for v in values:
            dl.reduce_value(v)
    return dl.close()


@pytest.fixture(scope="module")
def iteration(tmp_path_factory):
    """Map two splits with the CPU program (reference runtime), then hand the partial records to a reduce
    task of every build, grouped by key in the byte order the shuffle's sort would deliver them."""
    for path in BINARIES.values():
        if not os.path.exists(path):
            pytest.skip(f"{path} not built (needs /root/reference at build time)")
    tmp = tmp_path_factory.mktemp("reduce")
    k, d = 16, 2
    pts, cent = cases.blobs(6000, d, k, seed=77)
    cpath = str(tmp / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    halves = [pts[:2500], pts[2500:]]
    maps = ph.run_children([(BINARIES["cpu program, reference runtime"],
                             ph.kmeans_downlink(h, cpath, k, d, "binary", 1, 64), None) for h in halves])
    by_key = {}
    for proc, data, _ in maps:
        up = wire.parse_uplink(data)
        assert up.done
        for key, value in up.records():
            by_key.setdefault(key, []).append(value)
    groups = sorted(by_key.items())  # Text keys compare as bytes: "0", "1", "10", "11", ...
    runs = ph.run_children([(path, reduce_downlink(groups, k, d), None) for path in BINARIES.values()])
    return pts, cent, groups, dict(zip(BINARIES, runs))


def test_reduce_uplinks_are_byte_identical_across_programs_and_runtimes(iteration):
    _, _, _, runs = iteration
    ref = runs["cpu program, reference runtime"][1]
    assert wire.parse_uplink(ref).done and len(ref) > 100
    for name, (proc, data, _) in runs.items():
        assert proc.returncode == 1, (name, proc.stderr[-300:])   # `return runTask(...)`
        assert data == ref, f"{name}: reduce uplink differs from the CPU program on the reference runtime"


def test_new_centroids_are_fp32_of_sum_over_count(iteration):
    pts, cent, groups, runs = iteration
    k = cent.shape[0]
    d2 = ((pts[:, None, :] - cent[None, :, :]) ** 2).sum(axis=2, dtype=np.float32)
    assign = d2.argmin(axis=1)
    up = wire.parse_uplink(runs["gpu program, this runtime"][1])
    got = {int(key): value.split() for key, value in up.records()}
    assert sorted(got) == sorted(set(assign.tolist())) and [k_ for k_, _ in groups] == sorted(k_ for k_, _ in groups)
    for c, fields in got.items():
        members = pts[assign == c].astype(np.float64)
        assert int(fields[0]) == len(members)
        want = (members.sum(axis=0) / len(members)).astype(np.float32)
        have = np.array([np.float32(f.decode()) for f in fields[1:]], dtype=np.float32)   # "%.9g" round-trips FP32
        assert np.allclose(have, want, rtol=1e-6, atol=0), (c, have, want)


def test_reduce_rejects_a_malformed_partial(iteration):
    bad = reduce_downlink([(b"3", [b"2 1.0"])], 16, 2)   # d = 2 but one coordinate
    for name in ("cpu program, this runtime", "gpu program, this runtime"):
        proc, data, _ = ph.run_child(BINARIES[name], bad)
        assert proc.returncode == 0 and b"malformed partial record" in proc.stderr
        assert not wire.parse_uplink(data).done
