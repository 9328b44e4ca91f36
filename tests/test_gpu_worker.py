"""GPU: the persistent worker of a GPU map slot (csrc/worker/) -- map tasks that live for one task
each (the -gpubin under the Pipes protocol) hand their batches and file splits to it; the second
iteration's task for the same split maps from the worker's HBM cache."""
import os
import socket
import subprocess
import time

import numpy as np
import pytest

import cases
import pipes_harness as ph
from conftest import PKG, check_partials
from hgpumap import wire

pytestmark = pytest.mark.gpu

GPU_BIN = os.path.join(PKG, "kmeans-pipes-gpu")
WORKER = os.path.join(PKG, "hgm-worker")


@pytest.fixture()
def worker(tmp_path):
    sock = str(tmp_path / "gpu0.sock")
    proc = subprocess.Popen([WORKER, "--device", "0", "--socket", sock], stderr=subprocess.PIPE)
    for _ in range(600):  # context creation takes about a second
        if os.path.exists(sock):
            break
        assert proc.poll() is None, proc.stderr.read().decode()
        time.sleep(0.05)
    yield sock, proc
    if proc.poll() is None:
        proc.terminate()
        proc.wait(timeout=30)


def conf_for(cpath, k, d, sock, fmt="binary", cache=False):
    conf = {"kmeans.k": str(k), "kmeans.d": str(d), "kmeans.centroids": cpath, "kmeans.input.format": fmt,
            "kmeans.gpu.worker.socket": sock.replace("gpu0", "gpu%d"), "hadoop.pipes.java.recordreader": "true"}
    if cache:
        conf["kmeans.gpu.worker.cache"] = "true"
    return conf


def test_piped_task_through_the_worker(worker, oracle, tmp_path):
    sock, _ = worker
    pts, cent = cases.blobs(70001, 2, 128, seed=101)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf_for(cpath, 128, 2, sock))
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.BytesWritable")
          .run_map(b"ignored", 2, True))
    raw = pts.astype("<f4")
    for i in range(0, len(raw), 777):
        dl.map_item((i * 8).to_bytes(8, "big"), raw[i:i + 777].tobytes())
    for attempt in range(2):  # two tasks, one worker: the second finds a warm context
        t0 = time.perf_counter()
        proc, data, _ = ph.run_child(GPU_BIN, dl.close() if attempt == 0 else dl, args=("0",))
        wall = time.perf_counter() - t0
        up = wire.parse_uplink(data)
        assert up.done, proc.stderr
        counts, sums = wire.kmeans_partials([(key, v) for _, key, v in up.partitioned], 128, 2)
        o = oracle.map_combine(pts, cent)
        check_partials(sums, counts, o[0], o[1], o[2])
        assert up.counter("KMEANS", "POINTS") == len(pts)
    assert wall < 1.0, f"a task against a warm worker took {wall:.2f} s"


def test_file_splits_stay_in_the_workers_hbm_cache(worker, oracle, tmp_path):
    sock, _ = worker
    pts, cent0 = cases.blobs(90000, 2, 64, seed=103)
    data = str(tmp_path / "points.bin")
    pts.astype("<f4").tofile(data)
    cent = cent0
    splits = [(0, 40000 * 8), (40000 * 8, 50000 * 8)]
    for iteration in range(3):  # one job per iteration, as the reference runs K-means
        cpath = str(tmp_path / f"cent{iteration}.bin")
        cent.astype("<f4").tofile(cpath)
        conf = conf_for(cpath, 64, 2, sock, cache=True)
        conf["hadoop.pipes.java.recordreader"] = "false"
        total_counts = np.zeros(64, np.int64)
        total_sums = np.zeros((64, 2))
        hits = 0
        for off, length in splits:
            dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
                  .run_map(wire.file_split("file:" + data, off, length), 1, False))
            proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
            up = wire.parse_uplink(out)
            assert up.done, proc.stderr
            c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
            total_counts += c
            total_sums += s
            hits += up.counter("KMEANS", "GPU_CACHE_HITS") or 0
        o = oracle.map_combine(pts, cent)
        check_partials(total_sums, total_counts, o[0], o[1], o[2])
        assert hits == (0 if iteration == 0 else 2), (iteration, hits)
        nz = total_counts > 0
        cent = cent.copy()
        cent[nz] = (total_sums[nz] / total_counts[nz, None]).astype(np.float32)
    # a rewritten file is another split: no stale rows
    (pts + 1).astype("<f4").tofile(data)
    os.utime(data, (time.time() + 5, time.time() + 5))
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
          .run_map(wire.file_split("file:" + data, 0, 40000 * 8), 1, False))
    proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
    up = wire.parse_uplink(out)
    assert up.done and not up.counter("KMEANS", "GPU_CACHE_HITS")
    c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
    o = oracle.map_combine((pts + 1)[:40000], np.fromfile(cpath, "<f4").reshape(64, 2))
    check_partials(s, c, o[0], o[1], o[2])


def test_no_worker_no_task(tmp_path):
    """A task configured for a worker that is not there fails loudly; nothing falls back to the CPU."""
    cent = np.zeros((4, 2), np.float32)
    cpath = str(tmp_path / "cent.bin")
    cent.tofile(cpath)
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf_for(cpath, 4, 2, str(tmp_path / "gpu0.sock")))
          .set_input_types("k", "v").run_map(b"ignored", 1, True))
    dl.map_item(b"\0" * 8, np.zeros(2, "<f4").tobytes())
    proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    assert not wire.parse_uplink(out).done and b"cannot reach the GPU worker" in proc.stderr


def test_worker_errors_reach_the_task(worker, tmp_path):
    sock, proc = worker
    cent = np.zeros((4, 2), np.float32)
    cpath = str(tmp_path / "cent.bin")
    cent.tofile(cpath)
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf_for(cpath, 4, 2, sock))
          .set_input_types("k", "v").run_map(b"ignored", 1, True))
    dl.map_item(b"\0" * 8, b"\0" * 12)  # one and a half rows... the child's own check catches this one
    child, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    assert not wire.parse_uplink(out).done
    conf = conf_for(cpath, 4, 2, sock, fmt="text")
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v").run_map(b"ignored", 1, True))
    dl.map_item(b"\0" * 8, "1.0 oops")
    child, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    assert not wire.parse_uplink(out).done and b"GPU worker: text split" in child.stderr
    assert proc.poll() is None  # the worker survives a failed task


def test_a_batch_of_path_records_is_one_request_to_the_worker(worker, oracle, tmp_path):
    """Task batching: a map task whose records name several splits (kmeans.input.format=path, the NLineInputFormat
    route) sends them to the worker in ONE request; the second task finds all of them in the HBM cache, a rewritten
    file is staged again, and a new cache epoch empties the cache."""
    sock, _ = worker
    pts, cent = cases.blobs(60000, 2, 64, seed=113)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    files = []
    for i, part in enumerate(np.array_split(pts, 5)):
        path = str(tmp_path / f"part-{i}")
        part.astype("<f4").tofile(path)
        files.append(path)
    conf = conf_for(cpath, 64, 2, sock, fmt="path", cache=True)
    conf["kmeans.file.format"] = "binary"
    conf["kmeans.gpu.worker.cache.epoch"] = "data-set-A"

    def task():
        dl = (wire.Downlink().authenticate().start().set_job_conf(conf)
              .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text").run_map(b"split", 1, True))
        for i, path in enumerate(files):
            size = os.path.getsize(path)
            dl.map_item(i.to_bytes(8, "big"), path if i == 0 else f"{path}\t0\t{size}")
        proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
        up = wire.parse_uplink(out)
        assert up.done, proc.stderr
        c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
        return up, c, s

    o = oracle.map_combine(pts, cent)
    up, c, s = task()
    check_partials(s, c, o[0], o[1], o[2])
    assert up.counter("KMEANS", "GPU_CACHE_HITS") == 0 and up.counter("KMEANS", "GPU_BATCHES") == 1
    assert up.counter("KMEANS", "GPU_BYTES_STAGED") == pts.nbytes
    up, c, s = task()
    check_partials(s, c, o[0], o[1], o[2])
    assert up.counter("KMEANS", "GPU_CACHE_HITS") == 5 and up.counter("KMEANS", "GPU_BYTES_STAGED") == 0
    (pts[:12000] * 0.5).astype("<f4").tofile(files[0])  # same size, new content: mtime decides
    os.utime(files[0], (1, 1))
    up, c, s = task()
    o2 = oracle.map_combine(np.concatenate([pts[:12000] * 0.5, pts[12000:]]).astype(np.float32), cent)
    check_partials(s, c, o2[0], o2[1], o2[2])
    assert up.counter("KMEANS", "GPU_CACHE_HITS") == 4
    conf["kmeans.gpu.worker.cache.epoch"] = "data-set-B"
    up, c, s = task()
    check_partials(s, c, o2[0], o2[1], o2[2])
    assert up.counter("KMEANS", "GPU_CACHE_HITS") == 0


@pytest.mark.parametrize("fmt", ["text", "sequencefile"])
def test_text_and_sequencefile_splits_through_the_worker(worker, oracle, tmp_path, fmt):
    import seqfile

    sock, _ = worker
    pts, cent = cases.blobs(20000, 2, 64, seed=107)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    if fmt == "text":
        data = "".join("%.9g %.9g\n" % (x, y) for x, y in pts).encode()
        decode = lambda a, n: oracle.text_split_decode(data, a, n, 2)[0]  # noqa: E731
    else:
        data = seqfile.write(pts, 8, big_endian=True)
        decode = lambda a, n: oracle.seq_split_decode(data, a, n, 2, True)[0]  # noqa: E731
    path = str(tmp_path / ("points." + fmt))
    open(path, "wb").write(data)
    conf = conf_for(cpath, 64, 2, sock, fmt=fmt, cache=True)
    conf["hadoop.pipes.java.recordreader"] = "false"
    cut = len(data) // 2 + 5
    for run in range(2):
        total = 0
        for a, n in ((0, cut), (cut, len(data) - cut)):
            dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
                  .run_map(wire.file_split("file:" + path, a, n), 1, False))
            proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
            up = wire.parse_uplink(out)
            assert up.done, proc.stderr
            c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
            rows = decode(a, n)
            o = oracle.map_combine(rows, cent)
            check_partials(s, c, o[0], o[1], o[2])
            assert up.counter("KMEANS", "GPU_CACHE_HITS") == run
            total += int(c.sum())
        assert total == len(pts)
