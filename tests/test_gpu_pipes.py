"""GPU: the -gpubin (csrc/pipes/kmeans_pipes_gpu.cc) driven like the Java task drives it,
on this project's Pipes runtime and on the reference's, against the CPU program's records."""
import os

import numpy as np
import pytest

import cases
import pipes_harness as ph
from conftest import PKG, REF_DIR, check_partials
from hgpumap import wire

pytestmark = pytest.mark.gpu

GPU_BIN = os.path.join(PKG, "kmeans-pipes-gpu")
GPU_BIN_REFRT = os.path.join(REF_DIR, "kmeans-pipes-gpu-refrt")


def gpu_conf(cpath, k, d, fmt="binary", **extra):
    # piped input unless a test says otherwise: a job whose records come through Java's reader carries
    # hadoop.pipes.java.recordreader=true (pipes/Submitter.java:128-139; absent means false)
    conf = {"kmeans.k": str(k), "kmeans.d": str(d), "kmeans.centroids": cpath, "kmeans.input.format": fmt,
            "hadoop.pipes.java.recordreader": "true"}
    conf.update({key.replace("_", "."): str(v) for key, v in extra.items()})
    return conf


def piped_downlink(pts, conf, num_reduces, per_record):
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf)
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.BytesWritable")
          .run_map(b"ignored", num_reduces, True))
    raw = pts.astype("<f4")
    for i in range(0, len(raw), per_record):
        dl.map_item((i * 8).to_bytes(8, "big"), raw[i:i + per_record].tobytes())
    return dl.close()


def check_records(up, pts, cent, oracle, num_reduces):
    k, d = cent.shape
    recs = up.partitioned if num_reduces else [(0, key, v) for key, v in up.outputs]
    counts, sums = wire.kmeans_partials([(key, v) for _, key, v in recs], k, d)
    o_sums, o_counts, o_abs, _ = oracle.map_combine(pts, cent)
    check_partials(sums, counts, o_sums, o_counts, o_abs)
    keys = [key.decode() for _, key, _ in recs]
    assert keys == sorted(keys), "records must leave in CombineRunner (byte) key order"
    assert len(keys) == len(set(keys)) == int((o_counts > 0).sum()), "one combined record per non-empty cluster"
    if num_reduces:
        assert all(p == int(key) % num_reduces for p, key, _ in recs)
    assert up.counter("KMEANS", "POINTS") == len(pts)
    assert up.counter("KMEANS", "GPU_BATCHES") >= 1
    assert up.counter("KMEANS", "GPU_BYTES_STAGED") == pts.astype("<f4").nbytes  # every row crossed to the device once


@pytest.mark.parametrize("binary", [GPU_BIN, GPU_BIN_REFRT], ids=["own-runtime", "reference-runtime"])
def test_gpubin_piped_records(binary, oracle, tmp_path):
    if not os.path.exists(binary):
        pytest.skip(f"{binary} not built")
    pts, cent = cases.blobs(60001, 2, 128, seed=17)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    proc, data, _ = ph.run_child(binary, piped_downlink(pts, gpu_conf(cpath, 128, 2), 3, 1000), args=("0",))
    up = wire.parse_uplink(data)
    assert up.done and proc.returncode == 1, proc.stderr
    check_records(up, pts, cent, oracle, 3)


def test_gpubin_matches_cpubin_record_for_record(oracle, tmp_path):
    """Same job on the CPU program (reference runtime) and the GPU program: same keys, same
    partitions, same counts; sums within tolerance."""
    cpu = os.path.join(REF_DIR, "kmeans-pipes-cpu")
    if not os.path.exists(cpu):
        pytest.skip("oracle/_ref not built")
    pts, cent = cases.blobs(20000, 2, 128, seed=23)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    res = ph.run_children([
        (cpu, ph.kmeans_downlink(pts, cpath, 128, 2, "binary", 2, 1), None),
        (GPU_BIN, piped_downlink(pts, gpu_conf(cpath, 128, 2), 2, 500), None),
    ])
    cpu_up, gpu_up = wire.parse_uplink(res[0][1]), wire.parse_uplink(res[1][1])
    assert cpu_up.done and gpu_up.done, (res[0][0].stderr, res[1][0].stderr)
    assert [(p, k) for p, k, _ in cpu_up.partitioned] == [(p, k) for p, k, _ in gpu_up.partitioned]
    for (_, _, cv), (_, _, gv) in zip(cpu_up.partitioned, gpu_up.partitioned):
        cf, gf = cv.split(), gv.split()
        assert cf[0] == gf[0]
        np.testing.assert_allclose([float(x) for x in gf[1:]], [float(x) for x in cf[1:]], rtol=1e-9, atol=1e-6)


@pytest.mark.parametrize("key", ["false", None], ids=["key-false", "key-absent"])
def test_gpubin_cpp_record_reader_route_with_filesplit(oracle, tmp_path, key):
    """hadoop.pipes.java.recordreader=false, or absent -- Submitter.getIsJavaRecordReader defaults to false
    (pipes/Submitter.java:137-139), which is what the command in kmeans_pipes_gpu.cc's header submits: Java sends
    only the FileSplit with pipedInput = false; the child maps the byte range itself (hgm_map_file).  Two splits of
    one file = the whole file."""
    pts, cent = cases.blobs(50000, 2, 64, seed=29)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    data = str(tmp_path / "points.bin")
    pts.astype("<f4").tofile(data)
    conf = gpu_conf(cpath, 64, 2)
    if key is None:
        del conf["hadoop.pipes.java.recordreader"]
    else:
        conf["hadoop.pipes.java.recordreader"] = key
    total_counts = np.zeros(64, np.int64)
    total_sums = np.zeros((64, 2))
    for off, length in ((0, 20000 * 8), (20000 * 8, 30000 * 8)):
        dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
              .run_map(wire.file_split("file:" + data, off, length), 1, False))
        proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
        up = wire.parse_uplink(out)
        assert up.done, proc.stderr
        c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
        total_counts += c
        total_sums += s
        assert up.progress[-1] in (0.0, 1.0)
    o = oracle.map_combine(pts, cent)
    check_partials(total_sums, total_counts, o[0], o[1], o[2])


def test_gpubin_path_records_and_no_reduces(oracle, tmp_path):
    """The shipped NLineInputFormat config: each record names a file (conf/mapred-site.xml:13-22);
    with 0 reduces there is no partitioner and records leave as OUTPUT."""
    pts, cent = cases.blobs(30000, 2, 128, seed=31)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    files = []
    for i, part in enumerate(np.array_split(pts, 3)):
        path = str(tmp_path / f"part-{i}")
        part.astype("<f4").tofile(path)
        files.append(path)
    dl = (wire.Downlink().authenticate().start().set_job_conf(gpu_conf(cpath, 128, 2, "path"))
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text")
          .run_map(b"ignored", 0, True))
    dl.map_item((0).to_bytes(8, "big"), files[0])
    dl.map_item((1).to_bytes(8, "big"), f"{files[1]}\t0\t{os.path.getsize(files[1])}")
    dl.map_item((2).to_bytes(8, "big"), "file:" + files[2])
    proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    up = wire.parse_uplink(out)
    assert up.done and not up.partitioned, proc.stderr
    check_records(up, pts, cent, oracle, 0)


def test_gpubin_exports_recheck_counter(oracle, tmp_path):
    """the GPU program reports how many points the tensor-core path had to re-scan exactly
    (KMEANS/GPU_RECHECKS, registered like any Pipes counter: HadoopPipes.cc:346-359)."""
    pts, cent = cases.general_d(6000, 64, 256, seed=41)
    cent = cent.copy()
    cent[9] = cent[8]  # a duplicated centroid: every point of cluster 8 is a tie the scores cannot break
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    proc, data, _ = ph.run_child(GPU_BIN, piped_downlink(pts, gpu_conf(cpath, 256, 64), 1, 100), args=("0",))
    up = wire.parse_uplink(data)
    assert up.done, proc.stderr
    counts, _ = wire.kmeans_partials([(key, v) for _, key, v in up.partitioned], 256, 64)
    assert counts[9] == 0 and counts[8] > 0
    assert up.counter("KMEANS", "GPU_RECHECKS") >= counts[8]
    assert up.counter("KMEANS", "GPU_BYTES_STAGED") == pts.nbytes


def test_gpubin_fails_the_task_loudly(tmp_path):
    """No CPU fallback: a device that does not exist, or a malformed record, fails the task
    (no DONE), which is how Hadoop learns to re-run the split."""
    pts, cent = cases.blobs(100, 2, 8, seed=1)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    proc, out, _ = ph.run_child(GPU_BIN, piped_downlink(pts, gpu_conf(cpath, 8, 2), 1, 10), args=("63",))
    assert not wire.parse_uplink(out).done and proc.returncode == 0
    assert b"Hadoop Pipes Exception" in proc.stderr and b"device 63" in proc.stderr
    dl = (wire.Downlink().authenticate().start().set_job_conf(gpu_conf(cpath, 8, 2)).set_input_types("k", "v")
          .run_map(b"s", 1, True).map_item(b"k", b"\x00" * 12).close())
    proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
    assert not wire.parse_uplink(out).done and b"whole number of rows" in proc.stderr


def test_gpubin_text_records_match_the_cpubin(oracle, tmp_path):
    """kmeans.input.format=text on both programs: the CPU one parses every line with
    sscanf on the host, the GPU one batches the lines and decodes them on the device."""
    cpu = os.path.join(REF_DIR, "kmeans-pipes-cpu")
    if not os.path.exists(cpu):
        pytest.skip("oracle/_ref not built")
    pts, cent = cases.blobs(8000, 2, 128, seed=37)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    dl = ph.kmeans_downlink(pts, cpath, 128, 2, "text", 2)
    res = ph.run_children([(cpu, dl, None), (GPU_BIN, dl, None)])
    cpu_up, gpu_up = wire.parse_uplink(res[0][1]), wire.parse_uplink(res[1][1])
    assert cpu_up.done and gpu_up.done, (res[0][0].stderr, res[1][0].stderr)
    assert [(p, k) for p, k, _ in cpu_up.partitioned] == [(p, k) for p, k, _ in gpu_up.partitioned]
    for (_, _, cv), (_, _, gv) in zip(cpu_up.partitioned, gpu_up.partitioned):
        cf, gf = cv.split(), gv.split()
        assert cf[0] == gf[0]
        np.testing.assert_allclose([float(x) for x in gf[1:]], [float(x) for x in cf[1:]], rtol=1e-9, atol=1e-6)
    assert gpu_up.counter("KMEANS", "POINTS") == len(pts)


def test_gpubin_text_filesplits_cover_the_file_once(oracle, tmp_path):
    """C++ reader route over a text file cut at arbitrary bytes: LineRecordReader's rule gives
    every line to exactly one split."""
    pts, cent = cases.blobs(20000, 2, 64, seed=41)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    text = "".join("%.7g,%.7g\n" % (x, y) for x, y in pts).encode()
    data = str(tmp_path / "points.txt")
    open(data, "wb").write(text)
    conf = gpu_conf(cpath, 64, 2, "text")
    conf["hadoop.pipes.java.recordreader"] = "false"
    cuts = [0, 12345, len(text) // 2 + 3, len(text)]
    total_counts = np.zeros(64, np.int64)
    total_sums = np.zeros((64, 2))
    for a, b in zip(cuts, cuts[1:]):
        dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
              .run_map(wire.file_split("file:" + data, a, b - a), 1, False))
        proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
        up = wire.parse_uplink(out)
        assert up.done, proc.stderr
        c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
        rows, _ = oracle.text_split_decode(text, a, b - a, 2)
        assert c.sum() == len(rows) == up.counter("KMEANS", "POINTS")
        total_counts += c
        total_sums += s
    rows, _ = oracle.text_split_decode(text, 0, len(text), 2)
    o = oracle.map_combine(rows, cent)
    check_partials(total_sums, total_counts, o[0], o[1], o[2])


def test_gpubin_malformed_text_fails_the_task(tmp_path):
    cent = np.zeros((4, 2), np.float32)
    cpath = str(tmp_path / "cent.bin")
    cent.tofile(cpath)
    dl = (wire.Downlink().authenticate().start().set_job_conf(gpu_conf(cpath, 4, 2, "text"))
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text")
          .run_map(b"ignored", 1, True))
    dl.map_item((0).to_bytes(8, "big"), "1.0 2.0")
    dl.map_item((8).to_bytes(8, "big"), "3.0 4.0 5.0")
    proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    up = wire.parse_uplink(out)
    assert not up.done and b"whole number" in proc.stderr


def test_gpubin_sequencefile_filesplits(oracle, tmp_path):
    """C++ reader route over a SequenceFile<LongWritable, BytesWritable>: splits cut anywhere, records
    owned by sync entries (SequenceFileRecordReader.java:40-52, 71-84)."""
    import seqfile

    pts, cent = cases.blobs(30000, 2, 64, seed=43)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    data = seqfile.write(pts, 4, big_endian=True)
    path = str(tmp_path / "points.seq")
    open(path, "wb").write(data)
    conf = gpu_conf(cpath, 64, 2, "sequencefile")
    conf["hadoop.pipes.java.recordreader"] = "false"
    cuts = [0, len(data) // 3 + 11, len(data)]
    total_counts = np.zeros(64, np.int64)
    total_sums = np.zeros((64, 2))
    for a, b in zip(cuts, cuts[1:]):
        dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
              .run_map(wire.file_split("file:" + path, a, b - a), 1, False))
        proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
        up = wire.parse_uplink(out)
        assert up.done, proc.stderr
        c, s = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 64, 2)
        rows, _ = oracle.seq_split_decode(data, a, b - a, 2)
        assert c.sum() == len(rows)
        total_counts += c
        total_sums += s
    o = oracle.map_combine(pts, cent)
    check_partials(total_sums, total_counts, o[0], o[1], o[2])
