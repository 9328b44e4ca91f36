"""CPU: the -gpubin's side of the worker protocol (csrc/worker/worker_proto.hh) against a stand-in worker
written here -- framing, the `%d` device substitution, batches and file splits, error replies.  The
stand-in answers with the oracle's partials; the real worker is exercised by tests/test_gpu_worker.py."""
import os
import socket
import struct
import threading

import numpy as np
import pytest

import cases
import pipes_harness as ph
from conftest import PKG, check_partials
from hgpumap import wire

GPU_BIN = os.path.join(PKG, "kmeans-pipes-gpu")
MAGIC = 0x574D4748


class StandInWorker(threading.Thread):
    def __init__(self, path, oracle, fail_on_map=False):
        super().__init__(daemon=True)
        self.sock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self.sock.bind(path)
        self.sock.listen(1)
        self.oracle, self.fail = oracle, fail_on_map
        self.log = []

    def _read(self, conn, n):
        buf = b""
        while len(buf) < n:
            part = conn.recv(n - len(buf))
            if not part:
                raise EOFError
            buf += part
        return buf

    def _reply(self, conn, op, status, payload=b""):
        conn.sendall(struct.pack("<IIiIQ", MAGIC, op, status, 0, len(payload)) + payload)

    def _map(self, conn, op, rows, hit=0):
        if self.fail:
            return self._reply(conn, op, -1002, b"stand-in says no")
        sums, counts, _, _ = self.oracle.map_combine(rows, self.cent)
        head = struct.pack("<qfiiiqq", len(rows), 0.25, 1, hit, 0, 3, 0 if hit else rows.nbytes)  # MapReply, protocol 2
        self._reply(conn, op, 0, head + sums.astype("<f8").tobytes() + counts.astype("<i8").tobytes())

    def run(self):
        conn, _ = self.sock.accept()
        try:
            while True:
                magic, op, status, _, n = struct.unpack("<IIiIQ", self._read(conn, 24))
                assert magic == MAGIC and status == 0
                p = self._read(conn, n)
                self.log.append(op)
                if op == 1:
                    self._reply(conn, op, 0, struct.pack("<Ii", 2, 3))
                elif op == 2:
                    k, d = struct.unpack("<ii", p[:8])
                    self.cent = np.frombuffer(p[8:], "<f4").reshape(k, d).copy()
                    self._reply(conn, op, 0)
                elif op == 3:
                    fmt, = struct.unpack("<i", p[:4])
                    assert fmt == 0
                    self._map(conn, op, np.frombuffer(p[4:], "<f4").reshape(-1, self.cent.shape[1]))
                elif op == 4:
                    fmt, use_cache, off, length = struct.unpack("<iiqq", p[:24])
                    self.file_request = (fmt, use_cache, off, length, p[24:].decode())
                    raw = open(p[24:].decode(), "rb").read()[off:off + length]
                    self._map(conn, op, np.frombuffer(raw, "<f4").reshape(-1, self.cent.shape[1]), hit=use_cache)
                elif op == 8:  # MAP_FILES: a batch of splits in one request, one reply over all of them
                    fmt, use_cache, count = struct.unpack("<iii", p[:12])
                    at, rows = 12, []
                    self.batch = []
                    for _ in range(count):
                        off, length, plen = struct.unpack("<qqi", p[at:at + 20])
                        path = p[at + 20:at + 20 + plen].decode()
                        at += 20 + plen
                        raw = open(path, "rb").read()
                        rows.append(np.frombuffer(raw[off:] if length < 0 else raw[off:off + length], "<f4"))
                        self.batch.append((path, off, length))
                    self._map(conn, op, np.concatenate(rows).reshape(-1, self.cent.shape[1]), hit=0)
                elif op == 7:
                    self.epoch = p.decode()
                    self._reply(conn, op, 0)
                else:
                    self._reply(conn, op, -22, b"unknown")
        except EOFError:
            pass
        finally:
            conn.close()
            self.sock.close()


def job_conf(cpath, k, d, sock):
    return {"kmeans.k": str(k), "kmeans.d": str(d), "kmeans.centroids": cpath, "kmeans.input.format": "binary",
            "kmeans.gpu.worker.socket": sock, "hadoop.pipes.java.recordreader": "true"}  # piped input, as Submitter sets it for a Java reader


@pytest.mark.skipif(not os.path.exists(GPU_BIN), reason="kmeans-pipes-gpu not built")
def test_batches_and_file_splits_reach_the_worker(oracle, tmp_path):
    pts, cent = cases.blobs(5000, 2, 16, seed=111)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    sock = str(tmp_path / "gpu3.sock")
    w = StandInWorker(sock, oracle)
    w.start()
    dl = (wire.Downlink().authenticate().start().set_job_conf(job_conf(cpath, 16, 2, str(tmp_path / "gpu%d.sock")))
          .set_input_types("k", "v").run_map(b"ignored", 1, True))
    raw = pts.astype("<f4")
    for i in range(0, len(raw), 100):
        dl.map_item((i).to_bytes(8, "big"), raw[i:i + 100].tobytes())
    proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("3",))  # device 3 -> gpu3.sock
    up = wire.parse_uplink(out)
    assert up.done, proc.stderr
    counts, sums = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 16, 2)
    o = oracle.map_combine(pts, cent)
    check_partials(sums, counts, o[0], o[1], o[2])
    assert w.log == [1, 2, 3] and up.counter("KMEANS", "GPU_MAP_MICROS") == 250

    data = str(tmp_path / "points.bin")
    raw.tofile(data)
    w = StandInWorker(sock.replace("gpu3", "gpu0"), oracle)
    w.start()
    conf = job_conf(cpath, 16, 2, sock.replace("gpu3", "gpu0"))
    conf.update({"hadoop.pipes.java.recordreader": "false", "kmeans.gpu.worker.cache": "true"})
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
          .run_map(wire.file_split("file:" + data, 800, 8000), 1, False))
    proc, out, _ = ph.run_child(GPU_BIN, dl, args=("0",))
    up = wire.parse_uplink(out)
    assert up.done, proc.stderr
    assert w.file_request == (0, 1, 800, 8000, data)
    assert up.counter("KMEANS", "POINTS") == 1000 and up.counter("KMEANS", "GPU_CACHE_HITS") == 1


@pytest.mark.skipif(not os.path.exists(GPU_BIN), reason="kmeans-pipes-gpu not built")
def test_path_records_travel_as_one_batch(oracle, tmp_path):
    """kmeans.input.format=path with a worker: the task's records are gathered and sent in one MAP_FILES request
    (after the cache epoch and the centroids), whatever their number."""
    pts, cent = cases.blobs(3000, 2, 16, seed=115)
    cpath = str(tmp_path / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    files = []
    for i, part in enumerate(np.array_split(pts, 3)):
        path = str(tmp_path / f"part-{i}")
        part.astype("<f4").tofile(path)
        files.append(path)
    sock = str(tmp_path / "gpu0.sock")
    w = StandInWorker(sock, oracle)
    w.start()
    conf = job_conf(cpath, 16, 2, sock)
    conf.update({"kmeans.input.format": "path", "kmeans.file.format": "binary", "kmeans.gpu.worker.cache.epoch": "run-7"})
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf)
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text").run_map(b"split", 1, True))
    dl.map_item((0).to_bytes(8, "big"), files[0])
    dl.map_item((1).to_bytes(8, "big"), f"{files[1]}\t0\t{os.path.getsize(files[1])}")
    dl.map_item((2).to_bytes(8, "big"), "file:" + files[2])
    proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    up = wire.parse_uplink(out)
    assert up.done, proc.stderr
    counts, sums = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 16, 2)
    o = oracle.map_combine(pts, cent)
    check_partials(sums, counts, o[0], o[1], o[2])
    assert w.log == [1, 7, 2, 8] and w.epoch == "run-7"
    assert w.batch == [(files[0], 0, -1), (files[1], 0, os.path.getsize(files[1])), (files[2], 0, -1)]
    assert up.counter("KMEANS", "GPU_BATCHES") == 1 and up.counter("KMEANS", "GPU_RECHECKS") == 3


@pytest.mark.skipif(not os.path.exists(GPU_BIN), reason="kmeans-pipes-gpu not built")
def test_a_worker_error_fails_the_task(oracle, tmp_path):
    cent = np.zeros((4, 2), np.float32)
    cpath = str(tmp_path / "cent.bin")
    cent.tofile(cpath)
    sock = str(tmp_path / "w.sock")
    StandInWorker(sock, oracle, fail_on_map=True).start()
    dl = (wire.Downlink().authenticate().start().set_job_conf(job_conf(cpath, 4, 2, sock))
          .set_input_types("k", "v").run_map(b"ignored", 1, True))
    dl.map_item(b"\0" * 8, np.zeros(2, "<f4").tobytes())
    proc, out, _ = ph.run_child(GPU_BIN, dl.close(), args=("0",))
    assert not wire.parse_uplink(out).done and b"GPU worker: stand-in says no" in proc.stderr
