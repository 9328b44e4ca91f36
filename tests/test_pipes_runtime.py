"""CPU: this project's Pipes child runtime (csrc/pipes/pipes_runtime.cc) against the reference's
(c++/pipes/impl/HadoopPipes.cc, built into oracle/_ref).  The *-hgmrt binaries are the
REFERENCE's example sources compiled unchanged against hgm/pipes_api.hh -- source compatibility --
and their uplink must equal the reference runtime's byte for byte."""
import os
import time

import numpy as np
import pytest

import cases
import pipes_harness as ph
from conftest import GOLDEN, REF_DIR
from hgpumap import wire


def _bin(name):
    path = os.path.join(REF_DIR, name)
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (needs /root/reference at build time)")
    return path


@pytest.fixture(scope="module")
def replay(tmp_path_factory):
    tmp = tmp_path_factory.mktemp("rt")
    pts, cent = cases.blobs(5000, 2, 128, seed=123)
    cpath = str(tmp / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    data = tmp / "part0"
    data.write_text("\n".join(ph.TESTPIPES_INPUT) + "\n")
    nopipe_split = wire.vint(len(str(data))) + str(data).encode()  # Text.writeString, WordCountInputFormat.java:42-44
    nopipe = (wire.Downlink().authenticate().start()
              .set_job_conf({"mapred.task.partition": "0", "mapred.work.output.dir": "file:" + str(tmp / "out")})
              .set_input_types("k", "v").run_map(nopipe_split, 0, False))
    jobs = [
        (_bin("wordcount-part-hgmrt"), ph.wordcount_downlink(2), None),
        (_bin("wordcount-simple-hgmrt"), ph.wordcount_downlink(0), None),
        (_bin("kmeans-pipes-cpu-hgmrt"), ph.kmeans_downlink(pts, cpath, 128, 2, "binary", 1, 1, io_sort_mb=0), None),
        (_bin("kmeans-pipes-cpu"), ph.kmeans_downlink(pts, cpath, 128, 2, "binary", 1, 1, io_sort_mb=0), None),
        (_bin("wordcount-part-hgmrt"), wire.Downlink().authenticate().start().abort(), None),
        (_bin("wordcount-nopipe-hgmrt"), nopipe, None),
        (_bin("wordcount-part-hgmrt"), wire.Downlink().start(), None),  # no AUTHENTICATION_REQ first
        (_bin("wordcount-part-hgmrt"), wire.Downlink().authenticate().start(7), None),  # bad protocol version
    ]
    t0 = time.time()
    res = ph.run_children(jobs)
    return res, time.time() - t0


def test_reference_examples_on_this_runtime_give_the_golden_uplinks(replay):
    res, _ = replay
    assert res[0][1] == open(os.path.join(GOLDEN, "wordcount_part.uplink"), "rb").read()
    assert res[1][1] == open(os.path.join(GOLDEN, "wordcount_simple.uplink"), "rb").read()
    assert res[0][0].returncode == 1 and res[1][0].returncode == 1  # `return runTask(...)`


def test_kmeans_cpu_program_same_bytes_on_both_runtimes(replay):
    res, _ = replay
    assert res[2][1] == res[3][1] and len(res[2][1]) > 100000
    assert wire.parse_uplink(res[2][1]).done


def test_abort_missing_auth_and_bad_version_fail_the_task(replay):
    res, _ = replay
    proc, data, _ = res[4]
    assert proc.returncode == 0 and b"Hadoop Pipes Exception: Aborted by driver" in proc.stderr
    assert not wire.parse_uplink(data).done
    proc, data, _ = res[6]
    assert proc.returncode != 1 and b"received before authentication" in proc.stderr
    proc, data, _ = res[7]
    assert proc.returncode == 0 and b"Protocol version 7 not supported" in proc.stderr


def test_cpp_record_reader_route(replay):
    """hadoop.pipes.java.recordreader=false: the child reads the split itself
    (examples/pipes/impl/wordcount-nopipe.cc:67-113) and reports progress from the reader."""
    res, _ = replay
    proc, data, tmp = res[5]
    up = wire.parse_uplink(data)
    assert proc.returncode == 1 and up.done
    words = [k.decode() for k, _ in up.outputs]
    # wordcount-nopipe's reader drops the last line when the file ends with a newline-terminated record? no:
    assert words == " ".join(ph.TESTPIPES_INPUT).split()
    assert up.progress and 0.0 <= up.progress[-1] <= 1.0


def test_the_runtime_does_not_linger_after_done(replay):
    """The reference joins a ping thread that sleeps 5 s (HadoopPipes.cc:1049,1151); the batch
    above mixes one reference-runtime child with seven of ours, so it takes ~5 s, not ~40."""
    _, elapsed = replay
    assert elapsed < 9.0


def test_socket_mode_with_hmac_authentication():
    """The real transport: loopback socket + HMAC-SHA1 challenge/response
    (pipes/Application.java:197-211, HadoopPipes.cc:395-448), on both runtimes."""
    body = wire.Downlink().start().set_job_conf({"io.sort.mb": "1"}).set_input_types("k", "v").run_map(b"s", 2, True)
    off = 0
    for line in ph.TESTPIPES_INPUT:
        body.map_item(off.to_bytes(8, "big"), line)
        off += len(line) + 1
    body.close()
    golden = open(os.path.join(GOLDEN, "wordcount_part.uplink"), "rb").read()
    outs = {}
    for name in ("wordcount-part-hgmrt", "wordcount-part"):
        rc, uplink, err, expected = ph.run_child_over_socket(_bin(name), body)
        up = wire.parse_uplink(uplink)
        assert rc == 1 and up.done, err
        assert up.auth_response == expected.decode()
        outs[name] = uplink
    assert outs["wordcount-part-hgmrt"] == outs["wordcount-part"]
    # same records as the command-file golden, after the AUTHENTICATION_RESP message
    resp = wire.vint(wire.AUTHENTICATION_RESP) + wire.vint(28)
    assert outs["wordcount-part-hgmrt"].startswith(resp)
    assert outs["wordcount-part-hgmrt"][len(resp) + 28:] == golden


def test_socket_mode_rejects_a_wrong_digest():
    body = wire.Downlink().start().close()
    rc, uplink, err, _ = ph.run_child_over_socket(_bin("wordcount-part-hgmrt"), body, bad_digest=True)
    assert b"Server failed to authenticate" in err and rc not in (0, 1)
    assert uplink == b""
