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


@pytest.mark.parametrize("password,challenge", [
    (b"Jefe", b"what do ya want for nothing?"),                       # RFC 2202 case 2 (key shorter than a block)
    (b"\xaa" * 80, b"Test Using Larger Than Block-Size Key - Hash Key First"),  # RFC 2202 case 6 (key is hashed first)
    (b"k" * 64, b"c" * 55), (b"k" * 65, b"c" * 56), (b"x", b"c" * 119), (b"p" * 20, b"c" * 120),  # SHA-1 padding edges
])
def test_socket_mode_digest_is_hmac_sha1_for_any_key_and_challenge(password, challenge):
    """The runtime carries its own SHA-1/HMAC (no libcrypto): the AUTHENTICATION_RESP must equal
    base64(HMAC-SHA1(password, digest)) as Python's hmac computes it, for short and over-long keys and for
    message lengths on both sides of the padding boundary -- and the child must accept the server's digest."""
    body = wire.Downlink().start().close()
    rc, uplink, err, expected = ph.run_child_over_socket(_bin("wordcount-part-hgmrt"), body, password=password,
                                                         challenge=challenge)
    up = wire.parse_uplink(uplink)
    assert up.done, err
    assert up.auth_response == expected.decode()


def test_socket_mode_rejects_a_wrong_digest():
    body = wire.Downlink().start().close()
    rc, uplink, err, _ = ph.run_child_over_socket(_bin("wordcount-part-hgmrt"), body, bad_digest=True)
    assert b"Server failed to authenticate" in err and rc not in (0, 1)
    assert uplink == b""


# ---- the text protocol: a child started by hand, neither command port nor command file -------------
# (HadoopPipes.cc:1137-1139 -> TextProtocol :181-295, TextUpwardProtocol :122-179)

TEXT_MAP = (b"start\t0\nsetJobConf\t2\tmapred.reduce.tasks\t2\nsetInputTypes\tk\tv\nrunMap\tsplit\t2\ttrue\n"
            b"mapItem\t0\thello world hello\nmapItem\t12\tone two back\\slash  three \x01ctl caf\xc3\xa9\nclose\n")
TEXT_REDUCE = (b"start\t0\nsetJobConf\t2\tmapred.reduce.tasks\t1\nsetInputTypes\tk\tv\nrunReduce\t0\ttrue\n"
               b"reduceKey\tapple\nreduceValue\t1\nreduceValue\t41\nreduceKey\ttab\\there\nreduceValue\t7\nclose\n")
TEXT_CASES = {
    "map + combiner + partitioner": ("wordcount-part", TEXT_MAP),
    "reduce": ("wordcount-simple", TEXT_REDUCE),
    "odd setJobConf count": ("wordcount-simple", b"start\t0\nsetJobConf\t3\ta\tb\tc\nsetInputTypes\tk\tv\n"
                                                  b"runMap\ts\t0\ttrue\nmapItem\t0\tx y\nclose\n"),
    "abort": ("wordcount-part", b"start\t0\nabort\n"),
    "short command": ("wordcount-part", b"start\t0\nsetInputTypes\tk\n"),
    "long command": ("wordcount-part", b"start\t0\nsetInputTypes\tk\tv\textra\n"),
    "illegal command": ("wordcount-part", b"start\t0\nbogus\n"),
    "end of file": ("wordcount-part", b"start\t0\n"),
    "bad version": ("wordcount-part", b"start\t3\n"),
}


def _random_bytes_case():
    """400 map items of words over every byte but tab, newline and '%' (the reference hands the quoted string
    to fprintf as the format, HadoopPipes.cc:128-130): quoting of unprintables and high bytes must agree."""
    import random

    rng = random.Random(5)
    alphabet = [bytes([b]) for b in range(1, 256) if b not in (9, 10, 37)]
    lines = [b"start\t0\n", b"setJobConf\t2\tmapred.reduce.tasks\t0\n", b"setInputTypes\tk\tv\n", b"runMap\ts\t0\ttrue\n"]
    for i in range(400):
        words = [b"".join(rng.choice(alphabet) for _ in range(rng.randint(1, 6))).replace(b" ", b"_")
                 for _ in range(rng.randint(1, 5))]
        lines.append(b"mapItem\t%d\t" % i + b" ".join(words) + b"\n")
    return b"".join(lines) + b"close\n"


TEXT_CASES["random bytes"] = ("wordcount-simple", _random_bytes_case())


@pytest.fixture(scope="module")
def text_runs():
    """Every case on both runtimes at once (the reference joins its ping thread for 5 s per child)."""
    import subprocess

    procs = {}
    for name, (prog, stdin) in TEXT_CASES.items():
        for rt, binary in (("ref", _bin(prog)), ("ours", _bin(prog + "-hgmrt"))):
            p = subprocess.Popen([binary], env={}, stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            p.stdin.write(stdin)
            p.stdin.close()
            procs[name, rt] = p
    out = {}
    for key, p in procs.items():
        stdout, stderr = p.stdout.read(), p.stderr.read()
        p.wait(timeout=60)
        out[key] = (p.returncode, stdout, stderr)
    return out


@pytest.mark.parametrize("name", list(TEXT_CASES))
def test_text_protocol_same_bytes_and_outcome_as_the_reference_runtime(text_runs, name):
    ref, ours = text_runs[name, "ref"], text_runs[name, "ours"]
    assert ours[1] == ref[1], f"uplink text differs:\n{ours[1]!r}\n{ref[1]!r}"
    assert ours[0] == ref[0]  # `return runTask(...)`: 1 = done, 0 = Hadoop Pipes Exception
    failed = b"Hadoop Pipes Exception" in ref[2]
    assert (b"Hadoop Pipes Exception" in ours[2]) == failed
    if failed:  # same complaint; the reference appends its own file:line
        msg = ref[2].split(b"Hadoop Pipes Exception: ")[1].split(b" at /")[0].strip()
        assert msg in ours[2], (msg, ours[2])


def test_text_protocol_output_is_what_the_source_says(text_runs):
    rc, out, _ = text_runs["map + combiner + partitioner", "ours"]
    assert rc == 1 and out.endswith(b"done\n")
    assert b"parititionedOutput\t0\thello\t2\n" in out            # the reference's spelling (HadoopPipes.cc:145)
    assert b"parititionedOutput\t0\tback\\\\slash\t1\n" in out    # backslash quoted, StringUtils.cc:97-127
    assert b"parititionedOutput\t0\t\\01ctl\t1\n" in out          # unprintable byte as \\hex
    assert out.startswith(b"registerCounter\t0\tWORDCOUNT\tINPUT_WORDS\n")
    assert b"progress\t0.000000\n" in out
    rc, out, _ = text_runs["reduce", "ours"]
    assert b"output\tapple\t42\n" in out and b"output\ttab\\\\there\t7\n" in out
