"""CPU: pin the oracle and the wire codec to the reference's known answers."""
import os
import struct

import numpy as np
import pytest

import pipes_harness as ph
from conftest import GOLDEN
from hgpumap import wire

# Hadoop vint known answers: SURVEY.md section 8(c) (127 -> 7f, 128 -> 8f 80, -113 -> 87 70)
# plus the boundaries of the algorithm at SerialUtils.cc:182-211 / WritableUtils.writeVLong.
VINT_KATS = [
    (0, "00"), (1, "01"), (127, "7f"), (-1, "ff"), (-112, "90"),
    (128, "8f80"), (255, "8fff"), (256, "8e0100"), (65535, "8effff"), (65536, "8d010000"),
    (-113, "8770"), (-129, "8780"), (-257, "860100"),
    (2**31 - 1, "8c7fffffff"), (-2**31, "847fffffff"),
    (2**63 - 1, "887fffffffffffffff"), (-2**63, "807fffffffffffffff"),
]


@pytest.mark.parametrize("value,hexed", VINT_KATS)
def test_vint_known_answers(oracle, value, hexed):
    assert wire.vint(value).hex() == hexed
    assert oracle.vint(value).hex() == hexed
    assert wire.read_vint(bytes.fromhex(hexed), 0) == (value, len(hexed) // 2)
    assert oracle.read_vint(bytes.fromhex(hexed)) == (value, len(hexed) // 2)


def test_vint_python_matches_oracle_on_random_values(oracle):
    rng = np.random.default_rng(1)
    vals = np.concatenate([rng.integers(-2**62, 2**62, 2000), rng.integers(-70000, 70000, 2000),
                           rng.integers(-300, 300, 600)])
    for v in vals.tolist():
        enc = wire.vint(v)
        assert enc == oracle.vint(v)
        assert wire.read_vint(enc, 0) == (v, len(enc))


def test_truncated_vint_is_an_error(oracle):
    with pytest.raises(EOFError):
        wire.read_vint(bytes.fromhex("8e01"), 0)
    assert oracle.read_vint(bytes.fromhex("8e01"))[1] == -1


def test_xdr_float_is_big_endian(oracle):
    import ctypes as C
    buf = (C.c_uint8 * 4)()
    for f in (0.0, 1.0, -2.5, 0.1, 3.4e38):
        oracle.lib.hgo_xdr_float(C.c_float(f), buf)
        assert bytes(buf) == struct.pack(">f", f)


def test_to_float_follows_sscanf(oracle):
    # StringUtils.cc:49-56: sscanf("%f%c") must consume the whole field
    assert oracle.to_float("1.5") == np.float32(1.5)
    assert oracle.to_float("  -2e3") == np.float32(-2000.0)
    assert oracle.to_float("0.1") == np.float32(0.1)
    assert oracle.to_float("16777217") == np.float32(16777216.0)     # correctly rounded
    assert oracle.to_float("1.00000017881393432617187500001") == np.float32(1.0000002)  # just above a tie
    assert oracle.to_float("1.5x") is None
    assert oracle.to_float("") is None
    assert np.isinf(oracle.to_float("1e50"))


def test_golden_wordcount_part_uplink_matches_testpipes():
    """The reference's wordcount-part, replayed through the command-file harness, must give
    TestPipes.java:126-133 (everything in partition 0, combiner = reducer, keys sorted)."""
    up = wire.parse_uplink(open(os.path.join(GOLDEN, "wordcount_part.uplink"), "rb").read())
    assert up.done and not up.outputs
    assert all(p == 0 for p, _, _ in up.partitioned)
    text = "".join(f"{k.decode()}\t{v.decode()}\n" for _, k, v in up.partitioned)
    assert text == ph.FIXED_PARTITION_OUTPUT
    # TestPipes.java:188-195: at least one counter in group WORDCOUNT
    assert up.counter("WORDCOUNT", "INPUT_WORDS") == 57
    assert up.counter("WORDCOUNT", "OUTPUT_WORDS") == 44


def test_golden_wordcount_simple_uplink_matches_testpipes():
    """0 reduces: no combiner, no partitioner, records leave in map order (TestPipes.java:114-124)."""
    up = wire.parse_uplink(open(os.path.join(GOLDEN, "wordcount_simple.uplink"), "rb").read())
    assert up.done and not up.partitioned
    text = "".join(f"{k.decode()}\t{v.decode()}\n" for k, v in up.outputs)
    assert text == ph.NO_SORT_OUTPUT_IN_INPUT_ORDER


def test_downlink_roundtrip_framing():
    d = wire.Downlink().authenticate().start().set_job_conf({"a": "b"}).run_map(b"xyz", 2, True)
    d.map_item(b"\x00" * 8, "hello").close()
    raw = d.to_bytes()
    # AUTH_REQ "" ""  START 0  SET_JOB_CONF 2 "a" "b"  RUN_MAP "xyz" 2 1  MAP_ITEM k v  CLOSE
    assert raw == bytes.fromhex("0a0000" "0000" "0102" "0161" "0162" "03" "0378797a" "0201"
                                "0408" + "00" * 8 + "0568656c6c6f" "08")


def test_frame_fixed_map_items_equals_generic_path():
    rng = np.random.default_rng(0)
    keys = rng.integers(0, 256, size=(50, 8), dtype=np.uint8)
    vals = rng.integers(0, 256, size=(50, 8), dtype=np.uint8)
    d = wire.Downlink()
    for k, v in zip(keys, vals):
        d.map_item(k.tobytes(), v.tobytes())
    assert wire.frame_fixed_map_items(keys, vals) == d.to_bytes()


def test_file_split_roundtrip():
    s = wire.file_split("/data/part-00000", 1 << 33, 64 << 20)
    assert s[:2] == b"\x00\x10"  # UTF8.writeString: 2-byte big-endian length (FileSplit.java:81-85)
    assert wire.parse_file_split(s) == ("/data/part-00000", 1 << 33, 64 << 20)


@pytest.mark.parametrize("fname", ["kmeans_blobs_d2_k128.npz", "kmeans_text_d5_k7.npz"])
def test_oracle_restatement_matches_reference_runtime_golden(oracle, fname):
    """The C restatement vs what the K-means Pipes program produced INSIDE the reference
    runtime (golden generated by tests/golden/make_golden.py)."""
    g = np.load(os.path.join(GOLDEN, fname))
    sums, counts, abs_sums, assign = oracle.map_combine(g["points"], g["centroids"])
    np.testing.assert_array_equal(counts, g["counts"])
    np.testing.assert_array_equal(assign, g["assign"].astype(np.int32))
    np.testing.assert_allclose(sums, g["sums"], rtol=1e-12, atol=1e-9)
    assert int(g["points_counter"]) == g["points"].shape[0]
    # CombineRunner order: std::map<string,...> = byte-lexicographic keys (HadoopPipes.cc:660-706)
    keys = [str(k) for k in g["emitted_keys"]]
    assert keys == sorted(keys)
    k = g["centroids"].shape[0]
    nred = int(g["emitted_partitions"].max()) + 1
    assert all(int(p) == int(key) % max(nred, 1) or nred == 1 for p, key in zip(g["emitted_partitions"], g["emitted_keys"]))
    assert set(g["emitted_keys"].tolist()) == set(np.nonzero(g["counts"])[0].tolist())
    assert k >= len(keys)


def test_oracle_edge_case_semantics(oracle):
    g = np.load(os.path.join(GOLDEN, "kmeans_adversarial_2d.npz"))
    pts, cent = g["points"], g["centroids"]
    assign = oracle.assign(pts, cent)
    np.testing.assert_array_equal(assign, g["assign"].astype(np.int32))
    # duplicates 3/5/20: the lowest index wins; -0 centroid (10) never beats +0 (9)
    assert not np.isin(assign, [5, 20, 10]).any()
    # rows with a NaN coordinate make every distance NaN: nothing wins, index 0 stays
    nan_rows = np.isnan(pts).any(axis=1)
    assert nan_rows.any() and (assign[nan_rows] == 0).all()
    # +-inf rows: every distance is +inf (or NaN), never '<' +inf
    inf_rows = np.isinf(pts).any(axis=1) & ~nan_rows
    assert inf_rows.any() and (assign[inf_rows] == 0).all()
    # numpy restatement of the same scan, vectorised
    d2 = ((pts[:, None, 0] - cent[None, :, 0]) ** 2 + (pts[:, None, 1] - cent[None, :, 1]) ** 2).astype(np.float32)
    with np.errstate(invalid="ignore"):
        d2 = np.where(np.isnan(d2), np.inf, d2)
    ref = np.argmin(d2, axis=1)
    ref[np.isinf(d2.min(axis=1))] = 0
    np.testing.assert_array_equal(assign, ref)


def test_synth_generators_agree_bitwise(oracle):
    from hgpumap import synth
    c = synth.centres(synth.SEED_POINTS, 128, 2)
    assert np.array_equal(c.view(np.uint32), oracle.synth_centres(synth.SEED_POINTS, 128, 2).view(np.uint32))
    p = synth.perturb(synth.SEED_CENTROIDS, c)
    assert np.array_equal(p.view(np.uint32), oracle.synth_perturb(synth.SEED_CENTROIDS, c).view(np.uint32))
    x = synth.points(synth.SEED_POINTS, 12345, 4097, c)
    assert np.array_equal(x.view(np.uint32), oracle.synth_points(synth.SEED_POINTS, 12345, 4097, c).view(np.uint32))
    assert abs(float((x - c[oracle.assign(x, c)]).std()) - 1.0) < 0.05
    assert c.min() >= -100 and c.max() < 100
