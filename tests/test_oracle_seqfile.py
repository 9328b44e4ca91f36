"""CPU: the oracle's restatement of SequenceFile.Reader + SequenceFileRecordReader against files
written the way SequenceFile.Writer writes them (tests/seqfile.py).  The reference tree ships no
SequenceFile fixture, so the pinning is structural: round trip, and the split rule's defining
property -- consecutive splits deliver every record exactly once, whatever the cut points."""
import random

import numpy as np
import pytest

import seqfile


@pytest.mark.parametrize("big_endian", [True, False])
@pytest.mark.parametrize("rows_per_record", [1, 7, 300])
def test_round_trip_and_split_partition(oracle, big_endian, rows_per_record):
    rng = np.random.default_rng(3)
    pts = rng.normal(0, 50, size=(5000, 3)).astype(np.float32)
    data = seqfile.write(pts, rows_per_record, big_endian=big_endian, metadata=[("made-by", "tests/seqfile.py")])
    rows, nrec = oracle.seq_split_decode(data, 0, len(data), 3, big_endian)
    np.testing.assert_array_equal(rows.view(np.uint32), pts.view(np.uint32))
    assert nrec == -(-len(pts) // rows_per_record)
    r = random.Random(rows_per_record)
    cuts = sorted({0, len(data)} | {r.randrange(len(data)) for _ in range(9)})
    parts = [oracle.seq_split_decode(data, a, b - a, 3, big_endian)[0] for a, b in zip(cuts, cuts[1:])]
    np.testing.assert_array_equal(np.concatenate(parts).view(np.uint32), pts.view(np.uint32))


def test_every_cut_point_on_a_small_file(oracle):
    rng = np.random.default_rng(5)
    pts = rng.normal(0, 5, size=(900, 2)).astype(np.float32)
    data = seqfile.write(pts, 3)
    assert data.count(b"\xff\xff\xff\xff") >= 3  # several sync entries
    for cut in range(0, len(data) + 1, 7):
        left, _ = oracle.seq_split_decode(data, 0, cut, 2)
        right, _ = oracle.seq_split_decode(data, cut, len(data) - cut, 2)
        assert len(left) + len(right) == len(pts), cut
        np.testing.assert_array_equal(np.concatenate([left, right]).view(np.uint32), pts.view(np.uint32))


def test_rejects_what_the_reader_rejects(oracle):
    pts = np.zeros((10, 2), np.float32)
    good = seqfile.write(pts, 1)
    for bad in (b"SEQ\x05" + good[4:],                                            # older version
                seqfile.header(b"s" * 16, val_class="org.apache.hadoop.io.Text"),  # not BytesWritable
                seqfile.header(b"s" * 16, compressed=True),
                b"XXXX" + good[4:]):
        with pytest.raises(ValueError):
            oracle.seq_split_decode(bad, 0, len(bad), 2)
    with pytest.raises(ValueError):
        oracle.seq_split_decode(good, 0, len(good), 3)  # rows of 2 floats are not rows of 3
