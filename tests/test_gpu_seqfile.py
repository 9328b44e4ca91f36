"""GPU: SequenceFile splits walked on the device against the oracle's
restatement of SequenceFileRecordReader."""
import random

import numpy as np
import pytest

import cases
import seqfile
from conftest import check_partials
from hgpumap import capi

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("big_endian,rows_per_record", [(True, 1), (False, 5), (True, 1000)])
def test_sequencefile_splits_match_the_record_reader(oracle, tmp_path, big_endian, rows_per_record):
    fmt = capi.FMT_SEQ_F32BE if big_endian else capi.FMT_SEQ_F32LE
    pts, cent = cases.blobs(60000, 2, 128, seed=91)
    data = seqfile.write(pts, rows_per_record, big_endian=big_endian, metadata=[("k", "v")])
    path = str(tmp_path / "points.seq")
    open(path, "wb").write(data)
    r = random.Random(rows_per_record)
    cuts = sorted({0, len(data)} | {r.randrange(len(data)) for _ in range(6)})
    total = 0
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        sums, counts, _ = slot.map_file(path, 0, -1, fmt=fmt)   # the whole file
        o = oracle.map_combine(pts, cent)
        check_partials(sums, counts, o[0], o[1], o[2])
        for a, b in zip(cuts, cuts[1:]):
            rows, _ = oracle.seq_split_decode(data, a, b - a, 2, big_endian)
            sums, counts, _ = slot.map_file(path, a, b - a, fmt=fmt)
            assert slot.last_points == len(rows), (a, b)
            o = oracle.map_combine(rows, cent)
            check_partials(sums, counts, o[0], o[1], o[2])
            slot.cache_put_file(a, path, a, b - a, fmt=fmt)
            np.testing.assert_array_equal(slot.cache_read(a).view(np.uint32), rows.view(np.uint32))
            total += len(rows)
        assert total == len(pts)  # every record belongs to exactly one split


def test_ragged_records_and_empty_values(oracle, tmp_path):
    pts, cent = cases.general_d(4000, 5, 9, seed=93)
    data = seqfile.write(pts, lambda rng: int(rng.integers(0, 40)), big_endian=True, key_bytes=3)
    path = str(tmp_path / "ragged.seq")
    open(path, "wb").write(data)
    rows, _ = oracle.seq_split_decode(data, 0, len(data), 5)
    np.testing.assert_array_equal(rows.view(np.uint32), pts.view(np.uint32))
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        slot.cache_put_file(0, path, fmt=capi.FMT_SEQ_F32BE)
        np.testing.assert_array_equal(slot.cache_read(0).view(np.uint32), pts.view(np.uint32))
        assert slot.cache_map(0) == len(pts)


def test_bad_sequencefiles_fail_the_task(tmp_path):
    pts = np.arange(40, dtype=np.float32).reshape(20, 2)
    good = seqfile.write(pts, 1)
    cases_ = {
        "version": b"SEQ\x05" + good[4:],
        "BytesWritable": seqfile.header(b"s" * 16, val_class="org.apache.hadoop.io.Text"),
        "compressed": seqfile.header(b"s" * 16, compressed=True),
        "not a SequenceFile": b"XXXX" + good[4:],
    }
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(np.zeros((4, 2), np.float32))
        for what, data in cases_.items():
            path = str(tmp_path / "bad.seq")
            open(path, "wb").write(data)
            with pytest.raises(capi.HgmError) as err:
                slot.map_file(path, fmt=capi.FMT_SEQ_F32BE)
            assert err.value.code == -1002 and what in str(err.value), str(err.value)
        path = str(tmp_path / "good.seq")
        open(path, "wb").write(good)
        slot.set_centroids(np.zeros((4, 3), np.float32))      # rows of 2 floats are not rows of 3
        with pytest.raises(capi.HgmError) as err:
            slot.map_file(path, fmt=capi.FMT_SEQ_F32BE)
        assert "whole rows" in str(err.value)
        truncated = str(tmp_path / "cut.seq")
        open(truncated, "wb").write(good[:-5])                # the last record runs past the end
        slot.set_centroids(np.zeros((4, 2), np.float32))
        with pytest.raises(capi.HgmError) as err:
            slot.map_file(truncated, fmt=capi.FMT_SEQ_F32BE)
        assert "malformed" in str(err.value)
        with pytest.raises(capi.HgmError):                    # memory entry points cannot know the sync marker
            slot.map_split(good, fmt=capi.FMT_SEQ_F32BE)
