"""GPU: text splits decoded on the device -- field values bit-identical
to sscanf("%f") (HadoopUtils::toFloat), record framing as LineRecordReader does it."""
import os
import random

import numpy as np
import pytest

import cases
from conftest import check_partials
from hgpumap import capi
from test_text_parse_host import GRAMMAR, rand_decimal

pytestmark = pytest.mark.gpu


def render(points, rng, eol=b"\n", per_line=(1, 1), fmt="%.6f", seps=(b" ",), final_eol=True):
    """Points -> text lines; returns bytes."""
    out = []
    i = 0
    n, d = points.shape
    while i < n:
        m = rng.randint(*per_line)
        row = points[i:i + m].reshape(-1)
        parts = []
        for v in row:
            parts.append((fmt % v).encode())
            parts.append(rng.choice(seps))
        parts.pop()
        out.append(b"".join(parts))
        i += m
    text = eol.join(out)
    return text + eol if final_eol else text


def test_device_fields_equal_sscanf_bit_for_bit(oracle):
    rng = random.Random(5)
    fields = []
    while len(fields) < 200000:
        f = rand_decimal(rng).encode()
        if 0 < len(f) <= 127 and not any(c in f for c in b" ,\t\r\n"):
            fields.append(f)
    fields += [g.encode() for g in GRAMMAR if g and "\x00" not in g]
    obits, orc = oracle.to_float_many(fields)
    good = [f for f, r in zip(fields, orc) if r == 0]
    want = obits[orc == 0]
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(np.zeros((1, 1), np.float32))
        slot.cache_put(0, b"\n".join(good) + b"\n", fmt=capi.FMT_TEXT)
        got = slot.cache_read(0).reshape(-1).view(np.uint32)
        np.testing.assert_array_equal(got, want)
        # what sscanf rejects fails the call and names the byte
        bad = [f for f, r in zip(fields, orc) if r != 0]
        assert len(bad) > 1000
        for f in bad[:60] + [g.encode() for g in GRAMMAR if g and "\x00" not in g and oracle.to_float(g) is None]:
            text = b"1.5\n2.5\n" + f + b"\n3.5\n"
            with pytest.raises(capi.HgmError) as err:
                slot.cache_put(1, text, fmt=capi.FMT_TEXT)
            assert err.value.code == -1002 and "byte 8 " in str(err.value), (f, str(err.value))


@pytest.mark.parametrize("eol,per_line,seps,fmt,final_eol", [
    (b"\n", (1, 1), (b" ",), "%.6f", True),
    (b"\r\n", (1, 4), (b",", b"\t", b" ", b" , "), "%.9g", True),
    (b"\r", (2, 2), (b"\t",), "%e", False),
    (b"\n", (1, 3), (b"  ", b",,"), "%.3f", False),
])
def test_text_split_matches_the_cpu_mapper(oracle, eol, per_line, seps, fmt, final_eol):
    rng = random.Random(11)
    pts, cent = cases.blobs(30011, 2, 128, seed=71)
    text = render(pts, rng, eol, per_line, fmt, seps, final_eol)
    rows, _ = oracle.text_split_decode(text, 0, len(text), 2)
    assert len(rows) == len(pts)
    o = oracle.map_combine(rows, cent)
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        sums, counts, _ = slot.map_split(text, fmt=capi.FMT_TEXT)
        assert slot.last_points == len(pts)
        check_partials(sums, counts, o[0], o[1], o[2])
        slot.cache_put(3, text, fmt=capi.FMT_TEXT)
        np.testing.assert_array_equal(slot.cache_read(3).view(np.uint32), rows.view(np.uint32))


@pytest.mark.parametrize("eol", [b"\n", b"\r\n", b"\r"])
def test_file_splits_partition_the_lines_like_linerecordreader(oracle, tmp_path, eol):
    rng = random.Random(13)
    pts, cent = cases.blobs(5003, 3, 17, seed=73)
    text = render(pts, rng, eol, (1, 2), "%.5f", (b" ", b","), final_eol=rng.random() < 0.5)
    text = text.replace(eol, eol + eol, 3)  # a few empty lines
    path = str(tmp_path / "points.txt")
    open(path, "wb").write(text)
    size = len(text)
    cuts = sorted({0, size} | {rng.randrange(size) for _ in range(12)} | {size // 2, size // 2 + 1})
    total = 0
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        for a, b in zip(cuts, cuts[1:]):
            rows, _ = oracle.text_split_decode(text, a, b - a, 3)
            sums, counts, _ = slot.map_file(path, a, b - a, fmt=capi.FMT_TEXT)
            assert slot.last_points == len(rows), (a, b)
            o = oracle.map_combine(rows, cent)
            check_partials(sums, counts, o[0], o[1], o[2])
            total += slot.last_points
            slot.cache_put_file(a, path, a, b - a, fmt=capi.FMT_TEXT)
            np.testing.assert_array_equal(slot.cache_read(a).view(np.uint32), rows.view(np.uint32))
        assert total == len(pts)  # every line belongs to exactly one split
        # every byte position as a boundary, on a small file
        small = text[:text.rfind(eol, 0, 400) + len(eol)]
        open(path, "wb").write(small)
        for a in range(0, len(small) + 1):
            rows_l, _ = oracle.text_split_decode(small, 0, a, 3) if a else (np.zeros((0, 3), np.float32), 0)
            rows_r, _ = oracle.text_split_decode(small, a, len(small) - a, 3)
            n_l = slot.map_file(path, 0, a, fmt=capi.FMT_TEXT) and slot.last_points
            n_r = slot.map_file(path, a, len(small) - a, fmt=capi.FMT_TEXT) and slot.last_points
            assert (n_l, n_r) == (len(rows_l), len(rows_r)), a


def test_a_text_split_larger_than_the_staging_window(oracle):
    """40 MB of text: three 16 MiB windows with carried-over partial lines."""
    pts, cent = cases.blobs(2_000_000, 2, 128, seed=79)
    lines = np.char.add(np.char.add(np.char.mod("%.6f", pts[:, 0]), " "), np.char.mod("%.6f", pts[:, 1]))
    text = ("\n".join(lines.tolist()) + "\n").encode()
    assert len(text) > 34 * 2**20
    rows, _ = oracle.text_split_decode(text, 0, len(text), 2)
    o = oracle.map_combine(rows, cent)
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        sums, counts, _ = slot.map_split(text, fmt=capi.FMT_TEXT)
        assert slot.last_points == len(pts)
        check_partials(sums, counts, o[0], o[1], o[2])


def test_text_errors_fail_the_task(tmp_path):
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(np.zeros((4, 2), np.float32))
        for text, what in [(b"1 2\n3\n4 5\n", "whole number"), (b"1 2\n3 4 5", "whole number"),
                           (b"1 2\n" + b"9" * 128 + b" 1\n", "longer than 127"), (b"1 2\n3 4x\n", "byte 6 "),
                           (b"1;2\n", "byte 0 ")]:
            with pytest.raises(capi.HgmError) as err:
                slot.map_split(text, fmt=capi.FMT_TEXT)
            assert err.value.code == -1002 and what in str(err.value), str(err.value)
        sums, counts, _ = slot.map_split(b"", fmt=capi.FMT_TEXT)          # empty split
        assert counts.sum() == 0
        sums, counts, _ = slot.map_split(b"\n\n \n", fmt=capi.FMT_TEXT)   # only empty lines
        assert counts.sum() == 0
        sums, counts, _ = slot.map_split(b"1 2", fmt=capi.FMT_TEXT)       # no final line end
        assert counts.sum() == 1
