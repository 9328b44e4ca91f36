"""Writes uncompressed version-6 SequenceFiles the way SequenceFile.Writer does
(core/org/apache/hadoop/io/SequenceFile.java: header :920-943, sync entry :985-990,
checkAndWriteSync :1025-1030, append :1040-1075), for the tests of section 8(f)3."""
import struct

import numpy as np

SYNC_INTERVAL = 100 * 20


def vint(n: int) -> bytes:  # WritableUtils.writeVInt, small non-negative values only
    if n <= 127:
        return bytes([n])
    raw = n.to_bytes((n.bit_length() + 7) // 8, "big")
    return bytes([(-112 - len(raw)) & 0xFF]) + raw


def text(s: str) -> bytes:
    b = s.encode()
    return vint(len(b)) + b


def header(sync: bytes, key_class="org.apache.hadoop.io.LongWritable", val_class="org.apache.hadoop.io.BytesWritable",
           metadata=(), compressed=False, block=False, version=6) -> bytes:
    out = b"SEQ" + bytes([version]) + text(key_class) + text(val_class) + bytes([compressed, block])
    out += struct.pack(">i", len(metadata))
    for k, v in metadata:
        out += text(k) + text(v)
    return out + sync


def write(points: np.ndarray, rows_per_record, big_endian=True, seed=1, metadata=(), key_bytes=8) -> bytes:
    """points [n][d] -> file bytes; rows_per_record is an int or a callable(rng) -> int."""
    rng = np.random.default_rng(seed)
    sync = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    out = bytearray(header(sync, metadata=metadata))
    last_sync = 0
    n = len(points)
    raw = points.astype(">f4" if big_endian else "<f4")
    sizes = []  # rows per record; a callable may return 0 (an empty BytesWritable) now and then
    while sum(sizes) < n:
        m = rows_per_record if isinstance(rows_per_record, int) else int(rows_per_record(rng))
        sizes.append(max(0, min(m, n - sum(sizes))))
        assert len(sizes) < 4 * n + 16, "rows_per_record returns 0 too often"
    i = 0
    for rec, m in enumerate(sizes):
        if len(out) >= last_sync + SYNC_INTERVAL:          # checkAndWriteSync
            out += struct.pack(">i", -1) + sync
            last_sync = len(out)
        key = rec.to_bytes(key_bytes, "big")
        payload = raw[i:i + m].tobytes()
        value = struct.pack(">i", len(payload)) + payload   # BytesWritable.write
        out += struct.pack(">ii", len(key) + len(value), len(key)) + key + value
        i += m
    return bytes(out)
