"""Hadoop Pipes binary wire protocol, Java side restated in Python.

What the Java task writes to the C++ child (the *downlink*) and reads back (the
*uplink*).  Used to drive any Pipes child -- the reference's CPU programs or this
project's GPU bin -- without a JVM, through the child's command-file replay mode
(``hadoop.pipes.command.file``, reference src/c++/pipes/impl/HadoopPipes.cc:1130-1136;
the Java side records such files with ``hadoop.pipes.command-file.keep``,
src/mapred/org/apache/hadoop/mapred/pipes/Submitter.java:240-251).

Reference being mirrored:
  message codes   pipes/BinaryProtocol.java:66-84  ==  HadoopPipes.cc:296-300
  downlink calls  pipes/BinaryProtocol.java:262-347 (start, setJobConf, setInputTypes,
                  runMap, mapItem, runReduce, reduceKey, reduceValue, endOfInput, abort)
  framing         WritableUtils.writeVInt/VLong  ==  SerialUtils.cc:182-244;
                  strings/bytes = vint length + raw bytes (BinaryProtocol.java:349-368)
  uplink loop     pipes/BinaryProtocol.java:115-167
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Iterable, Iterator

# downlink
START, SET_JOB_CONF, SET_INPUT_TYPES, RUN_MAP, MAP_ITEM = 0, 1, 2, 3, 4
RUN_REDUCE, REDUCE_KEY, REDUCE_VALUE, CLOSE, ABORT, AUTHENTICATION_REQ = 5, 6, 7, 8, 9, 10
# uplink
OUTPUT, PARTITIONED_OUTPUT, STATUS, PROGRESS, DONE = 50, 51, 52, 53, 54
REGISTER_COUNTER, INCREMENT_COUNTER, AUTHENTICATION_RESP = 55, 56, 57

CURRENT_PROTOCOL_VERSION = 0  # BinaryProtocol.java:52


def vint(value: int) -> bytes:
    """Hadoop variable-length long (WritableUtils.writeVLong / SerialUtils.cc:182-211)."""
    if -112 <= value <= 127:
        return struct.pack("b", value)
    tag = -112
    if value < 0:
        value ^= -1
        tag = -120
    nbytes = (value.bit_length() + 7) // 8
    return struct.pack("b", tag - nbytes) + value.to_bytes(nbytes, "big")


def read_vint(buf: bytes | memoryview, pos: int) -> tuple[int, int]:
    """Decode one vint at ``pos``; returns (value, new_pos)."""
    first = struct.unpack_from("b", buf, pos)[0]
    if first >= -112:
        return first, pos + 1
    negative = first < -120
    nbytes = (-120 - first) if negative else (-112 - first)
    if pos + 1 + nbytes > len(buf):
        raise EOFError("truncated vint")
    value = int.from_bytes(bytes(buf[pos + 1:pos + 1 + nbytes]), "big")
    if negative:
        value ^= -1
    return value, pos + 1 + nbytes


def _blob(data: bytes | str) -> bytes:
    if isinstance(data, str):
        data = data.encode("utf-8")
    return vint(len(data)) + data


class Downlink:
    """Builds the byte stream a Java ``BinaryProtocol`` would send to the child."""

    def __init__(self) -> None:
        self._parts: list[bytes] = []

    # Application.java:197-211 sends the auth request first; with no job-token password
    # (command-file mode) the child accepts anything (HadoopPipes.cc:396-401).
    def authenticate(self, digest: str = "", challenge: str = "") -> "Downlink":
        self._parts += [vint(AUTHENTICATION_REQ), _blob(digest), _blob(challenge)]
        return self

    def start(self, version: int = CURRENT_PROTOCOL_VERSION) -> "Downlink":
        self._parts += [vint(START), vint(version)]
        return self

    def set_job_conf(self, conf: dict[str, str]) -> "Downlink":
        self._parts += [vint(SET_JOB_CONF), vint(2 * len(conf))]
        for key, value in conf.items():
            self._parts += [_blob(key), _blob(str(value))]
        return self

    def set_input_types(self, key_class: str, value_class: str) -> "Downlink":
        self._parts += [vint(SET_INPUT_TYPES), _blob(key_class), _blob(value_class)]
        return self

    def run_map(self, split: bytes, num_reduces: int, piped_input: bool) -> "Downlink":
        self._parts += [vint(RUN_MAP), _blob(split), vint(num_reduces), vint(1 if piped_input else 0)]
        return self

    def map_item(self, key: bytes | str, value: bytes | str) -> "Downlink":
        self._parts += [vint(MAP_ITEM), _blob(key), _blob(value)]
        return self

    def map_items_raw(self, framed: bytes) -> "Downlink":
        """Append pre-framed MAP_ITEM messages (see :func:`frame_fixed_map_items`)."""
        self._parts.append(framed)
        return self

    def run_reduce(self, partition: int, piped_output: bool) -> "Downlink":
        self._parts += [vint(RUN_REDUCE), vint(partition), vint(1 if piped_output else 0)]
        return self

    def reduce_key(self, key: bytes | str) -> "Downlink":
        self._parts += [vint(REDUCE_KEY), _blob(key)]
        return self

    def reduce_value(self, value: bytes | str) -> "Downlink":
        self._parts += [vint(REDUCE_VALUE), _blob(value)]
        return self

    def close(self) -> "Downlink":
        self._parts.append(vint(CLOSE))
        return self

    def abort(self) -> "Downlink":
        self._parts.append(vint(ABORT))
        return self

    def to_bytes(self) -> bytes:
        return b"".join(self._parts)

    def write(self, path: str) -> None:
        with open(path, "wb") as f:
            for part in self._parts:
                f.write(part)


def frame_fixed_map_items(keys: "object", values: "object") -> bytes:
    """Vectorised MAP_ITEM framing for many records of equal, short (<128 B) size.

    ``keys`` and ``values`` are 2-D uint8 numpy arrays, one row per record; each
    message is ``04 | len(key) key | len(value) value`` exactly as ``mapItem`` +
    ``writeObject`` produce for ``BytesWritable``-free raw payloads
    (BinaryProtocol.java:305-310, 349-368).
    """
    import numpy as np

    keys = np.ascontiguousarray(keys, dtype=np.uint8)
    values = np.ascontiguousarray(values, dtype=np.uint8)
    n, klen = keys.shape
    n2, vlen = values.shape
    if n != n2 or klen > 127 or vlen > 127:
        raise ValueError("frame_fixed_map_items needs equal row counts and records < 128 bytes")
    out = np.empty((n, 3 + klen + vlen), dtype=np.uint8)
    out[:, 0] = MAP_ITEM
    out[:, 1] = klen
    out[:, 2:2 + klen] = keys
    out[:, 2 + klen] = vlen
    out[:, 3 + klen:] = values
    return out.tobytes()


def file_split(path: str, start: int, length: int) -> bytes:
    """Serialised ``FileSplit``: UTF8.writeString(path) + long start + long length,
    all big-endian (src/mapred/org/apache/hadoop/mapred/FileSplit.java:81-85)."""
    raw = path.encode("utf-8")
    return struct.pack(">H", len(raw)) + raw + struct.pack(">qq", start, length)


def parse_file_split(split: bytes) -> tuple[str, int, int]:
    (n,) = struct.unpack_from(">H", split, 0)
    path = split[2:2 + n].decode("utf-8")
    start, length = struct.unpack_from(">qq", split, 2 + n)
    return path, start, length


@dataclass
class Uplink:
    """Everything a child said, in the order ``UplinkReaderThread`` would see it."""

    outputs: list[tuple[bytes, bytes]] = field(default_factory=list)
    partitioned: list[tuple[int, bytes, bytes]] = field(default_factory=list)
    statuses: list[str] = field(default_factory=list)
    progress: list[float] = field(default_factory=list)
    counters: dict[int, tuple[str, str]] = field(default_factory=dict)
    counter_values: dict[int, int] = field(default_factory=dict)
    auth_response: str | None = None
    done: bool = False

    def records(self) -> list[tuple[bytes, bytes]]:
        """Key/value pairs regardless of partitioning, in emission order is not
        preserved across the two kinds; callers that need order use one kind."""
        return self.outputs + [(k, v) for _, k, v in self.partitioned]

    def counter(self, group: str, name: str) -> int:
        for cid, (g, n) in self.counters.items():
            if g == group and n == name:
                return self.counter_values.get(cid, 0)
        raise KeyError((group, name))


def iter_uplink(buf: bytes) -> Iterator[tuple]:
    view = memoryview(buf)
    pos = 0

    def blob() -> bytes:
        nonlocal pos
        n, pos = read_vint(view, pos)
        if n < 0 or pos + n > len(view):
            raise EOFError("truncated uplink string")
        data = bytes(view[pos:pos + n])
        pos += n
        return data

    while pos < len(view):
        code, pos = read_vint(view, pos)
        if code == OUTPUT:
            key = blob()
            yield (OUTPUT, key, blob())
        elif code == PARTITIONED_OUTPUT:
            part, pos = read_vint(view, pos)
            key = blob()
            yield (PARTITIONED_OUTPUT, part, key, blob())
        elif code == STATUS:
            yield (STATUS, blob().decode("utf-8", "replace"))
        elif code == PROGRESS:
            (value,) = struct.unpack_from(">f", view, pos)  # XDR float, SerialUtils.cc:246-253
            pos += 4
            yield (PROGRESS, value)
        elif code == DONE:
            yield (DONE,)
        elif code == REGISTER_COUNTER:
            cid, pos = read_vint(view, pos)
            group = blob().decode("utf-8")
            yield (REGISTER_COUNTER, cid, group, blob().decode("utf-8"))
        elif code == INCREMENT_COUNTER:
            cid, pos = read_vint(view, pos)
            amount, pos = read_vint(view, pos)
            yield (INCREMENT_COUNTER, cid, amount)
        elif code == AUTHENTICATION_RESP:
            yield (AUTHENTICATION_RESP, blob().decode("ascii", "replace"))
        else:
            raise ValueError(f"Bad command code: {code}")  # BinaryProtocol.java:153-155


def parse_uplink(buf: bytes) -> Uplink:
    up = Uplink()
    for msg in iter_uplink(buf):
        code = msg[0]
        if up.done:
            raise ValueError("message after DONE")
        if code == OUTPUT:
            up.outputs.append((msg[1], msg[2]))
        elif code == PARTITIONED_OUTPUT:
            up.partitioned.append((msg[1], msg[2], msg[3]))
        elif code == STATUS:
            up.statuses.append(msg[1])
        elif code == PROGRESS:
            up.progress.append(msg[1])
        elif code == DONE:
            up.done = True
        elif code == REGISTER_COUNTER:
            up.counters[msg[1]] = (msg[2], msg[3])
        elif code == INCREMENT_COUNTER:
            up.counter_values[msg[1]] = up.counter_values.get(msg[1], 0) + msg[2]
        elif code == AUTHENTICATION_RESP:
            up.auth_response = msg[1]
    return up


def kmeans_partials(records: Iterable[tuple[bytes, bytes]], k: int, d: int):
    """Fold K-means partial records (key = decimal cluster id, value =
    ``"count s0 s1 ..."``) into (counts[k] int64, sums[k,d] float64).  A key may
    occur more than once (one record per combiner spill, HadoopPipes.cc:683-705)."""
    import numpy as np

    counts = np.zeros(k, dtype=np.int64)
    sums = np.zeros((k, d), dtype=np.float64)
    for key, value in records:
        c = int(key)
        fields = value.split()
        if len(fields) != d + 1:
            raise ValueError(f"partial record for cluster {c} has {len(fields)} fields, want {d + 1}")
        counts[c] += int(fields[0])
        sums[c] += [float(x) for x in fields[1:]]
    return counts, sums
