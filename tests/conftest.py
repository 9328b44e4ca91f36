"""Shared test plumbing.

`-m "not gpu"` : oracle vs golden vectors, wire codec, host logic, library load/exports.
`-m gpu`       : parity tests proper -- CUDA path through the C ABI vs the oracle.
Only tests (and smoke/bench baselines) touch oracle/; the product never does.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "hadoop-1.0.3-gpu_b200")
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (PKG, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run on the GPU box with -m gpu)")


def _have_gpu() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Oracle:
    """ctypes view of oracle/libkmeans_oracle.so (the CPU restatement)."""

    def __init__(self, path):
        lib = C.CDLL(path)
        self.lib = lib
        lib.hgo_nearest.restype = C.c_int32
        lib.hgo_vint_encode.restype = C.c_int32
        lib.hgo_vint_encode.argtypes = [C.c_int64, C.c_void_p]
        lib.hgo_vint_decode.restype = C.c_int32
        lib.hgo_vint_decode.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
        lib.hgo_to_float.restype = C.c_int32
        lib.hgo_to_float.argtypes = [C.c_char_p, C.POINTER(C.c_float)]
        lib.hgo_decode_text_record.restype = C.c_int32
        lib.hgo_decode_text_record.argtypes = [C.c_char_p, C.c_int64, C.c_void_p, C.c_int32]
        lib.hgo_map_combine.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.hgo_assign.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]
        lib.hgo_synth_centres.argtypes = [C.c_uint64, C.c_int32, C.c_int32, C.c_void_p]
        lib.hgo_synth_perturb.argtypes = [C.c_uint64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        lib.hgo_synth_points.argtypes = [C.c_uint64, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                         C.c_void_p, C.c_void_p]
        lib.hgo_xdr_float.argtypes = [C.c_float, C.c_void_p]

    @staticmethod
    def _p(a):
        return None if a is None else a.ctypes.data_as(C.c_void_p)

    def map_combine(self, pts, cent, want_assign=True):
        pts = np.ascontiguousarray(pts, dtype=np.float32)
        cent = np.ascontiguousarray(cent, dtype=np.float32)
        n = pts.shape[0] if pts.ndim == 2 else 0
        k, d = cent.shape
        sums = np.zeros((k, d), np.float64)
        counts = np.zeros(k, np.int64)
        abs_sums = np.zeros((k, d), np.float64)
        assign = np.zeros(n, np.int32) if want_assign else None
        self.lib.hgo_map_combine(self._p(pts), n, d, self._p(cent), k, self._p(sums), self._p(counts),
                                 self._p(abs_sums), self._p(assign))
        return sums, counts, abs_sums, assign

    def assign(self, pts, cent):
        pts = np.ascontiguousarray(pts, dtype=np.float32)
        cent = np.ascontiguousarray(cent, dtype=np.float32)
        out = np.zeros(pts.shape[0], np.int32)
        self.lib.hgo_assign(self._p(pts), pts.shape[0], cent.shape[1], self._p(cent), cent.shape[0], self._p(out))
        return out

    def vint(self, v):
        buf = (C.c_uint8 * 9)()
        n = self.lib.hgo_vint_encode(v, buf)
        return bytes(buf[:n])

    def read_vint(self, data):
        v = C.c_int64()
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        n = self.lib.hgo_vint_decode(buf, len(data), C.byref(v))
        return v.value, n

    def to_float(self, s):
        f = C.c_float()
        rc = self.lib.hgo_to_float(s.encode() if isinstance(s, str) else s, C.byref(f))
        return None if rc != 0 else np.float32(f.value)

    def to_float_many(self, fields):
        """fields: list of bytes -> (float32 bits as uint32 array, rc array)"""
        buf = b"".join(fields)
        off = np.zeros(len(fields) + 1, np.int64)
        np.cumsum([len(f) for f in fields], out=off[1:])
        out = np.zeros(len(fields), np.float32)
        rc = np.zeros(len(fields), np.int32)
        self.lib.hgo_to_float_many(buf, self._p(off), C.c_int64(len(fields)), self._p(out), self._p(rc))
        return out.view(np.uint32), rc

    def text_split_decode(self, file_bytes, off, length, d):
        """LineRecordReader + text record decode of FileSplit(off, length): (rows [n][d], records)."""
        cap = len(file_bytes) // 2 + 8
        out = np.empty(cap, np.float32)
        nrec = C.c_int64(0)
        self.lib.hgo_text_split_decode.restype = C.c_int64
        n = self.lib.hgo_text_split_decode(file_bytes, C.c_int64(len(file_bytes)), C.c_int64(off), C.c_int64(length),
                                           C.c_int32(d), self._p(out), C.c_int64(cap), C.byref(nrec))
        if n < 0:
            raise ValueError("record %d is malformed or ragged" % (-1 - n))
        return out[:n].reshape(-1, d).copy(), nrec.value

    def seq_split_decode(self, file_bytes, off, length, d, big_endian=True):
        """SequenceFileRecordReader + BytesWritable rows of FileSplit(off, length): (rows [n][d], records)."""
        cap = len(file_bytes) // 4 + 8
        out = np.empty(cap, np.float32)
        nrec = C.c_int64(0)
        self.lib.hgo_seq_split_decode.restype = C.c_int64
        n = self.lib.hgo_seq_split_decode(file_bytes, C.c_int64(len(file_bytes)), C.c_int64(off), C.c_int64(length),
                                          C.c_int32(d), C.c_int32(1 if big_endian else 0), self._p(out), C.c_int64(cap),
                                          C.byref(nrec))
        if n < 0:
            raise ValueError("record %d is malformed, or the header is not acceptable" % (-1 - n))
        return out[:n].reshape(-1, d).copy(), nrec.value

    def synth_centres(self, seed, k, d):
        out = np.empty((k, d), np.float32)
        self.lib.hgo_synth_centres(seed, k, d, self._p(out))
        return out

    def synth_perturb(self, seed, centres):
        out = np.empty_like(centres)
        self.lib.hgo_synth_perturb(seed, centres.shape[0], centres.shape[1], self._p(centres), self._p(out))
        return out

    def synth_points(self, seed, first, n, centres):
        out = np.empty((n, centres.shape[1]), np.float32)
        self.lib.hgo_synth_points(seed, first, n, centres.shape[1], centres.shape[0], self._p(centres), self._p(out))
        return out


@pytest.fixture(scope="session")
def oracle():
    so = os.path.join(ORACLE_DIR, "libkmeans_oracle.so")
    if not os.path.exists(so):
        subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, capture_output=True)
    return Oracle(so)


@pytest.fixture(scope="session")
def ref_bins():
    """The reference's own Pipes runtime + programs, compiled into oracle/_ref (travels to
    the GPU box prebuilt; rebuilt here when /root/reference is present)."""
    need = [os.path.join(REF_DIR, n) for n in ("wordcount-part", "wordcount-simple", "kmeans-pipes-cpu")]
    if not all(os.path.exists(p) for p in need):
        if os.path.isdir("/root/reference/hadoop-1.0.3/src/c++/pipes/impl"):
            subprocess.run(["make", "-C", ORACLE_DIR, "ref"], check=True, capture_output=True)
        else:
            pytest.skip("oracle/_ref not built and /root/reference absent")
    return {os.path.basename(p): p for p in need}


def check_partials(sums, counts, ref_sums, ref_counts, ref_abs, rtol=1e-5):
    """The parity bar of BASELINE.md section 2: counts bit-exact; sums within
    |s - s_ref| <= 1e-5 * max(|s_ref|, 1e-5 * sum|x|); NaN/inf must match in kind."""
    np.testing.assert_array_equal(counts, ref_counts)
    finite = np.isfinite(ref_sums)
    assert np.array_equal(np.isnan(sums), np.isnan(ref_sums)), "NaN pattern of sums differs"
    inf_mask = np.isinf(ref_sums)
    assert np.array_equal(sums[inf_mask], ref_sums[inf_mask]), "inf pattern of sums differs"
    tol = rtol * np.maximum(np.abs(ref_sums[finite]), 1e-5 * ref_abs[finite])
    err = np.abs(sums[finite] - ref_sums[finite])
    bad = err > tol
    assert not bad.any(), f"{bad.sum()} sums out of tolerance; worst err/tol = {np.max(err / np.maximum(tol, 1e-300))}"
