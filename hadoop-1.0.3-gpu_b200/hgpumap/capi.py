"""ctypes binding of libhgpumap.so -- one Python method per C-ABI entry point.

The class is called ``GpuMapSlot`` after what it is in the reference: one of the
TaskTracker's ``mapred.tasktracker.map.gpu.tasks.maximum`` GPU map slots, bound to
one device (mapred/TaskTracker.java:1428-1433, mapred/TaskTrackerStatus.java:536-551).
A slot maps input splits to combiner output (per-cluster counts and coordinate sums),
which is what the reference's Pipes mapper + combiner pair emits for a split
(c++/pipes/impl/HadoopPipes.cc:865-890 and :660-706).

No fallbacks: if the shared library is missing, ``load_library`` raises with the
build command; if there is no B200, ``GpuMapSlot(...)`` raises ``HgmError``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

FMT_F32LE, FMT_F32BE, FMT_TEXT, FMT_SEQ_F32BE, FMT_SEQ_F32LE = 0, 1, 2, 3, 4
KERNEL_AUTO, KERNEL_EXACT, KERNEL_GRID2D, KERNEL_TENSOR, KERNEL_GRID2D_SORT = 0, 1, 2, 3, 4
GRID2D_FORM_NONE, GRID2D_FORM_FIXED_POINT, GRID2D_FORM_SORT = 0, 1, 2
TENSOR_FORM_AUTO, TENSOR_FORM_SINGLE, TENSOR_FORM_PAIR = 0, 1, 2
TENSOR_FORM_PAIR_FUSED, TENSOR_FORM_PAIR_SEGMENTED = 3, 4  # pair kernel with the sums inside it / by the second pass

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


class HgmError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"hgpumap error {code}: {message}")
        self.code = code


def library_path() -> str:
    return os.path.join(_PKG_ROOT, "libhgpumap.so")


# name -> (restype, argtypes); the same list drives the symbol-export test
_c_ctx = C.c_void_p
_p = C.c_void_p
SIGNATURES = {
    "hgm_version": (C.c_int, []),
    "hgm_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "hgm_open": (C.c_int, [C.c_int, C.POINTER(_c_ctx)]),
    "hgm_close": (None, [_c_ctx]),
    "hgm_last_error": (C.c_char_p, [_c_ctx]),
    "hgm_set_centroids": (C.c_int, [_c_ctx, _p, C.c_int32, C.c_int32]),
    "hgm_set_kernel": (C.c_int, [_c_ctx, C.c_int32]),
    "hgm_last_kernel": (C.c_int, [_c_ctx, C.POINTER(C.c_int32)]),
    "hgm_last_tensor_phases": (C.c_int, [_c_ctx, C.POINTER(C.c_float)]),
    "hgm_cache_map_many": (C.c_int, [_c_ctx, C.POINTER(C.c_int64), C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "hgm_staged_bytes": (C.c_int, [_c_ctx, C.POINTER(C.c_int64)]),
    "hgm_set_tensor_form": (C.c_int, [C.c_int32]),
    "hgm_peer_handle": (C.c_int, [_c_ctx, C.c_int32, C.c_void_p]),
    "hgm_peer_open": (C.c_int, [_c_ctx, C.c_int32, C.c_void_p]),
    "hgm_peer_close": (C.c_int, [_c_ctx]),
    "hgm_checked_failures": (C.c_int, [C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "hgm_grid2d_form": (C.c_int, [_c_ctx, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "hgm_map_split": (C.c_int, [_c_ctx, _p, C.c_size_t, C.c_int32, _p, _p, _p, C.POINTER(C.c_int64)]),
    "hgm_map_file": (C.c_int, [_c_ctx, C.c_char_p, C.c_int64, C.c_int64, C.c_int32, _p, _p, _p,
                               C.POINTER(C.c_int64)]),
    "hgm_map_device": (C.c_int, [_c_ctx, _p, C.c_int64, _p]),
    "hgm_reset_partials": (C.c_int, [_c_ctx]),
    "hgm_read_partials": (C.c_int, [_c_ctx, _p, _p]),
    "hgm_partials_device": (C.c_int, [_c_ctx, C.POINTER(_p), C.POINTER(_p)]),
    "hgm_cache_put": (C.c_int, [_c_ctx, C.c_int64, _p, C.c_size_t, C.c_int32]),
    "hgm_cache_put_file": (C.c_int, [_c_ctx, C.c_int64, C.c_char_p, C.c_int64, C.c_int64, C.c_int32]),
    "hgm_cache_map": (C.c_int, [_c_ctx, C.c_int64, _p, C.POINTER(C.c_int64)]),
    "hgm_cache_drop": (C.c_int, [_c_ctx, C.c_int64]),
    "hgm_cache_map_all": (C.c_int, [_c_ctx, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "hgm_cache_info": (C.c_int, [_c_ctx, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "hgm_cache_read": (C.c_int, [_c_ctx, C.c_int64, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "hgm_sync": (C.c_int, [_c_ctx]),
    "hgm_stream": (C.c_int, [_c_ctx, C.POINTER(_p)]),
    "hgm_last_map_ms": (C.c_int, [_c_ctx, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "hgm_last_rechecks": (C.c_int, [_c_ctx, C.POINTER(C.c_int64)]),
    "hgm_launch_count": (C.c_int, [_c_ctx, C.POINTER(C.c_int64)]),
    "hgm_map_device_static": (C.c_int, [C.c_int, _p, C.c_int64, _p, C.c_int32, C.c_int32, _p, _p, _p, _p]),
    "hgm_device_alloc": (C.c_int, [_c_ctx, C.c_size_t, C.POINTER(_p)]),
    "hgm_device_free": (C.c_int, [_c_ctx, _p]),
    "hgm_host_alloc": (C.c_int, [_c_ctx, C.c_size_t, C.POINTER(_p)]),
    "hgm_host_free": (C.c_int, [_c_ctx, _p]),
    "hgm_memcpy_h2d": (C.c_int, [_c_ctx, _p, _p, C.c_size_t]),
    "hgm_memcpy_d2h": (C.c_int, [_c_ctx, _p, _p, C.c_size_t]),
    "hgm_comm_id": (C.c_int, [_p]),
    "hgm_comm_init": (C.c_int, [_c_ctx, C.c_int32, C.c_int32, _p]),
    "hgm_allreduce_partials": (C.c_int, [_c_ctx]),
    "hgm_synth_centres": (C.c_int, [C.c_uint64, C.c_int32, C.c_int32, _p]),
    "hgm_synth_perturb": (C.c_int, [C.c_uint64, C.c_int32, C.c_int32, _p, _p]),
    "hgm_synth_points_host": (C.c_int, [C.c_uint64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, _p, _p]),
    "hgm_synth_points_device": (C.c_int, [_c_ctx, C.c_uint64, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                          _p, _p]),
}


def load_library(path: str | None = None) -> C.CDLL:
    """dlopen libhgpumap.so and type every entry point.  Loading needs no GPU."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    path = path or library_path()
    if not os.path.exists(path):
        raise FileNotFoundError(
            f"{path} is not built; run `make -C {_PKG_ROOT}` (or __graft_entry__.build()). "
            "There is no CPU fallback for the GPU map path.")
    lib = C.CDLL(path)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _LIB = lib
    return lib


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class GpuMapSlot:
    """One GPU map slot = one ``hgm_ctx``."""

    def __init__(self, device: int = 0):
        self._lib = load_library()
        self._ctx = _c_ctx()
        rc = self._lib.hgm_open(int(device), C.byref(self._ctx))
        if rc != 0:
            raise HgmError(rc, self._lib.hgm_last_error(None).decode())
        self.device = int(device)
        self.k = 0
        self.d = 0

    # -- plumbing -------------------------------------------------------------------
    def _check(self, rc: int) -> None:
        if rc != 0:
            raise HgmError(rc, self._lib.hgm_last_error(self._ctx).decode())

    def close(self) -> None:
        if self._ctx:
            self._lib.hgm_close(self._ctx)
            self._ctx = _c_ctx()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass

    # -- job state --------------------------------------------------------------------
    def set_centroids(self, centroids: np.ndarray) -> None:
        c = np.ascontiguousarray(centroids, dtype=np.float32)
        if c.ndim != 2:
            raise ValueError("centroids must be [k, d]")
        self._check(self._lib.hgm_set_centroids(self._ctx, _ptr(c), c.shape[0], c.shape[1]))
        self.k, self.d = int(c.shape[0]), int(c.shape[1])

    def set_kernel(self, kernel: int) -> None:
        self._check(self._lib.hgm_set_kernel(self._ctx, kernel))

    def set_tensor_form(self, form: int) -> None:
        """Process-wide: which form of the tcgen05 score kernel runs (TENSOR_FORM_*)."""
        self._check(self._lib.hgm_set_tensor_form(form))

    def grid2d_form(self):
        """(form, cells per side) of the installed d == 2 index: GRID2D_FORM_*."""
        form, side = C.c_int32(0), C.c_int32(0)
        self._check(self._lib.hgm_grid2d_form(self._ctx, C.byref(form), C.byref(side)))
        return form.value, side.value

    def last_kernel(self) -> int:
        v = C.c_int32()
        self._check(self._lib.hgm_last_kernel(self._ctx, C.byref(v)))
        return v.value

    # -- the map task -----------------------------------------------------------------
    def map_split(self, split, fmt: int = FMT_F32LE, want_assign: bool = False):
        """Map one split held in host memory.  ``split`` is bytes-like or a numpy array
        (any dtype; its raw bytes are the records).  Returns (sums[k,d] f64,
        counts[k] i64, assign[n] i32 or None)."""
        buf = np.frombuffer(split, dtype=np.uint8) if not isinstance(split, np.ndarray) else split
        buf = np.ascontiguousarray(buf)
        nbytes = buf.nbytes
        sums = np.zeros((self.k, self.d), dtype=np.float64)
        counts = np.zeros(self.k, dtype=np.int64)
        assign = None
        if want_assign:
            if fmt == FMT_TEXT:
                raise ValueError("assignments of a text split need its point count; decode first")
            assign = np.empty(nbytes // (4 * max(self.d, 1)), dtype=np.int32)
        n = C.c_int64()
        self._check(self._lib.hgm_map_split(self._ctx, _ptr(buf) if nbytes else None, nbytes, fmt,
                                            _ptr(sums), _ptr(counts), _ptr(assign), C.byref(n)))
        self.last_points = n.value
        return sums, counts, assign

    def map_split_raw(self, host_ptr: int, nbytes: int, sums: np.ndarray, counts: np.ndarray,
                      fmt: int = FMT_F32LE, assign_ptr: int | None = None) -> int:
        """Pointer-level variant for pinned buffers owned by the caller (bench e2e leg)."""
        n = C.c_int64()
        self._check(self._lib.hgm_map_split(self._ctx, host_ptr, nbytes, fmt, _ptr(sums), _ptr(counts),
                                            assign_ptr, C.byref(n)))
        return n.value

    def map_file(self, path: str, offset: int = 0, length: int = -1, fmt: int = FMT_F32LE,
                 n_points_hint: int | None = None, want_assign: bool = False):
        sums = np.zeros((self.k, self.d), dtype=np.float64)
        counts = np.zeros(self.k, dtype=np.int64)
        assign = None
        if want_assign:
            size = length if length >= 0 else os.path.getsize(path) - offset
            assign = np.empty(n_points_hint or size // (4 * self.d), dtype=np.int32)
        n = C.c_int64()
        self._check(self._lib.hgm_map_file(self._ctx, path.encode(), offset, length, fmt, _ptr(sums),
                                           _ptr(counts), _ptr(assign), C.byref(n)))
        self.last_points = n.value
        return sums, counts, assign

    # -- HBM-resident split cache -------------------------------------------------------
    def cache_put(self, split_id: int, split, fmt: int = FMT_F32LE) -> None:
        buf = np.frombuffer(split, dtype=np.uint8) if not isinstance(split, np.ndarray) else split
        buf = np.ascontiguousarray(buf)
        self._check(self._lib.hgm_cache_put(self._ctx, split_id, _ptr(buf) if buf.nbytes else None, buf.nbytes, fmt))

    def cache_put_file(self, split_id: int, path: str, offset: int = 0, length: int = -1,
                       fmt: int = FMT_F32LE) -> None:
        self._check(self._lib.hgm_cache_put_file(self._ctx, split_id, path.encode(), offset, length, fmt))

    def cache_map(self, split_id: int, d_assign: int | None = None) -> int:
        n = C.c_int64()
        self._check(self._lib.hgm_cache_map(self._ctx, split_id, d_assign, C.byref(n)))
        return n.value

    def cache_map_all(self) -> tuple[int, int]:
        """Map every cached split; returns (points, launch groups)."""
        n, g = C.c_int64(), C.c_int32()
        self._check(self._lib.hgm_cache_map_all(self._ctx, C.byref(n), C.byref(g)))
        return n.value, g.value

    def cache_map_many(self, split_ids) -> tuple[int, int]:
        """Map the listed cached splits (neighbours in the arena share a launch); returns (points, launch groups)."""
        ids = (C.c_int64 * max(len(split_ids), 1))(*split_ids)
        n, g = C.c_int64(), C.c_int32()
        self._check(self._lib.hgm_cache_map_many(self._ctx, ids, len(split_ids), C.byref(n), C.byref(g)))
        return n.value, g.value

    def cache_read(self, split_id: int) -> np.ndarray:
        """The cached split's decoded rows, [n, d] float32."""
        n = C.c_int64()
        self._lib.hgm_cache_read(self._ctx, split_id, None, 0, C.byref(n))  # size query
        rows = np.empty(n.value, dtype=np.float32)
        self._check(self._lib.hgm_cache_read(self._ctx, split_id, _ptr(rows) if n.value else None, n.value, C.byref(n)))
        return rows.reshape(-1, self.d)

    def cache_drop(self, split_id: int = -1) -> None:
        self._check(self._lib.hgm_cache_drop(self._ctx, split_id))

    def cache_info(self) -> tuple[int, int]:
        n, b = C.c_int64(), C.c_int64()
        self._check(self._lib.hgm_cache_info(self._ctx, C.byref(n), C.byref(b)))
        return n.value, b.value

    # -- device-resident splits ---------------------------------------------------------
    def map_device(self, d_points: int, n_points: int, d_assign: int | None = None) -> None:
        self._check(self._lib.hgm_map_device(self._ctx, d_points, n_points, d_assign))

    def reset_partials(self) -> None:
        self._check(self._lib.hgm_reset_partials(self._ctx))

    def read_partials(self):
        sums = np.zeros((self.k, self.d), dtype=np.float64)
        counts = np.zeros(self.k, dtype=np.int64)
        self._check(self._lib.hgm_read_partials(self._ctx, _ptr(sums), _ptr(counts)))
        return sums, counts

    def partials_device(self) -> tuple[int, int]:
        s, c = _p(), _p()
        self._check(self._lib.hgm_partials_device(self._ctx, C.byref(s), C.byref(c)))
        return s.value, c.value

    def sync(self) -> None:
        self._check(self._lib.hgm_sync(self._ctx))

    def stream(self) -> int:
        p = _p()
        self._check(self._lib.hgm_stream(self._ctx, C.byref(p)))
        return p.value or 0

    def last_map_ms(self) -> tuple[float, int]:
        ms, n = C.c_float(), C.c_int32()
        self._check(self._lib.hgm_last_map_ms(self._ctx, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def last_tensor_phases(self):
        """(scores K2, re-check K3, sums K4) in ms for the last launch group of the tensor-core path."""
        ms = (C.c_float * 3)()
        self._check(self._lib.hgm_last_tensor_phases(self._ctx, ms))
        return [float(v) for v in ms]

    def staged_bytes(self) -> int:
        n = C.c_int64()
        self._check(self._lib.hgm_staged_bytes(self._ctx, C.byref(n)))
        return n.value

    def last_rechecks(self) -> int:
        n = C.c_int64()
        self._check(self._lib.hgm_last_rechecks(self._ctx, C.byref(n)))
        return n.value

    def launch_count(self) -> int:
        n = C.c_int64()
        self._check(self._lib.hgm_launch_count(self._ctx, C.byref(n)))
        return n.value

    def device_alloc(self, nbytes: int) -> int:
        p = _p()
        self._check(self._lib.hgm_device_alloc(self._ctx, nbytes, C.byref(p)))
        return p.value

    def device_free(self, ptr: int) -> None:
        self._check(self._lib.hgm_device_free(self._ctx, ptr))

    def host_alloc(self, nbytes: int) -> int:
        p = _p()
        self._check(self._lib.hgm_host_alloc(self._ctx, nbytes, C.byref(p)))
        return p.value

    def host_free(self, ptr: int) -> None:
        self._check(self._lib.hgm_host_free(self._ctx, ptr))

    def memcpy_h2d(self, d_dst: int, src: np.ndarray) -> None:
        src = np.ascontiguousarray(src)
        self._check(self._lib.hgm_memcpy_h2d(self._ctx, d_dst, _ptr(src), src.nbytes))

    def memcpy_d2h(self, dst: np.ndarray, d_src: int) -> None:
        self._check(self._lib.hgm_memcpy_d2h(self._ctx, _ptr(dst), d_src, dst.nbytes))

    # -- multi-GPU ----------------------------------------------------------------------
    @staticmethod
    def comm_id() -> bytes:
        lib = load_library()
        buf = C.create_string_buffer(128)
        rc = lib.hgm_comm_id(buf)
        if rc != 0:
            raise HgmError(rc, lib.hgm_last_error(None).decode())
        return buf.raw

    def comm_init(self, n_ranks: int, rank: int, comm_id: bytes) -> None:
        buf = C.create_string_buffer(bytes(comm_id), 128)
        self._check(self._lib.hgm_comm_init(self._ctx, n_ranks, rank, buf))

    def peer_handle(self, n_ranks: int) -> bytes:
        """This rank's inbox handle (64 bytes) for the peer-memory exchange of small partials."""
        buf = C.create_string_buffer(64)
        self._check(self._lib.hgm_peer_handle(self._ctx, n_ranks, buf))
        return buf.raw

    def peer_open(self, rank: int, handles) -> None:
        """handles: the n_ranks 64-byte handles in rank order."""
        blob = b"".join(bytes(h) for h in handles)
        buf = C.create_string_buffer(blob, len(blob))
        self._check(self._lib.hgm_peer_open(self._ctx, rank, buf))

    def peer_close(self) -> None:
        self._check(self._lib.hgm_peer_close(self._ctx))

    def allreduce_partials(self) -> None:
        self._check(self._lib.hgm_allreduce_partials(self._ctx))

    # -- synthetic workload -------------------------------------------------------------
    def synth_points_device(self, seed: int, first: int, n: int, centres: np.ndarray, d_out: int) -> None:
        c = np.ascontiguousarray(centres, dtype=np.float32)
        self._check(self._lib.hgm_synth_points_device(self._ctx, seed, first, n, c.shape[1], c.shape[0],
                                                      _ptr(c), d_out))


def map_device_static(device: int, d_points: int, n_points: int, centroids: np.ndarray,
                      d_assign: int | None = None, stream: int | None = None):
    """hgm_map_device_static: the d == 2 map task without any device allocation (the scratch
    is a static device array inside the library).  Returns (sums[k,2] f64, counts[k] i64)."""
    lib = load_library()
    c = np.ascontiguousarray(centroids, dtype=np.float32)
    k, d = c.shape
    sums = np.zeros((k, d), dtype=np.float64)
    counts = np.zeros(k, dtype=np.int64)
    rc = lib.hgm_map_device_static(device, d_points, n_points, _ptr(c), k, d, _ptr(sums), _ptr(counts),
                                   d_assign, stream)
    if rc != 0:
        raise HgmError(rc, lib.hgm_last_error(None).decode())
    return sums, counts


def synth_centres(seed: int, k: int, d: int) -> np.ndarray:
    lib = load_library()
    out = np.empty((k, d), dtype=np.float32)
    rc = lib.hgm_synth_centres(seed, k, d, _ptr(out))
    if rc != 0:
        raise HgmError(rc, lib.hgm_last_error(None).decode())
    return out


def synth_perturb(seed: int, centres: np.ndarray) -> np.ndarray:
    lib = load_library()
    c = np.ascontiguousarray(centres, dtype=np.float32)
    out = np.empty_like(c)
    rc = lib.hgm_synth_perturb(seed, c.shape[0], c.shape[1], _ptr(c), _ptr(out))
    if rc != 0:
        raise HgmError(rc, lib.hgm_last_error(None).decode())
    return out


def synth_points_host(seed: int, first: int, n: int, centres: np.ndarray) -> np.ndarray:
    lib = load_library()
    c = np.ascontiguousarray(centres, dtype=np.float32)
    out = np.empty((n, c.shape[1]), dtype=np.float32)
    rc = lib.hgm_synth_points_host(seed, first, n, c.shape[1], c.shape[0], _ptr(c), _ptr(out))
    if rc != 0:
        raise HgmError(rc, lib.hgm_last_error(None).decode())
    return out
