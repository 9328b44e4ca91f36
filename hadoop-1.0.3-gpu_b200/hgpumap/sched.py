"""ctypes binding of libhgsched.so (include/hgsched.h): the fork's hybrid CPU/GPU map
assignment rule (JobQueueTaskScheduler.assignTasks) and the one-TaskTracker simulator."""
from __future__ import annotations

import ctypes as C
import os

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX_DEVICES = 32
FIX_PENDING, FIX_PER_JOB_ACCEL, FIX_PRIOR_ACCEL = 1, 2, 4


class Job(C.Structure):
    _fields_ = [("desired_maps", C.c_int32), ("finished_maps", C.c_int32), ("running_maps", C.c_int32),
                ("cpu_mean_ms", C.c_int32), ("gpu_mean_ms", C.c_int32), ("has_gpu_executable", C.c_int32),
                ("prior_acceleration", C.c_double)]


class Tracker(C.Structure):
    _fields_ = [("max_cpu_slots", C.c_int32), ("max_gpu_slots", C.c_int32), ("running_cpu", C.c_int32),
                ("running_gpu", C.c_int32), ("busy_devices", C.c_uint32)]


class TaskStatus(C.Structure):
    _fields_ = [("is_map", C.c_int32), ("run_state", C.c_int32), ("cleanup_phase", C.c_int32),
                ("run_on_gpu", C.c_int32), ("gpu_device_id", C.c_int32)]


(STATE_RUNNING, STATE_SUCCEEDED, STATE_FAILED, STATE_UNASSIGNED, STATE_KILLED, STATE_COMMIT_PENDING,
 STATE_FAILED_UNCLEAN, STATE_KILLED_UNCLEAN) = range(8)


class Decision(C.Structure):
    _fields_ = [("n_cpu", C.c_int32), ("n_gpu", C.c_int32), ("gpu_device", C.c_int32 * MAX_DEVICES),
                ("cpu_pass_skipped", C.c_int32), ("pending_map_load", C.c_int32), ("acceleration", C.c_double)]


class SimParams(C.Structure):
    _fields_ = [("n_splits", C.c_int32), ("iterations", C.c_int32), ("cpu_slots", C.c_int32),
                ("gpu_slots", C.c_int32), ("cpu_task_ms", C.c_double), ("gpu_task_ms_first", C.c_double),
                ("gpu_task_ms_cached", C.c_double), ("task_overhead_ms", C.c_double), ("heartbeat_ms", C.c_double),
                ("out_of_band_heartbeat", C.c_int32), ("optional_scheduling", C.c_int32), ("fixes", C.c_int32),
                ("prior_acceleration", C.c_double), ("per_device_cache", C.c_int32), ("device_affinity", C.c_int32),
                ("task_jitter", C.c_double)]


class SimResult(C.Structure):
    _fields_ = [("makespan_ms", C.c_double), ("tasks_cpu", C.c_int64), ("tasks_gpu", C.c_int64),
                ("heartbeats", C.c_int64), ("cpu_pass_skips", C.c_int64), ("last_acceleration", C.c_double),
                ("points_per_sec", C.c_double), ("gpu_cache_hits", C.c_int64)]


_LIB = None


def load() -> C.CDLL:
    global _LIB
    if _LIB is None:
        path = os.path.join(_PKG_ROOT, "libhgsched.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} is not built; run `make -C {_PKG_ROOT}`")
        lib = C.CDLL(path)
        lib.hgs_assign.restype = C.c_int
        lib.hgs_assign.argtypes = [C.POINTER(Job), C.c_int32, C.POINTER(Tracker), C.c_int32, C.c_int32,
                                   C.c_int32, C.c_int32, C.POINTER(Decision)]
        lib.hgs_simulate.restype = C.c_int
        lib.hgs_simulate.argtypes = [C.POINTER(SimParams), C.POINTER(SimResult)]
        lib.hgs_tracker_from_reports.restype = C.c_int
        lib.hgs_tracker_from_reports.argtypes = [C.POINTER(TaskStatus), C.c_int32, C.c_int32, C.c_int32,
                                                 C.POINTER(Tracker)]
        lib.hgs_mean_map_time_ms.restype = C.c_int32
        lib.hgs_mean_map_time_ms.argtypes = [C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                                             C.c_int32, C.c_int32]
        lib.hgs_slots_new.restype = C.c_void_p
        lib.hgs_slots_new.argtypes = [C.c_int32, C.c_int32]
        lib.hgs_slots_delete.restype = None
        lib.hgs_slots_delete.argtypes = [C.c_void_p]
        lib.hgs_slots_acquire.restype = C.c_int32
        lib.hgs_slots_acquire.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
        lib.hgs_slots_release.restype = C.c_int32
        lib.hgs_slots_release.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
        lib.hgs_slots_release_task.restype = C.c_int32
        lib.hgs_slots_release_task.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
        lib.hgs_slots_notify.restype = None
        lib.hgs_slots_notify.argtypes = [C.c_void_p]
        lib.hgs_slots_snapshot.restype = C.c_int32
        lib.hgs_slots_snapshot.argtypes = [C.c_void_p, C.POINTER(Tracker)]
        lib.hgs_pick_split.restype = C.c_int32
        lib.hgs_pick_split.argtypes = [C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32]
        _LIB = lib
    return _LIB


def assign(jobs: list[Job], tracker: Tracker, num_trackers: int = 1, optional: bool = True, fixes: int = 0,
           want: int = 1 << 30) -> Decision:
    arr = (Job * max(len(jobs), 1))(*jobs)
    out = Decision()
    rc = load().hgs_assign(arr, len(jobs), C.byref(tracker), num_trackers, int(optional), fixes, want, C.byref(out))
    if rc != 0:
        raise ValueError(f"hgs_assign returned {rc}")
    return out


def simulate(**kw) -> SimResult:
    params = SimParams(**kw)
    out = SimResult()
    rc = load().hgs_simulate(C.byref(params), C.byref(out))
    if rc != 0:
        raise ValueError(f"hgs_simulate returned {rc}")
    return out


class SlotError(Exception):
    """An hgs_slots call that would break an assertion of TaskTracker.java (-34) or got a bad argument (-22)."""


class Slots:
    """The TaskTracker's free-slot counters (TaskLauncher, mapred/TaskTracker.java:2434-2629) + releaseSlot
    (:3397-3408) + the device of each GPU task (hgs_slots_*).  acquire blocks like the launcher thread does."""

    def __init__(self, max_cpu_slots: int, max_gpu_slots: int):
        self._lib = load()
        self._h = self._lib.hgs_slots_new(max_cpu_slots, max_gpu_slots)
        if not self._h:
            raise SlotError("bad slot counts")

    def close(self):
        if self._h:
            self._lib.hgs_slots_delete(self._h)
            self._h = None

    __del__ = close

    def _rc(self, rc: int) -> bool:
        if rc < 0:
            raise SlotError(rc)
        return rc == 0

    def acquire(self, on_gpu: bool, need: int = 1, device: int = -1, timeout_ms: int = -1, cancelled=None) -> bool:
        """True = slots taken; False = timed out or cancelled.  cancelled: a ctypes c_int32 another thread sets."""
        flag = C.byref(cancelled) if cancelled is not None else None
        return self._rc(self._lib.hgs_slots_acquire(self._h, int(on_gpu), need, device, timeout_ms, flag))

    def release(self, on_gpu: bool, n: int = 1, device: int = -1) -> None:
        self._rc(self._lib.hgs_slots_release(self._h, int(on_gpu), n, device))

    def release_task(self, on_gpu: bool, slot_taken: bool, n: int = 1, device: int = -1) -> bool:
        """TaskInProgress.releaseSlot; returns the new slotTaken (always False)."""
        taken = C.c_int32(int(slot_taken))
        self._rc(self._lib.hgs_slots_release_task(self._h, int(on_gpu), n, device, C.byref(taken)))
        return bool(taken.value)

    def notify(self) -> None:
        self._lib.hgs_slots_notify(self._h)

    def snapshot(self) -> Tracker:
        out = Tracker()
        self._rc(self._lib.hgs_slots_snapshot(self._h, C.byref(out)))
        return out


def pick_split(home_device: list[int], device: int, max_gpu_slots: int) -> int:
    """Index of the pending split to hand to GPU slot `device` (hgs_pick_split): one it already holds in
    HBM, else one nobody holds, else one of the device with the longest queue.  -1 if nothing is pending."""
    arr = (C.c_int32 * max(len(home_device), 1))(*home_device)
    rc = load().hgs_pick_split(arr, len(home_device), device, max_gpu_slots)
    if rc < -1:
        raise ValueError(f"hgs_pick_split returned {rc}")
    return rc


def tracker_from_reports(reports: list[TaskStatus], max_cpu_slots: int, max_gpu_slots: int) -> Tracker:
    """The heartbeat's slot accounting (TaskTrackerStatus.countCPUMapTasks / countGPUMapTasks /
    availableGPUDevices) over its task reports."""
    arr = (TaskStatus * max(len(reports), 1))(*reports)
    out = Tracker()
    rc = load().hgs_tracker_from_reports(arr, len(reports), max_cpu_slots, max_gpu_slots, C.byref(out))
    if rc != 0:
        raise ValueError(f"hgs_tracker_from_reports returned {rc}")
    return out


def mean_map_time_ms(start_ms: list[int], finish_ms: list[int], ran_on_gpu: list[bool], want_gpu: bool) -> int:
    """JobInProgress.getCPUMapTaskMeanTime / getGPUMapTaskMeanTime with their Java int arithmetic."""
    n = len(start_ms)
    a = (C.c_int64 * max(n, 1))(*start_ms)
    b = (C.c_int64 * max(n, 1))(*finish_ms)
    g = (C.c_int32 * max(n, 1))(*[int(x) for x in ran_on_gpu])
    return load().hgs_mean_map_time_ms(a, b, g, n, int(want_gpu))
