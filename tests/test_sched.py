"""CPU: the hybrid CPU/GPU assignment rule restated from JobQueueTaskScheduler.assignTasks
(mapred/JobQueueTaskScheduler.java:86-575).  The reference has no test for its GPU branch
(SURVEY.md section 4), so these pin the rule as the source states it, quirks included."""
import pytest

from hgpumap import sched
from hgpumap.sched import Job, Tracker


def job(desired=100, finished=0, running=0, cpu=0, gpu=0, exe=1):
    return Job(desired, finished, running, cpu, gpu, exe, 0.0)


def test_acceleration_factor_is_cpu_over_gpu_and_zero_until_both_are_known():
    tt = Tracker(3, 1, 0, 0, 0)
    assert sched.assign([job(cpu=0, gpu=50)], tt).acceleration == 0.0          # :175-177
    assert sched.assign([job(cpu=4000, gpu=0)], tt).acceleration == 0.0
    assert sched.assign([job(cpu=4000, gpu=50)], tt).acceleration == 80.0


def test_shipped_config_fills_three_cpu_slots_and_one_gpu_slot():
    """conf/mapred-site.xml:23-44: 3 CPU + 1 GPU map slots, optionalscheduling=true."""
    d = sched.assign([job()], Tracker(3, 1, 0, 0, 0))
    assert (d.n_cpu, d.n_gpu, d.gpu_device[0], d.cpu_pass_skipped) == (3, 1, 0, 0)
    d = sched.assign([job(running=3)], Tracker(3, 1, 2, 1, 0b1))
    assert (d.n_cpu, d.n_gpu) == (1, 0)


def test_cpu_pass_is_suppressed_near_the_tail_only_with_optional_scheduling():
    """:290-291: skip CPUs iff optional && max(pending,0) < accel * gpuSlots * trackers."""
    tt = Tracker(3, 1, 0, 0, 0)
    tail = job(desired=100, finished=85, running=0, cpu=4000, gpu=200)   # pending 15 < 20*1*1
    d = sched.assign([tail], tt, optional=True)
    assert d.cpu_pass_skipped == 1 and (d.n_cpu, d.n_gpu) == (0, 1) and d.pending_map_load == 15
    d = sched.assign([tail], tt, optional=False)
    assert d.cpu_pass_skipped == 0 and (d.n_cpu, d.n_gpu) == (3, 1)
    early = job(desired=100, finished=10, cpu=4000, gpu=200)             # pending 90 >= 20
    assert sched.assign([early], tt).cpu_pass_skipped == 0
    # more trackers raise the threshold: 15 < 20*1*1 but 90 < 20*1*8
    assert sched.assign([early], tt, num_trackers=8).cpu_pass_skipped == 1
    # before any GPU task has finished the factor is 0, so CPUs are never suppressed
    assert sched.assign([job(desired=5, finished=4, cpu=4000, gpu=0)], tt).cpu_pass_skipped == 0


def test_gpu_task_gets_the_first_free_device_id():
    """:354-361 with the bitmap of TaskTrackerStatus.availableGPUDevices (:536-551)."""
    d = sched.assign([job()], Tracker(0, 8, 0, 3, 0b00001011))
    assert d.n_gpu == 5 and list(d.gpu_device[:5]) == [2, 4, 5, 6, 7]
    d = sched.assign([job()], Tracker(0, 8, 0, 0, 0), want=2)
    assert d.n_gpu == 2 and list(d.gpu_device[:2]) == [0, 1]


def test_job_without_gpu_executable_gets_no_gpu_slot():
    d = sched.assign([job(exe=0)], Tracker(2, 2, 0, 0, 0))                # :342-347
    assert (d.n_cpu, d.n_gpu) == (2, 0)


def test_pending_load_quirk_and_its_fix():
    """:137-139 adds the RUNNING TOTAL of remaining maps for every job; the factor comes from
    the last job only (:143-144)."""
    jobs = [job(desired=10, finished=0, running=2, cpu=1000, gpu=100), job(desired=20, finished=5, running=1)]
    d = sched.assign(jobs, Tracker(1, 1, 0, 0, 0))
    assert d.pending_map_load == (10 - 2) + (10 + 15 - 1) and d.acceleration == 0.0
    d = sched.assign(jobs, Tracker(1, 1, 0, 0, 0), fixes=sched.FIX_PENDING | sched.FIX_PER_JOB_ACCEL)
    assert d.pending_map_load == (10 - 2) + (15 - 1) and d.acceleration == 10.0


def test_simulation_hybrid_beats_cpu_only_and_tail_rule_helps():
    base = dict(n_splits=120, iterations=1, cpu_task_ms=4000.0, gpu_task_ms_first=200.0, gpu_task_ms_cached=200.0,
                task_overhead_ms=0.0, heartbeat_ms=50.0, out_of_band_heartbeat=1, fixes=0, prior_acceleration=0.0)
    cpu_only = sched.simulate(cpu_slots=3, gpu_slots=0, optional_scheduling=0, **base)
    hybrid = sched.simulate(cpu_slots=3, gpu_slots=1, optional_scheduling=0, **base)
    tail = sched.simulate(cpu_slots=3, gpu_slots=1, optional_scheduling=1, **base)
    assert cpu_only.tasks_gpu == 0 and cpu_only.tasks_cpu == 120
    assert hybrid.makespan_ms < cpu_only.makespan_ms / 4
    assert tail.makespan_ms <= hybrid.makespan_ms and tail.cpu_pass_skips > 0
    assert hybrid.tasks_cpu + hybrid.tasks_gpu == 120 == tail.tasks_cpu + tail.tasks_gpu
    assert 15 < tail.last_acceleration <= 20  # integer-millisecond means like JobInProgress.java:527-565


def test_prior_acceleration_skips_the_cold_start():
    """Without a prior every job starts by filling all CPU slots (factor 0 until tasks of both
    kinds have finished); fed with a measured factor the rule suppresses them from the start."""
    tt = Tracker(3, 1, 0, 0, 0)
    fresh = Job(50, 0, 0, 0, 0, 1, 80.0)
    assert sched.assign([fresh], tt).cpu_pass_skipped == 0
    d = sched.assign([fresh], tt, fixes=sched.FIX_PRIOR_ACCEL)
    assert d.cpu_pass_skipped == 1 and d.acceleration == 80.0 and (d.n_cpu, d.n_gpu) == (0, 1)
    base = dict(n_splits=60, iterations=3, cpu_slots=8, gpu_slots=8, cpu_task_ms=14000.0, gpu_task_ms_first=1.2,
                gpu_task_ms_cached=0.05, task_overhead_ms=1.0, heartbeat_ms=3000.0, out_of_band_heartbeat=1,
                optional_scheduling=1)
    cold = sched.simulate(fixes=0, prior_acceleration=0.0, **base)
    warm = sched.simulate(fixes=sched.FIX_PRIOR_ACCEL, prior_acceleration=14000.0 / 2.2, **base)
    assert cold.tasks_cpu == 24 and cold.makespan_ms > 3 * 14000
    assert warm.tasks_cpu == 0 and warm.makespan_ms < 100


def test_simulation_rejects_bad_arguments():
    with pytest.raises(ValueError):
        sched.simulate(n_splits=1, iterations=0, cpu_slots=1, gpu_slots=1, cpu_task_ms=1, gpu_task_ms_first=1,
                       gpu_task_ms_cached=1, task_overhead_ms=0, heartbeat_ms=1, out_of_band_heartbeat=0,
                       optional_scheduling=0, fixes=0, prior_acceleration=0.0)


def test_hgsched_header_symbols_are_exported():
    import os
    import re

    from conftest import ROOT

    text = open(os.path.join(ROOT, "include", "hgsched.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = sorted(set(re.findall(r"\b(hgs_[a-z_]+)\s*\(", text)))
    assert names == ["hgs_assign", "hgs_mean_map_time_ms", "hgs_pick_split", "hgs_simulate", "hgs_slots_acquire",
                     "hgs_slots_delete", "hgs_slots_new", "hgs_slots_notify", "hgs_slots_release",
                     "hgs_slots_release_task", "hgs_slots_snapshot", "hgs_tracker_from_reports"]
    lib = sched.load()
    for name in names:
        assert hasattr(lib, name)


def test_pick_split_prefers_own_then_unheld_then_longest_queue():
    """Device-level locality below the reference's node locality (JobQueueTaskScheduler.java:350-361 stamps
    the first free device on whatever split it got)."""
    # (1) a split the device holds, wherever it sits in the pending list
    assert sched.pick_split([3, -1, 5, 2, 5], device=5, max_gpu_slots=8) == 2
    # (2) else a split nobody holds: it has to be staged somewhere anyway
    assert sched.pick_split([3, -1, 2, -1], device=5, max_gpu_slots=8) == 1
    # a home outside the tracker's slots (a worker that is gone) counts as held nowhere
    assert sched.pick_split([3, 9, 2], device=5, max_gpu_slots=8) == 1
    # (3) else steal from the device with the most pending resident splits; lowest id on a tie
    assert sched.pick_split([3, 2, 2, 3, 2], device=5, max_gpu_slots=8) == 1
    assert sched.pick_split([3, 2, 2, 3], device=5, max_gpu_slots=8) == 1
    assert sched.pick_split([3, 3, 2], device=5, max_gpu_slots=8) == 0
    # nothing pending, bad arguments
    assert sched.pick_split([], device=0, max_gpu_slots=1) == -1
    for dev, slots in ((8, 8), (-1, 8), (0, 0), (0, 33)):
        with pytest.raises(ValueError):
            sched.pick_split([0], device=dev, max_gpu_slots=slots)


def test_device_affine_picks_keep_iteration_two_in_the_hbm_caches():
    """128 splits on eight GPU slots, two iterations, task times jittered the way real tasks finish (a staged
    split takes three times a cached one): handing a freed slot the NEXT split (the reference) finds it
    cached on that device about one time in eight; hgs_pick_split finds nearly every split at home
    (measured on eight B200s: 123 of 128, profiles/r01_config3_job.md)."""
    import heapq
    import random

    n_splits, slots = 128, 8

    def run_iteration(home, affine, seed):
        rng = random.Random(seed)
        pending = list(range(n_splits))
        hits = 0
        now_home = dict(home)
        free_at = [(0.0, dev) for dev in range(slots)]
        heapq.heapify(free_at)
        while pending:
            t, dev = heapq.heappop(free_at)          # the slot that frees up first
            if affine:
                i = sched.pick_split([now_home.get(s, -1) for s in pending], dev, slots)
            else:
                i = 0
            s = pending.pop(i)
            hit = home.get(s, -1) == dev
            hits += hit
            now_home[s] = dev                        # the worker of `dev` holds it from now on
            heapq.heappush(free_at, (t + (1.0 if hit else 3.0) * rng.uniform(0.7, 1.3), dev))
        return hits, now_home

    _, home = run_iteration({}, affine=False, seed=1)        # iteration 1 stages everything
    hits_ref, _ = run_iteration(home, affine=False, seed=2)
    hits_aff, _ = run_iteration(home, affine=True, seed=2)
    assert hits_aff >= 0.9 * n_splits                         # misses only where a device ran out of its own early
    assert hits_ref <= 0.3 * n_splits


def test_simulation_per_device_cache_and_affinity():
    """One HBM cache per slot's worker: as the reference orders tasks (next split to the first free device)
    iteration 2 finds few splits at home and stages the rest again; with hgs_pick_split nearly all are hits.
    per_device_cache=0 keeps the older one-cache-per-node behaviour."""
    base = dict(n_splits=128, iterations=3, cpu_slots=0, gpu_slots=8, cpu_task_ms=11000.0, gpu_task_ms_first=6.8,
                gpu_task_ms_cached=2.2, task_overhead_ms=0.0, heartbeat_ms=3000.0, out_of_band_heartbeat=1,
                optional_scheduling=1, fixes=0, prior_acceleration=0.0, task_jitter=0.3)
    node = sched.simulate(**base)
    ref = sched.simulate(per_device_cache=1, device_affinity=0, **base)
    aff = sched.simulate(per_device_cache=1, device_affinity=1, **base)
    assert node.gpu_cache_hits == 0 and node.tasks_gpu == ref.tasks_gpu == aff.tasks_gpu == 3 * 128
    assert aff.gpu_cache_hits >= 0.9 * 2 * 128
    assert ref.gpu_cache_hits < aff.gpu_cache_hits
    assert aff.makespan_ms < ref.makespan_ms
    assert abs(aff.makespan_ms - node.makespan_ms) < 0.1 * node.makespan_ms
    # CPU slots take splits too and leave them where they were cached
    mixed = sched.simulate(per_device_cache=1, device_affinity=1, **{**base, "cpu_slots": 2, "cpu_task_ms": 50.0})
    assert mixed.tasks_cpu > 0 and mixed.tasks_cpu + mixed.tasks_gpu == 3 * 128


def test_tracker_accounting_from_task_reports():
    """TaskTrackerStatus.java:504-551: a map task holds its slot (a GPU task its device) while RUNNING,
    UNASSIGNED or cleaning up after an *_UNCLEAN end; reduces and finished tasks do not count."""
    TS = sched.TaskStatus
    reports = [
        TS(1, sched.STATE_RUNNING, 0, 0, 0),          # CPU map, running
        TS(1, sched.STATE_UNASSIGNED, 0, 1, 2),       # GPU map on device 2, launching
        TS(1, sched.STATE_SUCCEEDED, 0, 1, 0),        # finished GPU map: device 0 is free again
        TS(1, sched.STATE_FAILED_UNCLEAN, 1, 1, 5),   # GPU map in task cleanup: still holds device 5
        TS(1, sched.STATE_KILLED_UNCLEAN, 0, 1, 6),   # *_UNCLEAN but not in the cleanup phase: released
        TS(1, sched.STATE_COMMIT_PENDING, 0, 0, 0),   # not counted by the fork's getters (:524-528)
        TS(0, sched.STATE_RUNNING, 0, 0, 0),          # a reduce
    ]
    tt = sched.tracker_from_reports(reports, 3, 8)
    assert (tt.max_cpu_slots, tt.max_gpu_slots, tt.running_cpu, tt.running_gpu) == (3, 8, 1, 2)
    assert tt.busy_devices == (1 << 2) | (1 << 5)
    # the scheduler then hands out the free devices in id order
    d = sched.assign([job()], tt)
    assert d.n_cpu == 2 and list(d.gpu_device[:d.n_gpu]) == [0, 1, 3, 4, 6, 7]
    # a device id outside the slot range indexes past boolean[maxGPUMapSlots] in the Java (:546)
    with pytest.raises(ValueError):
        sched.tracker_from_reports([TS(1, sched.STATE_RUNNING, 0, 1, 8)], 3, 8)
    assert sched.tracker_from_reports([], 2, 0).busy_devices == 0


def test_mean_map_times_use_java_int_arithmetic():
    """JobInProgress.java:527-565: finished tasks of the wanted kind only, integer milliseconds, int sum."""
    t0 = 1_700_000_000_000
    start = [t0, t0 + 10, t0 + 20, t0 + 30, t0 + 40]
    finish = [t0 + 4000, t0 + 10 + 151, 0, t0 + 30 + 100, t0 + 40 + 5001]   # the third has not finished
    gpu = [False, True, True, True, False]
    assert sched.mean_map_time_ms(start, finish, gpu, want_gpu=False) == (4000 + 5001) // 2
    assert sched.mean_map_time_ms(start, finish, gpu, want_gpu=True) == (151 + 100) // 2      # 125, not 125.5
    assert sched.mean_map_time_ms([], [], [], want_gpu=True) == 0
    assert sched.mean_map_time_ms([t0], [0], [True], want_gpu=True) == 0
    # a GPU task shorter than a millisecond counts as 0 ms: the acceleration factor stays 0 (:175-177)
    assert sched.mean_map_time_ms([t0, t0], [t0, t0], [True, True], want_gpu=True) == 0
    # the int sum wraps: three tasks of 2^30 ms sum to -2^30 in a Java int
    big = 1 << 30
    assert sched.mean_map_time_ms([0, 0, 0], [big, big, big], [False] * 3, want_gpu=False) == -(1 << 30) // 3 + (1 if (-(1 << 30)) % 3 else 0)


# ---- the slot accounting unit (hgs_slots_*) -------------------------------------------------------------------

def test_slots_replay_one_heartbeat_of_launches_and_completions():
    """A TaskTracker with 2 CPU + 3 GPU map slots: the launcher takes slots in the order the heartbeat's actions
    arrive (TaskTracker.java:2526-2600), finishing tasks give them back (:2467-2485 via releaseSlot :3397-3408), and
    what the next heartbeat reports is what TaskTrackerStatus would count."""
    s = sched.Slots(2, 3)
    d = sched.assign([job(100, finished=0, running=0)], s.snapshot(), optional=False)
    assert (d.n_cpu, d.n_gpu) == (2, 3) and list(d.gpu_device[:3]) == [0, 1, 2]
    for _ in range(d.n_cpu):
        assert s.acquire(False, timeout_ms=0)
    for dev in d.gpu_device[:d.n_gpu]:
        assert s.acquire(True, device=dev, timeout_ms=0)
    t = s.snapshot()
    assert (t.running_cpu, t.running_gpu, t.busy_devices) == (2, 3, 0b111)
    assert sched.assign([job(100, running=5)], t, optional=False).n_gpu == 0  # nothing free, nothing handed out
    assert not s.acquire(True, device=1, timeout_ms=0) and not s.acquire(False, timeout_ms=0)
    assert s.release_task(True, slot_taken=True, device=1) is False          # GPU task on device 1 finishes
    t = s.snapshot()
    assert (t.running_gpu, t.busy_devices) == (2, 0b101)
    d = sched.assign([job(100, finished=1, running=4)], t, optional=False)
    assert d.n_gpu == 1 and d.gpu_device[0] == 1                              # the first free device id (:354-361)
    assert s.acquire(True, device=1, timeout_ms=0)
    # a task killed before the launcher got to it never took a slot: releaseSlot only wakes the launcher
    before = s.snapshot()
    assert s.release_task(False, slot_taken=False) is False
    after = s.snapshot()
    assert (before.running_cpu, before.running_gpu) == (after.running_cpu, after.running_gpu)
    s.close()


def test_slots_refuse_what_the_java_asserts_against():
    s = sched.Slots(1, 2)
    with pytest.raises(sched.SlotError):      # addFreeCPUSlots: assert (numCPUFreeSlots <= maxCPUSlots) :2472
        s.release(False)
    with pytest.raises(sched.SlotError):      # more slots than the tracker has: the launcher would wait for ever
        s.acquire(True, need=3, timeout_ms=0)
    assert s.acquire(True, device=0, timeout_ms=0)
    with pytest.raises(sched.SlotError):      # a second task stamped with a busy device (slot i <-> device i)
        s.acquire(True, device=0, timeout_ms=0)
    with pytest.raises(sched.SlotError):      # giving back a device nobody holds
        s.release(True, device=1)
    with pytest.raises(sched.SlotError):
        s.acquire(True, device=5, timeout_ms=0)
    s.release(True, device=0)
    assert s.snapshot().busy_devices == 0
    s.close()


def test_slots_launcher_blocks_until_a_task_finishes_or_is_killed():
    """The launcher thread waits on the counter (:2529-2553): a completion wakes it; a kill of the queued task
    (canBeLaunched() false) makes it give up without taking a slot."""
    import ctypes
    import threading
    import time

    s = sched.Slots(1, 1)
    assert s.acquire(True, device=0)
    got = []
    th = threading.Thread(target=lambda: got.append(s.acquire(True, device=0, timeout_ms=5000)))
    th.start()
    time.sleep(0.05)
    assert th.is_alive() and not got          # blocked: no free GPU slot
    s.release(True, device=0)
    th.join(timeout=5)
    assert got == [True] and s.snapshot().busy_devices == 1
    killed = ctypes.c_int32(0)
    th = threading.Thread(target=lambda: got.append(s.acquire(True, device=0, timeout_ms=5000, cancelled=killed)))
    th.start()
    time.sleep(0.05)
    killed.value = 1
    s.notify()                                # releaseSlot of a task that never took one: notifySlots
    th.join(timeout=5)
    assert got == [True, False] and s.snapshot().running_gpu == 1
    assert not s.acquire(False, need=1, timeout_ms=0) is False  # the CPU counter is untouched
    s.close()

