#!/usr/bin/env python
"""BASELINE config 3 as a JOB, JVM-free: K-means iterations over file splits, every map task a real Pipes
child driven through the protocol (command-file replay), placed on CPU and GPU map slots by the fork's
own hybrid rule (JobQueueTaskScheduler.assignTasks restated in libhgsched.so, `hgs_assign`), one call
per "heartbeat" = per task completion.  What the Java side would do and cannot do here (no JVM):

  every slot   one long-lived process (the per-slot child JVM of a TaskTracker with JVM reuse) that builds each task's
               downlink, starts its Pipes child and hands the partial records back
  GPU slot i   `kmeans-pipes-gpu <i>` child per task, C++ record reader route (FileSplit), talking to the
               slot's persistent worker (`hgm-worker`), file splits kept in its HBM cache
  CPU slot     the CPU K-means Pipes program (oracle/_ref/kmeans-pipes-cpu-hgmrt: the reference's
               example-style mapper+combiner on this repository's runtime, so that a task does not sit
               out the reference runtime's 5 s ping join), `kmeans.input.format=path`
  reduce       the partial records of all tasks are summed here and the centroids updated

It lives under tests/measure/ because the CPU slots run a program of oracle/.  The scheduler's inputs
(mean CPU / GPU task times) are the measured wall times of the tasks of this very job, as
JobInProgress keeps them (mapred/JobInProgress.java:527-565).

Task batching (--batch B): a GPU map task carries B splits instead of one -- `kmeans.input.format=path` records
`file \t offset \t length`, the shipped NLineInputFormat route (conf/mapred-site.xml:13-22) with
mapred.line.input.format.linespermap = B -- so the 2-3.5 ms it costs to start a Pipes child
(pipes/Application.java:156-194) is paid once per B splits; it is what JVM reuse
(mapred.job.reuse.jvm.num.tasks) does for the Java half of a task.  The slot bookkeeping is libhgsched's
hgs_slots_* (TaskLauncher + releaseSlot restated), the same unit hgs_simulate uses.

    python tests/measure/config3_job.py [--splits 24] [--split-mib 16] [--iterations 3] [--batch 1]
                                        [--cpu-slots 4] [--gpu-slots 1] [--optional 1] [--fixes 0]
                                        [--data-dir DIR] [--gen host|gpu]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = os.path.join(ROOT, "hadoop-1.0.3-gpu_b200")
sys.path[:0] = [PKG, os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402

from hgpumap import capi, sched, synth, wire  # noqa: E402

GPU_BIN = os.path.join(PKG, "kmeans-pipes-gpu")
CPU_BIN = os.path.join(ROOT, "oracle", "_ref", "kmeans-pipes-cpu-hgmrt")
WORKER = os.path.join(PKG, "hgm-worker")


class Task:
    """One map task = one Pipes child on a command file (hadoop.pipes.command.file).  Started without a thread or a
    pipe of its own -- the TaskTracker's launcher is a fork + exec, and so is this -- and reaped by the job loop."""

    def __init__(self, binary, downlink, args, tmp):
        fd, self.cmd = tempfile.mkstemp(prefix="task_", dir=tmp)
        os.close(fd)
        downlink.write(self.cmd)
        env = {"hadoop.pipes.command.file": self.cmd, "PATH": os.environ.get("PATH", "")}
        for key in ("LD_LIBRARY_PATH", "CUDA_VISIBLE_DEVICES"):
            if key in os.environ:
                env[key] = os.environ[key]
        self.err = open(self.cmd + ".err", "wb")
        self.t0 = time.perf_counter()
        self.proc = subprocess.Popen([binary, *args], env=env, stdout=subprocess.DEVNULL, stderr=self.err)

    def finish(self):
        """After the child has exited: (uplink, wall seconds)."""
        wall = time.perf_counter() - self.t0
        self.err.close()
        out = self.cmd + ".out"
        data = open(out, "rb").read() if os.path.exists(out) else b""
        up = wire.parse_uplink(data)
        if not up.done:
            raise RuntimeError(f"map task failed: {open(self.cmd + '.err', 'rb').read().decode()[-400:]}")
        for path in (self.cmd, out, self.cmd + ".err"):
            if os.path.exists(path):
                os.remove(path)
        return up, wall


def gpu_downlink(paths, cpath, k, d, sock, epoch):
    """One GPU map task over the splits `paths` (whole files): a FileSplit through the C++ reader for one split,
    path records (piped Text values, NLineInputFormat) for a batch."""
    conf = {"kmeans.k": str(k), "kmeans.d": str(d), "kmeans.centroids": cpath, "kmeans.gpu.worker.socket": sock,
            "kmeans.gpu.worker.cache": "true", "kmeans.gpu.worker.cache.epoch": epoch}
    if len(paths) == 1:
        conf.update({"kmeans.input.format": "binary", "hadoop.pipes.java.recordreader": "false"})
        return (wire.Downlink().authenticate().start().set_job_conf(conf).set_input_types("k", "v")
                .run_map(wire.file_split("file:" + paths[0], 0, os.path.getsize(paths[0])), 1, False))
    conf.update({"kmeans.input.format": "path", "kmeans.file.format": "binary", "hadoop.pipes.java.recordreader": "true"})
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf)
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text").run_map(b"split", 1, True))
    for i, path in enumerate(paths):
        dl.map_item(i.to_bytes(8, "big"), f"{path}\t0\t{os.path.getsize(path)}")
    return dl.close()


def cpu_downlink(path, cpath, k, d):
    conf = {"kmeans.k": str(k), "kmeans.d": str(d), "kmeans.centroids": cpath, "kmeans.input.format": "path",
            "io.sort.mb": "100"}
    dl = (wire.Downlink().authenticate().start().set_job_conf(conf)
          .set_input_types("org.apache.hadoop.io.LongWritable", "org.apache.hadoop.io.Text").run_map(b"split", 1, True))
    dl.map_item((0).to_bytes(8, "big"), path)
    return dl.close()


def slot_process(conn, tmp, k, d):
    """The long-lived process of one map slot -- the child JVM a TaskTracker keeps per slot when JVM reuse is on
    (mapred.job.reuse.jvm.num.tasks = -1).  For every task it is handed it does what PipesGPUMapRunner / PipesMapRunner
    does: build the downlink, start the Pipes child, collect its uplink; the partial records go back to the job as
    arrays.  The slots work side by side, as the JVMs of a TaskTracker do."""
    while True:
        job = conn.recv()
        if job is None:
            return
        kind, dev, paths, cpath, sock, epoch = job
        try:
            if kind == "gpu":
                task = Task(GPU_BIN, gpu_downlink(paths, cpath, k, d, sock, epoch), (str(dev),), tmp)
            else:
                task = Task(CPU_BIN, cpu_downlink(paths[0], cpath, k, d), (), tmp)
            task.proc.wait()
            up, wall = task.finish()
            counts, sums = wire.kmeans_partials([(key, v) for _, key, v in up.partitioned], k, d)
            conn.send((counts, sums, wall, up.counter("KMEANS", "GPU_CACHE_HITS") if kind == "gpu" else 0, None))
        except Exception as exc:  # noqa: BLE001 -- reported to the job, which fails
            conn.send((None, None, 0.0, 0, repr(exc)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--splits", type=int, default=24)
    ap.add_argument("--split-mib", type=int, default=16)
    ap.add_argument("--iterations", type=int, default=3)
    ap.add_argument("--cpu-slots", type=int, default=4)
    ap.add_argument("--gpu-slots", type=int, default=1)
    ap.add_argument("--optional", type=int, default=1, help="mapred.jobtracker.map.optionalscheduling")
    ap.add_argument("--fixes", type=int, default=0, help="HGS_FIX_* flags (4 = carry the acceleration factor over)")
    ap.add_argument("--affinity", type=int, default=0,
                    help="1: hand a GPU slot a split its worker already holds in HBM when there is one (the device-level "
                         "analogue of obtainNewNodeLocalMapTask; the reference picks splits without regard to devices)")
    ap.add_argument("--k", type=int, default=128)
    ap.add_argument("--out", default="")
    ap.add_argument("--batch", type=int, default=1, help="splits per GPU map task (path records; 1 = one FileSplit per task)")
    ap.add_argument("--data-dir", default="", help="where the split files go (default gpurun_out/job_tmp; use /dev/shm for 20 GB)")
    ap.add_argument("--devices", type=int, default=0, help="physical GPUs the slots are spread over (slot g -> device g %% N); "
                    "0 = one device per slot.  For trying the 8-slot job on a smaller box")
    ap.add_argument("--gen", choices=["host", "gpu"], default="host", help="generate the splits on the host or on GPU 0")
    args = ap.parse_args()
    k, d = args.k, 2
    tmp = args.data_dir or os.path.join(ROOT, "gpurun_out", "job_tmp")
    os.makedirs(tmp, exist_ok=True)
    # one long-lived process per map slot, forked before this process touches CUDA
    import multiprocessing as mp
    from multiprocessing.connection import wait as wait_ready

    slot_conns = {}
    slot_procs = []
    for name in [("gpu", g) for g in range(args.gpu_slots)] + [("cpu", i) for i in range(args.cpu_slots)]:
        here, there = mp.Pipe()
        proc = mp.Process(target=slot_process, args=(there, tmp, k, d), daemon=True)
        proc.start()
        slot_conns[name] = here
        slot_procs.append(proc)
    free_cpu = list(range(args.cpu_slots))
    centres = synth.centres(synth.SEED_POINTS, k, d)
    cent = synth.perturb(synth.SEED_CENTROIDS, centres)
    pts_per_split = args.split_mib * (1 << 20) // (4 * d)
    files = []
    t_gen = time.perf_counter()
    if args.gen == "gpu":  # the generator is bit-identical on host and device (csrc/synth.cu); the device is just faster
        with capi.GpuMapSlot(0) as gen:
            d_buf = gen.device_alloc(pts_per_split * d * 4)
            host = np.empty((pts_per_split, d), np.float32)
            for s in range(args.splits):
                gen.synth_points_device(synth.SEED_POINTS, s * pts_per_split, pts_per_split, centres, d_buf)
                gen.memcpy_d2h(host, d_buf)
                path = os.path.join(tmp, f"part-{s:05d}")
                host.tofile(path)
                files.append(path)
            gen.device_free(d_buf)
    else:
        for s in range(args.splits):
            path = os.path.join(tmp, f"part-{s:05d}")
            capi.synth_points_host(synth.SEED_POINTS, s * pts_per_split, pts_per_split, centres).tofile(path)
            files.append(path)
    n_points = args.splits * pts_per_split
    print(json.dumps({"generated_gb": n_points * d * 4 / 1e9, "seconds": time.perf_counter() - t_gen}), flush=True)

    workers, socks = [], []
    for g in range(args.gpu_slots):
        sock = os.path.join(tmp, f"gpu{g}.sock")
        if os.path.exists(sock):
            os.remove(sock)  # a stale socket of an earlier run would end the wait below too early
        device = g % args.devices if args.devices else g
        workers.append(subprocess.Popen([WORKER, "--device", str(device), "--socket", sock], stderr=subprocess.DEVNULL))
        socks.append(sock)
    for sock in socks:
        for _ in range(1200):
            if os.path.exists(sock):
                break
            time.sleep(0.05)

    report = {"workload": f"{args.splits} splits x {args.split_mib} MiB = {n_points:,} 2-D points, k={k}, "
                          f"{args.iterations} iterations; {args.gpu_slots} GPU + {args.cpu_slots} CPU map slots, "
                          f"optional={args.optional}, fixes={args.fixes}, affinity={args.affinity}, "
                          f"{args.batch} split(s) per GPU map task", "iterations": []}
    epoch = f"config3-{os.getpid()}"  # the workers' caches hold this data set only
    prior = 0.0
    cached_on = {}  # split -> device whose worker holds it

    def pick(dev, pending):
        if args.affinity:  # device-level locality: include/hgsched.h hgs_pick_split
            return pending.pop(sched.pick_split([cached_on.get(s, -1) for s in pending], dev, args.gpu_slots))
        return pending.pop(0)

    def pick_batch(dev, pending, want):
        """`want` splits for GPU slot `dev`.  Rule (1) of hgs_pick_split -- a split this device holds, lowest index first --
        applied in one pass while it lasts (the rule is applied per split by the library; asking it 38 times over 298
        pending splits from Python cost more than the tasks it was scheduling); the library decides the rest."""
        batch = []
        if args.affinity:
            mine = [s_ for s_ in pending if cached_on.get(s_, -1) == dev][:want]
            if mine:
                taken = set(mine)
                pending[:] = [s_ for s_ in pending if s_ not in taken]
                batch = mine
        while len(batch) < want and pending:
            batch.append(pick(dev, pending))
        return batch

    slots = sched.Slots(args.cpu_slots, args.gpu_slots)
    t_job = time.perf_counter()
    try:
        for it in range(args.iterations):
            cpath = os.path.join(tmp, f"centroids-{it}.bin")
            cent.astype("<f4").tofile(cpath)
            pending = list(range(args.splits))
            running = {}  # slot connection -> (kind, device or cpu slot, splits of the task)
            cpu_times, gpu_times = [], []
            sums = np.zeros((k, d))
            counts = np.zeros(k, np.int64)
            hits = skips = beats = 0
            finished_tasks = 0
            t_it = time.perf_counter()
            while pending or running:
                # one heartbeat: the tracker reports its slots, the rule answers with tasks to start.  A task is a
                # batch of splits; the rule counts tasks, as the JobTracker does
                tasks_left = (len(pending) + args.batch - 1) // args.batch
                job = sched.Job(finished_tasks + len(running) + tasks_left, finished_tasks, len(running),
                                int(np.mean(cpu_times)) if cpu_times else 0, int(max(np.mean(gpu_times), 1)) if gpu_times else 0,
                                1, prior)
                dec = sched.assign([job], slots.snapshot(), 1, bool(args.optional), args.fixes, want=tasks_left)
                beats += 1
                skips += dec.cpu_pass_skipped
                for i in range(dec.n_gpu):
                    if not pending:
                        break
                    dev = dec.gpu_device[i]
                    assert slots.acquire(True, device=dev, timeout_ms=0)  # the launcher: the rule only hands out free slots
                    batch = pick_batch(dev, pending, min(args.batch, len(pending)))
                    for s_ in batch:
                        cached_on[s_] = dev
                    conn = slot_conns[("gpu", dev)]
                    conn.send(("gpu", dev, [files[s_] for s_ in batch], cpath, socks[dev], epoch))
                    running[conn] = ("gpu", dev, batch)
                for _ in range(dec.n_cpu):
                    if not pending:
                        break
                    assert slots.acquire(False, timeout_ms=0)
                    s_ = pending.pop(0)
                    cpu_slot = free_cpu.pop()
                    conn = slot_conns[("cpu", cpu_slot)]
                    conn.send(("cpu", -1, [files[s_]], cpath, "", epoch))
                    running[conn] = ("cpu", cpu_slot, [s_])
                if not running:
                    raise RuntimeError("the scheduler assigned nothing although slots are free and maps are pending")
                for conn in wait_ready(list(running)):
                    kind, dev, batch = running.pop(conn)
                    c, sm, wall, task_hits, error = conn.recv()
                    if error:
                        raise RuntimeError(error)
                    slots.release_task(kind == "gpu", slot_taken=True, device=dev if kind == "gpu" else -1)  # releaseSlot
                    if kind == "cpu":
                        free_cpu.append(dev)
                    finished_tasks += 1
                    (cpu_times if kind == "cpu" else gpu_times).append(wall * 1e3)
                    counts += c
                    sums += sm
                    hits += task_hits
            assert counts.sum() == n_points, (counts.sum(), n_points)
            dt = time.perf_counter() - t_it
            if cpu_times and gpu_times:
                prior = float(np.mean(cpu_times) / max(np.mean(gpu_times), 1e-3))
            nz = counts > 0
            cent = cent.copy()
            cent[nz] = (sums[nz] / counts[nz, None]).astype(np.float32)
            rec = {"iteration": it + 1, "wall_s": dt, "points_per_s": n_points / dt, "cpu_tasks": len(cpu_times),
                   "gpu_tasks": len(gpu_times), "cpu_task_mean_ms": float(np.mean(cpu_times)) if cpu_times else None,
                   "gpu_task_mean_ms": float(np.mean(gpu_times)) if gpu_times else None, "gpu_cache_hits": hits,
                   "heartbeats": beats, "cpu_pass_skips": skips}
            report["iterations"].append(rec)
            print(json.dumps(rec))
    finally:
        for conn in slot_conns.values():
            try:
                conn.send(None)
            except (BrokenPipeError, OSError):
                pass
        for proc in slot_procs:
            proc.join(timeout=5)
            if proc.is_alive():
                proc.terminate()
        for w in workers:
            w.terminate()
        for path in files:
            os.remove(path)
    report["job_wall_s"] = time.perf_counter() - t_job
    print(json.dumps({"job_wall_s": report["job_wall_s"]}))
    if args.out:
        with open(args.out, "w") as f:
            f.write(json.dumps(report, indent=1) + "\n")


if __name__ == "__main__":
    main()
