#!/usr/bin/env python
"""bench.py -- K-means map throughput (points/s) on B200, the metric BASELINE.json names.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this framework
  python bench.py --impl reference [--gpus N] [--steps K] ...    # the reference's CPU path
  torchrun --nproc-per-node N ... bench.py --gpus N ...          # one rank per GPU

A "step" is one pass of the hot path (nearest-centroid assignment + combiner partial sums,
one K-means iteration's map phase) over one split-set of synthetic 2-D points:
BASELINE.json configs[1] = 200,000,000 points, d=2, k=128 per GPU ("weak" scaling: every
GPU slot maps its own 200 M-point shard; the k*d sums + k counts are then sum-allreduced
over NCCL).

  value     whole-job points/s with the points resident in HBM (CUDA events on the
            library's own stream; K steps between barriers; max over ranks)
  e2e       the same through the reference-facing C-ABI call hgm_map_split with PINNED HOST
            buffers: host->device copies of the whole split and the device->host read of
            the partials are inside the timed region
  roofline  the map kernel alone: algorithmic bytes (4*d per point)
            over its average CUDA-event duration, against MEASURED_PEAKS.json's HBM copy rate
  cpu_baseline  the reference-runtime CPU Pipes program (oracle/_ref) on a bounded sample
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "hadoop-1.0.3-gpu_b200")
sys.path[:0] = [PKG, os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402

METRIC = "kmeans_map_points_per_sec"
UNIT = "points/s"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--points", type=int, default=200_000_000, help="points per GPU (configs[1])")
    ap.add_argument("--k", "--clusters", dest="k", type=int, default=128)
    ap.add_argument("--d", "--dims", dest="d", type=int, default=2,
                    help="(under torchrun spell it --dims: its own parser takes a bare --d as an abbreviation)")
    ap.add_argument("--cpu-points", type=int, default=1 << 19, help="points per CPU process in the baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--kernel", type=int, default=0, help="force a kernel family (hgpumap.h HGM_KERNEL_*)")
    ap.add_argument("--tensor-form", choices=["auto", "single", "pair", "pair_fused", "pair_segmented"], default="auto")
    ap.add_argument("--no-extra", action="store_true", help="skip the `extra` lines (other inputs, configs 4 and 5)")
    return ap.parse_args()


def ncu_traffic(kernel_family, d, k, n):
    """(DRAM bytes per launch, what the capture says limits the kernel) for the dominant kernel, from the
    committed ncu --set full capture of exactly this workload (profiles/ncu_traffic.json), else (None, None)."""
    fam = kernel_family
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        entry = table[f"{fam}:d={d}:k={k}:points={n}"]
        return entry["bytes"], entry.get("limiter")
    except (OSError, KeyError, ValueError):
        return None, None


def kernel_name(capi, slot, family):
    """The dominant kernel of a family as it is named in the library (and in the ncu launch lists)."""
    if family in (capi.KERNEL_GRID2D, capi.KERNEL_GRID2D_SORT):
        return "grid2d_fx_kernel" if slot.grid2d_form()[0] == capi.GRID2D_FORM_FIXED_POINT else "grid2d_map_kernel"
    return {capi.KERNEL_EXACT: "exact_map_kernel", capi.KERNEL_TENSOR: "tensor_assign_kernel"}.get(family, "?")


def workload_name(args):
    return (f"K-means 1 iteration, {args.points:,} 2-D points per GPU, k={args.k}, single B200 GPU map slot"
            if args.d == 2 else f"K-means 1 iteration, {args.points:,} points per GPU, d={args.d}, k={args.k}")


# ---------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's own Pipes runtime, JVM-free
# ---------------------------------------------------------------------------------------

def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class CpuReference:
    """P concurrent Pipes children = P CPU map slots (mapred.tasktracker.map.cpu.tasks.maximum),
    each replaying a command file of `points` MAP_ITEM records through the reference's
    HadoopPipes.cc (or, if oracle/_ref is absent, the C port called in-process)."""

    def __init__(self, args, procs, points):
        from hgpumap import synth

        self.k, self.d, self.procs, self.points = args.k, args.d, procs, points
        self.bin = os.path.join(ROOT, "oracle", "_ref", "kmeans-pipes-cpu")
        self.kind = "reference" if os.path.exists(self.bin) else "port"
        centres = synth.centres(synth.SEED_POINTS, self.k, self.d)
        self.cent = synth.perturb(synth.SEED_CENTROIDS, centres)
        self.tmp = tempfile.mkdtemp(prefix="hgm_cpu_")
        self.splits = [synth.points(synth.SEED_POINTS, i * points, points, centres) for i in range(procs)]
        if self.kind == "reference":
            import pipes_harness as ph

            cpath = os.path.join(self.tmp, "centroids.bin")
            self.cent.astype("<f4").tofile(cpath)
            self.cmds = []
            for i, pts in enumerate(self.splits):
                path = os.path.join(self.tmp, f"cmd{i}")
                ph.kmeans_downlink(pts, cpath, self.k, self.d, "binary", 1, 1).write(path)
                self.cmds.append(path)
        else:
            import ctypes as C

            so = os.path.join(ROOT, "oracle", "libkmeans_oracle.so")
            if not os.path.exists(so):
                subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)
            self.lib = C.CDLL(so)
            self.procs = 1

    def sample(self):
        what = (f"{self.procs} concurrent Pipes children x {self.points:,} one-point MAP_ITEM records through the "
                "reference runtime (HadoopPipes.cc record pump + CombineRunner), map time excludes the runtime's "
                "5 s ping-thread join") if self.kind == "reference" else \
               f"C port of the mapper+combiner, 1 thread, {self.points:,} points, no Pipes protocol"
        return what

    def step(self):
        """Run one bounded sample; returns (points, seconds of map work = max over children)."""
        if self.kind == "port":
            import ctypes as C

            pts = self.splits[0]
            sums = np.zeros((self.k, self.d))
            counts = np.zeros(self.k, np.int64)
            p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
            t0 = time.perf_counter()
            self.lib.hgo_map_combine(p(pts), C.c_int64(len(pts)), self.d, p(self.cent), self.k, p(sums), p(counts), None, None)
            return len(pts), time.perf_counter() - t0
        children = []
        for i, cmd in enumerate(self.cmds):
            stats = os.path.join(self.tmp, f"stats{i}.json")
            env = {"hadoop.pipes.command.file": cmd, "HGM_ORACLE_STATS": stats, "PATH": os.environ.get("PATH", "")}
            children.append((subprocess.Popen([self.bin], env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE), stats))
        total, worst = 0, 0.0
        for proc, stats in children:
            _, err = proc.communicate(timeout=600)
            if proc.returncode != 1:  # runTask's `true` is the exit status (wordcount-simple.cc:63-66)
                raise RuntimeError(f"reference child failed: {err.decode()[-300:]}")
            rec = json.load(open(stats))
            total += rec["points"]
            worst = max(worst, rec["map_seconds"])
        return total, worst


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    ref = CpuReference(args, procs=cores, points=1 << 18)
    for _ in range(max(args.warmup, 3)):  # as many warm-up steps as the GPU arm takes (each is a bounded sample)
        ref.step()
    pts, secs = 0, 0.0
    for _ in range(args.steps):
        p, s = ref.step()
        pts += p
        secs += s
    value = pts / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "k": args.k, "d": args.d,
                   "sample_points_per_step": pts // args.steps},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": ref.procs, "kind": ref.kind, "sample": ref.sample()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_direct(args, threads, points_per_thread=1 << 18):
    """B-direct of BASELINE.md section 3: the mapper + combiner arithmetic called directly (the oracle's C
    restatement, no Pipes protocol, no per-record framing), one slice per host thread.  The arithmetic-only CPU
    rate, printed beside the protocol-bound B-pipes figure."""
    import ctypes as C
    from concurrent.futures import ThreadPoolExecutor

    from hgpumap import synth

    so = os.path.join(ROOT, "oracle", "libkmeans_oracle.so")
    if not os.path.exists(so):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)
    lib = C.CDLL(so)
    centres = synth.centres(synth.SEED_POINTS, args.k, args.d)
    cent = synth.perturb(synth.SEED_CENTROIDS, centres)
    slices = [synth.points(synth.SEED_POINTS, i * points_per_thread, points_per_thread, centres) for i in range(threads)]
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731

    def one(pts):
        sums = np.zeros((args.k, args.d))
        counts = np.zeros(args.k, np.int64)
        lib.hgo_map_combine(p(pts), C.c_int64(len(pts)), args.d, p(cent), args.k, p(sums), p(counts), None, None)
        return int(counts.sum())

    with ThreadPoolExecutor(threads) as pool:
        list(pool.map(one, slices[:threads]))  # warm
        t0 = time.perf_counter()
        done = sum(pool.map(one, slices))      # ctypes releases the GIL: the threads run on separate cores
        dt = time.perf_counter() - t0
    return {"value": done / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{threads} threads x {points_per_thread:,} points through the C restatement of the mapper + combiner "
                      "(oracle/kmeans_oracle.c), called directly: no Pipes protocol"}


def other_points(torch, dev, kind, n, centres, seed=20261108):
    """Inputs a map task does not choose (d = 2): `uniform` = points uniform over the centroids' box + 20 %, ~13 %
    of them in multi-candidate cells of the index and a few outside it; `skew` = 60 % of the points in ONE cluster,
    the rest spread over the others."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    c = torch.as_tensor(centres, device=dev)
    if kind in ("uniform", "uniform_in"):
        lo, hi = c.min(0).values, c.max(0).values
        pad = (0.2 if kind == "uniform" else 0.0) * (hi - lo)
        return (torch.rand(n, c.shape[1], device=dev, generator=g) * (hi - lo + 2 * pad) + (lo - pad)).float()
    if kind == "skew":
        ids = torch.randint(0, c.shape[0], (n,), device=dev, generator=g)
        ids[torch.rand(n, device=dev, generator=g) < 0.6] = 5 % c.shape[0]
        return (c[ids] + torch.randn(n, c.shape[1], device=dev, generator=g)).float()
    raise ValueError(kind)


# ---------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------

class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def wait_first(self, timeout=5.0):
        t0 = time.perf_counter()
        while self.proc and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def stop(self, t_lo=None, t_hi=None):
        """Summarise the samples taken in [t_lo, t_hi] (host perf_counter times)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for stamp, row in self.rows:
            if (t_lo is not None and stamp < t_lo) or (t_hi is not None and stamp > t_hi + 0.05):
                continue
            parts = [p.strip() for p in row.split(",")]
            if len(parts) != 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[2:]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------
# this framework
# ---------------------------------------------------------------------------------------

def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def tf32_peak_tflops(torch, dev):
    """cuBLAS TF32 GEMM rate measured here and now, the way the driver measured the bf16 peak
    (8192^3, best of 5, CUDA events): the denominator for the kind::tf32 tensor-core path."""
    n = 8192
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(n, n, device=dev, dtype=torch.float32)
        b = torch.randn(n, n, device=dev, dtype=torch.float32)
        for _ in range(2):
            torch.matmul(a, b)
        best = float("inf")
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def bind_to_gpu_numa_node(local_rank):
    """One process per GPU: run on, and allocate pinned memory from, the NUMA node the GPU hangs off
    (sysfs local_cpulist of its PCI function), so that the e2e copies do not cross a socket interconnect.
    On this pool's boxes (one virtual socket, numa_node = -1) it changes nothing, and the e2e rate is the platform's
    own: a plain pinned cudaMemcpyAsync sweep (tools/h2d_sweep.py, profiles/r02_h2d_sweep.md) gives 55.5 GB/s per GPU
    at 1 and 2 GPUs, 28.8 at 4 (115 aggregate) and 23.2-35.6 at 8 (235 aggregate); hgm_map_split reaches 55 / 55 / 28.7 /
    23.3 (max over ranks = the slowest GPU).
    Returns a description of the placement or None."""
    try:
        import pynvml

        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:  # 00000000:1b:00.0 -> 0000:1b:00.0
            bus = bus[4:]
        base = f"/sys/bus/pci/devices/{bus}"
        cpulist = open(base + "/local_cpulist").read().strip()
        node = open(base + "/numa_node").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"numa_node": int(node), "cpus": len(cpus), "pci": bus}
    except Exception:  # no NVML, no sysfs entry, a single-node box: leave the placement alone
        return None


def timed_steps(torch, dist, slot, stream, ptr, n, steps, world, warm=3):
    """`steps` passes of reset -> map -> (allreduce) over n HBM-resident points: (ms per step, map ms per step),
    CUDA events on the library's stream, barrier on both sides, max over ranks."""
    def one():
        slot.reset_partials()
        slot.map_device(ptr, n)
        if world > 1:
            slot.allreduce_partials()

    for _ in range(warm):
        one()
    slot.sync()
    k_a = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    k_b = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0.record(stream)
    for i in range(steps):
        slot.reset_partials()
        k_a[i].record(stream)
        slot.map_device(ptr, n)
        k_b[i].record(stream)
        if world > 1:
            slot.allreduce_partials()
    ev1.record(stream)
    slot.sync()
    if world > 1:
        dist.barrier()
    t = torch.tensor([ev0.elapsed_time(ev1) / steps, statistics.mean(a.elapsed_time(b) for a, b in zip(k_a, k_b))],
                     dtype=torch.float64, device=torch.device("cuda", torch.cuda.current_device()))
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def run_extras(args, torch, dist, slot, stream, dev, rank, world, points, centres, peak):
    """The lines the headline does not show (same JSON line, key `extra`): configs[1]'s kernel on inputs a map task
    does not choose, the round-1 combiner beside the new one, and BASELINE configs 4 and 5 at a per-GPU shard that
    fits the step budget, with their kernels timed apart and the cost of the exact re-check at 0 / 0.1 / 1 / 10 %."""
    from hgpumap import capi, synth

    extra = []
    n, k = args.points, args.k
    steps = max(3, min(args.steps, 10))

    def d2_line(name, family=capi.KERNEL_AUTO):
        slot.set_kernel(family)
        ms, kms = timed_steps(torch, dist, slot, stream, points.data_ptr(), n, steps, world)
        _, counts = slot.read_partials()
        assert int(counts.sum()) == n * world
        extra.append({"workload": name, "points_per_gpu": n, "d": 2, "k": k, "value": world * n / (ms * 1e-3),
                      "unit": UNIT, "ms_per_step": ms, "kernel": kernel_name(capi, slot, slot.last_kernel()),
                      "kernel_ms": kms, "roofline_frac_hbm": n * 8 / (kms * 1e-3) / 1e9 / peak})
        slot.set_kernel(capi.KERNEL_AUTO)

    if args.d == 2:
        d2_line("configs[1] input, counting-sort combiner (the round-1 kernel, HGM_KERNEL_GRID2D_SORT)", capi.KERNEL_GRID2D_SORT)
        points.view(n, 2).copy_(other_points(torch, dev, "uniform", n, centres, seed=20261108 + rank))
        d2_line("uniform points over the centroids' box + 20 % (~13 % in multi-candidate cells, some outside the index)")
        points.view(n, 2).copy_(other_points(torch, dev, "skew", n, centres, seed=20261108 + rank))
        d2_line("skewed points: one cluster holds 60 % of every tile")
        d2_line("skewed points, counting-sort combiner", capi.KERNEL_GRID2D_SORT)

    # ---- BASELINE configs 4 and 5: the tcgen05 path ------------------------------------------------------
    tf32 = tf32_peak_tflops(torch, dev)
    try:
        bf16 = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        bf16 = None
    for d, kk, m in ((64, 1024, 40_000_000), (128, 4096, 8_000_000)):
        cen = synth.centres(synth.SEED_POINTS, kk, d)
        cent = synth.perturb(synth.SEED_CENTROIDS, cen)
        buf = torch.empty(m * d, dtype=torch.float32, device=dev)
        slot.set_centroids(cent)
        slot.synth_points_device(synth.SEED_POINTS, rank * m, m, cen, buf.data_ptr())
        entry = None
        fused = (args.tensor_form == "pair_fused" or  # launch_tensor_map's rule
                 (args.tensor_form in ("auto", "pair") and d > 32 and kk >= (1024 if d <= 64 else 2048)))
        for dup in (0, max(1, kk // 1000), kk // 100, kk // 10):  # duplicated centroids: their clusters' points are ties
            c2 = cent.copy()
            if dup:
                c2[kk - dup:] = c2[:dup]
                slot.set_centroids(c2)
            ms, kms = timed_steps(torch, dist, slot, stream, buf.data_ptr(), m, max(3, steps // 2), world)
            _, counts = slot.read_partials()
            assert int(counts.sum()) == m * world
            phases = slot.last_tensor_phases()
            flagged = slot.last_rechecks()
            if dup == 0:
                tf = 2.0 * d * kk * m / (kms * 1e-3) / 1e12
                entry = {"workload": f"BASELINE configs[{3 if d == 64 else 4}] shape: d={d}, k={kk}, {m:,} points per GPU",
                         "points_per_gpu": m, "d": d, "k": kk, "value": world * m / (ms * 1e-3), "unit": UNIT,
                         "ms_per_step": ms, "kernel": "tensor_assign_pair_kernel" if d > 32 else "tensor_assign_kernel",
                         "sums": "inside K2 (decided rows leave as FP64 reductions from shared memory)" if fused else "K4 segmented second pass",
                         "kernel_ms": {"map_call": kms, "K2_scores_tcgen05": phases[0], "K3_exact_recheck": phases[1],
                                       "K4_counts_and_rechecked_rows" if fused else "K4_segmented_sums": phases[2]},
                         "rechecked_points": flagged, "achieved_tflops": tf,
                         "roofline_frac_tf32_measured_in_run": tf / tf32, "tf32_tflops_cublas_in_run": tf32,
                         # cuBLAS's 8192^3 TF32 GEMM is a measured yardstick, not the pipe's ceiling (the kernel can pass it:
                         # it also multiplies the |c|^2 K-block, (d + 32) / d more flops than counted here); the nominal
                         # dense TF32 rate (half of the 2.25 PFLOP/s bf16 figure) is the other denominator
                         "roofline_frac_tf32_nominal": tf / 1125.0, "tf32_tflops_nominal": 1125.0,
                         "roofline_frac_bf16_sustained": tf / bf16 if bf16 else None, "bf16_tflops_sustained": bf16,
                         "recheck_sensitivity": []}
            else:
                entry["recheck_sensitivity"].append(
                    {"duplicated_centroids": dup, "rechecked_points": flagged, "rechecked_frac": flagged / m,
                     "value": world * m / (ms * 1e-3), "K3_exact_recheck_ms": phases[1], "map_call_ms": kms})
        if fused and args.tensor_form == "auto":  # the same call with the sums made by K4's second pass over the rows
            slot.set_centroids(cent)
            slot.set_tensor_form(capi.TENSOR_FORM_PAIR_SEGMENTED)
            ms, kms = timed_steps(torch, dist, slot, stream, buf.data_ptr(), m, max(3, steps // 2), world)
            phases = slot.last_tensor_phases()
            slot.set_tensor_form(capi.TENSOR_FORM_AUTO)
            entry["same_call_with_K4_second_pass"] = {
                "value": world * m / (ms * 1e-3), "map_call_ms": kms, "K2_scores_tcgen05": phases[0],
                "K3_exact_recheck": phases[1], "K4_segmented_sums": phases[2]}
        extra.append(entry)
        del buf
    return extra


def run_ours(args):
    import torch
    import torch.distributed as dist

    from hgpumap import capi, synth

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run")
    # N = 1 keeps the whole box for the cpu_baseline leg; with one rank per GPU each rank moves next to its GPU
    placement = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()  # nvidia-smi takes a while to come up; it is read only around the timed region
    n, k, d = args.points, args.k, args.d
    centres = synth.centres(synth.SEED_POINTS, k, d)
    centroids = synth.perturb(synth.SEED_CENTROIDS, centres)
    slot = capi.GpuMapSlot(local_rank)
    if args.kernel:
        slot.set_kernel(args.kernel)
    slot.set_tensor_form({"auto": capi.TENSOR_FORM_AUTO, "single": capi.TENSOR_FORM_SINGLE,
                          "pair": capi.TENSOR_FORM_PAIR, "pair_fused": capi.TENSOR_FORM_PAIR_FUSED,
                          "pair_segmented": capi.TENSOR_FORM_PAIR_SEGMENTED}[args.tensor_form])
    slot.set_centroids(centroids)
    points = torch.empty(n * d, dtype=torch.float32, device=dev)  # device memory: plumbing only
    slot.synth_points_device(synth.SEED_POINTS, rank * n, n, centres, points.data_ptr())
    peer_path = False
    if world > 1:
        box = [capi.GpuMapSlot.comm_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        slot.comm_init(world, rank, box[0])
        # small partials (the d = 2 shapes): exchange through peer memory instead of NCCL's launch latency
        try:
            handle, peer_ok = slot.peer_handle(world), 1
        except capi.HgmError:
            handle, peer_ok = b"\0" * 64, 0
        handles = [None] * world
        dist.all_gather_object(handles, handle)
        if peer_ok:
            try:
                slot.peer_open(rank, handles)
            except capi.HgmError:
                peer_ok = 0
        flag = torch.tensor([peer_ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        peer_path = bool(flag.item())
        if not peer_path:
            slot.peer_close()
    stream = torch.cuda.ExternalStream(slot.stream(), device=dev)

    def step():
        slot.reset_partials()
        slot.map_device(points.data_ptr(), n)
        if world > 1:
            slot.allreduce_partials()

    for _ in range(max(args.warmup, 3)):
        step()
    slot.sync()
    sums, counts = slot.read_partials()
    assert int(counts.sum()) == n * world, f"partials count {int(counts.sum())} != {n * world}"
    kernel_family = slot.last_kernel()

    # ---- value: K steps, HBM-resident, CUDA events on the library's stream -------------
    k_a = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    k_b = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if rank == 0:
        sampler.wait_first()
    # The K timed steps last only milliseconds, less than nvidia-smi's sampling period, so
    # the clocks are sampled over ~1 s of the identical load run back to back with (and
    # including) the timed region; these extra steps are untimed warm-up.  Every rank runs the
    # SAME number of them (a step holds a collective): the count is agreed on first.
    t_probe = time.perf_counter()
    for _ in range(8):
        step()
    slot.sync()
    per_step = torch.tensor([(time.perf_counter() - t_probe) / 8], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(per_step, op=dist.ReduceOp.MAX)
    load_steps = int(min(4000, max(16, 1.0 / max(per_step.item(), 1e-5))))
    t_load0 = time.perf_counter()
    for _ in range(load_steps):
        step()
    slot.sync()
    launches0 = slot.launch_count()
    barrier()
    ev0.record(stream)
    for i in range(args.steps):
        slot.reset_partials()
        k_a[i].record(stream)
        slot.map_device(points.data_ptr(), n)
        k_b[i].record(stream)
        if world > 1:
            slot.allreduce_partials()
    ev1.record(stream)
    slot.sync()
    barrier()
    clocks = sampler.stop(t_load0 + 0.1, time.perf_counter()) if rank == 0 else None
    launches = slot.launch_count() - launches0
    elapsed_ms = ev0.elapsed_time(ev1)
    kernel_ms = statistics.mean(a.elapsed_time(b) for a, b in zip(k_a, k_b))
    t = torch.tensor([elapsed_ms, kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms, kernel_ms = t.tolist()
    value = world * n * args.steps / (elapsed_ms * 1e-3)

    # ---- e2e: the C-ABI call a Pipes child / JNI stub makes, host buffers ----------------
    e2e = None
    if not args.no_e2e:
        host = torch.empty(n * d, dtype=torch.float32, pin_memory=True)
        host.copy_(points)
        torch.cuda.synchronize()
        h_sums = np.zeros((k, d))
        h_counts = np.zeros(k, np.int64)
        nbytes = n * d * 4

        def e2e_step():
            slot.map_split_raw(host.data_ptr(), nbytes, h_sums, h_counts)
            if world > 1:
                slot.allreduce_partials()
                return slot.read_partials()
            return h_sums, h_counts

        for _ in range(3):
            e2e_step()
        e2e_steps = max(3, min(args.steps, 10))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            s_, c_ = e2e_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        assert int(c_.sum()) == n * world
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = t.item()
        e2e = {"value": world * n * e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": k * d * 8 + k * 8, "steps": e2e_steps, "ms_per_step": 1e3 * dt / e2e_steps,
               "api": "hgm_map_split (pinned host split -> chunked async H2D -> kernel -> partials D2H)"}
        del host

    peak, peak_src = hbm_peak()
    main_kernel = kernel_name(capi, slot, kernel_family)
    extra = None
    main_rechecks = slot.last_rechecks() if kernel_family == capi.KERNEL_TENSOR else 0
    if not args.no_extra and d == 2:  # every rank takes part: the steps hold a collective
        extra = run_extras(args, torch, dist, slot, stream, dev, rank, world, points, centres, peak)
    if rank == 0:
        achieved = n * d * 4 / (kernel_ms * 1e-3) / 1e9
        fam = main_kernel
        traffic, limiter = ncu_traffic(fam, d, k, n)
        if kernel_family == capi.KERNEL_TENSOR:
            # GEMM-shaped: algorithmic flops = 2*d*k per point, against the
            # TF32 tensor-pipe rate measured in this run; kernel_ms spans the score kernel and the
            # (normally empty) exact re-check launch that follows it
            tf32 = tf32_peak_tflops(torch, dev)
            achieved_tf = 2.0 * d * k * n / (kernel_ms * 1e-3) / 1e12
            roofline = {"bound": "tensor", "achieved": achieved_tf, "peak": tf32, "unit": "TFLOP/s",
                        "frac": achieved_tf / tf32, "traffic": traffic, "kernel": fam, "kernel_ms": kernel_ms,
                        "peak_source": "measured in this run: cuBLAS TF32 8192^3, best of 5 (the path runs "
                                       "tcgen05 kind::tf32; MEASURED_PEAKS.json holds only the bf16 rate)",
                        "algorithmic_flops_per_point": 2 * d * k, "rechecked_points_last_step": main_rechecks}
        else:
            roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": traffic, "kernel": fam, "kernel_ms": kernel_ms, "peak_source": peak_src,
                        "algorithmic_bytes_per_point": 4 * d}
            if limiter:  # what the committed ncu capture of this workload shows the kernel is bound by
                roofline["limiter"] = limiter
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args), "points_per_gpu": n, "k": k, "d": d, "kernel": fam,
                       "l2": f"inputs ({n * d * 4 / 1e6:.0f} MB per GPU) larger than the 126 MB L2; no flush needed",
                       "parallelism": (f"{world} GPU map slot(s), split-sharded; partials of k*d sums + k counts combined by "
                                       + ("a one-CTA exchange through peer memory over NVLink (hgm_peer_open; NCCL for "
                                          "partials over 64 KiB)" if peer_path else "one NCCL allreduce"))
                       if world > 1 else "1 GPU map slot",
                       "host_placement": placement},
            "roofline": roofline,
            "clocks": clocks,
            "gpu_launches": launches,
        }
        if e2e:
            line["e2e"] = e2e
        if extra is not None:
            line["extra"] = extra
        if world == 1 and not args.no_cpu_baseline:
            cores = min(host_cores(), 32)
            ref = CpuReference(args, procs=cores, points=args.cpu_points)
            p, s = ref.step()
            line["cpu_baseline"] = {"value": p / s, "unit": UNIT, "cores": ref.procs, "kind": ref.kind,
                                    "sample": ref.sample(),
                                    "direct": cpu_direct(args, cores)}  # the arithmetic alone, no protocol (B-direct)
        print(json.dumps(line), flush=True)
    slot.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
