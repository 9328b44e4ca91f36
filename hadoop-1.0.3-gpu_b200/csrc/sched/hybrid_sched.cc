// hybrid_sched.cc -- the fork's hybrid CPU/GPU map assignment rule and a one-TaskTracker
// simulator around it (include/hgsched.h has the reference citations).
#include "../../../include/hgsched.h"

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <new>
#include <queue>
#include <vector>

namespace {

// JobQueueTaskScheduler.java:134-177: statistics over the running jobs
struct QueueStats {
  int pending = 0;
  int cpu_mean = 0, gpu_mean = 0;
  double accel = 0.0;
};

QueueStats queue_stats(const hgs_job *jobs, int n, int fixes) {
  QueueStats s;
  int remaining = 0;
  long long cpu_sum = 0, gpu_sum = 0;
  int cpu_jobs = 0, gpu_jobs = 0;
  for (int i = 0; i < n; ++i) {
    const hgs_job &j = jobs[i];
    if (fixes & HGS_FIX_PENDING) {
      s.pending += j.desired_maps - j.finished_maps - j.running_maps;
    } else {
      remaining += j.desired_maps - j.finished_maps;  // :137
      s.pending += remaining - j.running_maps;        // :139 adds the RUNNING TOTAL, not this job's share
    }
    s.cpu_mean = j.cpu_mean_ms;  // :143-144 the last running job wins
    s.gpu_mean = j.gpu_mean_ms;
    if (j.cpu_mean_ms) { cpu_sum += j.cpu_mean_ms; ++cpu_jobs; }
    if (j.gpu_mean_ms) { gpu_sum += j.gpu_mean_ms; ++gpu_jobs; }
  }
  if (fixes & HGS_FIX_PER_JOB_ACCEL) {
    s.cpu_mean = cpu_jobs ? (int)(cpu_sum / cpu_jobs) : 0;
    s.gpu_mean = gpu_jobs ? (int)(gpu_sum / gpu_jobs) : 0;
  }
  s.accel = (s.cpu_mean == 0 || s.gpu_mean == 0) ? 0.0 : (double)s.cpu_mean / (double)s.gpu_mean;  // :175-177
  if (s.accel == 0.0 && (fixes & HGS_FIX_PRIOR_ACCEL))
    for (int i = 0; i < n; ++i)
      if (jobs[i].prior_acceleration > 0.0) s.accel = jobs[i].prior_acceleration;
  return s;
}

}  // namespace

extern "C" int hgs_assign(const hgs_job *jobs, int32_t n_jobs, const hgs_tracker *tt, int32_t num_trackers,
                          int32_t optional, int32_t fixes, int32_t want, hgs_decision *out) {
  if (!jobs || !tt || !out || n_jobs < 0 || num_trackers < 1 || tt->max_gpu_slots > HGS_MAX_DEVICES) return -22;
  *out = hgs_decision{};
  const QueueStats s = queue_stats(jobs, n_jobs, fixes);
  out->acceleration = s.accel;
  out->pending_map_load = s.pending;
  int avail_cpu = tt->max_cpu_slots - tt->running_cpu;  // :257-258
  int avail_gpu = tt->max_gpu_slots - tt->running_gpu;
  if (avail_cpu < 0) avail_cpu = 0;
  if (avail_gpu < 0) avail_gpu = 0;
  int left = std::max(want, 0);

  // CPU pass (:290-326): skipped near the tail of the job when GPUs alone finish sooner
  const bool skip_cpu = optional && std::max(s.pending, 0) < s.accel * tt->max_gpu_slots * num_trackers;
  out->cpu_pass_skipped = skip_cpu ? 1 : 0;
  if (!skip_cpu) {
    const int take = std::min(avail_cpu, left);
    out->n_cpu = take;
    left -= take;
  }
  // GPU pass (:334-387): always; a job without a GPU executable stops the pass (`break`)
  bool gpu_ok = n_jobs > 0;
  for (int i = 0; i < n_jobs && i < 1; ++i) gpu_ok = jobs[i].has_gpu_executable != 0;
  if (gpu_ok) {
    uint32_t busy = tt->busy_devices;
    const int scan = tt->max_gpu_slots;  // node-local pick scans trackerGPUMapCapacity ids (:354)
    const int take = std::min(avail_gpu, left);
    for (int t = 0; t < take; ++t) {
      int dev = 0;  // Task's default GPUDeviceId when no free id is found
      for (int j = 0; j < scan; ++j)
        if (!(busy & (1u << j))) { dev = j; busy |= 1u << j; break; }
      out->gpu_device[out->n_gpu++] = dev;
    }
  }
  return 0;
}

extern "C" int hgs_tracker_from_reports(const hgs_task_status *reports, int32_t n, int32_t max_cpu_slots,
                                        int32_t max_gpu_slots, hgs_tracker *out) {
  if (!out || n < 0 || (n > 0 && !reports) || max_cpu_slots < 0 || max_gpu_slots < 0 ||
      max_gpu_slots > HGS_MAX_DEVICES)
    return -22;
  *out = hgs_tracker{max_cpu_slots, max_gpu_slots, 0, 0, 0u};
  for (int i = 0; i < n; ++i) {
    const hgs_task_status &ts = reports[i];
    const bool cleanup = ts.cleanup_phase && (ts.run_state == HGS_STATE_FAILED_UNCLEAN ||
                                              ts.run_state == HGS_STATE_KILLED_UNCLEAN);  // TaskStatus.java:242-246
    const bool holds_slot = ts.run_state == HGS_STATE_RUNNING || ts.run_state == HGS_STATE_UNASSIGNED || cleanup;
    if (!ts.is_map || !holds_slot) continue;
    if (!ts.run_on_gpu) { ++out->running_cpu; continue; }
    if (ts.gpu_device_id < 0 || ts.gpu_device_id >= max_gpu_slots) return -22;
    ++out->running_gpu;
    out->busy_devices |= 1u << ts.gpu_device_id;
  }
  return 0;
}

extern "C" int32_t hgs_mean_map_time_ms(const int64_t *start_ms, const int64_t *finish_ms, const int32_t *ran_on_gpu,
                                        int32_t n, int32_t want_gpu) {
  if (n < 0 || (n > 0 && (!start_ms || !finish_ms || !ran_on_gpu))) return -22;
  uint32_t total = 0;  // Java int arithmetic: two's-complement wrap, done unsigned here to stay defined
  int32_t tasks = 0;
  for (int i = 0; i < n; ++i) {
    if (finish_ms[i] == 0 || (ran_on_gpu[i] != 0) != (want_gpu != 0)) continue;
    total += (uint32_t)(uint64_t)(finish_ms[i] - start_ms[i]);  // totalTime += (int)(long difference)
    ++tasks;
  }
  if (tasks == 0) return 0;
  return (int32_t)total / tasks;  // Java's int division truncates toward zero, like C's
}

extern "C" int32_t hgs_pick_split(const int32_t *home, int32_t n, int32_t device, int32_t max_gpu_slots) {
  if (n < 0 || (n > 0 && !home) || max_gpu_slots < 1 || max_gpu_slots > HGS_MAX_DEVICES || device < 0 ||
      device >= max_gpu_slots)
    return -22;
  if (n == 0) return -1;
  int first_unheld = -1;
  int queued[HGS_MAX_DEVICES] = {};  // pending resident splits per device
  int first_of[HGS_MAX_DEVICES];
  for (int j = 0; j < HGS_MAX_DEVICES; ++j) first_of[j] = -1;
  for (int i = 0; i < n; ++i) {
    const int h = home[i];
    if (h == device) return i;  // (1)
    if (h < 0 || h >= max_gpu_slots) {  // held nowhere (or by a slot that no longer exists)
      if (first_unheld < 0) first_unheld = i;
    } else {
      if (first_of[h] < 0) first_of[h] = i;
      ++queued[h];
    }
  }
  if (first_unheld >= 0) return first_unheld;  // (2)
  int victim = -1;
  for (int j = 0; j < max_gpu_slots; ++j)
    if (queued[j] > 0 && (victim < 0 || queued[j] > queued[victim])) victim = j;
  return first_of[victim];  // (3): every pending split is held by some other device
}

// ---- TaskLauncher's free-slot counters + releaseSlot (mapred/TaskTracker.java:2434-2629, 3397-3408) -----------
struct hgs_slots {
  std::mutex lock;
  std::condition_variable freed;  // the Java has one monitor per counter; one condition serves both here
  int32_t max_cpu, max_gpu, free_cpu, free_gpu;
  uint32_t busy_devices = 0;
};

extern "C" hgs_slots *hgs_slots_new(int32_t max_cpu_slots, int32_t max_gpu_slots) {
  if (max_cpu_slots < 0 || max_gpu_slots < 0 || max_gpu_slots > HGS_MAX_DEVICES) return nullptr;
  hgs_slots *s = new (std::nothrow) hgs_slots();
  if (!s) return nullptr;
  s->max_cpu = s->free_cpu = max_cpu_slots;  // TaskLauncher(...) :2445-2453
  s->max_gpu = s->free_gpu = max_gpu_slots;
  return s;
}
extern "C" void hgs_slots_delete(hgs_slots *s) { delete s; }

extern "C" int32_t hgs_slots_acquire(hgs_slots *s, int32_t on_gpu, int32_t need, int32_t device, int32_t timeout_ms,
                                     const volatile int32_t *cancelled) {
  if (!s || need < 1) return -22;
  if (device >= 0 && (!on_gpu || device >= s->max_gpu)) return -22;
  if (need > (on_gpu ? s->max_gpu : s->max_cpu)) return -34;  // the launcher would wait for ever
  std::unique_lock<std::mutex> hold(s->lock);
  int32_t &free_slots = on_gpu ? s->free_gpu : s->free_cpu;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms < 0 ? 0 : timeout_ms);
  while (free_slots < need) {  // :2529, :2568
    if (cancelled && *cancelled) return 1;  // !tip.canBeLaunched(): killed while still in the launcher queue
    if (timeout_ms == 0) return 1;
    if (timeout_ms < 0) s->freed.wait(hold);
    else if (s->freed.wait_until(hold, deadline) == std::cv_status::timeout && free_slots < need) return 1;
  }
  if (cancelled && *cancelled) return 1;
  if (device >= 0) {
    if (s->busy_devices & (1u << device)) return -34;  // two tasks on one device: slot i <-> device i is broken
    s->busy_devices |= 1u << device;
  }
  free_slots -= need;  // :2560, :2597; the Java asserts the result >= 0, which the loop above guarantees
  return 0;
}

extern "C" int32_t hgs_slots_release(hgs_slots *s, int32_t on_gpu, int32_t n, int32_t device) {
  if (!s || n < 1) return -22;
  if (device >= 0 && (!on_gpu || device >= s->max_gpu)) return -22;
  std::lock_guard<std::mutex> hold(s->lock);
  int32_t &free_slots = on_gpu ? s->free_gpu : s->free_cpu;
  if (free_slots + n > (on_gpu ? s->max_gpu : s->max_cpu)) return -34;  // assert (free <= max) :2472, :2482
  if (device >= 0) {
    if (!(s->busy_devices & (1u << device))) return -34;
    s->busy_devices &= ~(1u << device);
  }
  free_slots += n;
  s->freed.notify_all();
  return 0;
}

extern "C" int32_t hgs_slots_release_task(hgs_slots *s, int32_t on_gpu, int32_t n, int32_t device, int32_t *slot_taken) {
  if (!s || !slot_taken) return -22;
  if (*slot_taken) {  // :3398-3407
    const int32_t rc = hgs_slots_release(s, on_gpu, n, device);
    if (rc == 0) *slot_taken = 0;
    return rc;
  }
  hgs_slots_notify(s);  // "wake up the launcher. it may be waiting to block slots for this task."
  return 0;
}

extern "C" void hgs_slots_notify(hgs_slots *s) {
  if (!s) return;
  std::lock_guard<std::mutex> hold(s->lock);
  s->freed.notify_all();
}

extern "C" int32_t hgs_slots_snapshot(hgs_slots *s, hgs_tracker *out) {
  if (!s || !out) return -22;
  std::lock_guard<std::mutex> hold(s->lock);
  out->max_cpu_slots = s->max_cpu;
  out->max_gpu_slots = s->max_gpu;
  out->running_cpu = s->max_cpu - s->free_cpu;
  out->running_gpu = s->max_gpu - s->free_gpu;
  out->busy_devices = s->busy_devices;
  return 0;
}

extern "C" int hgs_simulate(const hgs_sim_params *p, hgs_sim_result *out) {
  if (!p || !out || p->n_splits < 0 || p->iterations < 1 || p->cpu_slots < 0 || p->gpu_slots < 0 ||
      p->gpu_slots > HGS_MAX_DEVICES || p->heartbeat_ms <= 0 || !(p->task_jitter >= 0.0 && p->task_jitter < 1.0))
    return -22;
  // task-time scale in [1 - jitter, 1 + jitter]: an integer hash, the same on every run
  auto jitter = [p](int it, int task) {
    uint32_t h = (uint32_t)it * 0x9E3779B9u ^ ((uint32_t)task + 0x7F4A7C15u) * 0x85EBCA6Bu;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
    return 1.0 + p->task_jitter * ((double)h / 2147483648.0 - 1.0);
  };
  *out = hgs_sim_result{};
  struct Running { double finish; bool gpu; int dev; double start; };
  auto later = [](const Running &a, const Running &b) { return a.finish > b.finish; };
  double now = 0.0;
  std::vector<int32_t> home((size_t)p->n_splits, -1);  // device whose HBM cache holds the split
  std::vector<int32_t> pending, pending_home;
  for (int it = 0; it < p->iterations; ++it) {
    std::priority_queue<Running, std::vector<Running>, decltype(later)> running(later);
    int finished = 0, started = 0;
    pending.resize((size_t)p->n_splits);
    for (int i = 0; i < p->n_splits; ++i) pending[(size_t)i] = i;
    long long cpu_time = 0, gpu_time = 0;  // integer milliseconds like JobInProgress sums them
    int cpu_done = 0, gpu_done = 0;
    // the TaskTracker's slot accounting: the same unit a live tracker would use (never blocks here: the scheduler
    // only hands out what the heartbeat reported free)
    std::unique_ptr<hgs_slots, void (*)(hgs_slots *)> slots(hgs_slots_new(p->cpu_slots, p->gpu_slots), hgs_slots_delete);
    if (!slots) return -22;
    double next_hb = now;
    const double gpu_ms = it == 0 ? p->gpu_task_ms_first : p->gpu_task_ms_cached;
    while (finished < p->n_splits) {
      // advance to the next event: a heartbeat, or (out-of-band) a task completion
      double t_next = next_hb;
      if (!running.empty() && running.top().finish <= t_next) {
        const Running r = running.top();
        running.pop();
        now = std::max(now, r.finish);
        ++finished;
        const long long took = (long long)(r.finish - r.start);  // finishTime - startTime, ms
        if (r.gpu) { gpu_time += took; ++gpu_done; }
        else { cpu_time += took; ++cpu_done; }
        int32_t taken = 1;
        if (hgs_slots_release_task(slots.get(), r.gpu, 1, r.gpu ? r.dev : -1, &taken) != 0) return -34;
        if (p->out_of_band_heartbeat) next_hb = now;
        continue;
      }
      now = std::max(now, t_next);
      next_hb = now + p->heartbeat_ms;
      ++out->heartbeats;
      hgs_tracker tt;
      hgs_slots_snapshot(slots.get(), &tt);
      hgs_job job{p->n_splits, finished, tt.running_cpu + tt.running_gpu, cpu_done ? (int)(cpu_time / cpu_done) : 0,
                  gpu_done ? (int)(gpu_time / gpu_done) : 0, 1, p->prior_acceleration};
      hgs_decision d;
      hgs_assign(&job, 1, &tt, 1, p->optional_scheduling, p->fixes, p->n_splits - started, &d);
      out->cpu_pass_skips += d.cpu_pass_skipped;
      out->last_acceleration = d.acceleration;
      for (int i = 0; i < d.n_cpu; ++i) {
        pending.erase(pending.begin());  // a CPU slot takes the next split; it stays where it was cached
        running.push(Running{now + p->task_overhead_ms + p->cpu_task_ms * jitter(it, started), false, -1, now});
        if (hgs_slots_acquire(slots.get(), 0, 1, -1, 0, nullptr) != 0) return -34;  // the rule handed out a busy slot
        ++started; ++out->tasks_cpu;
      }
      for (int i = 0; i < d.n_gpu; ++i) {
        const int dev = d.gpu_device[i];
        size_t at = 0;
        if (p->per_device_cache && p->device_affinity) {
          pending_home.resize(pending.size());
          for (size_t j = 0; j < pending.size(); ++j) pending_home[j] = home[(size_t)pending[j]];
          at = (size_t)hgs_pick_split(pending_home.data(), (int32_t)pending_home.size(), dev, p->gpu_slots);
        }
        const int32_t split = pending[at];
        pending.erase(pending.begin() + (long)at);
        double ms = gpu_ms;
        if (p->per_device_cache) {
          const bool hit = home[(size_t)split] == dev;
          ms = hit ? p->gpu_task_ms_cached : p->gpu_task_ms_first;
          out->gpu_cache_hits += hit;
          home[(size_t)split] = dev;
        }
        running.push(Running{now + p->task_overhead_ms + ms * jitter(it, started), true, dev, now});
        if (hgs_slots_acquire(slots.get(), 1, 1, dev, 0, nullptr) != 0) return -34;  // ... or a busy device
        ++started; ++out->tasks_gpu;
      }
      if (running.empty() && started >= p->n_splits) break;
    }
  }
  out->makespan_ms = now;
  return 0;
}
