/* hgsched.h -- C ABI of libhgsched.so: the fork's hybrid CPU/GPU map scheduling rule, restated.
 *
 * The reference decides, on every TaskTracker heartbeat, how many map tasks go to CPU slots
 * and how many to GPU slots (and on which device):
 *   JobQueueTaskScheduler.assignTasks   mapred/org/apache/hadoop/mapred/JobQueueTaskScheduler.java:86-575
 *     job statistics                    :134-169   (pendingMapLoad, mean CPU / GPU map times)
 *     acceleration factor               :175-177   cpuMean / gpuMean, 0 if either is 0
 *     CPU pass (suppressed near the tail) :290-326 skipped iff optionalscheduling &&
 *                                                  max(pending,0) < accel * gpuSlots * trackers
 *     GPU pass, device = first free id  :334-387
 *   mean times (integer ms, finished tasks only)   mapred/JobInProgress.java:527-565
 *   free device bitmap                             mapred/TaskTrackerStatus.java:536-551
 *   slot counters the launcher blocks on           mapred/TaskTracker.java:2434-2629
 * Java cannot be built in this image, so the rule lives here in C++ behind a C ABI (a JNI or
 * test harness binds it); hgs_simulate drives it over one TaskTracker for BASELINE config 3.
 */
#ifndef HGSCHED_H
#define HGSCHED_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define HGS_MAX_DEVICES 32

/* One running job as assignTasks sees it (JobInProgress getters). */
typedef struct {
  int32_t desired_maps;       /* desiredMaps()                                   */
  int32_t finished_maps;      /* finishedMaps()                                  */
  int32_t running_maps;       /* runningMaps()                                   */
  int32_t cpu_mean_ms;        /* getCPUMapTaskMeanTime(): int ms, 0 if none yet  */
  int32_t gpu_mean_ms;        /* getGPUMapTaskMeanTime()                         */
  int32_t has_gpu_executable; /* hadoop.pipes.gpu.executable is set (:342-347)   */
  double prior_acceleration;  /* HGS_FIX_PRIOR_ACCEL: factor measured earlier (previous iteration /
                                 job, or a calibration run), used until this job has finished
                                 tasks of both kinds; 0 = none                  */
} hgs_job;

/* One TaskTracker heartbeat (TaskTrackerStatus). */
typedef struct {
  int32_t max_cpu_slots;   /* mapred.tasktracker.map.cpu.tasks.maximum */
  int32_t max_gpu_slots;   /* mapred.tasktracker.map.gpu.tasks.maximum */
  int32_t running_cpu;     /* countCPUMapTasks()                       */
  int32_t running_gpu;     /* countGPUMapTasks()                       */
  uint32_t busy_devices;   /* bit j set = device j runs a GPU map      */
} hgs_tracker;

/* One entry of a heartbeat's task reports (mapred/TaskStatus.java: runState :46-47, phase :43,
 * runOnGPU / GPUDeviceId :66-67, 390-395 -- the two fields the fork added to the wire format, :417-418). */
#define HGS_STATE_RUNNING 0
#define HGS_STATE_SUCCEEDED 1
#define HGS_STATE_FAILED 2
#define HGS_STATE_UNASSIGNED 3
#define HGS_STATE_KILLED 4
#define HGS_STATE_COMMIT_PENDING 5
#define HGS_STATE_FAILED_UNCLEAN 6
#define HGS_STATE_KILLED_UNCLEAN 7
typedef struct {
  int32_t is_map;         /* getIsMap()                                               */
  int32_t run_state;      /* HGS_STATE_*                                              */
  int32_t cleanup_phase;  /* phase == CLEANUP (inTaskCleanupPhase needs *_UNCLEAN too) */
  int32_t run_on_gpu;     /* runOnGPU()                                               */
  int32_t gpu_device_id;  /* GPUDeviceId()                                            */
} hgs_task_status;

/* The GPU-slot accounting of a heartbeat: countCPUMapTasks / countGPUMapTasks / availableGPUDevices of
 * mapred/TaskTrackerStatus.java:504-551 over the task reports.  A map task occupies its slot (and a GPU
 * task its device: slot i <-> device i) while RUNNING, UNASSIGNED, or in the cleanup phase of a
 * *_UNCLEAN state (:524-528, 541-545).  Fills max_*_slots, running_cpu, running_gpu and busy_devices of
 * *out.  Returns 0; -22 on a bad argument or a GPU task whose device id is outside 0..max_gpu_slots-1
 * (the Java indexes a boolean[maxGPUMapSlots] with it, :546, and would throw). */
int hgs_tracker_from_reports(const hgs_task_status *reports, int32_t n_reports, int32_t max_cpu_slots,
                             int32_t max_gpu_slots, hgs_tracker *out);

/* get{CPU,GPU}MapTaskMeanTime of mapred/JobInProgress.java:527-565: over the map task reports whose
 * finish time is not 0 and whose successful attempt ran on the wanted kind of slot, the mean of
 * finish - start in milliseconds -- summed in a Java `int` (the compound assignment narrows each long
 * difference and the sum wraps at 2^31) and divided as integers; 0 when no such task has finished.
 * These two numbers are all the scheduler knows about task speed (hgs_job.cpu_mean_ms / gpu_mean_ms). */
int32_t hgs_mean_map_time_ms(const int64_t *start_ms, const int64_t *finish_ms, const int32_t *ran_on_gpu,
                             int32_t n_reports, int32_t want_gpu);

/* ---- the TaskTracker's slot accounting as a unit -------------------------------------------------------------
 * TaskLauncher (mapred/TaskTracker.java:2434-2629): two counters of free slots, CPU and GPU, that the launcher
 * thread blocks on before it starts a task (:2526-2600: wait while free < needed, then subtract, assert >= 0) and
 * that a finishing task gives back (addFree{CPU,GPU}Slots :2467-2485: add, assert <= maximum, notifyAll);
 * TaskInProgress.releaseSlot (:3397-3408) gives a slot back only if the task had taken one (slotTaken), else it
 * only wakes the launcher.  On top of the reference: the device of a GPU task is tracked (slot i <-> device i,
 * mapred/TaskTrackerStatus.java:536-551), and taking a device that is already taken is an error -- the state the
 * device-affinity patch (patches/) exists to prevent.  Thread-safe; acquire blocks like the launcher does.
 * Return codes: 0 done; 1 not done (timeout, or cancelled while waiting = the task was killed in the queue,
 * :2535-2541); -22 bad argument; -34 the call would break an assertion of the Java (free > maximum on release,
 * more slots needed than exist, device already busy / not busy). */
typedef struct hgs_slots hgs_slots;
hgs_slots *hgs_slots_new(int32_t max_cpu_slots, int32_t max_gpu_slots);
void hgs_slots_delete(hgs_slots *s);
/* Wait until `need` slots of the kind are free and take them.  timeout_ms: < 0 wait without limit, 0 do not wait.
 * cancelled (may be NULL) is read every time the waiter wakes: non-zero = give up without taking anything.
 * device: for a GPU task the device id it was assigned (0..max_gpu_slots-1, marked busy), -1 for none. */
int32_t hgs_slots_acquire(hgs_slots *s, int32_t on_gpu, int32_t need, int32_t device, int32_t timeout_ms,
                          const volatile int32_t *cancelled);
/* addFreeCPUSlots / addFreeGPUSlots; device as given to acquire (its bit is cleared) or -1. */
int32_t hgs_slots_release(hgs_slots *s, int32_t on_gpu, int32_t n, int32_t device);
/* TaskInProgress.releaseSlot: gives the slots back iff *slot_taken, which it clears; otherwise only wakes waiters. */
int32_t hgs_slots_release_task(hgs_slots *s, int32_t on_gpu, int32_t n, int32_t device, int32_t *slot_taken);
/* notifySlots: wake every waiter (it re-reads its `cancelled`). */
void hgs_slots_notify(hgs_slots *s);
/* What a heartbeat reports: fills max_*_slots, running_cpu, running_gpu (slots in use) and busy_devices. */
int32_t hgs_slots_snapshot(hgs_slots *s, hgs_tracker *out);

/* Behaviour switches.  All 0 = the reference as written, quirks included. */
#define HGS_FIX_PENDING 1     /* pendingMapLoad per job instead of the running total (:139)        */
#define HGS_FIX_PER_JOB_ACCEL 2 /* factor from every job's times, not only the last job's (:143-144) */
#define HGS_FIX_PRIOR_ACCEL 4 /* start from a measured factor instead of 0 ("fed by measured per-task
                                 GPU acceleration"): without it every job begins by filling all CPU
                                 slots, and a GPU task shorter than 1 ms (integer-ms means,
                                 JobInProgress.java:527-565) keeps the factor at 0 forever */
/* Not modelled: data locality.  Every task is handed out by the node-local branch (:350-366), so the
 * non-local branch's shorter device-id scan (:373, availableGPUMapSlots instead of the capacity) and
 * its one-task-per-heartbeat `break scheduleGPUMaps` never apply on the single node simulated here. */

typedef struct {
  int32_t n_cpu;                         /* map tasks handed to CPU slots                    */
  int32_t n_gpu;                         /* map tasks handed to GPU slots                    */
  int32_t gpu_device[HGS_MAX_DEVICES];   /* device id stamped on each GPU task (setGPUDeviceId) */
  int32_t cpu_pass_skipped;              /* 1 if the optional-scheduling rule suppressed CPUs */
  int32_t pending_map_load;
  double acceleration;
} hgs_decision;

/* One assignTasks call.  `want` = how many more map tasks the job queue could hand out (the
 * obtainNew*MapTask calls return null beyond it); all tasks are taken as node-local.
 * Returns 0, or -22 on a bad argument. */
int hgs_assign(const hgs_job *jobs, int32_t n_jobs, const hgs_tracker *tracker, int32_t num_trackers,
               int32_t optional_scheduling, int32_t fixes, int32_t want, hgs_decision *out);

/* Device-level locality for GPU map slots.  The reference stamps the first free device id on whatever
 * task obtainNewNodeLocalMapTask returned (mapred/JobQueueTaskScheduler.java:350-361): it knows which NODE
 * holds a split's blocks, not which DEVICE of the node holds its rows in HBM.  With one HBM split cache per
 * device (hgm_cache_*, one hgm-worker per slot) iteration 2's task for a split then lands on another device
 * seven times out of eight and is staged again (measured: 4.9 vs 9.4 G points/s on eight slots,
 * profiles/r01_config3_job.md).  This is the pick one level below node locality, to be called with the
 * device id hgs_assign chose: home_device[i] is the device whose cache holds pending split i, or -1.
 * Order of preference: (1) a split this device holds; (2) a split no device holds (it has to be staged
 * somewhere); (3) a split of the device with the most pending resident splits -- that device would finish
 * last, so its queue is the one to shorten (lowest device id on a tie, lowest index within it).
 * Returns the index into home_device[], -1 if n_pending == 0, -22 on a bad argument. */
int32_t hgs_pick_split(const int32_t *home_device, int32_t n_pending, int32_t device, int32_t max_gpu_slots);

/* Discrete-event run of `iterations` K-means map phases of `n_splits` map tasks each on ONE
 * TaskTracker with the given slots, driven by hgs_assign at every heartbeat. */
typedef struct {
  int32_t n_splits, iterations;
  int32_t cpu_slots, gpu_slots;
  double cpu_task_ms;        /* one split on one CPU slot                              */
  double gpu_task_ms_first;  /* one split on a GPU slot, iteration 1 (staged from host) */
  double gpu_task_ms_cached; /* iterations >= 2 (split resident in HBM)                */
  double task_overhead_ms;   /* per-task launch cost (child JVM + Pipes child + handshake) */
  double heartbeat_ms;       /* TaskTracker heartbeat period                           */
  int32_t out_of_band_heartbeat; /* 1 = also heartbeat when a task finishes            */
  int32_t optional_scheduling, fixes;
  double prior_acceleration; /* handed to every job when HGS_FIX_PRIOR_ACCEL is set */
  int32_t per_device_cache;  /* 0: every split is resident from iteration 2 on, whichever slot maps it
                                (one cache per node); 1: a split is resident on the device that mapped it
                                last (one HBM cache per slot's worker) -- a task on another device takes
                                gpu_task_ms_first again and moves the split there */
  int32_t device_affinity;   /* with per_device_cache: 0 = next split to the first free device, as the
                                reference does; 1 = hgs_pick_split chooses the split for the device */
  double task_jitter;        /* 0..1: every task time is scaled by 1 +- task_jitter (a fixed hash of
                                iteration and task number, so runs repeat); without it equal task times
                                hand device j the same splits every iteration by accident */
} hgs_sim_params;

typedef struct {
  double makespan_ms;
  int64_t tasks_cpu, tasks_gpu, heartbeats, cpu_pass_skips;
  double last_acceleration;
  double points_per_sec;     /* filled by the caller's points-per-split if > 0 */
  int64_t gpu_cache_hits;    /* GPU tasks that found their split resident (per_device_cache only) */
} hgs_sim_result;

int hgs_simulate(const hgs_sim_params *params, hgs_sim_result *out);

#ifdef __cplusplus
}
#endif
#endif
