/* hgpumap.h -- C ABI of libhgpumap.so, the B200 (sm_100a) K-means map-task runtime.
 *
 * This is the drop-in boundary for the one path the millecker/hadoop-1.0.3-gpu fork
 * accelerates: the K-means *map* function (nearest-centroid assignment + combiner
 * partial sums) run as a GPU map task.  In the reference that path is reached through
 * C++ virtuals and a socket protocol, not a C ABI:
 *
 *   Java task  --PipesGPUMapRunner.run-->  child process  --HadoopPipes::runTask-->
 *   Mapper::map(MapContext&) per record, Reducer::reduce(ReduceContext&) as combiner
 *
 * so each entry point below names the reference interface it stands under
 * (paths relative to /root/reference/hadoop-1.0.3/src).  Two hosts bind this header:
 * the Pipes child `kmeans-pipes-gpu` (the -gpubin; csrc/pipes/kmeans_pipes_gpu.cc) and a JNI
 * stub inside the task JVM.
 *
 * Conventions: plain C, no C++ or torch types; every function returns 0 or a negative
 * HGM_E* code and records a message readable with hgm_last_error(); the caller owns
 * every host buffer; one hgm_ctx per GPU map slot, used by one thread at a time
 * (mapred/TaskTracker.java:2526-2562 gives a slot to one task at a time); there is NO
 * CPU fallback: without a usable sm_100 device hgm_open fails and the task must fail
 * (the scheduler re-runs the split, mapred/JobInProgress.java:3146-3159).
 */
#ifndef HGPUMAP_H
#define HGPUMAP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hgm_ctx hgm_ctx;

#define HGM_OK 0
#define HGM_EINVAL (-22)   /* bad argument or call order */
#define HGM_ENOMEM (-12)   /* host or device allocation failed */
#define HGM_ENODEV (-19)   /* no such device / not an sm_100 GPU */
#define HGM_EIO (-5)       /* file or stream error */
#define HGM_ECUDA (-1000)  /* CUDA runtime error; text in hgm_last_error */
#define HGM_ENCCL (-1001)  /* NCCL missing or failed */
#define HGM_EFORMAT (-1002)/* malformed input record */

/* Record formats of a split (JobConf kmeans.input.format; the reference's map input is
 * one std::string per record, c++/pipes/api/hadoop/Pipes.hh:89). */
#define HGM_FMT_F32LE 0 /* rows of d little-endian IEEE float32 (primary format)           */
#define HGM_FMT_F32BE 1 /* rows of d big-endian float32 (DataOutput.writeFloat order)       */
#define HGM_FMT_TEXT 2  /* TextInputFormat lines (LF, CR or CR LF ends), each a whole number */
                        /* of points; fields split at space, comma, tab as splitString does */
                        /* (c++/utils/impl/StringUtils.cc:80-95); a field means what        */
                        /* HadoopUtils::toFloat (:49-56, sscanf "%f") returns for it, bit   */
                        /* for bit, and what toFloat rejects fails the call (HGM_EFORMAT).  */
                        /* Decoded on the device.  hgm_map_file / hgm_cache_put_file frame  */
                        /* (offset, length) the way LineRecordReader does (mapred/          */
                        /* LineRecordReader.java:79-95,137-157): the first line belongs to  */
                        /* the previous split unless offset == 0, the last one is read to   */
                        /* its end.  Assignments are not returned for text (pass NULL).     */
#define HGM_FMT_SEQ_F32BE 3 /* an uncompressed version-6 SequenceFile (core/org/apache/hadoop/io/  */
#define HGM_FMT_SEQ_F32LE 4 /* SequenceFile.java) whose values are BytesWritable holding whole rows  */
                            /* of d float32, big-endian (DataOutput.writeFloat) or little-endian;    */
                            /* keys are skipped.  File entry points only (hgm_map_file,              */
                            /* hgm_cache_put_file): (offset, length) is framed the way               */
                            /* SequenceFileRecordReader does (mapred/SequenceFileRecordReader.java:  */
                            /* 40-52, 71-84): from the first sync entry at or after offset to the    */
                            /* first one at or after offset+length; records are walked on the device */
                            /* in parallel between sync entries.  Compressed files: HGM_EFORMAT.     */

/* Kernel families (hgm_set_kernel); AUTO picks by shape. */
#define HGM_KERNEL_AUTO 0
#define HGM_KERNEL_EXACT 1   /* brute-force FP32 scan, any d and k                          */
#define HGM_KERNEL_GRID2D 2  /* d == 2: exact candidate pruning through a cell index; the    */
                             /* combiner is the fixed-point form where it applies (k <= 254, */
                             /* coordinates within 2^+-100), else the counting-sort form     */
#define HGM_KERNEL_TENSOR 3  /* tcgen05 distance GEMM + FP32 re-check of near ties          */
#define HGM_KERNEL_GRID2D_SORT 4 /* d == 2 cell index with the counting-sort combiner, any k <= 1024 */

/* ---- library / device ------------------------------------------------------------- */

/* ABI version of this header (major*1000 + minor); 1001 = 1.1: entry points added since 1.0, none changed. */
int hgm_version(void);

/* Number of CUDA devices visible; the TaskTracker sizes
 * mapred.tasktracker.map.gpu.tasks.maximum from it (mapred/TaskTracker.java:1428-1433). */
int hgm_device_count(int *count);

/* Open the GPU map slot on `device_id`.  Stands for the device id the Java side appends
 * to the child's argv (mapred/pipes/Application.java:179-181; slot i <-> device i,
 * mapred/TaskTrackerStatus.java:536-551).  Creates the CUDA context, one stream and the
 * pinned staging ring once; keep the ctx alive across splits. */
int hgm_open(int device_id, hgm_ctx **out);
void hgm_close(hgm_ctx *ctx);

/* Last error text of `ctx`, or of the calling thread when ctx is NULL (hgm_open failed).
 * The Pipes adapter rethrows it as HadoopUtils::Error (c++/utils/api/hadoop/SerialUtils.hh:
 * 29-63) so Hadoop sees a failed task. */
const char *hgm_last_error(hgm_ctx *ctx);

/* ---- job state ---------------------------------------------------------------------- */

/* Install the k x d FP32 centroid table (row-major, host memory) for the following map
 * calls and zero the partial sums.  Stands for what a Mapper constructor reads from the
 * JobConf (c++/pipes/impl/HadoopPipes.cc:801 createMapper; JobConf access Pipes.hh:46-54).
 * Builds the kernel-side index (cell table for d == 2) on the device. */
int hgm_set_centroids(hgm_ctx *ctx, const float *centroids, int32_t k, int32_t d);

/* Force a kernel family (tests and profiling); HGM_EINVAL if the shape does not fit it.
 * Changing to or from HGM_KERNEL_GRID2D_SORT rebuilds the d == 2 index of the installed table
 * (the two combiners keep different table sizes in shared memory). */
int hgm_set_kernel(hgm_ctx *ctx, int32_t kernel);
/* Which family the last map call actually ran (one of HGM_KERNEL_*; HGM_KERNEL_GRID2D_SORT only
 * when that form was forced). */
int hgm_last_kernel(hgm_ctx *ctx, int32_t *kernel);
/* Which form of the tcgen05 score kernel runs (process-wide; parity tests of both forms and comparison runs):
 * AUTO picks by d (one CTA per 256-point tile for d <= 32, a cta_group::2 pair above). */
#define HGM_TENSOR_FORM_AUTO 0
#define HGM_TENSOR_FORM_SINGLE 1
#define HGM_TENSOR_FORM_PAIR 2
#define HGM_TENSOR_FORM_PAIR_FUSED 3      /* pair kernel, per-cluster sums always made inside it */
#define HGM_TENSOR_FORM_PAIR_SEGMENTED 4  /* pair kernel, per-cluster sums always by the segmented reduction (K4) */
int hgm_set_tensor_form(int32_t form);
/* Diagnostics build only (`make CHECKED=1`): how many shared-memory indices / centroid ids of the d == 2 kernels
 * missed their bound since the library was loaded, and the source line of the first.  HGM_EINVAL in the product build. */
int hgm_checked_failures(int32_t *count, int32_t *first_line);
/* d == 2 only: which combiner the installed index drives -- 0 none (no index), 1 fixed-point
 * (csrc/kernels_grid2d_fx.cu), 2 counting sort (csrc/kernels_grid2d.cu) -- and its table side G. */
int hgm_grid2d_form(hgm_ctx *ctx, int32_t *form, int32_t *cells_per_side);

/* ---- the map task ------------------------------------------------------------------- */

/* One map task over one input split held in HOST memory: stage to HBM through pinned
 * buffers (async, chunked, overlapped with the kernel), decode records on the device,
 * assign every point to its nearest centroid and leave the combiner's output.
 *   sums_kxd[k*d]  FP64 per-cluster coordinate sums      (required)
 *   counts_k[k]    int64 per-cluster point counts        (required)
 *   assign_or_null int32 cluster index per point, or NULL
 *   n_points_out   number of points decoded, or NULL
 * Replaces, for a whole split at once: Mapper::map per record (Pipes.hh:158-161, called
 * from HadoopPipes.cc:884-885), the CombineRunner buffer (HadoopPipes.cc:660-706) and the
 * combiner Reducer::reduce (Pipes.hh:166-169, called from HadoopPipes.cc:700-702).
 * Results are for THIS split only (the ctx partials are reset first). */
int hgm_map_split(hgm_ctx *ctx, const void *host_bytes, size_t nbytes, int32_t fmt,
                  double *sums_kxd, int64_t *counts_k, int32_t *assign_or_null,
                  int64_t *n_points_out);

/* Same, reading the split from a local file: the C++ RecordReader route
 * (hadoop.pipes.java.recordreader=false; HadoopPipes.cc:793-799, example
 * examples/pipes/impl/wordcount-nopipe.cc:67-113; FileSplit = path, start, length,
 * mapred/FileSplit.java:81-85).  length < 0 means "to end of file". */
int hgm_map_file(hgm_ctx *ctx, const char *path, int64_t offset, int64_t length, int32_t fmt,
                 double *sums_kxd, int64_t *counts_k, int32_t *assign_or_null,
                 int64_t *n_points_out);

/* ---- device-resident splits (iterations >= 2 of a job, and the kernel-only bench) ---- */

/* Map n points already in HBM (row-major FP32, 8-byte aligned at least) and ADD the
 * result into the ctx's device partials.  Asynchronous on the ctx stream.
 * d_assign_or_null is a device pointer to n int32, or NULL. */
int hgm_map_device(hgm_ctx *ctx, const float *d_points, int64_t n_points,
                   int32_t *d_assign_or_null);
/* Zero the device partials (asynchronous). */
int hgm_reset_partials(hgm_ctx *ctx);
/* Copy the device partials to host memory (synchronises the ctx stream). */
int hgm_read_partials(hgm_ctx *ctx, double *sums_kxd, int64_t *counts_k);
/* Device addresses of the partials, laid out [k*d doubles][k int64] contiguously, for a
 * caller that runs its own collective (torch.distributed) on them. */
int hgm_partials_device(hgm_ctx *ctx, double **d_sums, int64_t **d_counts);
/* The HBM-resident split cache: iteration 1 of a job stages each split once and keeps the
 * decoded rows on the device; iterations 2.. map them from HBM (BASELINE config 3: 10
 * iterations over the same 20 GB).  It is a cache only -- the unit of recovery stays the
 * split (the reference re-runs a failed map from its input, mapred/JobInProgress.java:
 * 3146-3159): a put that does not fit returns HGM_ENOMEM and the caller maps the split with
 * hgm_map_split instead; hgm_close drops everything, and so does hgm_set_centroids with a
 * different d.
 *   hgm_cache_put       stage + decode a split held in host memory, keep it under split_id
 *                       (replacing an older entry of that id)
 *   hgm_cache_put_file  same from a FileSplit byte range (mapred/FileSplit.java:81-85)
 *   hgm_cache_map       hgm_map_device on the cached rows: ADDS into the partials, async
 *   hgm_cache_map_all   map every cached split; splits that lie back to back in the cache's
 *                       arena (2 GiB blocks, filled in put order) share one launch, so an
 *                       iteration over a resident input costs a handful of launches
 *   hgm_cache_map_many  the same for a list of split ids (a task that names several splits); an id that is
 *                       listed twice is mapped twice
 *   hgm_cache_drop      forget one split, or all of them with split_id < 0 (an arena block
 *                       goes back to the driver when its last split is dropped)
 *   hgm_cache_info      number of cached splits and their bytes
 *   hgm_cache_read      copy a cached split's decoded rows (LE float32) back to the host:
 *                       rows_out holds cap_floats floats; *n_floats_out = floats the split
 *                       has (HGM_EINVAL, nothing copied, if that is more than cap_floats) */
int hgm_cache_put(hgm_ctx *ctx, int64_t split_id, const void *host_bytes, size_t nbytes, int32_t fmt);
int hgm_cache_put_file(hgm_ctx *ctx, int64_t split_id, const char *path, int64_t offset, int64_t length,
                       int32_t fmt);
int hgm_cache_map(hgm_ctx *ctx, int64_t split_id, int32_t *d_assign_or_null, int64_t *n_points_out);
int hgm_cache_map_all(hgm_ctx *ctx, int64_t *n_points_out, int32_t *n_launch_groups_out);
int hgm_cache_map_many(hgm_ctx *ctx, const int64_t *split_ids, int32_t n_ids, int64_t *n_points_out,
                       int32_t *n_launch_groups_out);
int hgm_cache_drop(hgm_ctx *ctx, int64_t split_id);
int hgm_cache_info(hgm_ctx *ctx, int64_t *n_splits, int64_t *bytes);
int hgm_cache_read(hgm_ctx *ctx, int64_t split_id, float *rows_out, int64_t cap_floats, int64_t *n_floats_out);

/* Wait for everything queued on the ctx stream. */
int hgm_sync(hgm_ctx *ctx);
/* The ctx's compute stream as a cudaStream_t, so a host that has its own CUDA binding can
 * order work or record events against the map kernels (bench.py times with CUDA events
 * on exactly this stream). */
int hgm_stream(hgm_ctx *ctx, void **cuda_stream);
/* Device time of the last map call's kernels in milliseconds (CUDA events on the ctx
 * stream; valid after hgm_sync / hgm_read_partials), and how many kernels it launched. */
int hgm_last_map_ms(hgm_ctx *ctx, float *ms, int32_t *launches);
/* How many points of the last map call on the tensor-core path had a top-2 score margin
 * inside the TF32 rounding bound and were re-scanned with the exact FP32 arithmetic
 * (0 for the other kernel families).  Synchronises the ctx stream. */
int hgm_last_rechecks(hgm_ctx *ctx, int64_t *points);
/* CUDA-event times (ms) of the last launch group of the tensor-core path, by phase: [0] scores + assignment on
 * tcgen05 (K2), [1] exact re-check of the flagged points (K3), [2] per-cluster sums by segmented reduction (K4).
 * Synchronises on the group's last event. */
int hgm_last_tensor_phases(hgm_ctx *ctx, float *ms3);
/* Bytes of input splits this ctx has copied host -> device since hgm_open (a split mapped from the HBM cache adds
 * nothing).  The Pipes program exports the per-task difference as counter KMEANS/GPU_BYTES_STAGED and the
 * re-checks as KMEANS/GPU_RECHECKS (mechanism: c++/pipes/impl/HadoopPipes.cc:346-359). */
int hgm_staged_bytes(hgm_ctx *ctx, int64_t *bytes);
/* Total kernels this ctx has launched since hgm_open. */
int hgm_launch_count(hgm_ctx *ctx, int64_t *launches);

/* The d == 2 map task for a caller that is not allowed to allocate device memory -- a slot
 * running inside an HBM budget someone else owns (a TaskTracker's per-slot arena, a JVM or
 * pool allocator that holds every byte), where cudaMalloc is off limits and its device-wide
 * synchronisation would stall the other slots.  It has no counterpart in the reference.
 * Nothing is allocated: the cell index, the centroid copy and the partial sums live in one
 * static device array of about 1.22 MiB that every GPU loading this library holds until the
 * process exits.  Limits: d must be 2, 1 <= k <= 1024, and the centroid table must be
 * indexable (finite, |coordinate| <= 1e18); otherwise HGM_EINVAL with an explanation.
 * d_points (n_points rows, 8-byte aligned) and d_assign_or_null are device pointers;
 * centroids (k*2 floats), sums_kxd and counts_k are host pointers; cuda_stream is a
 * cudaStream_t or NULL.  The call returns after the results are in the host arrays.
 * All calls on one GPU share that array, which keeps the previous call's contents (the call
 * clears what it accumulates into): calls that overlap on a GPU race -- run them one at a
 * time.  The first call on a device also looks up the array's device address; make that
 * first call before starting any stream capture.  Every call waits for cuda_stream before it
 * returns (sums_kxd and counts_k are host memory and are complete then) and leaves the caller's
 * current device as it found it. */
int hgm_map_device_static(int device_id, const float *d_points, int64_t n_points,
                          const float *centroids, int32_t k, int32_t d, double *sums_kxd,
                          int64_t *counts_k, int32_t *d_assign_or_null, void *cuda_stream);

/* Device / pinned-host memory for hosts that have no CUDA runtime of their own (JNI). */
int hgm_device_alloc(hgm_ctx *ctx, size_t nbytes, void **d_ptr);
int hgm_device_free(hgm_ctx *ctx, void *d_ptr);
int hgm_host_alloc(hgm_ctx *ctx, size_t nbytes, void **h_ptr);
int hgm_host_free(hgm_ctx *ctx, void *h_ptr);
int hgm_memcpy_h2d(hgm_ctx *ctx, void *d_dst, const void *h_src, size_t nbytes);
int hgm_memcpy_d2h(hgm_ctx *ctx, void *h_dst, const void *d_src, size_t nbytes);

/* ---- multi-GPU: one ctx per device, partials combined over NVLink -------------------- */

/* Replaces the intra-box part of the map->reduce shuffle of per-task partial records
 * (mapred/TaskTracker.java:4050 MapOutputServlet) by ONE sum-allreduce of the
 * [k*d] FP64 sums + [k] int64 counts across the GPU slots of a box.
 * hgm_comm_id fills 128 bytes on rank 0, which the host distributes (JobConf / file /
 * torch.distributed); every rank then calls hgm_comm_init and, per iteration,
 * hgm_allreduce_partials.  NCCL is loaded with dlopen at hgm_comm_id/hgm_comm_init time;
 * HGM_ENCCL if it is absent. */
int hgm_comm_id(void *id128);
int hgm_comm_init(hgm_ctx *ctx, int32_t n_ranks, int32_t rank, const void *id128);
int hgm_allreduce_partials(hgm_ctx *ctx);
/* Optional, for the small partials of the d == 2 shapes (up to 64 KiB: k*(d+1)*8 bytes), where NCCL's launch
 * latency (23-38 us) shows against a 0.45 ms map kernel: an exchange through peer memory.  Every rank calls
 * hgm_peer_handle (allocates its inbox, fills 64 bytes), the host gathers the n_ranks handles in rank order (the
 * way it distributes the NCCL id) and every rank calls hgm_peer_open with all of them.  From then on
 * hgm_allreduce_partials of partials that fit uses ONE one-CTA kernel on the ctx stream that stores the partials
 * into every rank's inbox over NVLink and adds the n_ranks slots up in rank order (bit-identical on every rank);
 * larger partials, or a ctx without hgm_peer_open, go through NCCL as before.  All ranks are processes of one box
 * with peer access between their GPUs (CUDA IPC); a failure to map a peer is reported and leaves the NCCL path. */
int hgm_peer_handle(hgm_ctx *ctx, int32_t n_ranks, void *handle64);
int hgm_peer_open(hgm_ctx *ctx, int32_t rank, const void *handles_n_x_64);
/* Back to NCCL for every size.  If hgm_peer_open failed on ANY rank, every rank must call this (the exchange is
 * collective: a rank left on the peer path would wait for the others until its kernel gives up). */
int hgm_peer_close(hgm_ctx *ctx);

/* ---- synthetic workload of the benchmark -------------------- */

/* Integer-hash mixture-of-k-blobs generator; bit-identical on host and device.
 * Not part of the reference (it publishes no workload); used by bench.py and tests. */
int hgm_synth_centres(uint64_t seed, int32_t k, int32_t d, float *centres_out);
int hgm_synth_perturb(uint64_t seed, int32_t k, int32_t d, const float *centres,
                      float *centroids_out);
int hgm_synth_points_host(uint64_t seed, int64_t first, int64_t n, int32_t d, int32_t k,
                          const float *centres, float *points_out);
int hgm_synth_points_device(hgm_ctx *ctx, uint64_t seed, int64_t first, int64_t n, int32_t d,
                            int32_t k, const float *centres_host, float *d_points_out);

#ifdef __cplusplus
}
#endif
#endif /* HGPUMAP_H */
