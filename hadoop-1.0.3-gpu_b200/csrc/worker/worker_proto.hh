// worker_proto.hh -- wire format between a short-lived map task (the -gpubin child that Hadoop forks per
// task, pipes/Application.java:156-194) and the persistent GPU worker of its slot (hgm-worker).
//
// Why: a Pipes child lives for one task; creating a CUDA context costs ~1 s, mapping a 64 MiB split
// 0.3 ms (profiles/r01_config1_run.md).  The worker owns the context, the
// pinned staging buffers and the HBM split cache of ONE device = one GPU map slot
// (mapred/TaskTrackerStatus.java:536-551: slot i <-> device i) and serves one task at a time over a
// unix-domain socket; the child keeps speaking the Pipes protocol to Java and forwards batches here.
// No counterpart in the reference (its children are expected to own the GPU themselves).
//
// All integers little-endian.  Request and reply share the header; a reply with status != 0 carries
// the error text of hgm_last_error as payload and the task fails (no CPU fallback).
#pragma once
#include <stdint.h>

namespace hgmw {

constexpr uint32_t kMagic = 0x574D4748u;  // "HGMW"
constexpr uint32_t kVersion = 2;  // 2: MapReply carries re-checks and staged bytes; kDropCache

enum Op : uint32_t {
  kHello = 1,         // payload: uint32 version                         -> reply payload: uint32 version, int32 device
  kSetCentroids = 2,  // payload: int32 k, int32 d, float[k*d]           -> empty
  kMapBytes = 3,      // payload: int32 fmt, bytes of the split/batch    -> MapReply + sums + counts
  kMapFile = 4,       // payload: int32 fmt, int32 use_cache, int64 offset, int64 length, path bytes -> same
  kStats = 5,         // empty                                           -> WorkerStats
  kShutdown = 6,      // empty                                           -> empty, then the worker exits
  kMapFiles = 8,      // payload: int32 fmt, int32 use_cache, int32 n, then n x (int64 offset, int64 length, int32 path
                      //   bytes, path) -> MapReply + sums + counts over all n splits; cache_hit = how many were resident
  kDropCache = 7,     // payload: epoch bytes (may be empty)                 -> empty; a non-empty epoch that differs from the
                      //   worker's current one, or an empty payload, drops every cached split (a new data set)
};

struct Header {
  uint32_t magic;
  uint32_t op;       // request: Op; reply: the request's Op
  int32_t status;    // request: 0; reply: 0 or an HGM_E* code
  uint32_t reserved;
  uint64_t payload;  // bytes that follow
};

struct MapReply {      // followed by k*d doubles (sums) and k int64 (counts)
  int64_t n_points;
  float map_ms;        // CUDA-event time of the map kernels
  int32_t launches;
  int32_t cache_hit;   // the split came from the HBM cache
  int32_t reserved;
  int64_t rechecks;      // points the tensor-core path re-scanned exactly (hgm_last_rechecks)
  int64_t staged_bytes;  // bytes this call copied host -> device (0 on a cache hit)
};

struct WorkerStats {
  int64_t tasks, map_calls, cache_hits, cache_misses, cached_splits, cached_bytes, kernel_launches, cache_evictions;
};

}  // namespace hgmw
