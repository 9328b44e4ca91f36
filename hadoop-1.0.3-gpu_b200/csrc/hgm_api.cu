// hgm_api.cu -- the C ABI of libhgpumap.so (include/hgpumap.h): context, staging, dispatch.
//
// One hgm_ctx is one GPU map slot (mapred/TaskTracker.java:2526-2562): a CUDA context on
// one device, a compute stream, two copy streams with pinned bounce buffers, the centroid
// table with its kernel-side index, and the device partials [k*d FP64 sums][k int64 counts].
// A split is streamed through HBM in chunks: while chunk i is being mapped, chunk i+1 is
// crossing PCIe; nothing is ever computed on the host.
#include <errno.h>
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include <unistd.h>

#include <algorithm>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "hgm_internal.h"

using namespace hgm;

namespace {
thread_local char g_err[512] = "";

constexpr size_t kChunkBytes = size_t(64) << 20;  // HBM chunk per pipeline stage
constexpr int kStages = 3;
constexpr size_t kBounceBytes = size_t(16) << 20;  // pinned bounce for pageable callers
constexpr size_t kArenaBytes = size_t(2) << 30;    // split-cache arena block
// Rows one launch group of the scan / tensor paths handles: their scratch (assignments, re-check list, cluster
// order: 12 bytes per row) is allocated ONCE per job at this size and longer inputs are mapped piece by piece --
// no cudaFree/cudaMalloc (a device-wide synchronisation under the other slots' kernels) inside a map call.
constexpr size_t kScratchRows = size_t(1) << 26;
}  // namespace

struct hgm_ctx {
  int device = -1;
  int sm_count = 0;
  size_t smem_optin = 0;
  cudaStream_t stream = nullptr;          // kernels
  cudaStream_t copy[kStages] = {};        // H2D / D2H
  cudaEvent_t ev_h2d[kStages] = {}, ev_free[kStages] = {}, ev_t0 = nullptr, ev_t1 = nullptr;
  cudaEvent_t ev_phase[4] = {};  // tensor path: before K2, after K2 (scores), after K3 (re-check), after K4 (sums)
  bool phases_open = false;
  void *d_chunk[kStages] = {};            // kChunkBytes each
  int32_t *d_assign_chunk[kStages] = {};  // lazily allocated
  size_t assign_chunk_cap = 0;
  void *h_bounce[kStages] = {};
  cudaEvent_t ev_bounce[kStages] = {};
  void *d_text[kStages] = {};             // staged text pieces (kBounceBytes each), lazily allocated
  unsigned int *d_tblocks[kStages] = {};  // field counts / offsets per 4 KiB of text
  TextStats *d_tstats[kStages] = {};
  TextStats *h_tstats = nullptr;          // pinned, kStages entries

  int k = 0, d = 0;
  std::vector<float> h_cent;    // host copy of the installed table (the d == 2 index is planned from it)
  float *d_cent = nullptr;      // k*d floats (+ pad to 16 B)
  void *d_partials = nullptr;   // [k*d double][k int64]
  bool grid_ok = false;
  Grid2D grid{};
  uint16_t *d_table = nullptr;  // G*G
  uint16_t *d_cand = nullptr;   // G*G*8
  Grid2D outer_grid{};          // the coarse index around the fine one (fixed-point form); G == 0: none
  uint16_t *d_otable = nullptr, *d_ocand = nullptr;
  bool tensor_ok = false;       // tcgen05 path prepared for this centroid table
  float *d_caug = nullptr;      // [k_pad][d+32] augmented centroid matrix
  unsigned int *d_tflags = nullptr;  // [0] max |c| bits, [1] re-check counter
  int32_t *d_tassign = nullptr, *d_tlist = nullptr;  // scratch: assignments / re-check list
  int32_t *d_torder = nullptr;       // point indices grouped by cluster (K4)
  unsigned int *d_tmeta = nullptr;   // [lcounts k][offsets k+1][cursor k][items k+1]
  size_t tscratch_rows = 0;
  unsigned long long last_rechecks = 0;
  int forced = HGM_KERNEL_AUTO, last_kernel = HGM_KERNEL_AUTO;
  int last_launches = 0;
  int64_t total_launches = 0;
  int64_t staged_bytes = 0;  // split bytes copied host -> device since hgm_open
  bool timing_open = false;
  void *nccl_comm = nullptr;
  void *peer_state = nullptr;  // hgm_comm.cu: the peer-memory exchange of small partials (hgm_peer_open)
  // HBM-resident splits by id.  Rows live in arena blocks (one cudaMalloc per kArenaBytes, not per
  // split) and are appended back to back, so that neighbours can be mapped by ONE launch
  // (hgm_cache_map_all); a block is returned to the driver when its last split is dropped.
  struct ArenaBlock { unsigned char *base; size_t cap, used; int live; };
  struct CachedSplit { void *d_rows; int64_t n_points; size_t nbytes; int block; };
  std::vector<ArenaBlock> arena;
  std::unordered_map<int64_t, CachedSplit> cache;
  char err[512] = "";
};

namespace {

int fail(hgm_ctx *ctx, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  snprintf(g_err, sizeof g_err, "%s", buf);
  if (ctx) snprintf(ctx->err, sizeof ctx->err, "%s", buf);
  return code;
}

#define CU(ctx, call)                                                                         \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(ctx, HGM_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),     \
                  __FILE__, __LINE__);                                                        \
  } while (0)

#define NEED_CTX(ctx)                                             \
  do {                                                            \
    if (!(ctx)) return fail(nullptr, HGM_EINVAL, "ctx is NULL");  \
    cudaError_t e_ = cudaSetDevice((ctx)->device);                \
    if (e_ != cudaSuccess)                                        \
      return fail(ctx, HGM_ECUDA, "cudaSetDevice(%d): %s", (ctx)->device, cudaGetErrorString(e_)); \
  } while (0)

double *sums_ptr(hgm_ctx *c) { return static_cast<double *>(c->d_partials); }
unsigned long long *counts_ptr(hgm_ctx *c) {
  return reinterpret_cast<unsigned long long *>(sums_ptr(c) + (size_t)c->k * c->d);
}
size_t partial_bytes(const hgm_ctx *c) { return (size_t)c->k * c->d * 8 + (size_t)c->k * 8; }

void free_job(hgm_ctx *c) {
  cudaFree(c->d_cent); c->d_cent = nullptr;
  cudaFree(c->d_partials); c->d_partials = nullptr;
  cudaFree(c->d_table); c->d_table = nullptr;
  cudaFree(c->d_cand); c->d_cand = nullptr;
  cudaFree(c->d_otable); c->d_otable = nullptr;
  cudaFree(c->d_ocand); c->d_ocand = nullptr;
  c->outer_grid = Grid2D{};
  cudaFree(c->d_caug); c->d_caug = nullptr;
  cudaFree(c->d_tflags); c->d_tflags = nullptr;
  cudaFree(c->d_tmeta); c->d_tmeta = nullptr;  // sized by k; the row scratch (d_tassign, d_tlist, d_torder) outlives jobs
  c->tensor_ok = false;
  c->grid_ok = false;
  c->k = c->d = 0;
  c->h_cent.clear();
}

// ---- split-cache arena ----------------------------------------------------------------------
void *arena_alloc(hgm_ctx *c, size_t nbytes, int *block_out, cudaError_t *err) {
  const size_t need = nbytes ? nbytes : 16;
  *err = cudaSuccess;
  for (int b = (int)c->arena.size() - 1; b >= 0; --b) {  // newest first: appends stay contiguous
    hgm_ctx::ArenaBlock &blk = c->arena[b];
    if (blk.base && blk.cap - blk.used >= need) {
      void *p = blk.base + blk.used;
      blk.used += (need + 15) & ~size_t(15);  // rows of whole 16-byte groups stay back to back
      if (blk.used > blk.cap) blk.used = blk.cap;
      blk.live += 1;
      *block_out = b;
      return p;
    }
    break;  // only the newest block is appended to
  }
  void *base = nullptr;
  size_t cap = std::max(kArenaBytes, need);
  cudaError_t e = cudaMalloc(&base, cap);
  if (e != cudaSuccess && cap > need) {  // HBM is nearly full: take what this split needs
    cudaGetLastError();
    cap = need;
    e = cudaMalloc(&base, cap);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    *err = e;
    return nullptr;
  }
  c->arena.push_back(hgm_ctx::ArenaBlock{static_cast<unsigned char *>(base), cap, std::min(cap, (need + 15) & ~size_t(15)), 1});
  *block_out = (int)c->arena.size() - 1;
  return base;
}

void arena_release(hgm_ctx *c, int block) {
  hgm_ctx::ArenaBlock &blk = c->arena[block];
  if (--blk.live == 0) {
    cudaFree(blk.base);
    blk.base = nullptr;
    blk.cap = blk.used = 0;
  }
}

void cache_clear(hgm_ctx *c) {
  for (auto &blk : c->arena) cudaFree(blk.base);
  c->arena.clear();
  c->cache.clear();
}

// Choose the d == 2 cell grid for a centroid table; false if the table is not indexable
// (non-finite or astronomically large coordinates) and the exact scan must be used.
bool plan_grid2d(const float *cent, int k, size_t smem_optin, bool sort_form, Grid2D *out, Grid2D *outer_out = nullptr,
                 size_t table_cap = 0) {
  if (k > 1024) return false;
  double lo[2] = {INFINITY, INFINITY}, hi[2] = {-INFINITY, -INFINITY};
  for (int c = 0; c < k; ++c)
    for (int j = 0; j < 2; ++j) {
      const double v = cent[2 * c + j];
      if (!std::isfinite(v) || fabs(v) > 1e18) return false;
      lo[j] = std::min(lo[j], v);
      hi[j] = std::max(hi[j], v);
    }
  float s[2], B[2], fs[2] = {0.f, 0.f};
  int32_t b[2];
  // geometry of a G x G grid over the centroids' bounding box + 12.5 % margin; fs = the fixed-point
  // scale of kernels_grid2d_fx.cu for that grid (0 where the coordinates' magnitude rules it out)
  auto geometry = [&](int G, double margin_frac) -> bool {
    for (int j = 0; j < 2; ++j) {
      double ext = hi[j] - lo[j];
      const double mag = std::max(fabs(lo[j]), fabs(hi[j]));
      if (!(ext > mag * 1e-4)) ext = std::max(mag * 1e-4, 1e-30);  // degenerate axis
      const double margin = margin_frac * ext;
      const double origin = lo[j] - margin;
      const double scale = (double)(G - 1) / (ext + 2 * margin);
      const float sf = (float)scale;
      if (!(sf > 0.f) || !std::isfinite(sf)) return false;
      const double off = -floor(origin * (double)sf + 0.5);  // RN(origin*s) + b == 0
      if (fabs(off) > 4.0e6) return false;                   // keep b + 2^23 exact in FP32
      s[j] = sf;
      b[j] = (int32_t)off;
      B[j] = (float)(off + 8388608.0);
      // a point of an interior cell ix in [0, G-2] has |x*s + b - ix| <= 1/2 (one FMA, exponent of 2^23):
      // |x| <= reach.  The scale puts 2^21 above it.
      const double reach = std::max(fabs(off) + 0.5, fabs((double)G - 1.5 - off)) / (double)sf * 1.000001;
      const int e = ilogb(reach) + 1;  // 2^e > reach
      fs[j] = (reach > 0 && e >= -100 && e <= 100) ? (float)ldexp(1.0, 21 - e) : 0.f;
    }
    return true;
  };
  if (outer_out) *outer_out = Grid2D{};
  int G = sort_form ? 0 : grid2d_fx_table_size(k, smem_optin, table_cap);
  int shape = kShapeFx;
  if (G > 0 && outer_out && geometry(kOuterCells, 1.5)) {  // the coarse index: four times the box on each axis
    outer_out->sx = s[0]; outer_out->sy = s[1];
    outer_out->Bx = B[0]; outer_out->By = B[1];
    outer_out->bx = b[0]; outer_out->by = b[1];
    outer_out->G = kOuterCells;
  }
  if (G > 0 && !(geometry(G, 0.125) && fs[0] > 0.f && fs[1] > 0.f)) G = 0;
  if (G == 0) {  // the counting-sort form
    if (outer_out) *outer_out = Grid2D{};
    if (!grid2d_plan(k, smem_optin, &G, &shape) || !geometry(G, 0.125)) return false;
    fs[0] = fs[1] = 0.f;
  }
  out->sx = s[0]; out->sy = s[1];
  out->Bx = B[0]; out->By = B[1];
  out->bx = b[0]; out->by = b[1];
  out->G = G;
  out->shape = shape;
  out->fsx = fs[0]; out->fsy = fs[1];
  return true;
}

// (re)build the d == 2 cell index of the installed table for the form the ctx is set to
int build_grid_index(hgm_ctx *c);

int pick_kernel(hgm_ctx *c) {
  if (c->forced == HGM_KERNEL_EXACT) return HGM_KERNEL_EXACT;
  if (c->forced == HGM_KERNEL_GRID2D) return c->grid_ok ? HGM_KERNEL_GRID2D : -1;
  if (c->forced == HGM_KERNEL_GRID2D_SORT) return c->grid_ok && c->grid.shape != kShapeFx ? HGM_KERNEL_GRID2D_SORT : -1;
  if (c->forced == HGM_KERNEL_TENSOR) return c->tensor_ok ? HGM_KERNEL_TENSOR : -1;
  if (c->grid_ok) return HGM_KERNEL_GRID2D;
  // the GEMM-shaped path pays off once a point meets a few thousand centroid coordinates
  // below these the scan is cheaper than a padded tile (measured, 50 M points: d=16/k=16 12.5 ms tensor vs 16.7 exact --
  // the generic-d scan pays d FP64 atomics per point; d=8/k=64 15.6 vs 6.6 -- the register scan does not)
  if (c->tensor_ok && (long long)c->k * c->d >= (c->d > 8 ? 256 : 2048)) return HGM_KERNEL_TENSOR;
  return HGM_KERNEL_EXACT;
}

// scratch of the scan + segmented-sum paths (assignments, re-check list, cluster order), once per job
int ensure_scan_scratch(hgm_ctx *c) {
  if (c->tscratch_rows == 0) {
    cudaError_t e = cudaMalloc(&c->d_tlist, kScratchRows * 4);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_tassign, kScratchRows * 4);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_torder, kScratchRows * 8);  // 8: the re-check keeps a 64-bit key per flagged row there
    if (e != cudaSuccess) {
      cudaGetLastError();
      cudaFree(c->d_tlist); cudaFree(c->d_tassign); cudaFree(c->d_torder);
      c->d_tlist = c->d_tassign = c->d_torder = nullptr;
      return fail(c, e == cudaErrorMemoryAllocation ? HGM_ENOMEM : HGM_ECUDA, "scratch for the scan path: %s", cudaGetErrorString(e));
    }
    c->tscratch_rows = kScratchRows;
  }
  if (!c->d_tmeta) CU(c, cudaMalloc(&c->d_tmeta, tensor_meta_words(c->k) * 4));
  return HGM_OK;
}

int map_piece_async(hgm_ctx *c, int kern, const float *d_pts, int64_t n, int32_t *d_assign);

// d_tflags: [0] bits of max |c|, [1] points the last launch group re-checked, [2..3] the same summed over a map call
__global__ void add_rechecks_kernel(unsigned int *flags) {
  *reinterpret_cast<unsigned long long *>(flags + 2) += flags[1];
}

// queue the map of n device-resident points on ctx->stream; adds into the partials
int map_points_async(hgm_ctx *c, const float *d_pts, int64_t n, int32_t *d_assign) {
  if (n == 0) return HGM_OK;
  const int kern = pick_kernel(c);
  if (kern < 0) return fail(c, HGM_EINVAL, "forced kernel %d does not fit k=%d d=%d", c->forced, c->k, c->d);
  if (kern == HGM_KERNEL_GRID2D || kern == HGM_KERNEL_GRID2D_SORT) return map_piece_async(c, kern, d_pts, n, d_assign);
  const int rc = ensure_scan_scratch(c);  // a no-op after hgm_set_centroids unless the kernel was forced later
  if (rc != HGM_OK) return rc;
  for (int64_t first = 0; first < n; first += (int64_t)kScratchRows) {
    const int64_t m = std::min<int64_t>((int64_t)kScratchRows, n - first);
    const int prc = map_piece_async(c, kern, d_pts + (size_t)first * c->d, m, d_assign ? d_assign + first : nullptr);
    if (prc != HGM_OK) return prc;
  }
  return HGM_OK;
}

int map_piece_async(hgm_ctx *c, int kern, const float *d_pts, int64_t n, int32_t *d_assign) {
  cudaError_t e;
  int launches = 1;
  if (kern == HGM_KERNEL_TENSOR) {
    if (reinterpret_cast<uintptr_t>(d_pts) & 15)
      return fail(c, HGM_EINVAL, "the tensor path needs 16-byte aligned points");
    e = launch_tensor_map(d_pts, n, c->d, c->d_cent, c->k, c->d_caug, c->d_tflags, d_assign ? d_assign : c->d_tassign,
                          c->d_tlist, c->d_tflags + 1, c->d_tmeta, c->d_torder, sums_ptr(c), counts_ptr(c), c->sm_count,
                          c->smem_optin, c->stream, c->ev_phase, &launches);
    c->phases_open = e == cudaSuccess;
    if (e == cudaSuccess) {
      add_rechecks_kernel<<<1, 1, 0, c->stream>>>(c->d_tflags);
      e = cudaGetLastError();
    }
    launches += 1;  // add_rechecks_kernel
  } else if (kern == HGM_KERNEL_GRID2D || kern == HGM_KERNEL_GRID2D_SORT) {
    const Grid2DOuter outer{c->outer_grid, c->d_otable, c->d_ocand};
    e = launch_grid2d_map(d_pts, n, c->d_cent, c->k, c->grid, c->d_table, c->d_cand, sums_ptr(c),
                          counts_ptr(c), d_assign, c->sm_count, c->smem_optin, c->stream, &outer);
  } else {
    e = launch_exact_map(d_pts, n, c->d, c->d_cent, c->k, d_assign ? d_assign : c->d_tassign, c->d_tmeta, c->d_torder,
                         sums_ptr(c), counts_ptr(c), c->sm_count, c->smem_optin, c->stream);
    launches = 4 + (c->d - 1) / 128;
  }
  if (e != cudaSuccess) return fail(c, HGM_ECUDA, "map kernel launch failed: %s", cudaGetErrorString(e));
  c->last_kernel = kern;
  c->last_launches += launches;
  c->total_launches += launches;
  return HGM_OK;
}

int build_grid_index(hgm_ctx *c) {
  c->grid_ok = false;
  if (c->d != 2) return HGM_OK;
  Grid2D g;
  Grid2D outer;
  if (!plan_grid2d(c->h_cent.data(), c->k, c->smem_optin, c->forced == HGM_KERNEL_GRID2D_SORT, &g, &outer)) return HGM_OK;
  if (!c->d_table || g.G != c->grid.G) {
    cudaFree(c->d_table); cudaFree(c->d_cand);
    c->d_table = nullptr; c->d_cand = nullptr;
    CU(c, cudaMalloc(&c->d_table, grid2d_table_bytes(g.G)));
    CU(c, cudaMalloc(&c->d_cand, (size_t)g.G * g.G * kMaxCand * 2));
  }
  c->grid = g;
  cudaError_t e = launch_grid2d_build(c->d_cent, c->k, g, c->d_table, c->d_cand, c->stream);
  if (e != cudaSuccess) return fail(c, HGM_ECUDA, "grid2d build launch: %s", cudaGetErrorString(e));
  c->total_launches += 1;
  c->outer_grid = outer;
  if (outer.G != 0) {
    if (!c->d_otable) {
      CU(c, cudaMalloc(&c->d_otable, grid2d_table_bytes(kOuterCells)));
      CU(c, cudaMalloc(&c->d_ocand, (size_t)kOuterCells * kOuterCells * kMaxCand * 2));
    }
    e = launch_grid2d_build(c->d_cent, c->k, outer, c->d_otable, c->d_ocand, c->stream);
    if (e != cudaSuccess) return fail(c, HGM_ECUDA, "grid2d outer build launch: %s", cudaGetErrorString(e));
    c->total_launches += 1;
  }
  c->grid_ok = true;
  return HGM_OK;
}

bool is_pinned(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

}  // namespace

namespace hgm {
void comm_destroy(void *comm);
void peer_destroy(void *state);
void **ctx_peer_slot(hgm_ctx *c) { return &c->peer_state; }
void ctx_count_launches(hgm_ctx *c, int n) { c->last_launches += n; c->total_launches += n; }
int ctx_shape(hgm_ctx *c, int *device, int *k, int *d, cudaStream_t *stream, void ***comm_slot) {
  *device = c->device; *k = c->k; *d = c->d; *stream = c->stream; *comm_slot = &c->nccl_comm;
  return 0;
}
int ctx_fail(hgm_ctx *c, int code, const char *msg) { return fail(c, code, "%s", msg); }
// for entry points that live in other translation units and have no ctx
int report_error(int code, const char *msg) { return fail(nullptr, code, "%s", msg); }
bool plan_grid2d_for(const float *cent, int k, size_t smem_optin, Grid2D *out, size_t table_cap) {
  return plan_grid2d(cent, k, smem_optin, false, out, nullptr, table_cap);
}
}  // namespace hgm

extern "C" {

int hgm_version(void) { return 1001; }  // 1.1: + peer exchange, grid2d form / tensor form, phases, staged bytes

int hgm_device_count(int *count) {
  if (!count) return fail(nullptr, HGM_EINVAL, "count is NULL");
  CU(nullptr, cudaGetDeviceCount(count));
  return HGM_OK;
}

const char *hgm_last_error(hgm_ctx *ctx) { return ctx ? ctx->err : g_err; }

int hgm_open(int device_id, hgm_ctx **out) {
  HGM_RANGE("hgm_open");
  if (!out) return fail(nullptr, HGM_EINVAL, "out is NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, HGM_ENODEV, "no CUDA device: %s (there is no CPU fallback in a GPU map slot)",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device_id < 0 || device_id >= ndev)
    return fail(nullptr, HGM_ENODEV, "device %d out of range (have %d)", device_id, ndev);
  cudaDeviceProp prop;
  CU(nullptr, cudaGetDeviceProperties(&prop, device_id));
  if (prop.major != 10)
    return fail(nullptr, HGM_ENODEV, "device %d is sm_%d%d; this library carries sm_100a code only",
                device_id, prop.major, prop.minor);
  CU(nullptr, cudaSetDevice(device_id));
  hgm_ctx *c = new (std::nothrow) hgm_ctx();
  if (!c) return fail(nullptr, HGM_ENOMEM, "out of host memory");
  c->device = device_id;
  c->sm_count = prop.multiProcessorCount;
  c->smem_optin = prop.sharedMemPerBlockOptin;
#define CUO(call)                                                                            \
  do {                                                                                       \
    cudaError_t e2_ = (call);                                                                \
    if (e2_ != cudaSuccess) {                                                                \
      fail(nullptr, HGM_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e2_));             \
      hgm_close(c);                                                                          \
      return HGM_ECUDA;                                                                      \
    }                                                                                        \
  } while (0)
  CUO(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CUO(cudaEventCreate(&c->ev_t0));
  CUO(cudaEventCreate(&c->ev_t1));
  for (int i = 0; i < 4; ++i) CUO(cudaEventCreate(&c->ev_phase[i]));
  for (int i = 0; i < kStages; ++i) {
    CUO(cudaStreamCreateWithFlags(&c->copy[i], cudaStreamNonBlocking));
    CUO(cudaEventCreateWithFlags(&c->ev_h2d[i], cudaEventDisableTiming));
    CUO(cudaEventCreateWithFlags(&c->ev_free[i], cudaEventDisableTiming));
    CUO(cudaEventCreateWithFlags(&c->ev_bounce[i], cudaEventDisableTiming));
  }
#undef CUO
  *out = c;
  return HGM_OK;
}

void hgm_close(hgm_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->peer_state) { hgm::peer_destroy(c->peer_state); c->peer_state = nullptr; }
  if (c->nccl_comm) { hgm::comm_destroy(c->nccl_comm); c->nccl_comm = nullptr; }
  cache_clear(c);
  free_job(c);
  cudaFree(c->d_tassign); cudaFree(c->d_tlist); cudaFree(c->d_torder);
  for (int i = 0; i < kStages; ++i) {
    if (c->copy[i]) { cudaStreamSynchronize(c->copy[i]); cudaStreamDestroy(c->copy[i]); }
    if (c->ev_h2d[i]) cudaEventDestroy(c->ev_h2d[i]);
    if (c->ev_free[i]) cudaEventDestroy(c->ev_free[i]);
    if (c->ev_bounce[i]) cudaEventDestroy(c->ev_bounce[i]);
    cudaFree(c->d_chunk[i]);
    cudaFree(c->d_assign_chunk[i]);
    if (c->h_bounce[i]) cudaFreeHost(c->h_bounce[i]);
    cudaFree(c->d_text[i]);
    cudaFree(c->d_tblocks[i]);
    cudaFree(c->d_tstats[i]);
  }
  if (c->h_tstats) cudaFreeHost(c->h_tstats);
  if (c->ev_t0) cudaEventDestroy(c->ev_t0);
  if (c->ev_t1) cudaEventDestroy(c->ev_t1);
  for (int i = 0; i < 4; ++i) if (c->ev_phase[i]) cudaEventDestroy(c->ev_phase[i]);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int hgm_set_centroids(hgm_ctx *c, const float *centroids, int32_t k, int32_t d) {
  HGM_RANGE("hgm_set_centroids");
  NEED_CTX(c);
  if (!centroids || k <= 0 || d <= 0) return fail(c, HGM_EINVAL, "need centroids, k > 0, d > 0");
  if ((int64_t)k * d > (int64_t(1) << 28)) return fail(c, HGM_EINVAL, "k*d too large");
  CU(c, cudaStreamSynchronize(c->stream));
  if (k != c->k || d != c->d) {
    if (d != c->d) cache_clear(c);  // cached rows are rows of the old d
    free_job(c);
    const size_t cent_bytes = (((size_t)k * d * 4) + 15) & ~size_t(15);
    // the shape is recorded only once both buffers exist: a failed allocation leaves the ctx without a job
    cudaError_t ea = cudaMalloc(&c->d_cent, cent_bytes);
    if (ea == cudaSuccess) ea = cudaMalloc(&c->d_partials, (size_t)k * d * 8 + (size_t)k * 8);
    if (ea != cudaSuccess) {
      cudaGetLastError();
      free_job(c);
      return fail(c, ea == cudaErrorMemoryAllocation ? HGM_ENOMEM : HGM_ECUDA, "device memory for k=%d d=%d: %s", k, d,
                  cudaGetErrorString(ea));
    }
    c->k = k; c->d = d;
    CU(c, cudaMemsetAsync(c->d_cent, 0, cent_bytes, c->stream));
  }
  CU(c, cudaMemcpyAsync(c->d_cent, centroids, (size_t)k * d * 4, cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaMemsetAsync(c->d_partials, 0, partial_bytes(c), c->stream));
  c->h_cent.assign(centroids, centroids + (size_t)k * d);
  {
    const int rc = build_grid_index(c);
    if (rc != HGM_OK) return rc;
  }
  c->tensor_ok = false;
  if (tensor_shape_ok(k, d, c->smem_optin)) {
    if (!c->d_caug) {
      CU(c, cudaMalloc(&c->d_caug, tensor_caug_floats(k, d) * 4));
      CU(c, cudaMalloc(&c->d_tflags, 16));
    }
    cudaError_t e = launch_tensor_prep(c->d_cent, k, d, c->d_caug, c->d_tflags, c->stream);
    if (e != cudaSuccess) return fail(c, HGM_ECUDA, "tensor prep launch: %s", cudaGetErrorString(e));
    c->total_launches += 1;
    c->tensor_ok = true;
  }
  if (!c->grid_ok || c->forced == HGM_KERNEL_EXACT || c->forced == HGM_KERNEL_TENSOR) {
    const int rc = ensure_scan_scratch(c);
    if (rc != HGM_OK) return rc;
  }
  // the caller's centroid buffer may be reused as soon as we return
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}

int hgm_set_kernel(hgm_ctx *c, int32_t kernel) {
  NEED_CTX(c);
  if (kernel < HGM_KERNEL_AUTO || kernel > HGM_KERNEL_GRID2D_SORT) return fail(c, HGM_EINVAL, "unknown kernel %d", kernel);
  const bool was_sort = c->forced == HGM_KERNEL_GRID2D_SORT;
  c->forced = kernel;
  if (was_sort != (kernel == HGM_KERNEL_GRID2D_SORT) && c->d == 2 && !c->h_cent.empty()) {
    CU(c, cudaStreamSynchronize(c->stream));  // maps in flight still read the old index
    const int rc = build_grid_index(c);
    if (rc != HGM_OK) return rc;
    CU(c, cudaStreamSynchronize(c->stream));
  }
  return HGM_OK;
}

// Bound misses counted by a `make CHECKED=1` build of the d == 2 kernels (always 0, line 0 in the product build).
int hgm_checked_failures(int32_t *count, int32_t *first_line) {
  unsigned int a[2] = {0, 0}, b[2] = {0, 0};
#ifdef HGM_CHECKED
  cudaDeviceSynchronize();
  grid2d_checked_failures(a);
  grid2d_fx_checked_failures(b);
#endif
  if (count) *count = (int32_t)(a[0] + b[0]);
  if (first_line) *first_line = (int32_t)(a[0] ? a[1] : b[1]);
#ifdef HGM_CHECKED
  return HGM_OK;
#else
  return fail(nullptr, HGM_EINVAL, "this library was not built with CHECKED=1");
#endif
}

int hgm_set_tensor_form(int32_t form) {
  if (form < HGM_TENSOR_FORM_AUTO || form > HGM_TENSOR_FORM_PAIR_SEGMENTED) return fail(nullptr, HGM_EINVAL, "unknown tensor form %d", form);
  tensor_set_form(form);
  return HGM_OK;
}

int hgm_grid2d_form(hgm_ctx *c, int32_t *form, int32_t *cells_per_side) {
  if (!c || !form) return fail(c, HGM_EINVAL, "NULL argument");
  *form = !c->grid_ok ? 0 : (c->grid.shape == kShapeFx ? 1 : 2);
  if (cells_per_side) *cells_per_side = c->grid_ok ? c->grid.G : 0;
  return HGM_OK;
}

int hgm_last_kernel(hgm_ctx *c, int32_t *kernel) {
  if (!c || !kernel) return fail(c, HGM_EINVAL, "NULL argument");
  *kernel = c->last_kernel;
  return HGM_OK;
}

int hgm_reset_partials(hgm_ctx *c) {
  NEED_CTX(c);
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  CU(c, cudaMemsetAsync(c->d_partials, 0, partial_bytes(c), c->stream));
  return HGM_OK;
}

int hgm_map_device(hgm_ctx *c, const float *d_points, int64_t n, int32_t *d_assign) {
  HGM_RANGE("hgm_map_device");
  NEED_CTX(c);
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (n < 0 || (n > 0 && !d_points)) return fail(c, HGM_EINVAL, "bad points argument");
  if (reinterpret_cast<uintptr_t>(d_points) & 7) return fail(c, HGM_EINVAL, "d_points must be 8-byte aligned");
  c->last_launches = 0;
  if (c->d_tflags) cudaMemsetAsync(c->d_tflags + 2, 0, 8, c->stream);  // re-check total of this map call
  CU(c, cudaEventRecord(c->ev_t0, c->stream));
  const int rc = map_points_async(c, d_points, n, d_assign);
  if (rc != HGM_OK) return rc;
  CU(c, cudaEventRecord(c->ev_t1, c->stream));
  c->timing_open = true;
  return HGM_OK;
}

int hgm_read_partials(hgm_ctx *c, double *sums, int64_t *counts) {
  HGM_RANGE("hgm_read_partials");
  NEED_CTX(c);
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (!sums || !counts) return fail(c, HGM_EINVAL, "sums and counts are required");
  CU(c, cudaMemcpyAsync(sums, sums_ptr(c), (size_t)c->k * c->d * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(counts, counts_ptr(c), (size_t)c->k * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}

int hgm_partials_device(hgm_ctx *c, double **d_sums, int64_t **d_counts) {
  if (!c || !c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (d_sums) *d_sums = sums_ptr(c);
  if (d_counts) *d_counts = reinterpret_cast<int64_t *>(counts_ptr(c));
  return HGM_OK;
}

int hgm_sync(hgm_ctx *c) {
  NEED_CTX(c);
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}

int hgm_stream(hgm_ctx *c, void **cuda_stream) {
  if (!c || !cuda_stream) return fail(c, HGM_EINVAL, "NULL argument");
  *cuda_stream = c->stream;
  return HGM_OK;
}

int hgm_last_map_ms(hgm_ctx *c, float *ms, int32_t *launches) {
  NEED_CTX(c);
  if (!c->timing_open) return fail(c, HGM_EINVAL, "no map call has been timed yet");
  CU(c, cudaEventSynchronize(c->ev_t1));
  if (ms) CU(c, cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
  if (launches) *launches = c->last_launches;
  return HGM_OK;
}

int hgm_last_rechecks(hgm_ctx *c, int64_t *points) {
  NEED_CTX(c);
  if (!points) return fail(c, HGM_EINVAL, "points is NULL");
  *points = 0;
  if (c->last_kernel != HGM_KERNEL_TENSOR || !c->d_tflags) return HGM_OK;
  unsigned long long n = 0;
  CU(c, cudaMemcpyAsync(&n, c->d_tflags + 2, 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  *points = (int64_t)n;
  return HGM_OK;
}

int hgm_last_tensor_phases(hgm_ctx *c, float *ms3) {
  NEED_CTX(c);
  if (!ms3) return fail(c, HGM_EINVAL, "ms3 is NULL");
  if (!c->phases_open) return fail(c, HGM_EINVAL, "the tensor-core path has not run on this ctx");
  CU(c, cudaEventSynchronize(c->ev_phase[3]));
  for (int i = 0; i < 3; ++i) CU(c, cudaEventElapsedTime(&ms3[i], c->ev_phase[i], c->ev_phase[i + 1]));
  return HGM_OK;
}

int hgm_staged_bytes(hgm_ctx *c, int64_t *bytes) {
  if (!c || !bytes) return fail(c, HGM_EINVAL, "NULL argument");
  *bytes = c->staged_bytes;
  return HGM_OK;
}

int hgm_launch_count(hgm_ctx *c, int64_t *launches) {
  if (!c || !launches) return fail(c, HGM_EINVAL, "NULL argument");
  *launches = c->total_launches;
  return HGM_OK;
}

int hgm_device_alloc(hgm_ctx *c, size_t nbytes, void **d_ptr) {
  NEED_CTX(c);
  if (!d_ptr) return fail(c, HGM_EINVAL, "d_ptr is NULL");
  cudaError_t e = cudaMalloc(d_ptr, nbytes ? nbytes : 1);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(c, HGM_ENOMEM, "cudaMalloc(%zu): %s", nbytes, cudaGetErrorString(e)); }
  return HGM_OK;
}
int hgm_device_free(hgm_ctx *c, void *d_ptr) {
  NEED_CTX(c);
  CU(c, cudaStreamSynchronize(c->stream));
  CU(c, cudaFree(d_ptr));
  return HGM_OK;
}
int hgm_host_alloc(hgm_ctx *c, size_t nbytes, void **h_ptr) {
  NEED_CTX(c);
  if (!h_ptr) return fail(c, HGM_EINVAL, "h_ptr is NULL");
  cudaError_t e = cudaHostAlloc(h_ptr, nbytes ? nbytes : 1, cudaHostAllocDefault);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(c, HGM_ENOMEM, "cudaHostAlloc(%zu): %s", nbytes, cudaGetErrorString(e)); }
  return HGM_OK;
}
int hgm_host_free(hgm_ctx *c, void *h_ptr) {
  NEED_CTX(c);
  CU(c, cudaFreeHost(h_ptr));
  return HGM_OK;
}
int hgm_memcpy_h2d(hgm_ctx *c, void *d_dst, const void *h_src, size_t nbytes) {
  NEED_CTX(c);
  CU(c, cudaMemcpyAsync(d_dst, h_src, nbytes, cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}
int hgm_memcpy_d2h(hgm_ctx *c, void *h_dst, const void *d_src, size_t nbytes) {
  NEED_CTX(c);
  CU(c, cudaMemcpyAsync(h_dst, d_src, nbytes, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}

// ---- the streamed map task ---------------------------------------------------------------

static int ensure_pipeline(hgm_ctx *c, bool want_assign, size_t rows_per_chunk, bool want_bounce) {
  for (int i = 0; i < kStages; ++i) {
    if (!c->d_chunk[i]) CU(c, cudaMalloc(&c->d_chunk[i], kChunkBytes));
    if (want_bounce && !c->h_bounce[i]) CU(c, cudaHostAlloc(&c->h_bounce[i], kBounceBytes, cudaHostAllocDefault));
  }
  if (want_assign && c->assign_chunk_cap < rows_per_chunk) {
    for (int i = 0; i < kStages; ++i) {
      cudaFree(c->d_assign_chunk[i]);
      c->d_assign_chunk[i] = nullptr;
      CU(c, cudaMalloc(&c->d_assign_chunk[i], rows_per_chunk * 4));
    }
    c->assign_chunk_cap = rows_per_chunk;
  }
  return HGM_OK;
}

// Byte source of a split: host memory or a file region read through the bounce buffers.
struct SplitSource {
  const unsigned char *mem = nullptr;
  FILE *file = nullptr;
  int64_t file_base = 0;  // file offset of the split's first byte
  size_t nbytes = 0;
  bool pinned = false;
};

// Host side of the staging: bytes [src_off, src_off + nbytes) of the split into a pinned bounce buffer.
// One thread copies (memcpy, or pread from the page cache) at ~10 GB/s, a fifth of what the H2D link
// takes; slices of a piece are therefore filled by a few threads at once (HGM_COPY_THREADS, default 8).
static int copy_threads() {
  static const int n = [] {
    const char *v = getenv("HGM_COPY_THREADS");
    int t = v ? atoi(v) : 8;
    const int hw = (int)std::thread::hardware_concurrency();
    if (hw > 0 && t > hw) t = hw;
    return t < 1 ? 1 : (t > 16 ? 16 : t);
  }();
  return n;
}
static bool stage_bytes(const SplitSource &src, size_t src_off, void *dst, size_t nbytes) {
  auto slice = [&](size_t lo, size_t hi, bool *ok) {
    unsigned char *out = static_cast<unsigned char *>(dst);
    if (!src.file) {
      memcpy(out + lo, src.mem + src_off + lo, hi - lo);
      return;
    }
    const int fd = fileno(src.file);
    while (lo < hi) {
      const ssize_t got = pread(fd, out + lo, hi - lo, (off_t)(src.file_base + (int64_t)(src_off + lo)));
      if (got <= 0) {
        if (got < 0 && errno == EINTR) continue;
        *ok = false;  // short file or I/O error
        return;
      }
      lo += (size_t)got;
    }
  };
  const int threads = nbytes >= (size_t(2) << 20) ? copy_threads() : 1;
  bool ok[16];
  for (int t = 0; t < threads; ++t) ok[t] = true;
  if (threads == 1) {
    slice(0, nbytes, &ok[0]);
    return ok[0];
  }
  std::thread pool[15];
  const size_t step = ((nbytes / threads) + 4095) & ~size_t(4095);
  int started = 0;
  for (int t = 1; t < threads; ++t) {
    const size_t lo = std::min(nbytes, step * t), hi = std::min(nbytes, step * (t + 1));
    if (lo < hi) pool[started++] = std::thread(slice, lo, hi, &ok[t]);
  }
  slice(0, std::min(nbytes, step), &ok[0]);
  for (int t = 0; t < started; ++t) pool[t].join();
  for (int t = 0; t < threads; ++t)
    if (!ok[t]) return false;
  return true;
}

static int map_stream(hgm_ctx *c, SplitSource &src, int32_t fmt, double *sums, int64_t *counts,
                      int32_t *assign, int64_t *n_points_out) {
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (!sums || !counts) return fail(c, HGM_EINVAL, "sums_kxd and counts_k are required");
  if (fmt != HGM_FMT_F32LE && fmt != HGM_FMT_F32BE)
    return fail(c, HGM_EFORMAT, "record format %d is not supported by this build", fmt);
  const size_t row = (size_t)c->d * 4;
  if (src.nbytes % row != 0)
    return fail(c, HGM_EFORMAT, "split of %zu bytes is not a whole number of %zu-byte rows", src.nbytes, row);
  const int64_t n = (int64_t)(src.nbytes / row);
  if (n_points_out) *n_points_out = n;
  size_t rows_per_chunk = kChunkBytes / row;
  rows_per_chunk &= ~size_t(8191);  // whole tiles of every kernel, keeps 16-byte alignment
  if (rows_per_chunk == 0) rows_per_chunk = kChunkBytes / row;
  if (rows_per_chunk == 0) return fail(c, HGM_EINVAL, "a row of d=%d floats exceeds the staging chunk", c->d);
  const bool bounce = !src.pinned;
  int rc = ensure_pipeline(c, assign != nullptr, rows_per_chunk, bounce);
  if (rc != HGM_OK) return rc;

  c->last_launches = 0;
  if (c->d_tflags) cudaMemsetAsync(c->d_tflags + 2, 0, 8, c->stream);  // re-check total of this map call
  CU(c, cudaMemsetAsync(c->d_partials, 0, partial_bytes(c), c->stream));
  CU(c, cudaEventRecord(c->ev_t0, c->stream));
  int64_t done = 0;
  int stage = 0;
  int bounce_i = 0;
  bool stage_used[kStages] = {};
  bool bounce_used[kStages] = {};
  while (done < n) {
    const int64_t m = std::min<int64_t>((int64_t)rows_per_chunk, n - done);
    const size_t bytes = (size_t)m * row;
    cudaStream_t cs = c->copy[stage];
    // the chunk buffer is free once the kernel that read it (and its assign D2H) is done
    if (stage_used[stage]) CU(c, cudaStreamWaitEvent(cs, c->ev_free[stage], 0));
    if (!bounce) {
      CU(c, cudaMemcpyAsync(c->d_chunk[stage], src.mem + (size_t)done * row, bytes, cudaMemcpyHostToDevice, cs));
      c->staged_bytes += (int64_t)bytes;
    } else {
      size_t off = 0;
      while (off < bytes) {
        const size_t piece = std::min(kBounceBytes, bytes - off);
        if (bounce_used[bounce_i]) CU(c, cudaEventSynchronize(c->ev_bounce[bounce_i]));
        if (!stage_bytes(src, (size_t)done * row + off, c->h_bounce[bounce_i], piece))
          return fail(c, HGM_EIO, "short read from split file: %s", strerror(errno));
        CU(c, cudaMemcpyAsync(static_cast<unsigned char *>(c->d_chunk[stage]) + off, c->h_bounce[bounce_i], piece,
                              cudaMemcpyHostToDevice, cs));
        c->staged_bytes += (int64_t)piece;
        CU(c, cudaEventRecord(c->ev_bounce[bounce_i], cs));
        bounce_used[bounce_i] = true;
        bounce_i = (bounce_i + 1) % kStages;
        off += piece;
      }
    }
    CU(c, cudaEventRecord(c->ev_h2d[stage], cs));
    CU(c, cudaStreamWaitEvent(c->stream, c->ev_h2d[stage], 0));
    if (fmt == HGM_FMT_F32BE) {  // K5: decode on the device, in place
      cudaError_t e = launch_bswap32(static_cast<uint32_t *>(c->d_chunk[stage]), m * c->d, c->stream);
      if (e != cudaSuccess) return fail(c, HGM_ECUDA, "decode launch: %s", cudaGetErrorString(e));
      c->last_launches += 1;
      c->total_launches += 1;
    }
    int32_t *d_asg = assign ? c->d_assign_chunk[stage] : nullptr;
    rc = map_points_async(c, static_cast<const float *>(c->d_chunk[stage]), m, d_asg);
    if (rc != HGM_OK) return rc;
    if (assign)
      CU(c, cudaMemcpyAsync(assign + done, d_asg, (size_t)m * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaEventRecord(c->ev_free[stage], c->stream));
    stage_used[stage] = true;
    stage = (stage + 1) % kStages;
    done += m;
  }
  CU(c, cudaEventRecord(c->ev_t1, c->stream));
  c->timing_open = true;
  CU(c, cudaMemcpyAsync(sums, sums_ptr(c), (size_t)c->k * c->d * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(counts, counts_ptr(c), (size_t)c->k * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}

// ---- HBM-resident split cache -------------------------------------------------------------

static void cache_erase(hgm_ctx *c, int64_t id) {
  auto it = c->cache.find(id);
  if (it == c->cache.end()) return;
  arena_release(c, it->second.block);
  c->cache.erase(it);
}

// ---- text splits: LineRecordReader framing on the host, decode on the device ----------------

static int ensure_text(hgm_ctx *c) {
  for (int i = 0; i < kStages; ++i) {
    if (!c->d_text[i]) CU(c, cudaMalloc(&c->d_text[i], kBounceBytes));
    if (!c->d_tblocks[i]) CU(c, cudaMalloc(&c->d_tblocks[i], text_block_count(kBounceBytes) * sizeof(unsigned int)));
    if (!c->d_tstats[i]) CU(c, cudaMalloc(&c->d_tstats[i], sizeof(TextStats)));
  }
  if (!c->h_tstats) CU(c, cudaHostAlloc(&c->h_tstats, kStages * sizeof(TextStats), cudaHostAllocDefault));
  return HGM_OK;
}

// Bytes LineReader.readLine (core/org/apache/hadoop/util/LineReader.java:106-167) consumes from
// file position pos: through the first LF, or the first CR and an LF right behind it; to EOF if
// neither shows up.
static int64_t line_bytes_from(FILE *f, int64_t pos) {
  if (fseeko(f, (off_t)pos, SEEK_SET) != 0) return 0;
  int64_t n = 0;
  for (;;) {
    const int ch = fgetc(f);
    if (ch == EOF) return n;
    ++n;
    if (ch == '\n') return n;
    if (ch == '\r') {
      const int nx = fgetc(f);
      return nx == '\n' ? n + 1 : n;
    }
  }
}

// The byte range of the records LineRecordReader hands out for FileSplit(off, len)
// (mapred/org/apache/hadoop/mapred/LineRecordReader.java:79-95 and :137-157): a split that does
// not start at 0 gives its first (partial or whole) line to its predecessor -- found by reading
// one line from off-1 -- and every line that STARTS before off+len is read to its end.
static void text_split_bounds(FILE *f, int64_t off, int64_t len, int64_t *start_out, int64_t *stop_out) {
  fseeko(f, 0, SEEK_END);
  const int64_t size = (int64_t)ftello(f);
  int64_t end = len < 0 ? size : off + len;
  if (end > size) end = size;
  int64_t start = off;
  if (off != 0) start = (off - 1) + line_bytes_from(f, off - 1);
  if (start > size) start = size;
  int64_t stop = start;
  if (start < end) stop = (end - 1) + line_bytes_from(f, end - 1);
  *start_out = start;
  *stop_out = stop;
}

// Where to cut a staged window so that it ends at a line end (0 = no line end inside).
static size_t text_cut(const unsigned char *buf, size_t n) {
  for (size_t p = n; p > 0; --p)
    if (buf[p - 1] == '\n') return p;
  for (size_t p = n > 0 ? n - 1 : 0; p > 0; --p)  // CR-only files; a CR in the last byte may have its LF outside
    if (buf[p - 1] == '\r') return p;
  return 0;
}

static int text_error(hgm_ctx *c, const TextStats &st, int64_t origin) {
  const unsigned code = (unsigned)(st.first_error & 15u);
  const long long pos = (long long)(st.first_error >> 4) + origin;
  switch (code) {
    case HGM_TEXT_MALFORMED:
      return fail(c, HGM_EFORMAT, "text split: the field at byte %lld is not a number sscanf(\"%%f\") accepts", pos);
    case HGM_TEXT_RAGGED:
      return fail(c, HGM_EFORMAT, "text split: the line ending at byte %lld does not hold a whole number of d=%d points",
                  pos, c->d);
    case HGM_TEXT_LONG:
      return fail(c, HGM_EFORMAT, "text split: the field at byte %lld is longer than 127 bytes", pos);
    default:
      return fail(c, HGM_ECUDA, "text split: decoder ran out of room at byte %lld", pos);
  }
}

static int map_text_stream(hgm_ctx *c, SplitSource &src, double *sums, int64_t *counts, int32_t *assign,
                           int64_t *n_points_out) {
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (!sums || !counts) return fail(c, HGM_EINVAL, "sums_kxd and counts_k are required");
  if (assign)
    return fail(c, HGM_EINVAL, "assignments of a text split are not returned (the caller cannot size the array); "
                               "cache the split and map it from HBM instead");
  int rc = ensure_pipeline(c, false, 1, true);
  if (rc == HGM_OK) rc = ensure_text(c);
  if (rc != HGM_OK) return rc;

  c->last_launches = 0;
  if (c->d_tflags) cudaMemsetAsync(c->d_tflags + 2, 0, 8, c->stream);  // re-check total of this map call
  CU(c, cudaMemsetAsync(c->d_partials, 0, partial_bytes(c), c->stream));
  CU(c, cudaEventRecord(c->ev_t0, c->stream));
  size_t remaining = src.nbytes, src_pos = 0;  // src_pos: source offset of the next fresh byte
  const unsigned char *carry = nullptr;
  size_t carry_len = 0;
  int64_t points = 0;
  int stage = 0;
  bool used[kStages] = {};
  while (remaining > 0 || carry_len > 0) {
    cudaStream_t cs = c->copy[stage];
    unsigned char *buf = static_cast<unsigned char *>(c->h_bounce[stage]);
    if (used[stage]) CU(c, cudaEventSynchronize(c->ev_bounce[stage]));
    if (carry_len) memmove(buf, carry, carry_len);  // the unfinished line of the previous window
    const int64_t origin = (int64_t)src_pos - (int64_t)carry_len;
    const size_t fresh = std::min(remaining, kBounceBytes - carry_len);
    if (fresh && !stage_bytes(src, src_pos, buf + carry_len, fresh))
      return fail(c, HGM_EIO, "short read from split file: %s", strerror(errno));
    src_pos += fresh;
    remaining -= fresh;
    const size_t n = carry_len + fresh;
    const size_t cut = remaining == 0 ? n : text_cut(buf, n);
    if (cut == 0) return fail(c, HGM_EFORMAT, "text split: a line longer than %zu bytes at byte %lld", kBounceBytes, (long long)origin);
    carry = buf + cut;
    carry_len = n - cut;

    CU(c, cudaMemcpyAsync(c->d_text[stage], buf, cut, cudaMemcpyHostToDevice, cs));
    c->staged_bytes += (int64_t)cut;
    CU(c, cudaEventRecord(c->ev_bounce[stage], cs));
    if (used[stage]) CU(c, cudaStreamWaitEvent(cs, c->ev_free[stage], 0));  // rows still being mapped
    cudaError_t e = launch_text_decode(static_cast<const unsigned char *>(c->d_text[stage]), cut, c->d,
                                       static_cast<float *>(c->d_chunk[stage]), kChunkBytes / 4, c->d_tblocks[stage],
                                       c->d_tstats[stage], cs);
    if (e != cudaSuccess) return fail(c, HGM_ECUDA, "text decode launch: %s", cudaGetErrorString(e));
    c->last_launches += 3;
    c->total_launches += 3;
    CU(c, cudaMemcpyAsync(&c->h_tstats[stage], c->d_tstats[stage], sizeof(TextStats), cudaMemcpyDeviceToHost, cs));
    CU(c, cudaEventRecord(c->ev_h2d[stage], cs));
    CU(c, cudaStreamSynchronize(cs));  // only this copy stream: the previous piece keeps mapping meanwhile
    const TextStats st = c->h_tstats[stage];
    if (st.first_error != ~0ull) return text_error(c, st, origin);
    if (st.n_fields % (unsigned long long)c->d != 0)
      return fail(c, HGM_EFORMAT, "text split: the line ending at byte %lld does not hold a whole number of d=%d points",
                  (long long)(origin + (int64_t)cut), c->d);
    const int64_t m = (int64_t)(st.n_fields / (unsigned long long)c->d);
    CU(c, cudaStreamWaitEvent(c->stream, c->ev_h2d[stage], 0));
    rc = map_points_async(c, static_cast<const float *>(c->d_chunk[stage]), m, nullptr);
    if (rc != HGM_OK) return rc;
    CU(c, cudaEventRecord(c->ev_free[stage], c->stream));
    used[stage] = true;
    stage = (stage + 1) % kStages;
    points += m;
  }
  if (n_points_out) *n_points_out = points;
  CU(c, cudaEventRecord(c->ev_t1, c->stream));
  c->timing_open = true;
  CU(c, cudaMemcpyAsync(sums, sums_ptr(c), (size_t)c->k * c->d * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(counts, counts_ptr(c), (size_t)c->k * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return HGM_OK;
}

// A whole text split decoded into exactly-sized rows for the HBM cache.
static int cache_store_text(hgm_ctx *c, int64_t split_id, SplitSource &src) {
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called (d is unknown)");
  if (split_id < 0) return fail(c, HGM_EINVAL, "split_id must be >= 0");
  if (src.nbytes >= (size_t(1) << 33)) return fail(c, HGM_EINVAL, "a text split must be smaller than 8 GiB");
  int rc = ensure_pipeline(c, false, 1, true);
  if (rc != HGM_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  cache_erase(c, split_id);
  unsigned char *d_text = nullptr;
  unsigned int *d_blocks = nullptr;
  TextStats *d_stats = nullptr;
  float *d_rows = nullptr;
  int block = -1;
  TextStats st = {0ull, ~0ull};
  auto cleanup = [&](int code) {
    cudaFree(d_text);
    cudaFree(d_blocks);
    cudaFree(d_stats);
    if (code != HGM_OK && block >= 0) arena_release(c, block);
    return code;
  };
  auto no_room = [&](cudaError_t e) {
    cudaGetLastError();
    return cleanup(fail(c, HGM_ENOMEM, "split %lld (%zu bytes of text) does not fit the HBM cache: %s",
                        (long long)split_id, src.nbytes, cudaGetErrorString(e)));
  };
  cudaError_t e = cudaMalloc(&d_text, src.nbytes + 16);
  if (e != cudaSuccess) return no_room(e);
  e = cudaMalloc(&d_blocks, (text_block_count(src.nbytes) + 1) * sizeof(unsigned int));
  if (e != cudaSuccess) return no_room(e);
  e = cudaMalloc(&d_stats, sizeof(TextStats));
  if (e != cudaSuccess) return no_room(e);
  size_t off = 0;
  int bi = 0;
  while (off < src.nbytes) {
    const size_t piece = std::min(kBounceBytes, src.nbytes - off);
    if (cudaEventSynchronize(c->ev_bounce[bi]) != cudaSuccess) cudaGetLastError();
    if (!stage_bytes(src, off, c->h_bounce[bi], piece))
      return cleanup(fail(c, HGM_EIO, "short read from split file: %s", strerror(errno)));
    e = cudaMemcpyAsync(d_text + off, c->h_bounce[bi], piece, cudaMemcpyHostToDevice, c->stream);
    c->staged_bytes += (int64_t)piece;
    if (e == cudaSuccess) e = cudaEventRecord(c->ev_bounce[bi], c->stream);
    if (e != cudaSuccess) return cleanup(fail(c, HGM_ECUDA, "cache copy: %s", cudaGetErrorString(e)));
    bi = (bi + 1) % kStages;
    off += piece;
  }
  // pass 1 counts the fields so that the rows can be allocated exactly; pass 2 parses
  for (int pass = 0; pass < 2; ++pass) {
    e = launch_text_decode(d_text, src.nbytes, c->d, pass ? d_rows : nullptr, pass ? (size_t)st.n_fields : 0, d_blocks,
                           d_stats, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&st, d_stats, sizeof st, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return cleanup(fail(c, HGM_ECUDA, "text decode: %s", cudaGetErrorString(e)));
    c->total_launches += pass ? 3 : 2;
    if (pass == 0) {
      if (st.n_fields % (unsigned long long)c->d != 0)
        return cleanup(fail(c, HGM_EFORMAT, "text split: %llu numbers are not a whole number of d=%d points",
                            st.n_fields, c->d));
      d_rows = static_cast<float *>(arena_alloc(c, (size_t)st.n_fields * 4, &block, &e));
      if (!d_rows) return no_room(e);
    } else if (st.first_error != ~0ull) {
      return cleanup(text_error(c, st, 0));
    }
  }
  c->cache[split_id] = hgm_ctx::CachedSplit{d_rows, (int64_t)(st.n_fields / (unsigned long long)c->d),
                                            (size_t)st.n_fields * 4, block};
  return cleanup(HGM_OK);
}

// ---- SequenceFile splits: header and sync search on the host, record walk on the device ------

struct SeqHeader {
  int64_t header_end = 0;  // position of the first record
  SeqSync sync{};
};

// WritableUtils.readVInt (same coding as c++/utils/impl/SerialUtils.cc:217-244)
static bool seq_read_vint(const unsigned char *&p, const unsigned char *end, int64_t *out) {
  if (p >= end) return false;
  const int8_t first = (int8_t)*p++;
  if (first >= -112) { *out = first; return true; }
  const bool negative = first < -120;
  const int len = negative ? -(first + 120) : -(first + 112);
  if (p + len > end) return false;
  int64_t v = 0;
  for (int i = 0; i < len; ++i) v = (v << 8) | *p++;
  *out = negative ? ~v : v;
  return true;
}
static bool seq_read_text(const unsigned char *&p, const unsigned char *end, std::string *out) {
  int64_t len = 0;
  if (!seq_read_vint(p, end, &len) || len < 0 || p + len > end) return false;
  if (out) out->assign(reinterpret_cast<const char *>(p), (size_t)len);
  p += len;
  return true;
}

// SequenceFile.Reader.init (core/org/apache/hadoop/io/SequenceFile.java:1506-1571): version 6,
// Text class names, the two compression flags, metadata, 16 sync bytes.
static int seq_read_header(hgm_ctx *c, FILE *f, SeqHeader *h) {
  std::vector<unsigned char> buf(64 * 1024);
  fseeko(f, 0, SEEK_SET);
  const size_t got = fread(buf.data(), 1, buf.size(), f);
  const unsigned char *p = buf.data(), *end = p + got;
  if (got < 4 || p[0] != 'S' || p[1] != 'E' || p[2] != 'Q') return fail(c, HGM_EFORMAT, "not a SequenceFile");
  if (p[3] != 6) return fail(c, HGM_EFORMAT, "SequenceFile version %d is not supported (6 is)", (int)p[3]);
  p += 4;
  std::string key_class, val_class;
  if (!seq_read_text(p, end, &key_class) || !seq_read_text(p, end, &val_class))
    return fail(c, HGM_EFORMAT, "SequenceFile header is truncated");
  if (val_class != "org.apache.hadoop.io.BytesWritable")
    return fail(c, HGM_EFORMAT, "SequenceFile values are %s; rows must be BytesWritable", val_class.c_str());
  if (p + 2 > end) return fail(c, HGM_EFORMAT, "SequenceFile header is truncated");
  if (p[0] != 0 || p[1] != 0) return fail(c, HGM_EFORMAT, "compressed SequenceFiles are not supported");
  p += 2;
  if (p + 4 > end) return fail(c, HGM_EFORMAT, "SequenceFile header is truncated");
  const int64_t meta = ((int64_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3];
  p += 4;
  for (int64_t i = 0; i < 2 * meta; ++i)
    if (!seq_read_text(p, end, nullptr)) return fail(c, HGM_EFORMAT, "SequenceFile metadata is truncated or over 64 KiB");
  if (p + 16 > end) return fail(c, HGM_EFORMAT, "SequenceFile header is truncated");
  memcpy(h->sync.marker, p, 16);
  h->header_end = (int64_t)(p + 16 - buf.data());
  return HGM_OK;
}

// SequenceFile.Reader.sync(position) (SequenceFile.java:2221-2246): the first sync entry whose
// escape starts at or after position, the end of the file if there is none.
static int64_t seq_sync_at_or_after(FILE *f, int64_t size, const SeqHeader &h, int64_t position) {
  if (position + 20 >= size) return size;
  std::vector<unsigned char> buf(256 * 1024 + 16);
  int64_t at = position + 4;  // where a marker may start
  size_t keep = 0;
  fseeko(f, (off_t)at, SEEK_SET);
  for (;;) {
    const size_t got = fread(buf.data() + keep, 1, buf.size() - keep, f);
    const size_t have = keep + got;
    if (have < 16) return size;
    for (size_t i = 0; i + 16 <= have; ++i)
      if (buf[i] == h.sync.marker[0] && memcmp(buf.data() + i, h.sync.marker, 16) == 0) return at + (int64_t)i - 4;
    if (got == 0) return size;
    keep = 15;  // a marker may straddle the buffers
    memmove(buf.data(), buf.data() + have - keep, keep);
    at += (int64_t)(have - keep);
  }
}

// Byte range of the records SequenceFileRecordReader delivers for FileSplit(off, len)
// (mapred/org/apache/hadoop/mapred/SequenceFileRecordReader.java:40-52, 71-84): from the first sync
// entry at or after off (the header's end for the split that contains it) up to the first sync
// entry at or after off + len.
static int seq_split_bounds(hgm_ctx *c, FILE *f, int64_t off, int64_t len, SeqHeader *h, int64_t *start, int64_t *stop) {
  const int rc = seq_read_header(c, f, h);
  if (rc != HGM_OK) return rc;
  fseeko(f, 0, SEEK_END);
  const int64_t size = (int64_t)ftello(f);
  int64_t end = len < 0 ? size : off + len;
  if (end > size) end = size;
  *start = off > h->header_end ? seq_sync_at_or_after(f, size, *h, off) : h->header_end;
  *stop = end > h->header_end ? seq_sync_at_or_after(f, size, *h, end) : h->header_end;
  if (*start >= end) *stop = *start;  // "more = start < end"
  if (*stop < *start) *stop = *start;
  return HGM_OK;
}

// Stage the byte range, walk it on the device, leave the rows in *d_rows (an arena slot when block_out is
// given, else a cudaMalloc the caller frees).  *n_floats <- floats decoded.
static int decode_seq(hgm_ctx *c, FILE *f, int64_t start, int64_t stop, const SeqHeader &h, bool big_endian,
                      float **d_rows, int64_t *n_floats, int *block_out) {
  const size_t nbytes = (size_t)(stop - start);
  *d_rows = nullptr;
  *n_floats = 0;
  if (nbytes >= (size_t(1) << 33)) return fail(c, HGM_EINVAL, "a SequenceFile split must be smaller than 8 GiB");
  int rc = ensure_pipeline(c, false, 1, true);
  if (rc != HGM_OK) return rc;
  CU(c, cudaStreamSynchronize(c->stream));
  unsigned char *d_bytes = nullptr;
  unsigned int *d_blocks = nullptr, *d_seg = nullptr;
  long long *d_pos = nullptr;
  TextStats *d_stats = nullptr;
  TextStats st = {0ull, ~0ull};
  int block = -1;
  auto cleanup = [&](int code) {
    cudaFree(d_bytes); cudaFree(d_blocks); cudaFree(d_seg); cudaFree(d_pos); cudaFree(d_stats);
    if (code != HGM_OK && *d_rows) {
      if (block >= 0) arena_release(c, block); else cudaFree(*d_rows);
      *d_rows = nullptr;
    }
    return code;
  };
  auto no_room = [&](cudaError_t e) {
    cudaGetLastError();
    return cleanup(fail(c, HGM_ENOMEM, "SequenceFile split of %zu bytes does not fit the device: %s", nbytes,
                        cudaGetErrorString(e)));
  };
  cudaError_t e = cudaMalloc(&d_bytes, nbytes + 32);
  if (e == cudaSuccess) e = cudaMalloc(&d_blocks, (seq_block_count(nbytes) + 1) * sizeof(unsigned int));
  if (e == cudaSuccess) e = cudaMalloc(&d_pos, (nbytes / 20 + 2) * sizeof(long long));
  if (e == cudaSuccess) e = cudaMalloc(&d_stats, sizeof(TextStats));
  if (e != cudaSuccess) return no_room(e);
  SplitSource src;
  src.file = f;
  src.file_base = start;
  src.nbytes = nbytes;
  size_t off = 0;
  int bi = 0;
  while (off < nbytes) {
    const size_t piece = std::min(kBounceBytes, nbytes - off);
    if (cudaEventSynchronize(c->ev_bounce[bi]) != cudaSuccess) cudaGetLastError();
    if (!stage_bytes(src, off, c->h_bounce[bi], piece))
      return cleanup(fail(c, HGM_EIO, "short read from split file: %s", strerror(errno)));
    e = cudaMemcpyAsync(d_bytes + off, c->h_bounce[bi], piece, cudaMemcpyHostToDevice, c->stream);
    c->staged_bytes += (int64_t)piece;
    if (e == cudaSuccess) e = cudaEventRecord(c->ev_bounce[bi], c->stream);
    if (e != cudaSuccess) return cleanup(fail(c, HGM_ECUDA, "split copy: %s", cudaGetErrorString(e)));
    bi = (bi + 1) % kStages;
    off += piece;
  }
  auto fetch = [&]() {
    cudaError_t er = cudaMemcpyAsync(&st, d_stats, sizeof st, cudaMemcpyDeviceToHost, c->stream);
    if (er == cudaSuccess) er = cudaStreamSynchronize(c->stream);
    return er;
  };
  e = launch_seq_find_syncs(d_bytes, nbytes, h.sync, d_blocks, d_pos, d_stats, c->stream);
  if (e == cudaSuccess) e = fetch();
  if (e != cudaSuccess) return cleanup(fail(c, HGM_ECUDA, "SequenceFile sync search: %s", cudaGetErrorString(e)));
  const long long n_sync = (long long)st.n_fields;
  c->total_launches += 3;
  e = cudaMalloc(&d_seg, (size_t)(n_sync + 2) * sizeof(unsigned int));
  if (e != cudaSuccess) return no_room(e);
  e = launch_seq_walk(d_bytes, nbytes, d_pos, n_sync, c->d, big_endian, d_seg, nullptr, 0, d_stats, 0, c->sm_count, c->stream);
  if (e == cudaSuccess) e = fetch();
  if (e != cudaSuccess) return cleanup(fail(c, HGM_ECUDA, "SequenceFile record walk: %s", cudaGetErrorString(e)));
  c->total_launches += 2;
  if (st.first_error != ~0ull) {
    const long long pos = (long long)(st.first_error >> 4) + start;
    return cleanup(fail(c, HGM_EFORMAT, (st.first_error & 15u) == HGM_TEXT_RAGGED
                            ? "SequenceFile split: the record at byte %lld does not hold whole rows of d=%d float32"
                            : "SequenceFile split: the record at byte %lld is malformed (d=%d)", pos, c->d));
  }
  const size_t floats = (size_t)st.n_fields;
  if (block_out) {
    *d_rows = static_cast<float *>(arena_alloc(c, floats * 4, &block, &e));
    if (!*d_rows) return no_room(e);
  } else {
    e = cudaMalloc(d_rows, floats ? floats * 4 : 16);
    if (e != cudaSuccess) { *d_rows = nullptr; return no_room(e); }
  }
  if (floats) {
    e = launch_seq_walk(d_bytes, nbytes, d_pos, n_sync, c->d, big_endian, d_seg, *d_rows, floats, d_stats, 1, c->sm_count, c->stream);
    if (e == cudaSuccess) e = fetch();
    if (e != cudaSuccess) return cleanup(fail(c, HGM_ECUDA, "SequenceFile record copy: %s", cudaGetErrorString(e)));
    c->total_launches += 1;
    if (st.first_error != ~0ull) return cleanup(fail(c, HGM_ECUDA, "SequenceFile record copy ran out of room"));
  }
  *n_floats = (int64_t)floats;
  if (block_out) *block_out = block;
  return cleanup(HGM_OK);
}

static int map_seq_file(hgm_ctx *c, const char *path, int64_t offset, int64_t length, bool big_endian, double *sums,
                        int64_t *counts, int32_t *assign, int64_t *n_points_out) {
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (!sums || !counts) return fail(c, HGM_EINVAL, "sums_kxd and counts_k are required");
  if (assign) return fail(c, HGM_EINVAL, "assignments of a SequenceFile split are not returned (cache it instead)");
  if (!path || offset < 0) return fail(c, HGM_EINVAL, "bad path or offset");
  FILE *f = fopen(path, "rb");
  if (!f) return fail(c, HGM_EIO, "cannot open %s: %s", path, strerror(errno));
  SeqHeader h;
  int64_t start = 0, stop = 0, n_floats = 0;
  float *d_rows = nullptr;
  int rc = seq_split_bounds(c, f, offset, length, &h, &start, &stop);
  if (rc == HGM_OK) rc = decode_seq(c, f, start, stop, h, big_endian, &d_rows, &n_floats, nullptr);
  fclose(f);
  if (rc != HGM_OK) return rc;
  const int64_t n = n_floats / c->d;
  if (n_points_out) *n_points_out = n;
  c->last_launches = 0;
  if (c->d_tflags) cudaMemsetAsync(c->d_tflags + 2, 0, 8, c->stream);  // re-check total of this map call
  cudaError_t e = cudaMemsetAsync(c->d_partials, 0, partial_bytes(c), c->stream);
  if (e == cudaSuccess) e = cudaEventRecord(c->ev_t0, c->stream);
  if (e == cudaSuccess) rc = map_points_async(c, d_rows, n, nullptr);
  if (e == cudaSuccess && rc == HGM_OK) {
    e = cudaEventRecord(c->ev_t1, c->stream);
    c->timing_open = true;
    if (e == cudaSuccess) e = cudaMemcpyAsync(sums, sums_ptr(c), (size_t)c->k * c->d * 8, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(counts, counts_ptr(c), (size_t)c->k * 8, cudaMemcpyDeviceToHost, c->stream);
  }
  const cudaError_t es = cudaStreamSynchronize(c->stream);
  cudaFree(d_rows);
  if (rc != HGM_OK) return rc;
  if (e != cudaSuccess || es != cudaSuccess)
    return fail(c, HGM_ECUDA, "SequenceFile map: %s", cudaGetErrorString(e != cudaSuccess ? e : es));
  return HGM_OK;
}

static int cache_seq_file(hgm_ctx *c, int64_t split_id, const char *path, int64_t offset, int64_t length, bool big_endian) {
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called (d is unknown)");
  if (split_id < 0) return fail(c, HGM_EINVAL, "split_id must be >= 0");
  if (!path || offset < 0) return fail(c, HGM_EINVAL, "bad path or offset");
  FILE *f = fopen(path, "rb");
  if (!f) return fail(c, HGM_EIO, "cannot open %s: %s", path, strerror(errno));
  SeqHeader h;
  int64_t start = 0, stop = 0, n_floats = 0;
  float *d_rows = nullptr;
  int block = -1;
  {
    const cudaError_t es = cudaStreamSynchronize(c->stream);
    if (es != cudaSuccess) {
      fclose(f);
      return fail(c, HGM_ECUDA, "cudaStreamSynchronize: %s", cudaGetErrorString(es));
    }
  }
  cache_erase(c, split_id);
  int rc = seq_split_bounds(c, f, offset, length, &h, &start, &stop);
  if (rc == HGM_OK) rc = decode_seq(c, f, start, stop, h, big_endian, &d_rows, &n_floats, &block);
  fclose(f);
  if (rc != HGM_OK) return rc;
  c->cache[split_id] = hgm_ctx::CachedSplit{d_rows, n_floats / c->d, (size_t)n_floats * 4, block};
  return HGM_OK;
}

// Opens path and frames the split: exact bytes for binary rows, LineRecordReader's rule for text.
static int open_split_file(hgm_ctx *c, const char *path, int64_t offset, int64_t length, int32_t fmt, SplitSource *src) {
  if (!path || offset < 0) return fail(c, HGM_EINVAL, "bad path or offset");
  FILE *f = fopen(path, "rb");
  if (!f) return fail(c, HGM_EIO, "cannot open %s: %s", path, strerror(errno));
  int64_t start = offset, stop;
  if (fmt == HGM_FMT_TEXT) {
    text_split_bounds(f, offset, length, &start, &stop);
  } else {
    if (length < 0) {
      fseeko(f, 0, SEEK_END);
      length = std::max<int64_t>((int64_t)ftello(f) - offset, 0);
    }
    stop = offset + length;
  }
  if (fseeko(f, (off_t)start, SEEK_SET) != 0) {
    fclose(f);
    return fail(c, HGM_EIO, "cannot seek %s to %lld", path, (long long)start);
  }
  src->file = f;
  src->file_base = start;
  src->nbytes = (size_t)(stop - start);
  src->pinned = false;
  return HGM_OK;
}

int hgm_map_split(hgm_ctx *c, const void *host_bytes, size_t nbytes, int32_t fmt, double *sums,
                  int64_t *counts, int32_t *assign, int64_t *n_points_out) {
  HGM_RANGE("hgm_map_split");
  NEED_CTX(c);
  if (nbytes > 0 && !host_bytes) return fail(c, HGM_EINVAL, "host_bytes is NULL");
  SplitSource src;
  src.mem = static_cast<const unsigned char *>(host_bytes);
  src.nbytes = nbytes;
  src.pinned = nbytes > 0 && is_pinned(host_bytes);
  if (fmt == HGM_FMT_TEXT) return map_text_stream(c, src, sums, counts, assign, n_points_out);
  if (fmt == HGM_FMT_SEQ_F32BE || fmt == HGM_FMT_SEQ_F32LE)
    return fail(c, HGM_EFORMAT, "SequenceFile splits are read from the file (its header holds the sync marker): use hgm_map_file");
  return map_stream(c, src, fmt, sums, counts, assign, n_points_out);
}

int hgm_map_file(hgm_ctx *c, const char *path, int64_t offset, int64_t length, int32_t fmt,
                 double *sums, int64_t *counts, int32_t *assign, int64_t *n_points_out) {
  HGM_RANGE("hgm_map_file");
  NEED_CTX(c);
  if (fmt == HGM_FMT_SEQ_F32BE || fmt == HGM_FMT_SEQ_F32LE)
    return map_seq_file(c, path, offset, length, fmt == HGM_FMT_SEQ_F32BE, sums, counts, assign, n_points_out);
  SplitSource src;
  int rc = open_split_file(c, path, offset, length, fmt, &src);
  if (rc != HGM_OK) return rc;
  rc = fmt == HGM_FMT_TEXT ? map_text_stream(c, src, sums, counts, assign, n_points_out)
                           : map_stream(c, src, fmt, sums, counts, assign, n_points_out);
  fclose(src.file);
  if (rc != HGM_OK) {
    for (int i = 0; i < kStages; ++i) cudaStreamSynchronize(c->copy[i]);
    cudaStreamSynchronize(c->stream);
  }
  return rc;
}

static int cache_store(hgm_ctx *c, int64_t split_id, SplitSource &src, int32_t fmt) {
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called (d is unknown)");
  if (split_id < 0) return fail(c, HGM_EINVAL, "split_id must be >= 0");
  if (fmt != HGM_FMT_F32LE && fmt != HGM_FMT_F32BE)
    return fail(c, HGM_EFORMAT, "record format %d is not supported by this build", fmt);
  const size_t row = (size_t)c->d * 4;
  if (src.nbytes % row != 0)
    return fail(c, HGM_EFORMAT, "split of %zu bytes is not a whole number of %zu-byte rows", src.nbytes, row);
  CU(c, cudaStreamSynchronize(c->stream));
  cache_erase(c, split_id);
  int block = -1;
  cudaError_t e = cudaSuccess;
  void *d_rows = arena_alloc(c, src.nbytes, &block, &e);
  if (!d_rows)
    return fail(c, HGM_ENOMEM, "split %lld (%zu bytes) does not fit the HBM cache: %s", (long long)split_id,
                src.nbytes, cudaGetErrorString(e));
  int rc = HGM_OK;
  if (src.file || !src.pinned) {  // files and pageable memory: through the pinned bounce buffers
    rc = ensure_pipeline(c, false, 1, true);
    size_t off = 0;
    int bi = 0;
    while (rc == HGM_OK && off < src.nbytes) {
      const size_t piece = std::min(kBounceBytes, src.nbytes - off);
      if (cudaEventSynchronize(c->ev_bounce[bi]) != cudaSuccess) cudaGetLastError();
      if (!stage_bytes(src, off, c->h_bounce[bi], piece)) {
        rc = fail(c, HGM_EIO, "short read from split file: %s", strerror(errno));
        break;
      }
      e = cudaMemcpyAsync(static_cast<unsigned char *>(d_rows) + off, c->h_bounce[bi], piece, cudaMemcpyHostToDevice,
                          c->stream);
      c->staged_bytes += (int64_t)piece;
      if (e == cudaSuccess) e = cudaEventRecord(c->ev_bounce[bi], c->stream);
      if (e != cudaSuccess) rc = fail(c, HGM_ECUDA, "cache copy: %s", cudaGetErrorString(e));
      bi = (bi + 1) % kStages;
      off += piece;
    }
  } else if (src.nbytes) {
    e = cudaMemcpyAsync(d_rows, src.mem, src.nbytes, cudaMemcpyHostToDevice, c->stream);
    c->staged_bytes += (int64_t)src.nbytes;
    if (e != cudaSuccess) rc = fail(c, HGM_ECUDA, "cache copy: %s", cudaGetErrorString(e));
  }
  if (rc == HGM_OK && fmt == HGM_FMT_F32BE) {
    e = launch_bswap32(static_cast<uint32_t *>(d_rows), (int64_t)(src.nbytes / 4), c->stream);
    if (e != cudaSuccess) rc = fail(c, HGM_ECUDA, "decode launch: %s", cudaGetErrorString(e));
    else c->total_launches += 1;
  }
  if (rc == HGM_OK && cudaStreamSynchronize(c->stream) != cudaSuccess) rc = fail(c, HGM_ECUDA, "cache copy failed");
  if (rc != HGM_OK) {
    arena_release(c, block);
    return rc;
  }
  c->cache[split_id] = hgm_ctx::CachedSplit{d_rows, (int64_t)(src.nbytes / row), src.nbytes, block};
  return HGM_OK;
}

int hgm_cache_put(hgm_ctx *c, int64_t split_id, const void *host_bytes, size_t nbytes, int32_t fmt) {
  HGM_RANGE("hgm_cache_put");
  NEED_CTX(c);
  if (nbytes > 0 && !host_bytes) return fail(c, HGM_EINVAL, "host_bytes is NULL");
  SplitSource src;
  src.mem = static_cast<const unsigned char *>(host_bytes);
  src.nbytes = nbytes;
  src.pinned = nbytes > 0 && is_pinned(host_bytes);
  if (fmt == HGM_FMT_TEXT) return cache_store_text(c, split_id, src);
  if (fmt == HGM_FMT_SEQ_F32BE || fmt == HGM_FMT_SEQ_F32LE)
    return fail(c, HGM_EFORMAT, "SequenceFile splits are read from the file (its header holds the sync marker): use hgm_cache_put_file");
  return cache_store(c, split_id, src, fmt);
}

int hgm_cache_put_file(hgm_ctx *c, int64_t split_id, const char *path, int64_t offset, int64_t length, int32_t fmt) {
  HGM_RANGE("hgm_cache_put_file");
  NEED_CTX(c);
  if (fmt == HGM_FMT_SEQ_F32BE || fmt == HGM_FMT_SEQ_F32LE)
    return cache_seq_file(c, split_id, path, offset, length, fmt == HGM_FMT_SEQ_F32BE);
  SplitSource src;
  int rc = open_split_file(c, path, offset, length, fmt, &src);
  if (rc != HGM_OK) return rc;
  rc = fmt == HGM_FMT_TEXT ? cache_store_text(c, split_id, src) : cache_store(c, split_id, src, fmt);
  fclose(src.file);
  return rc;
}

int hgm_cache_map(hgm_ctx *c, int64_t split_id, int32_t *d_assign, int64_t *n_points_out) {
  HGM_RANGE("hgm_cache_map");
  NEED_CTX(c);
  auto it = c->cache.find(split_id);
  if (it == c->cache.end()) return fail(c, HGM_EINVAL, "split %lld is not in the cache", (long long)split_id);
  if (n_points_out) *n_points_out = it->second.n_points;
  return hgm_map_device(c, static_cast<const float *>(it->second.d_rows), it->second.n_points, d_assign);
}

// map the given cached splits; those that sit back to back in an arena block go out as one launch
static int map_cached(hgm_ctx *c, std::vector<const hgm_ctx::CachedSplit *> &order, int64_t *n_points_out,
                      int32_t *n_launch_groups_out);

int hgm_cache_map_all(hgm_ctx *c, int64_t *n_points_out, int32_t *n_launch_groups_out) {
  HGM_RANGE("hgm_cache_map_all");
  NEED_CTX(c);
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  std::vector<const hgm_ctx::CachedSplit *> order;
  order.reserve(c->cache.size());
  for (auto &kv : c->cache) order.push_back(&kv.second);
  return map_cached(c, order, n_points_out, n_launch_groups_out);
}

int hgm_cache_map_many(hgm_ctx *c, const int64_t *split_ids, int32_t n_ids, int64_t *n_points_out,
                       int32_t *n_launch_groups_out) {
  HGM_RANGE("hgm_cache_map_many");
  NEED_CTX(c);
  if (!c->d_partials) return fail(c, HGM_EINVAL, "hgm_set_centroids has not been called");
  if (n_ids < 0 || (n_ids > 0 && !split_ids)) return fail(c, HGM_EINVAL, "bad split id list");
  std::vector<const hgm_ctx::CachedSplit *> order;
  order.reserve((size_t)n_ids);
  for (int32_t i = 0; i < n_ids; ++i) {
    auto it = c->cache.find(split_ids[i]);
    if (it == c->cache.end()) return fail(c, HGM_EINVAL, "split %lld is not in the cache", (long long)split_ids[i]);
    order.push_back(&it->second);
  }
  return map_cached(c, order, n_points_out, n_launch_groups_out);
}

static int map_cached(hgm_ctx *c, std::vector<const hgm_ctx::CachedSplit *> &order, int64_t *n_points_out,
                      int32_t *n_launch_groups_out) {
  std::sort(order.begin(), order.end(),
            [](const hgm_ctx::CachedSplit *a, const hgm_ctx::CachedSplit *b) { return a->d_rows < b->d_rows; });
  int64_t total = 0;
  int32_t groups = 0;
  size_t i = 0;
  // one timed region over all the launches (hgm_last_map_ms)
  c->last_launches = 0;
  if (c->d_tflags) cudaMemsetAsync(c->d_tflags + 2, 0, 8, c->stream);
  CU(c, cudaEventRecord(c->ev_t0, c->stream));
  while (i < order.size()) {  // splits that sit back to back in an arena block go out as one launch
    const unsigned char *begin = static_cast<const unsigned char *>(order[i]->d_rows);
    const unsigned char *end = begin + order[i]->nbytes;
    int64_t n = order[i]->n_points;
    size_t j = i + 1;
    while (j < order.size() && order[j]->block == order[i]->block && order[j]->d_rows == end) {
      end += order[j]->nbytes;
      n += order[j]->n_points;
      ++j;
    }
    if (n > 0) {
      const int rc = map_points_async(c, reinterpret_cast<const float *>(begin), n, nullptr);
      if (rc != HGM_OK) return rc;
      ++groups;
    }
    total += n;
    i = j;
  }
  CU(c, cudaEventRecord(c->ev_t1, c->stream));
  c->timing_open = true;
  if (n_points_out) *n_points_out = total;
  if (n_launch_groups_out) *n_launch_groups_out = groups;
  return HGM_OK;
}

int hgm_cache_read(hgm_ctx *c, int64_t split_id, float *rows_out, int64_t cap_floats, int64_t *n_floats_out) {
  NEED_CTX(c);
  auto it = c->cache.find(split_id);
  if (it == c->cache.end()) return fail(c, HGM_EINVAL, "split %lld is not in the cache", (long long)split_id);
  const int64_t n = (int64_t)(it->second.nbytes / 4);
  if (n_floats_out) *n_floats_out = n;
  if (n > cap_floats || (n > 0 && !rows_out))
    return fail(c, HGM_EINVAL, "split %lld holds %lld floats, the buffer %lld", (long long)split_id, (long long)n,
                (long long)cap_floats);
  if (n > 0) {
    CU(c, cudaMemcpyAsync(rows_out, it->second.d_rows, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
  }
  return HGM_OK;
}

int hgm_cache_drop(hgm_ctx *c, int64_t split_id) {
  NEED_CTX(c);
  CU(c, cudaStreamSynchronize(c->stream));
  if (split_id >= 0) {
    cache_erase(c, split_id);
  } else {
    cache_clear(c);
  }
  return HGM_OK;
}

int hgm_cache_info(hgm_ctx *c, int64_t *n_splits, int64_t *bytes) {
  if (!c) return fail(nullptr, HGM_EINVAL, "ctx is NULL");
  int64_t total = 0;
  for (auto &kv : c->cache) total += (int64_t)kv.second.nbytes;
  if (n_splits) *n_splits = (int64_t)c->cache.size();
  if (bytes) *bytes = total;
  return HGM_OK;
}

// ---- synthetic workload ------------------------------------------------------------------

int hgm_synth_centres(uint64_t seed, int32_t k, int32_t d, float *out) {
  if (!out || k <= 0 || d <= 0) return fail(nullptr, HGM_EINVAL, "bad argument");
  synth_centres_host(seed, k, d, out);
  return HGM_OK;
}
int hgm_synth_perturb(uint64_t seed, int32_t k, int32_t d, const float *centres, float *out) {
  if (!out || !centres || k <= 0 || d <= 0) return fail(nullptr, HGM_EINVAL, "bad argument");
  synth_perturb_host(seed, k, d, centres, out);
  return HGM_OK;
}
int hgm_synth_points_host(uint64_t seed, int64_t first, int64_t n, int32_t d, int32_t k,
                          const float *centres, float *out) {
  if (!out || !centres || k <= 0 || d <= 0 || n < 0) return fail(nullptr, HGM_EINVAL, "bad argument");
  synth_points_host(seed, first, n, d, k, centres, out);
  return HGM_OK;
}
int hgm_synth_points_device(hgm_ctx *c, uint64_t seed, int64_t first, int64_t n, int32_t d, int32_t k,
                            const float *centres_host, float *d_out) {
  NEED_CTX(c);
  if (!centres_host || (!d_out && n > 0) || k <= 0 || d <= 0 || n < 0) return fail(c, HGM_EINVAL, "bad argument");
  float *d_centres = nullptr;
  CU(c, cudaMalloc(&d_centres, (size_t)k * d * 4));
  cudaError_t e = cudaMemcpyAsync(d_centres, centres_host, (size_t)k * d * 4, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = launch_synth_points(seed, first, n, d, k, d_centres, d_out, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  cudaFree(d_centres);
  if (e != cudaSuccess) return fail(c, HGM_ECUDA, "synth: %s", cudaGetErrorString(e));
  c->total_launches += 1;
  return HGM_OK;
}

}  // extern "C"
