// hgm_comm.cu -- the one collective of the path: a sum-allreduce of the per-GPU partials
// [k*d FP64 sums][k int64 counts] across the GPU map slots of a box, over NVLink/NVSwitch.
// It stands in for the intra-box leg of the reference's map->reduce shuffle of partial
// records (mapred/TaskTracker.java:4050 MapOutputServlet).  The
// payload is 3 KB (d=2, k=128) to 4.2 MB (d=128, k=4096): latency-bound, so NCCL's own
// kernels are used as they are, loaded with dlopen so that single-GPU hosts need no NCCL.
// ONE collective per iteration: the int64 counts are turned into FP64 in place (exact below 2^53, and a
// count cannot exceed the number of points), the whole buffer goes through one ncclAllReduce of k*(d+1)
// doubles, and the counts are turned back.  Round 1 issued two grouped collectives (FP64 sums, int64
// counts): 20 / 25 / 36 us per step at 2 / 4 / 8 GPUs for 3 KB, the limiter of the 1 -> 8 scaling curve.
#include <dlfcn.h>

#include "hgm_internal.h"

// hgm_ctx internals needed here (kept in hgm_api.cu); accessors avoid exposing the struct.
extern "C" int hgm_partials_device(hgm_ctx *, double **, int64_t **);
namespace hgm {
int ctx_shape(hgm_ctx *c, int *device, int *k, int *d, cudaStream_t *stream, void ***comm_slot);
int ctx_fail(hgm_ctx *c, int code, const char *msg);
void **ctx_peer_slot(hgm_ctx *c);
void ctx_count_launches(hgm_ctx *c, int n);
void peer_destroy(void *state);
}  // namespace hgm

namespace {

struct NcclId { char internal[128]; };
typedef void *NcclComm;
typedef int (*fn_get_id)(NcclId *);
typedef int (*fn_init_rank)(NcclComm *, int, NcclId, int);
typedef int (*fn_allreduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_void)(void);
typedef int (*fn_destroy)(NcclComm);
typedef const char *(*fn_errstr)(int);

struct Nccl {
  void *lib = nullptr;
  fn_get_id get_id = nullptr;
  fn_init_rank init_rank = nullptr;
  fn_allreduce allreduce = nullptr;
  fn_void group_start = nullptr, group_end = nullptr;
  fn_destroy destroy = nullptr;
  fn_errstr errstr = nullptr;
  char why[256] = "";
} g_nccl;

constexpr int kNcclFloat64 = 8, kNcclSum = 0;  // nccl.h enum values (stable ABI)

bool load_nccl() {
  if (g_nccl.lib) return true;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char *n : names) {
    g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib) break;
  }
  if (!g_nccl.lib) {
    snprintf(g_nccl.why, sizeof g_nccl.why, "cannot load libnccl.so.2: %s", dlerror());
    return false;
  }
  g_nccl.get_id = (fn_get_id)dlsym(g_nccl.lib, "ncclGetUniqueId");
  g_nccl.init_rank = (fn_init_rank)dlsym(g_nccl.lib, "ncclCommInitRank");
  g_nccl.allreduce = (fn_allreduce)dlsym(g_nccl.lib, "ncclAllReduce");
  g_nccl.group_start = (fn_void)dlsym(g_nccl.lib, "ncclGroupStart");
  g_nccl.group_end = (fn_void)dlsym(g_nccl.lib, "ncclGroupEnd");
  g_nccl.destroy = (fn_destroy)dlsym(g_nccl.lib, "ncclCommDestroy");
  g_nccl.errstr = (fn_errstr)dlsym(g_nccl.lib, "ncclGetErrorString");
  if (!g_nccl.get_id || !g_nccl.init_rank || !g_nccl.allreduce || !g_nccl.group_start ||
      !g_nccl.group_end || !g_nccl.destroy || !g_nccl.errstr) {
    snprintf(g_nccl.why, sizeof g_nccl.why, "libnccl is missing an expected symbol");
    dlclose(g_nccl.lib);
    g_nccl.lib = nullptr;
    return false;
  }
  return true;
}

// counts <-> FP64 in place; one small CTA each side of the collective, on the same stream
__global__ void counts_to_f64_kernel(long long *counts, int k) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < k) reinterpret_cast<double *>(counts)[i] = (double)counts[i];
}
__global__ void f64_to_counts_kernel(long long *counts, int k) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < k) counts[i] = __double2ll_rn(reinterpret_cast<double *>(counts)[i]);
}

// ---- small partials: one-shot exchange through peer memory --------------------------------------------------
// For the d == 2 shapes the payload is a few KB and the map kernel takes under half a millisecond, so the
// collective's LATENCY is what the 1 -> 8 GPU curve loses (NCCL: 23 / 28 / 38 us per step at 2 / 4 / 8 GPUs
// against a 0.45 ms kernel).  Every rank owns an inbox of [2 parities][n_ranks] slots in its own HBM, mapped into
// the other ranks' address spaces with CUDA IPC.  One CTA per rank, right behind the map kernel on the same
// stream: it stores this rank's partials into slot [parity][rank] of EVERY inbox (plain stores over NVLink),
// fences, raises its flag in every inbox, waits for the n_ranks flags of its own inbox, and adds the slots up in
// rank order -- the same order on every rank, so every rank ends with bit-identical sums -- into its partials.
// Two parities: a rank can be at most one exchange ahead of the slowest (it needs that rank's flag of the
// current exchange to finish), so the slot it writes next is never one a peer is still reading.  The wait is
// bounded: a peer that never arrives ends as a trapped kernel and a failed task, never as a hung GPU.
// Plain loads and stores only; nothing here is a TMA copy.
constexpr int kPeerMaxRanks = 16;
constexpr size_t kPeerMaxPayload = 64 * 1024;  // bytes of partials the one-CTA exchange takes; beyond it NCCL is used

struct PeerState {
  int n_ranks = 0, rank = 0;
  size_t slot_bytes = 0;          // payload capacity of a slot
  unsigned char *inbox = nullptr; // own: [flags 2 * kPeerMaxRanks * 8 B][2][n_ranks][slot_bytes]
  unsigned char *peers[kPeerMaxRanks] = {};  // every rank's inbox in THIS process's address space (own included)
  unsigned long long epoch = 0;
};
struct PeerArgs {
  unsigned char *peers[kPeerMaxRanks];
  int n_ranks, rank;
  unsigned long long epoch;
  size_t slot_bytes;
};
__host__ __device__ inline size_t peer_flag_bytes() { return (size_t)2 * kPeerMaxRanks * 8; }
__host__ __device__ inline size_t peer_inbox_bytes(int n_ranks, size_t slot_bytes) {
  return peer_flag_bytes() + (size_t)2 * n_ranks * slot_bytes;
}

__global__ void __launch_bounds__(1024) peer_allreduce_kernel(PeerArgs a, double *sums, long long *counts, int n_sums, int k) {
  const int parity = (int)(a.epoch & 1ull);
  const int words = n_sums + k;  // 8-byte words of payload: sums then counts, as they lie in the partials
  const unsigned long long *mine = reinterpret_cast<const unsigned long long *>(sums);
  for (int r = 0; r < a.n_ranks; ++r) {  // my partials -> slot [parity][rank] of every inbox
    unsigned long long *slot = reinterpret_cast<unsigned long long *>(
        a.peers[r] + peer_flag_bytes() + ((size_t)parity * a.n_ranks + a.rank) * a.slot_bytes);
    for (int i = threadIdx.x; i < words; i += blockDim.x) slot[i] = mine[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < a.n_ranks) {  // flag [parity][rank] of inbox threadIdx.x := epoch
    volatile unsigned long long *flag =
        reinterpret_cast<volatile unsigned long long *>(a.peers[threadIdx.x]) + parity * kPeerMaxRanks + a.rank;
    *flag = a.epoch;
    __threadfence_system();
    // ... and wait for rank threadIdx.x's flag in MY inbox
    volatile unsigned long long *want =
        reinterpret_cast<volatile unsigned long long *>(a.peers[a.rank]) + parity * kPeerMaxRanks + threadIdx.x;
    const long long t0 = clock64();
    while (*want < a.epoch) {
      if (clock64() - t0 > 4000000000ll) __trap();  // ~2 s: the peer is gone
      __nanosleep(100);
    }
    __threadfence_system();
  }
  __syncthreads();
  const unsigned char *base = a.peers[a.rank] + peer_flag_bytes() + (size_t)parity * a.n_ranks * a.slot_bytes;
  for (int i = threadIdx.x; i < words; i += blockDim.x) {
    if (i < n_sums) {
      double s = 0.0;
      for (int r = 0; r < a.n_ranks; ++r) s += reinterpret_cast<const volatile double *>(base + (size_t)r * a.slot_bytes)[i];
      sums[i] = s;
    } else {
      long long s = 0;
      for (int r = 0; r < a.n_ranks; ++r) s += reinterpret_cast<const volatile long long *>(base + (size_t)r * a.slot_bytes)[i];
      counts[i - n_sums] = s;
    }
  }
}

int nccl_fail(hgm_ctx *c, const char *what, int rc) {
  char buf[256];
  snprintf(buf, sizeof buf, "%s: %s", what, g_nccl.errstr ? g_nccl.errstr(rc) : "?");
  return hgm::ctx_fail(c, HGM_ENCCL, buf);
}

}  // namespace

extern "C" {

int hgm_comm_id(void *id128) {
  if (!id128) return hgm::ctx_fail(nullptr, HGM_EINVAL, "id128 is NULL");
  if (!load_nccl()) return hgm::ctx_fail(nullptr, HGM_ENCCL, g_nccl.why);
  NcclId id;
  const int rc = g_nccl.get_id(&id);
  if (rc != 0) return nccl_fail(nullptr, "ncclGetUniqueId", rc);
  memcpy(id128, &id, sizeof id);
  return HGM_OK;
}

int hgm_comm_init(hgm_ctx *c, int32_t n_ranks, int32_t rank, const void *id128) {
  int dev, k, d;
  cudaStream_t st;
  void **slot;
  if (!c || !id128 || n_ranks < 1 || rank < 0 || rank >= n_ranks)
    return hgm::ctx_fail(c, HGM_EINVAL, "bad communicator arguments");
  if (!load_nccl()) return hgm::ctx_fail(c, HGM_ENCCL, g_nccl.why);
  hgm::ctx_shape(c, &dev, &k, &d, &st, &slot);
  cudaSetDevice(dev);
  if (*slot) { g_nccl.destroy((NcclComm)*slot); *slot = nullptr; }
  NcclId id;
  memcpy(&id, id128, sizeof id);
  NcclComm comm = nullptr;
  const int rc = g_nccl.init_rank(&comm, n_ranks, id, rank);
  if (rc != 0) return nccl_fail(c, "ncclCommInitRank", rc);
  *slot = comm;
  return HGM_OK;
}

int hgm_peer_handle(hgm_ctx *c, int32_t n_ranks, void *handle64) {
  if (!c || !handle64 || n_ranks < 1 || n_ranks > kPeerMaxRanks) return hgm::ctx_fail(c, HGM_EINVAL, "bad peer arguments");
  int dev, k, d;
  cudaStream_t st;
  void **nccl;
  hgm::ctx_shape(c, &dev, &k, &d, &st, &nccl);
  cudaSetDevice(dev);
  void **slot = hgm::ctx_peer_slot(c);
  if (*slot) { hgm::peer_destroy(*slot); *slot = nullptr; }
  PeerState *p = new PeerState();
  p->n_ranks = n_ranks;
  p->slot_bytes = kPeerMaxPayload;
  const size_t bytes = peer_inbox_bytes(n_ranks, p->slot_bytes);
  cudaError_t e = cudaMalloc(&p->inbox, bytes);
  if (e == cudaSuccess) e = cudaMemset(p->inbox, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p->inbox);
  if (e != cudaSuccess) {
    cudaGetLastError();
    cudaFree(p->inbox);
    delete p;
    return hgm::ctx_fail(c, HGM_ECUDA, cudaGetErrorString(e));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the handle travels as 64 bytes");
  memcpy(handle64, &h, 64);
  *slot = p;
  return HGM_OK;
}

int hgm_peer_open(hgm_ctx *c, int32_t rank, const void *handles) {
  if (!c || !handles) return hgm::ctx_fail(c, HGM_EINVAL, "bad peer arguments");
  PeerState *p = static_cast<PeerState *>(*hgm::ctx_peer_slot(c));
  if (!p) return hgm::ctx_fail(c, HGM_EINVAL, "hgm_peer_handle has not been called");
  if (rank < 0 || rank >= p->n_ranks) return hgm::ctx_fail(c, HGM_EINVAL, "rank out of range");
  int dev, k, d;
  cudaStream_t st;
  void **nccl;
  hgm::ctx_shape(c, &dev, &k, &d, &st, &nccl);
  cudaSetDevice(dev);
  p->rank = rank;
  for (int r = 0; r < p->n_ranks; ++r) {
    if (r == rank) { p->peers[r] = p->inbox; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const unsigned char *>(handles) + (size_t)r * 64, 64);
    void *ptr = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      char msg[200];
      snprintf(msg, sizeof msg, "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
      for (int q = 0; q < r; ++q)
        if (q != rank && p->peers[q]) { cudaIpcCloseMemHandle(p->peers[q]); p->peers[q] = nullptr; }
      p->peers[rank] = nullptr;
      return hgm::ctx_fail(c, HGM_ECUDA, msg);
    }
    p->peers[r] = static_cast<unsigned char *>(ptr);
  }
  p->epoch = 0;
  return HGM_OK;
}

int hgm_peer_close(hgm_ctx *c) {
  if (!c) return hgm::ctx_fail(nullptr, HGM_EINVAL, "ctx is NULL");
  void **slot = hgm::ctx_peer_slot(c);
  if (*slot) {
    int dev, k, d;
    cudaStream_t st;
    void **nccl;
    hgm::ctx_shape(c, &dev, &k, &d, &st, &nccl);
    cudaSetDevice(dev);
    cudaStreamSynchronize(st);
    hgm::peer_destroy(*slot);
    *slot = nullptr;
  }
  return HGM_OK;
}

int hgm_allreduce_partials(hgm_ctx *c) {
  HGM_RANGE("hgm_allreduce_partials");
  int dev, k, d;
  cudaStream_t st;
  void **slot;
  if (!c) return hgm::ctx_fail(nullptr, HGM_EINVAL, "ctx is NULL");
  hgm::ctx_shape(c, &dev, &k, &d, &st, &slot);
  double *sums = nullptr;
  int64_t *counts = nullptr;
  int rc = hgm_partials_device(c, &sums, &counts);
  if (rc != HGM_OK) return rc;
  cudaSetDevice(dev);
  PeerState *p = static_cast<PeerState *>(*hgm::ctx_peer_slot(c));
  if (p && p->peers[p->rank] && ((size_t)k * d + (size_t)k) * 8 <= p->slot_bytes) {  // small partials: peer memory
    PeerArgs a{};
    for (int r = 0; r < p->n_ranks; ++r) a.peers[r] = p->peers[r];
    a.n_ranks = p->n_ranks;
    a.rank = p->rank;
    a.epoch = ++p->epoch;
    a.slot_bytes = p->slot_bytes;
    peer_allreduce_kernel<<<1, 1024, 0, st>>>(a, sums, reinterpret_cast<long long *>(counts), k * d, k);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return hgm::ctx_fail(c, HGM_ECUDA, cudaGetErrorString(e));
    hgm::ctx_count_launches(c, 1);
    return HGM_OK;
  }
  if (!*slot) return hgm::ctx_fail(c, HGM_EINVAL, "hgm_comm_init has not been called");
  // the counts follow the sums in one buffer (hgm_api.cu: [k*d FP64][k int64])
  if (reinterpret_cast<int64_t *>(sums + (size_t)k * d) != counts)
    return hgm::ctx_fail(c, HGM_EINVAL, "partials are not one contiguous buffer");
  const int blocks = (k + 255) / 256;
  counts_to_f64_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<long long *>(counts), k);
  const int n = g_nccl.allreduce(sums, sums, (size_t)k * (d + 1), kNcclFloat64, kNcclSum, (NcclComm)*slot, st);
  // turned back whatever the collective said: the partials must not be left half converted
  f64_to_counts_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<long long *>(counts), k);
  if (n != 0) return nccl_fail(c, "ncclAllReduce(partials)", n);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return hgm::ctx_fail(c, HGM_ECUDA, cudaGetErrorString(e));
  hgm::ctx_count_launches(c, 2);  // the two conversions; NCCL's own kernel is not ours to count
  return HGM_OK;
}

}  // extern "C"

namespace hgm {
void peer_destroy(void *state) {
  PeerState *p = static_cast<PeerState *>(state);
  if (!p) return;
  for (int r = 0; r < p->n_ranks; ++r)
    if (r != p->rank && p->peers[r] && p->peers[r] != p->inbox) cudaIpcCloseMemHandle(p->peers[r]);
  cudaFree(p->inbox);
  delete p;
}
void comm_destroy(void *comm) {
  if (comm && g_nccl.destroy) g_nccl.destroy((NcclComm)comm);
}
}  // namespace hgm
