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

int hgm_allreduce_partials(hgm_ctx *c) {
  int dev, k, d;
  cudaStream_t st;
  void **slot;
  if (!c) return hgm::ctx_fail(nullptr, HGM_EINVAL, "ctx is NULL");
  hgm::ctx_shape(c, &dev, &k, &d, &st, &slot);
  if (!*slot) return hgm::ctx_fail(c, HGM_EINVAL, "hgm_comm_init has not been called");
  double *sums = nullptr;
  int64_t *counts = nullptr;
  int rc = hgm_partials_device(c, &sums, &counts);
  if (rc != HGM_OK) return rc;
  cudaSetDevice(dev);
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
  return HGM_OK;
}

}  // extern "C"

namespace hgm {
void comm_destroy(void *comm) {
  if (comm && g_nccl.destroy) g_nccl.destroy((NcclComm)comm);
}
}  // namespace hgm
