// hgm_static.cu -- hgm_map_device_static: the d == 2 map task for callers that must not
// allocate device memory.
//
// Who that is: a GPU map slot that runs under a device-memory budget it does not own.  A
// TaskTracker that has carved HBM into per-slot arenas for the resident split cache, or a
// host runtime that keeps every byte in its own pool (a JVM off-heap manager, an RMM-style
// arena), hands the task its input and output buffers and forbids cudaMalloc -- both because
// the bytes are spoken for and because cudaMalloc/cudaFree synchronise the whole device
// while other slots' kernels are running.  hgm_open + hgm_set_centroids allocate (index,
// partials, staging); this entry point does not: everything the kernels need beyond the
// caller's buffers lives in ONE file-scope device array, `g_scratch` below.
//
// Consequences a caller must know:
//  * every GPU that loads libhgpumap.so carries the array (kScratchBytes, ~1.22 MiB) until
//    the process exits, whether this entry point is used or not;
//  * the job is limited to what fits: d == 2, k <= kStaticMaxK, and a centroid table the cell
//    index can hold (finite, |coordinate| <= 1e18); anything else is refused with a message;
//  * all calls on a device share the array and it keeps the last call's contents.  The
//    partial sums in it are cleared at the start of each call; the index is rebuilt by each
//    call before it is read.  Two calls of this process that overlap on one GPU (issued from
//    different threads or on different streams) race on it: issue them one after the other;
//  * the first call on a device looks the array's address up (cudaGetSymbolAddress) and
//    reads the device attributes; make that first call outside any stream capture;
//  * every call copies the centroids in and the results out of pageable host memory and waits for
//    `cuda_stream` before it returns: the results are in the caller's host buffers then.  The
//    caller's current device is put back before returning.
#include <math.h>

#include <mutex>

#include "hgm_internal.h"

namespace hgm {
// defined in hgm_api.cu: the same planner hgm_set_centroids uses
bool plan_grid2d_for(const float *cent, int k, size_t smem_optin, Grid2D *out, size_t table_cap);
int report_error(int code, const char *msg);
}  // namespace hgm

namespace {

constexpr int kStaticMaxK = 1024;
constexpr int kStaticMaxG = 256;
constexpr size_t kTableBytes = ((size_t)kStaticMaxG * kStaticMaxG * 3 + 32 + 255) & ~size_t(255);  // 192 KiB: 16-bit table + 8-bit copy
constexpr size_t kCandBytes = (size_t)kStaticMaxG * kStaticMaxG * hgm::kMaxCand * 2;   // 1 MiB
constexpr size_t kCentBytes = (size_t)kStaticMaxK * 2 * 4;                             // 8 KiB
constexpr size_t kPartialBytes = (size_t)kStaticMaxK * (2 * 8 + 8);                    // 24 KiB
constexpr size_t kScratchBytes = kTableBytes + kCandBytes + kCentBytes + kPartialBytes;

// The scratch.  In host code this NAME is only a handle for the runtime, not a device
// address: the address each device gave it is fetched below and that pointer is what the
// kernels receive.
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

struct DeviceSetup {
  unsigned char *scratch = nullptr;  // device address of g_scratch on this device
  int sm_count = 0;
  size_t smem_optin = 0;
  bool ready = false;
};
constexpr int kMaxDevices = 64;
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_lock;

// once per device, with that device current
int setup_device(int device, DeviceSetup **out) {
  std::lock_guard<std::mutex> hold(g_setup_lock);
  DeviceSetup &s = g_setup[device];
  if (!s.ready) {
    void *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      char msg[256];
      snprintf(msg, sizeof msg, "cudaGetSymbolAddress(g_scratch) on device %d: %s", device,
               cudaGetErrorString(e));
      return hgm::report_error(HGM_ECUDA, msg);
    }
    int sms = 0, smem = 0, major = 0;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess ||
        (e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)) != cudaSuccess ||
        (e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device)) != cudaSuccess) {
      char msg[256];
      snprintf(msg, sizeof msg, "cudaDeviceGetAttribute on device %d: %s", device, cudaGetErrorString(e));
      return hgm::report_error(HGM_ECUDA, msg);
    }
    if (major != 10) return hgm::report_error(HGM_ENODEV, "this library carries sm_100a code only");
    s.scratch = static_cast<unsigned char *>(ptr);
    s.sm_count = sms;
    s.smem_optin = (size_t)smem;
    s.ready = true;
  }
  *out = &s;
  return HGM_OK;
}

}  // namespace

extern "C" int hgm_map_device_static(int device_id, const float *d_points, int64_t n_points,
                                     const float *centroids, int32_t k, int32_t d, double *sums_kxd,
                                     int64_t *counts_k, int32_t *d_assign_or_null, void *cuda_stream) {
  using namespace hgm;
  if (d != 2)
    return report_error(HGM_EINVAL, "hgm_map_device_static handles d == 2 only (its scratch is a fixed-size "
                                    "2-D cell index); use hgm_open/hgm_map_device for other shapes");
  if (k < 1 || k > kStaticMaxK) {
    char msg[200];
    snprintf(msg, sizeof msg, "hgm_map_device_static: k = %d is outside 1..%d, the most its static scratch holds",
             k, kStaticMaxK);
    return report_error(HGM_EINVAL, msg);
  }
  if (!centroids || !sums_kxd || !counts_k || n_points < 0 || (n_points > 0 && !d_points))
    return report_error(HGM_EINVAL, "hgm_map_device_static: NULL argument");
  if (reinterpret_cast<uintptr_t>(d_points) & 7)
    return report_error(HGM_EINVAL, "hgm_map_device_static: d_points must be 8-byte aligned");
  if (device_id < 0 || device_id >= kMaxDevices) return report_error(HGM_ENODEV, "device id out of range");
  int caller_device = -1;
  cudaGetDevice(&caller_device);
  cudaError_t e = cudaSetDevice(device_id);
  if (e != cudaSuccess) return report_error(HGM_ENODEV, cudaGetErrorString(e));
  struct RestoreDevice {  // leave the caller's current device as it was, on every path out
    int device;
    ~RestoreDevice() { if (device >= 0) cudaSetDevice(device); }
  } restore{caller_device};
  DeviceSetup *dev = nullptr;
  int rc = setup_device(device_id, &dev);
  if (rc != HGM_OK) return rc;

  Grid2D grid;
  if (!plan_grid2d_for(centroids, k, dev->smem_optin, &grid, kTableBytes) || grid.G > kStaticMaxG ||
      grid2d_table_bytes(grid.G) > kTableBytes + (grid.shape == kShapeFx ? 0 : (size_t)grid.G * grid.G))
    return report_error(HGM_EINVAL, "hgm_map_device_static: this centroid table cannot be indexed (non-finite or "
                                    "huge coordinates); use hgm_open/hgm_map_device, which fall back to the exact scan");

  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  uint16_t *table = reinterpret_cast<uint16_t *>(dev->scratch);
  uint16_t *cand = reinterpret_cast<uint16_t *>(dev->scratch + kTableBytes);
  float *cent = reinterpret_cast<float *>(dev->scratch + kTableBytes + kCandBytes);
  double *sums = reinterpret_cast<double *>(dev->scratch + kTableBytes + kCandBytes + kCentBytes);
  unsigned long long *counts = reinterpret_cast<unsigned long long *>(sums + 2 * (size_t)k);

#define HGM_TRY(call)                                                        \
  do {                                                                       \
    cudaError_t e_ = (call);                                                 \
    if (e_ != cudaSuccess) {                                                 \
      char msg_[256];                                                        \
      snprintf(msg_, sizeof msg_, "%s: %s", #call, cudaGetErrorString(e_));  \
      return report_error(HGM_ECUDA, msg_);                                  \
    }                                                                        \
  } while (0)
  // the shared scratch still holds the previous call's sums and index: start from zero where
  // the kernels accumulate (partials; the padded tail of the centroid copy the TMA reads)
  HGM_TRY(cudaMemsetAsync(cent, 0, kCentBytes + kPartialBytes, stream));
  HGM_TRY(cudaMemcpyAsync(cent, centroids, (size_t)k * 2 * 4, cudaMemcpyHostToDevice, stream));
  HGM_TRY(launch_grid2d_build(cent, k, grid, table, cand, stream));
  HGM_TRY(launch_grid2d_map(d_points, n_points, cent, k, grid, table, cand, sums, counts, d_assign_or_null,
                            dev->sm_count, dev->smem_optin, stream));
  HGM_TRY(cudaMemcpyAsync(sums_kxd, sums, (size_t)k * 2 * 8, cudaMemcpyDeviceToHost, stream));
  HGM_TRY(cudaMemcpyAsync(counts_k, counts, (size_t)k * 8, cudaMemcpyDeviceToHost, stream));
  HGM_TRY(cudaStreamSynchronize(stream));
#undef HGM_TRY
  return HGM_OK;
}
