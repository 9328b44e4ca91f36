// kernels_exact.cu -- the brute-force FP32 map kernel: any d, any k, no index.
//
// This is the reference scan itself (oracle/kmeans_oracle.c hgo_nearest states the same
// contract on the CPU): one thread per point walks every centroid in ascending index,
// distance = FP32 sum_j (x_j - c_j)^2 with j ascending and no FMA contraction, strict '<'
// from (+inf, 0).  It fills the map-body slot Mapper::map (c++/pipes/api/hadoop/Pipes.hh:
// 158-161) and the combiner slot Reducer::reduce (Pipes.hh:166-169) for shapes the
// specialised kernels do not take, and it is the re-check kernel behind them.
//
// The per-(point, centroid) sum must run over j sequentially to round like the scan does,
// so the parallel axes are points (threads) and, in registers, 8 centroids per thread at a time;
// centroids stream through shared memory in transposed chunks and are read as warp-wide broadcasts.
#include "hgm_internal.h"

namespace hgm {
namespace {

constexpr int kThreads = 128;
constexpr int kSmemHistLimit = 8192;  // clusters counted in shared memory per CTA (else straight to global)

// Scan only: assign[p] and the per-cluster counts of this launch.  The sums follow as a segmented
// reduction (launch_segmented_sums, kernels_tensor.cu).  The first version added every point with d FP64
// atomics -- CAS loops in shared memory, REDs in global: with few clusters they all collide (d=5, k=2:
// 10 ms per 100 M points against 6.4 ms at k=64), with wide rows they dominate (d=300, k=9: 58 ms per 4 M).
template <int D>
__global__ void __launch_bounds__(kThreads)
    exact_map_kernel(const float *__restrict__ pts, long long n, int d_rt,
                     const float *__restrict__ cent, int k, int kc, int smem_hist,
                     unsigned int *__restrict__ lcounts, int32_t *__restrict__ assign) {
  const int d = D > 0 ? D : d_rt;
  extern __shared__ __align__(16) unsigned char raw[];
  // layout: [hist u32 k] (if smem_hist) | centroid chunk | xT d*kThreads (D==0)
  unsigned int *hist_s = reinterpret_cast<unsigned int *>(raw);
  float *cent_s = reinterpret_cast<float *>(  // 16-byte aligned: read as float4
      (reinterpret_cast<uintptr_t>(hist_s + (smem_hist ? k : 0)) + 15) & ~uintptr_t(15));
  float *xT = cent_s + (size_t)((kc + 7) & ~7) * d;
  const int t = threadIdx.x;

  if (smem_hist)
    for (int i = t; i < k; i += kThreads) hist_s[i] = 0u;
  __syncthreads();

  const long long ntiles = (n + kThreads - 1) / kThreads;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long p = tile * kThreads + t;
    const bool live = p < n;
    float xr[D > 0 ? D : 1];
    if (D > 0) {
#pragma unroll
      for (int j = 0; j < D; ++j) xr[j] = live ? pts[p * D + j] : 0.f;
    } else {
      // stage this tile's rows transposed: xT[j][thread]; coalesced global reads
      const long long base = tile * kThreads * (long long)d;
      const long long lim = n * (long long)d;
      for (int i = t; i < kThreads * d; i += kThreads) {
        const long long gidx = base + i;
        const float v = gidx < lim ? pts[gidx] : 0.f;
        xT[(i % d) * kThreads + (i / d)] = v;
      }
    }
    float best = INFINITY;
    int arg = 0;
    // d > 8: centroids stream through shared memory in chunks stored TRANSPOSED (cT[j][c], c padded to 8): a
    // thread scores 8 centroids at a time -- per coordinate one x (one LDS) and two
    // broadcast LDS.128 feed 24 FP32 instructions, instead of one LDS per 3 as a row-major chunk costs.
    // Every centroid keeps its own accumulator and sees j = 0..d-1 in order, so the rounding is the
    // scan's; the 8 results are then compared in ascending index with strict '<'.
    const int kcp = (kc + 7) & ~7;
    for (int c0 = 0; c0 < k; c0 += kc) {
      const int m = min(kc, k - c0);
      __syncthreads();  // previous chunk fully consumed (and xT staged)
      if constexpr (D > 0) {
        // d <= 8: the point sits in registers and the scan is bound by FP32 issue (3 unfused
        // instructions per coordinate): row-major chunk, one centroid at a time (measured faster here)
        for (int i = t; i < m * d; i += kThreads) cent_s[i] = cent[(size_t)c0 * d + i];
        __syncthreads();
        for (int c = 0; c < m; ++c) {
          float a = 0.0f;
#pragma unroll
          for (int j = 0; j < D; ++j) a = __fadd_rn(a, sqdiff(xr[j], cent_s[c * D + j]));
          if (a < best) { best = a; arg = c0 + c; }
        }
      } else {
      for (int i = t; i < m * d; i += kThreads) cent_s[(i % d) * kcp + (i / d)] = cent[(size_t)c0 * d + i];
      __syncthreads();
      for (int c = 0; c < m; c += 8) {
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = 0.0f;
        for (int j = 0; j < d; ++j) {
          const float x = xT[j * kThreads + t];
          const float4 u = *reinterpret_cast<const float4 *>(cent_s + j * kcp + c);
          const float4 v = *reinterpret_cast<const float4 *>(cent_s + j * kcp + c + 4);
          a[0] = __fadd_rn(a[0], sqdiff(x, u.x)); a[1] = __fadd_rn(a[1], sqdiff(x, u.y));
          a[2] = __fadd_rn(a[2], sqdiff(x, u.z)); a[3] = __fadd_rn(a[3], sqdiff(x, u.w));
          a[4] = __fadd_rn(a[4], sqdiff(x, v.x)); a[5] = __fadd_rn(a[5], sqdiff(x, v.y));
          a[6] = __fadd_rn(a[6], sqdiff(x, v.z)); a[7] = __fadd_rn(a[7], sqdiff(x, v.w));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (c + i < m && a[i] < best) { best = a[i]; arg = c0 + c + i; }
      }
      }
    }
    if (live) {
      assign[p] = arg;
      atomicAdd(smem_hist ? &hist_s[arg] : &lcounts[arg], 1u);
    }
    __syncthreads();  // xT is rewritten by the next tile
  }
  if (smem_hist) {
    __syncthreads();
    for (int i = t; i < k; i += kThreads)
      if (hist_s[i]) atomicAdd(&lcounts[i], hist_s[i]);
  }
}

__global__ void bswap32_kernel(uint32_t *w, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    w[i] = __byte_perm(w[i], 0, 0x0123);
}

}  // namespace

cudaError_t launch_exact_map(const float *d_pts, int64_t n, int d, const float *d_cent, int k, int32_t *d_assign,
                             unsigned int *d_meta, int32_t *d_order, double *d_sums, unsigned long long *d_counts,
                             int sm_count, size_t smem_optin, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (!d_assign || n > 0x7FFFFFFFll) return cudaErrorInvalidValue;
  const int smem_hist = k <= kSmemHistLimit ? 1 : 0;
  const size_t hist_bytes = smem_hist ? (((size_t)k * 4 + 15) & ~size_t(15)) : 0;
  const bool reg_path = d <= 8;
  const size_t xt_bytes = reg_path ? 0 : (size_t)d * kThreads * 4;
  if (hist_bytes + xt_bytes + (size_t)d * 32 > smem_optin) return cudaErrorInvalidValue;
  size_t room = smem_optin - hist_bytes - xt_bytes;
  if (room > 64 * 1024) room = 64 * 1024;  // keep a few CTAs per SM
  int kc = (int)(room / ((size_t)d * 4)) & ~7;  // the chunk is stored transposed, 8 centroids per step
  if (kc < 8) return cudaErrorInvalidValue;
  if (kc > k) kc = k;
  const size_t smem = hist_bytes + (size_t)((kc + 7) & ~7) * d * 4 + xt_bytes;

  void (*kern)(const float *, long long, int, const float *, int, int, int, unsigned int *, int32_t *) = nullptr;
  switch (reg_path ? d : 0) {
    case 1: kern = exact_map_kernel<1>; break;
    case 2: kern = exact_map_kernel<2>; break;
    case 3: kern = exact_map_kernel<3>; break;
    case 4: kern = exact_map_kernel<4>; break;
    case 5: kern = exact_map_kernel<5>; break;
    case 6: kern = exact_map_kernel<6>; break;
    case 7: kern = exact_map_kernel<7>; break;
    case 8: kern = exact_map_kernel<8>; break;
    default: kern = exact_map_kernel<0>; break;
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(d_meta, 0, (size_t)k * 4, s);  // the launch's per-cluster counts
  if (e != cudaSuccess) return e;
  const long long ntiles = (n + kThreads - 1) / kThreads;
  int per_sm = (int)(smem_optin / (smem + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  long long grid = (long long)sm_count * per_sm;
  if (grid > ntiles) grid = ntiles;
  kern<<<(int)grid, kThreads, smem, s>>>(d_pts, (long long)n, d, d_cent, k, kc, smem_hist, d_meta, d_assign);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  return launch_segmented_sums(d_pts, n, d, k, d_assign, d_meta, d_order, d_sums, d_counts, sm_count, s);
}

cudaError_t launch_bswap32(uint32_t *d_words, int64_t n_words, cudaStream_t s) {
  if (n_words <= 0) return cudaSuccess;
  long long blocks = (n_words + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  bswap32_kernel<<<(int)blocks, 256, 0, s>>>(d_words, (long long)n_words);
  return cudaGetLastError();
}

}  // namespace hgm
