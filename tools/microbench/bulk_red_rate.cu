// Two ways for ONE warp to add rows of 64 FP32 values to an FP64 table in L2 (the fused sums of K2's decider warps):
//   0  a warp per row, two REDG.E.ADD.F64 per lane-row (what csrc/kernels_tensor.cu does)
//   1  rows converted into a shared-memory staging buffer, one cp.reduce.async.bulk (.add.f64, 512 B) per row
//      issued by an elected lane: UBLKRED, the TMA engine does the adds
// Cycles per row per warp, with 2 warps per SM, on a few SMs (the warp's own issue cost) and on all 148 (L2 bound).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_red_rate bulk_red_rate.cu && ./bulk_red_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(64, 1) rows_kernel(double *__restrict__ sums, const float *__restrict__ src, int rows_per_warp,
                                                     int k, long long *cycles) {
  __shared__ float tile[2][64][64];               // "the point tile": 64 rows per warp
  __shared__ __align__(128) double stage[2][2][8][64];  // [warp][buffer][row][coordinate]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 2 * 64 * 64; i += 64) (&tile[0][0][0])[i] = src[i];
  __syncthreads();
  unsigned h = blockIdx.x * 977u + warp * 131u + 7u;
  const long long t0 = clock64();
  int buf = 0;
  for (int r0 = 0; r0 < rows_per_warp; r0 += 8) {
    unsigned cl[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) { h = h * 1664525u + 1013904223u; cl[q] = (h >> 8) % (unsigned)k; }
    if (MODE == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float *row = tile[warp][(r0 + q) & 63];
        double *out = sums + (size_t)cl[q] * 64 + lane;
        atomicAdd(out, (double)row[lane]);
        atomicAdd(out + 32, (double)row[lane + 32]);
      }
    } else {
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // the buffer used two rounds ago has been read
      __syncwarp();
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float *row = tile[warp][(r0 + q) & 63];
        stage[warp][buf][q][lane] = (double)row[lane];
        stage[warp][buf][q][lane + 32] = (double)row[lane + 32];
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
#pragma unroll
        for (int q = 0; q < 8; ++q)
          asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], 512;" ::"l"(
                           sums + (size_t)cl[q] * 64),
                       "r"(smem_u32(&stage[warp][buf][q][0]))
                       : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      __syncwarp();
      buf ^= 1;
    }
  }
  if (MODE == 1) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  const long long t1 = clock64();
  if (lane == 0) cycles[blockIdx.x * 2 + warp] = t1 - t0;
}

template <int MODE>
void run(const char *name, int grid, int rows_per_warp, int k) {
  double *sums; float *src; long long *cyc;
  cudaMalloc(&sums, (size_t)k * 64 * 8); cudaMemset(sums, 0, (size_t)k * 64 * 8);
  cudaMalloc(&src, 2 * 64 * 64 * 4); cudaMemset(src, 0, 2 * 64 * 64 * 4);
  cudaMalloc(&cyc, 148 * 2 * 8);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9f;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(a);
    rows_kernel<MODE><<<grid, 64>>>(sums, src, rows_per_warp, k, cyc);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  long long h[2]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  printf("%-28s grid %3d k=%4d: %7.3f ms, %6.1f cycles per row per warp, %6.1f G adds/s  (%s)\n", name, grid, k, best,
         (double)h[0] / rows_per_warp, (double)grid * 2 * rows_per_warp * 64 / best * 1e-6, cudaGetErrorString(cudaGetLastError()));
  cudaFree(sums); cudaFree(src); cudaFree(cyc);
}

int main() {
  const int rows = 135168;  // 40 M rows / 148 SMs / 2 warps
  for (int grid : {8, 148}) {
    run<0>("REDG.F64 x 2 per row", grid, rows, 1024);
    run<1>("UBLKRED 512 B per row", grid, rows, 1024);
  }
  run<0>("REDG.F64 x 2 per row", 148, rows, 16);
  run<1>("UBLKRED 512 B per row", 148, rows, 16);
  return 0;
}
