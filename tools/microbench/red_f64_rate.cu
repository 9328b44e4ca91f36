// How fast can the SMs push FP64 reductions into a k x d table that lives in L2?  The question behind fusing the
// per-cluster sums (K4) into the assignment kernel (K2): every decided row of a point tile would leave as d
// RED.ADD.F64, a warp per row, 256 contiguous bytes per warp instruction, cluster = effectively random.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_f64_rate red_f64_rate.cu && ./red_f64_rate
// Prints ms for 40 M rows x 64 columns into 1024 clusters (config 4's shard) with W warps per SM doing nothing else,
// and with rows whose values come from global memory (the 10.2 GB the fused form would not read again is left out).
#include <cstdio>
#include <cuda_runtime.h>

template <int D>
__global__ void red_rows(double *__restrict__ sums, long long rows, int k, unsigned seed) {
  const int lane = threadIdx.x & 31;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long r = warp; r < rows; r += nwarps) {
    unsigned h = (unsigned)r * 2654435761u + seed;
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    const int c = (int)(h % (unsigned)k);
    double *row = sums + (size_t)c * D;
    const double v = (double)(int)(h & 1023u) * 0.125;
#pragma unroll
    for (int j = 0; j < D / 32; ++j) atomicAdd(row + j * 32 + lane, v + j);  // result unused: RED.E.ADD.F64
  }
}

template <int D>
void run(int warps, long long rows, int k) {
  double *sums;
  cudaMalloc(&sums, (size_t)k * D * 8);
  cudaMemset(sums, 0, (size_t)k * D * 8);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9f;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(a);
    red_rows<D><<<148, warps * 32>>>(sums, rows, k, 17u + rep);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (ms < best) best = ms;
  }
  printf("d=%3d k=%4d rows=%lld, %2d warps per SM: %8.3f ms  = %6.1f G adds/s  (%s)\n", D, k, rows, warps, best,
         (double)rows * D / best * 1e-6, cudaGetErrorString(cudaGetLastError()));
  cudaFree(sums);
}

int main() {
  for (int w : {2, 4, 8, 16}) run<64>(w, 40000000LL, 1024);
  for (int w : {2, 8}) run<128>(w, 8000000LL, 4096);
  for (int w : {2, 8}) run<64>(w, 40000000LL, 16);  // few clusters: same-address pressure
  return 0;
}
