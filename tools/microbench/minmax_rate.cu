// Issue rate of the min/max family on sm_100a: cycles per warp-instruction per SM sub-partition.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o minmax_rate minmax_rate.cu && ./minmax_rate
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 4096

template <int OP>
__global__ void rate_kernel(float *out, long long *cycles, float seed) {
  float v[CHAINS];
  unsigned u[CHAINS];
  for (int i = 0; i < CHAINS; ++i) { v[i] = seed + threadIdx.x + i; u[i] = __float_as_uint(v[i]); }
  float w = seed * 3.0f, z = seed * 5.0f;
  unsigned uw = __float_as_uint(w), uz = __float_as_uint(z);
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (OP == 0) v[i] = fminf(v[i], w);                                   // FMNMX
      if (OP == 1) asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(w), "f"(z));  // FMNMX3
      if (OP == 2) asm volatile("min.s32 %0, %0, %1;" : "+r"(u[i]) : "r"(uw));             // VIMNMX
      if (OP == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(uw), "r"(uz));  // LOP3
      if (OP == 4) asm volatile("min.u32 %0, %0, %1;" : "+r"(u[i]) : "r"(uw));
      if (OP == 5) u[i] = (unsigned)min(min((int)u[i], (int)uw), (int)uz);                  // VIMNMX3 if the compiler fuses
      if (OP == 6) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(w), "f"(z));  // FFMA
      if (OP == 7) asm volatile("{.reg .b32 t; min.f16x2 t, %0, %1; mov.b32 %0, t;}" : "+r"(u[i]) : "r"(uw));  // HMNMX2
      if (OP == 8) asm volatile("{.reg .b32 t; min.bf16x2 t, %0, %1; mov.b32 %0, t;}" : "+r"(u[i]) : "r"(uw));
      if (OP == 9) asm volatile("{.reg .b32 t; min.s16x2 t, %0, %1; mov.b32 %0, t;}" : "+r"(u[i]) : "r"(uw));   // VIMNMX.S16x2
    }
    w += 1.0f; uw += 1;
  }
  const long long t1 = clock64();
  float acc = 0;
  for (int i = 0; i < CHAINS; ++i) acc += v[i] + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char *name, int threads) {
  float *out; long long *cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  rate_kernel<OP><<<148, threads>>>(out, cyc, 1.5f);
  rate_kernel<OP><<<148, threads>>>(out, cyc, 1.5f);
  long long h[148];
  cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaGetLastError();
  const double warps_per_smsp = threads / 32 / 4.0;
  const double per = (double)h[0] / ((double)ITERS * CHAINS * warps_per_smsp);
  printf("%-12s threads=%4d  cycles per warp-instruction per SMSP = %.2f  (%s)\n", name, threads, per, cudaGetErrorString(e));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int threads : {128, 512, 1024}) {
    run<0>("FMNMX", threads); run<1>("FMNMX3", threads); run<2>("min.s32", threads); run<4>("min.u32", threads);
    run<5>("min3.s32", threads); run<3>("LOP3", threads); run<6>("FFMA", threads); run<7>("min.f16x2", threads);
    run<8>("min.bf16x2", threads); run<9>("min.s16x2", threads);
  }
  return 0;
}
