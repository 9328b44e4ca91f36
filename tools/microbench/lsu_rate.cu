// Shared-memory pipe rates on sm_100a for the access patterns the d == 2 map kernel can choose between:
// cycles per warp-instruction per SM, one CTA of 768 threads per SM (the shape of the kernel).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lsu_rate lsu_rate.cu && ./lsu_rate
// Patterns:
//   0  ATOMS.ADD (no result), bank = lane ("lane-private" accumulator word), random row
//   1  ATOMS.ADD with result, same addressing
//   2  ATOMS.ADD (no result), random word (the rank counter of the counting sort)
//   3  LDS.U8 from a random byte of a 36 KiB table
//   4  STS.64 to a random 8-byte slot (the scatter of the counting sort)
//   5  STS.64, bank = lane
//   6  LDS.U8 random + 5 x pattern 0 (the lane-private fixed-point combiner, memory side only)
//   7  ATOMS.ADD (no result), 16 private lanes (lane & 15): two lanes per bank pair
//   8  pattern 6 with four of the five atomics only
//   9  pattern 6 with the table byte fetched through the texture unit (tex1Dfetch from linear global memory)
//  10  pattern 6 with the table byte loaded from global memory through L1 (ld.global.nc.u8)
//  11  tex1Dfetch alone;  12  ld.global.nc.u8 alone
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 2048
#define UNROLL 8

template <int OP>
__global__ void __launch_bounds__(768, 1) rate_kernel(unsigned *out, long long *cycles, unsigned seed, int rows,
                                                      cudaTextureObject_t tex, const unsigned char *__restrict__ gtable) {
  extern __shared__ __align__(16) unsigned char smem[];
  unsigned *const words = reinterpret_cast<unsigned *>(smem);           // rows * 32 words
  unsigned char *const table = smem + (size_t)rows * 128;               // 36864 bytes
  for (int i = threadIdx.x; i < rows * 32 + 36864 / 4; i += blockDim.x) words[i] = i * 2654435761u >> 24;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31;
  unsigned s = seed + threadIdx.x * 747796405u + blockIdx.x * 2891336453u;
  unsigned acc = 0;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
    unsigned r[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) { s = s * 1664525u + 1013904223u; r[u] = s >> 8; }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const unsigned row = r[u] % (unsigned)rows;
      if (OP == 0) atomicAdd(&words[row * 32 + lane], r[u]);
      if (OP == 1) acc += atomicAdd(&words[row * 32 + lane], r[u]);
      if (OP == 2) atomicAdd(&words[r[u] % (unsigned)(rows * 32)], 1u);
      if (OP == 3) acc += table[r[u] % 36864u];
      if (OP == 4) reinterpret_cast<uint2 *>(words)[r[u] % (unsigned)(rows * 16)] = make_uint2(r[u], s);
      if (OP == 5) reinterpret_cast<uint2 *>(words)[row * 16 + (lane & 15)] = make_uint2(r[u], s);
      if (OP == 6 || OP == 8) {
        const unsigned c = table[r[u] % 36864u] & 127u;  // "cluster" from the table: dependent address like the real thing
        unsigned *p = &words[c * 5 * 32 + lane];
        atomicAdd(p, r[u]); atomicAdd(p + 32, s); atomicAdd(p + 64, r[u] ^ s); atomicAdd(p + 96, r[u] + s);
        if (OP == 6) atomicAdd(p + 128, 1u);
      }
      if (OP == 7) atomicAdd(&words[row * 16 + (lane & 15)], r[u]);
      if (OP == 9 || OP == 10) {
        const unsigned cell = r[u] % 65536u;
        const unsigned c = (OP == 9 ? (unsigned)tex1Dfetch<unsigned char>(tex, (int)cell) : (unsigned)__ldg(gtable + cell)) & 127u;
        unsigned *p = &words[c * 5 * 32 + lane];
        atomicAdd(p, r[u]); atomicAdd(p + 32, s); atomicAdd(p + 64, r[u] ^ s); atomicAdd(p + 96, r[u] + s);
        atomicAdd(p + 128, 1u);
      }
      if (OP == 11) acc += tex1Dfetch<unsigned char>(tex, (int)(r[u] % 65536u));
      if (OP == 12) acc += __ldg(gtable + r[u] % 65536u);
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char *name, int per_iter) {
  unsigned *out; long long *cyc;
  cudaMalloc(&out, 148 * 768 * 4); cudaMalloc(&cyc, 148 * 8);
  const int rows = 640;  // 128 clusters x 5 words, 80 KiB
  const size_t smem = (size_t)rows * 128 + 36864;
  static unsigned char *gtable = nullptr;
  static cudaTextureObject_t tex = 0;
  if (!gtable) {
    cudaMalloc(&gtable, 65536);
    unsigned char h[65536];
    for (int i = 0; i < 65536; ++i) h[i] = (unsigned char)((i * 2654435761u) >> 24);
    cudaMemcpy(gtable, h, sizeof h, cudaMemcpyHostToDevice);
    cudaResourceDesc rd{};
    rd.resType = cudaResourceTypeLinear;
    rd.res.linear.devPtr = gtable;
    rd.res.linear.desc = cudaCreateChannelDesc<unsigned char>();
    rd.res.linear.sizeInBytes = 65536;
    cudaTextureDesc td{};
    td.readMode = cudaReadModeElementType;
    cudaCreateTextureObject(&tex, &rd, &td, nullptr);
  }
  cudaFuncSetAttribute(rate_kernel<OP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int rep = 0; rep < 2; ++rep) rate_kernel<OP><<<148, 768, smem>>>(out, cyc, 12345u, rows, tex, gtable);
  long long h[148];
  cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaGetLastError();
  const double warp_instr = (double)ITERS * UNROLL * 24 * per_iter;
  printf("%-44s %6.2f cycles per warp-instruction per SM   %7.2f per 32 points  (%s)\n", name,
         (double)h[0] / warp_instr, (double)h[0] / ((double)ITERS * UNROLL * 24), cudaGetErrorString(e));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("ATOMS.ADD no result, bank = lane", 1);
  run<1>("ATOMS.ADD with result, bank = lane", 1);
  run<7>("ATOMS.ADD no result, 16 private lanes", 1);
  run<2>("ATOMS.ADD no result, random word", 1);
  run<3>("LDS.U8 random byte of 36 KiB", 1);
  run<4>("STS.64 random slot", 1);
  run<5>("STS.64 bank = lane", 1);
  run<6>("LDS.U8 random + 5 ATOMS bank = lane", 6);
  run<8>("LDS.U8 random + 4 ATOMS bank = lane", 5);
  run<9>("tex1Dfetch u8 (64 KiB) + 5 ATOMS bank = lane", 6);
  run<10>("ld.global.nc.u8 (64 KiB) + 5 ATOMS bank = lane", 6);
  run<11>("tex1Dfetch u8 random byte of 64 KiB", 1);
  run<12>("ld.global.nc.u8 random byte of 64 KiB", 1);
  return 0;
}
