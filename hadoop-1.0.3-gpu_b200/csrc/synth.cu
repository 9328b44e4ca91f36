// synth.cu -- the benchmark's synthetic workload: a mixture of k
// unit-variance blobs around centres uniform in [-100, 100)^d.  The reference publishes no
// workload, so this is the project's own; all randomness is integer hashing and every
// float operation is a single correctly rounded IEEE op (no FMA), so host and device emit
// the same bits for the same (seed, point index) and splits can be generated anywhere.
#include "hgm_internal.h"

namespace hgm {

__host__ __device__ inline uint64_t mix64(uint64_t seed, uint64_t ctr) {
  uint64_t z = seed + (ctr + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// Irwin-Hall(8): sum of eight 16-bit uniforms, centred; variance 8*(65536^2-1)/12
__host__ __device__ inline int32_t ih8(uint64_t a, uint64_t b) {
  int32_t s = 0;
  for (int i = 0; i < 4; ++i) {
    s += (int32_t)((a >> (16 * i)) & 0xFFFF);
    s += (int32_t)((b >> (16 * i)) & 0xFFFF);
  }
  return s - 262140;
}

__host__ __device__ inline float mul_rn(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fmul_rn(a, b);
#else
  volatile float r = a * b;  // volatile: keep the product from fusing into a later add
  return r;
#endif
}
__host__ __device__ inline float add_rn(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fadd_rn(a, b);
#else
  volatile float r = a + b;
  return r;
#endif
}

constexpr float kIh8Scale = 0x1.3988e2p-16f;  // 1/sqrt(8*(65536^2-1)/12)
constexpr float kCentreStep = 0x1.9p-17f;     // 200 * 2^-24

__host__ __device__ inline float synth_coord(uint64_t seed, uint64_t i, uint64_t stride, int j,
                                             float centre) {
  const int32_t v = ih8(mix64(seed, i * stride + 1 + 2 * (uint64_t)j),
                        mix64(seed, i * stride + 2 + 2 * (uint64_t)j));
  return add_rn(centre, mul_rn((float)v, kIh8Scale));
}

__host__ __device__ inline uint32_t synth_component(uint64_t seed, uint64_t i, uint64_t stride, int k) {
  const uint64_t h = mix64(seed, i * stride);
  return (uint32_t)(((h >> 32) * (uint64_t)k) >> 32);
}

void synth_centres_host(uint64_t seed, int k, int d, float *out) {
  for (int64_t i = 0; i < (int64_t)k * d; ++i) {
    const uint32_t u = (uint32_t)(mix64(seed, (uint64_t)i) >> 40);
    out[i] = add_rn(mul_rn((float)u, kCentreStep), -100.0f);
  }
}

void synth_perturb_host(uint64_t seed, int k, int d, const float *centres, float *out) {
  for (int64_t i = 0; i < (int64_t)k * d; ++i) {
    const int32_t v = ih8(mix64(seed, 2 * (uint64_t)i), mix64(seed, 2 * (uint64_t)i + 1));
    out[i] = add_rn(centres[i], mul_rn(mul_rn((float)v, kIh8Scale), 0.1f));
  }
}

void synth_points_host(uint64_t seed, int64_t first, int64_t n, int d, int k, const float *centres,
                       float *out) {
  const uint64_t stride = 2 * (uint64_t)d + 1;
  for (int64_t r = 0; r < n; ++r) {
    const uint64_t i = (uint64_t)(first + r);
    const uint32_t comp = synth_component(seed, i, stride, k);
    for (int j = 0; j < d; ++j)
      out[(size_t)r * d + j] = synth_coord(seed, i, stride, j, centres[(size_t)comp * d + j]);
  }
}

namespace {
__global__ void synth_points_kernel(uint64_t seed, long long first, long long n, int d, int k,
                                    const float *__restrict__ centres, float *__restrict__ out) {
  const uint64_t stride = 2 * (uint64_t)d + 1;
  const long long total = n * d;
  const long long step = (long long)gridDim.x * blockDim.x;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += step) {
    const long long r = e / d;
    const int j = (int)(e - r * d);
    const uint64_t i = (uint64_t)(first + r);
    const uint32_t comp = synth_component(seed, i, stride, k);
    out[e] = synth_coord(seed, i, stride, j, centres[(size_t)comp * d + j]);
  }
}
}  // namespace

cudaError_t launch_synth_points(uint64_t seed, int64_t first, int64_t n, int d, int k,
                                const float *d_centres, float *d_out, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  long long blocks = (n * d + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  synth_points_kernel<<<(int)blocks, 256, 0, s>>>(seed, first, n, d, k, d_centres, d_out);
  return cudaGetLastError();
}

}  // namespace hgm
