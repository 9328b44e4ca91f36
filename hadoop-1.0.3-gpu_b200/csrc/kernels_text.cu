// kernels_text.cu -- K6: text records -> rows of FP32, on the device.
//
// What is replaced: the per-field sscanf("%f") of HadoopUtils::toFloat
// (c++/utils/impl/StringUtils.cc:49-56) behind splitString(" ,\t") (:80-95) that a CPU
// Pipes mapper runs on every TextInputFormat line.  Same results, bit for bit: decimal and
// hexadecimal literals are rounded to nearest-even exactly (denormals, overflow to inf,
// "inf"/"infinity"/"nan", "1e" and "0x1p" with empty exponents as glibc's scanf takes
// them); anything scanf would stop short of is an error, not a guess.
//
// Three launches per staged piece of text:
//   text_count_kernel   field starts per 4 KiB block
//   text_scan_kernel    exclusive scan of the block counts (one CTA), total -> stats
//   text_parse_kernel   one thread per 16 bytes; every field start parses its field and
//                       writes out[index]; every line end checks "whole points per line"
// HBM traffic: the text twice (second pass mostly L2), 4 bytes out per field.
#include "hgm_internal.h"
#include "text_parse.inl"

namespace hgm {
namespace {

constexpr int kTextThreads = 256;
constexpr int kBytesPerThread = 16;
constexpr int kTextBlockBytes = kTextThreads * kBytesPerThread;  // 4096
constexpr int kMaxField = 127;  // longest field the reference's decoders take (oracle: char field[128])

__device__ __forceinline__ bool is_sep(unsigned c) {  // splitString's " ,\t" plus the line ends
  return c == ' ' || c == ',' || c == '\t' || c == '\n' || c == '\r';
}
__device__ __forceinline__ bool is_eol(unsigned c) { return c == '\n' || c == '\r'; }

// bit j set <=> byte (base + j) starts a field; prev = the byte before base (a separator at 0)
__device__ __forceinline__ unsigned start_mask(const unsigned char *b, int valid, unsigned prev, unsigned *eol_mask) {
  unsigned m = 0, e = 0;
  bool prev_sep = is_sep(prev);
#pragma unroll
  for (int j = 0; j < kBytesPerThread; ++j) {
    if (j < valid) {
      const bool s = is_sep(b[j]);
      if (!s && prev_sep) m |= 1u << j;
      if (is_eol(b[j])) e |= 1u << j;
      prev_sep = s;
    }
  }
  *eol_mask = e;
  return m;
}

__device__ __forceinline__ void load16(const unsigned char *text, long long base, long long n, unsigned char *b,
                                       int *valid, unsigned *prev) {
  const long long left = n - base;
  *valid = left >= kBytesPerThread ? kBytesPerThread : (left > 0 ? (int)left : 0);
  if (*valid == kBytesPerThread) {
    const uint4 v = *reinterpret_cast<const uint4 *>(text + base);  // text buffers are 16-byte aligned
    memcpy(b, &v, 16);
  } else {
    for (int j = 0; j < kBytesPerThread; ++j) b[j] = j < *valid ? text[base + j] : (unsigned char)' ';
  }
  *prev = (base > 0 && *valid > 0) ? text[base - 1] : (unsigned)' ';
}

__global__ void __launch_bounds__(kTextThreads) text_count_kernel(const unsigned char *__restrict__ text, long long n,
                                                                   unsigned int *__restrict__ block_counts) {
  __shared__ unsigned int warp_sum[kTextThreads / 32];
  const long long base = ((long long)blockIdx.x * kTextThreads + threadIdx.x) * kBytesPerThread;
  unsigned char b[kBytesPerThread];
  int valid;
  unsigned prev, eol;
  load16(text, base, n, b, &valid, &prev);
  unsigned cnt = __popc(start_mask(b, valid, prev, &eol));
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = 0;
    for (int w = 0; w < kTextThreads / 32; ++w) t += warp_sum[w];
    block_counts[blockIdx.x] = t;
  }
}

// in place: block_counts[i] <- sum of block_counts[0..i), stats->n_fields <- total
__global__ void __launch_bounds__(1024) text_scan_kernel(unsigned int *__restrict__ block_counts, long long nb,
                                                         TextStats *__restrict__ stats) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long first = 0; first < nb; first += 1024) {
    const long long i = first + threadIdx.x;
    const unsigned long long v = i < nb ? block_counts[i] : 0ull;
    unsigned long long inc = v;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      const unsigned long long w = warp_tot[lane];
      unsigned long long winc = w;
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += t;
      }
      warp_tot[lane] = winc - w;  // exclusive over warps
    }
    __syncthreads();
    const unsigned long long excl = carry_s + warp_tot[warp] + inc - v;
    // offsets are stored as 32 bits: one staged piece holds < 2^32 fields (16 MiB pieces; whole-split
    // decodes are bounded by the launcher)
    if (i < nb) block_counts[i] = (unsigned int)excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) stats->n_fields = carry_s;
}

__device__ void report(TextStats *stats, unsigned code, long long pos) {
  // lowest position wins, so the message names the first bad byte whatever the schedule
  const unsigned long long key = ((unsigned long long)pos << 4) | code;
  atomicMin(&stats->first_error, key);
}

__global__ void __launch_bounds__(kTextThreads) text_parse_kernel(const unsigned char *__restrict__ text, long long n,
                                                                   int d, const unsigned int *__restrict__ block_offsets,
                                                                   float *__restrict__ out, long long out_cap,
                                                                   TextStats *__restrict__ stats) {
  __shared__ unsigned int warp_tot[kTextThreads / 32];
  const long long base = ((long long)blockIdx.x * kTextThreads + threadIdx.x) * kBytesPerThread;
  unsigned char b[kBytesPerThread];
  int valid;
  unsigned prev, eol;
  load16(text, base, n, b, &valid, &prev);
  const unsigned starts = start_mask(b, valid, prev, &eol);
  const unsigned cnt = __popc(starts);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned inc = cnt;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  unsigned before = block_offsets[blockIdx.x] + inc - cnt;
  for (int w = 0; w < warp; ++w) before += warp_tot[w];

  // "whole points per line": the CPU mapper asserts n % d == 0 per record (oracle/kmeans_pipes_cpu.cc)
  unsigned e = eol;
  while (e) {
    const int j = __ffs(e) - 1;
    e &= e - 1;
    const unsigned fields_so_far = before + __popc(starts & ((1u << j) - 1));
    if (fields_so_far % (unsigned)d != 0) report(stats, HGM_TEXT_RAGGED, base + j);
  }
  unsigned m = starts;
  unsigned idx = before;
  while (m) {
    const int j = __ffs(m) - 1;
    m &= m - 1;
    const long long p0 = base + j;
    int len = 0;
    while (p0 + len < n && len <= kMaxField && !is_sep(text[p0 + len])) ++len;
    uint32_t bits = 0;
    bool trouble = false;
    if (len > kMaxField) report(stats, HGM_TEXT_LONG, p0);
    else if (!hgm_text::parse_field(text + p0, len, &bits, &trouble)) report(stats, HGM_TEXT_MALFORMED, p0);
    else if (trouble) report(stats, HGM_TEXT_INTERNAL, p0);
    if ((long long)idx < out_cap) out[idx] = __uint_as_float(bits);
    else report(stats, HGM_TEXT_INTERNAL, p0);
    ++idx;
  }
}

}  // namespace

// in place: values[i] <- sum of values[0..i), stats->n_fields <- total (also used by kernels_seq.cu)
cudaError_t launch_u32_scan(unsigned int *d_values, long long n, TextStats *d_stats, cudaStream_t s) {
  text_scan_kernel<<<1, 1024, 0, s>>>(d_values, n, d_stats);
  return cudaGetLastError();
}

size_t text_block_count(size_t nbytes) { return (nbytes + kTextBlockBytes - 1) / kTextBlockBytes; }

cudaError_t launch_text_decode(const unsigned char *d_text, size_t nbytes, int d, float *d_out, size_t out_cap,
                               unsigned int *d_block_counts, TextStats *d_stats, cudaStream_t s) {
  const TextStats init = {0ull, ~0ull};
  cudaError_t e = cudaMemcpyAsync(d_stats, &init, sizeof init, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess || nbytes == 0) return e;
  const long long nb = (long long)text_block_count(nbytes);
  if (nb > 0x7fffffffLL) return cudaErrorInvalidValue;
  text_count_kernel<<<(unsigned)nb, kTextThreads, 0, s>>>(d_text, (long long)nbytes, d_block_counts);
  text_scan_kernel<<<1, 1024, 0, s>>>(d_block_counts, nb, d_stats);
  if (d_out)  // a count-only pass leaves the parse out
    text_parse_kernel<<<(unsigned)nb, kTextThreads, 0, s>>>(d_text, (long long)nbytes, d, d_block_counts, d_out,
                                                            (long long)out_cap, d_stats);
  return cudaGetLastError();
}

}  // namespace hgm
