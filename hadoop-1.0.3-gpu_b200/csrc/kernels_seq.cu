// kernels_seq.cu -- K7: SequenceFile records -> rows of FP32, on the device.
//
// Input: the bytes of an uncompressed version-6 SequenceFile between two sync entries (or the
// header's end and a sync entry), i.e. what SequenceFileRecordReader hands out for one FileSplit
// (mapred/org/apache/hadoop/mapred/SequenceFileRecordReader.java:40-52, 71-84); values are
// BytesWritable (4-byte big-endian length + bytes, core/org/apache/hadoop/io/BytesWritable.java:
// 158-161) holding rows of d float32.  Record framing is a chain (length, key length, key, value:
// SequenceFile.java:1949-1990), so the unit of parallelism is what the format itself provides for
// splitting: the sync entries the writer drops every >= 2000 bytes (SequenceFile.java:1025-1030).
//   seq_mark_kernel     sync entries (escape + 16-byte marker) starting in each 4 KiB block, counted
//   text_scan_kernel    (shared with K6) exclusive scan of the counts
//   seq_list_kernel     their byte positions, in file order
//   seq_walk_kernel     one warp per segment between two sync entries walks its records: pass 1 counts
//                       the floats per segment, pass 2 (after a scan) copies them out, byte-swapped
//                       if the rows are big-endian
#include "hgm_internal.h"

namespace hgm {
namespace {

constexpr int kSeqThreads = 256;
constexpr int kSeqBytesPerThread = 16;
constexpr int kSeqBlockBytes = kSeqThreads * kSeqBytesPerThread;

__device__ __forceinline__ bool sync_at(const unsigned char *b, long long n, long long p, const unsigned char *sync16) {
  if (p + 20 > n) return false;
  if ((b[p] & b[p + 1] & b[p + 2] & b[p + 3]) != 0xFFu) return false;
  for (int j = 0; j < 16; ++j)
    if (b[p + 4 + j] != sync16[j]) return false;
  return true;
}

// mode 0: counts per block; mode 1: positions at offsets[block] + rank inside the block
__global__ void __launch_bounds__(kSeqThreads) seq_mark_kernel(const unsigned char *__restrict__ b, long long n,
                                                                SeqSync sync, unsigned int *__restrict__ block_counts,
                                                                long long *__restrict__ positions, int mode) {
  __shared__ unsigned int warp_tot[kSeqThreads / 32];
  const long long base = ((long long)blockIdx.x * kSeqThreads + threadIdx.x) * kSeqBytesPerThread;
  unsigned mask = 0;
  for (int j = 0; j < kSeqBytesPerThread; ++j)
    if (sync_at(b, n, base + j, sync.marker)) mask |= 1u << j;
  const unsigned cnt = __popc(mask);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned inc = cnt;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  unsigned before = inc - cnt, total = 0;
  for (int w = 0; w < kSeqThreads / 32; ++w) {
    if (w < warp) before += warp_tot[w];
    total += warp_tot[w];
  }
  if (mode == 0) {
    if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
  } else {
    unsigned idx = block_counts[blockIdx.x] + before;
    while (mask) {
      const int j = __ffs(mask) - 1;
      mask &= mask - 1;
      positions[idx++] = base + j;
    }
  }
}

__device__ __forceinline__ long long be32(const unsigned char *p) {
  return (int)(((unsigned)p[0] << 24) | ((unsigned)p[1] << 16) | ((unsigned)p[2] << 8) | p[3]);
}

__device__ void seq_report(TextStats *stats, unsigned code, long long pos) {
  atomicMin(&stats->first_error, ((unsigned long long)pos << 4) | code);
}

// segment s = bytes [positions[s-1] + 20, positions[s])  (s = 0 starts at 0, the last ends at n)
// pass 0: seg_floats[s] <- floats in the segment; pass 1: copy them to out[seg_offsets[s] ...]
__global__ void seq_walk_kernel(const unsigned char *__restrict__ b, long long n, const long long *__restrict__ positions,
                                long long n_sync, int d, int big_endian, unsigned int *__restrict__ seg_floats,
                                float *__restrict__ out, long long out_cap, TextStats *__restrict__ stats, int pass) {
  const int lane = threadIdx.x & 31;
  const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long s = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; s <= n_sync; s += warps) {
    long long pos = s == 0 ? 0 : positions[s - 1] + 20;
    const long long end = s == n_sync ? n : positions[s];
    long long floats = 0;
    long long dst = pass ? (long long)seg_floats[s] : 0;  // pass 1: the scanned offsets
    while (pos < end) {  // every lane walks the same chain (broadcast loads), the payload copy is shared
      if (pos + 8 > end) { if (lane == 0) seq_report(stats, HGM_TEXT_MALFORMED, pos); break; }
      const long long length = be32(b + pos), key_len = be32(b + pos + 4);
      const long long val = pos + 8 + key_len, val_len = length - key_len;
      if (length < 0 || key_len < 0 || val_len < 4 || pos + 8 + length > end) {
        if (lane == 0) seq_report(stats, HGM_TEXT_MALFORMED, pos);
        break;
      }
      const long long bytes = be32(b + val);
      if (bytes != val_len - 4 || bytes % (4 * (long long)d) != 0) {
        if (lane == 0) seq_report(stats, HGM_TEXT_RAGGED, pos);
        break;
      }
      const long long m = bytes >> 2;
      if (pass) {
        const unsigned char *src = b + val + 4;
        for (long long i = lane; i < m; i += 32) {
          const unsigned char *q = src + 4 * i;
          const unsigned u = big_endian ? (((unsigned)q[0] << 24) | ((unsigned)q[1] << 16) | ((unsigned)q[2] << 8) | q[3])
                                        : (((unsigned)q[3] << 24) | ((unsigned)q[2] << 16) | ((unsigned)q[1] << 8) | q[0]);
          if (dst + i < out_cap) out[dst + i] = __uint_as_float(u);
          else seq_report(stats, HGM_TEXT_INTERNAL, pos);
        }
      }
      dst += m;
      floats += m;
      pos += 8 + length;
    }
    if (!pass && lane == 0) {
      if (floats > 0xFFFFFFFFll) seq_report(stats, HGM_TEXT_INTERNAL, pos);
      seg_floats[s] = (unsigned int)floats;
    }
  }
}

}  // namespace

// text_scan_kernel lives in kernels_text.cu
cudaError_t launch_u32_scan(unsigned int *d_values, long long n, TextStats *d_stats, cudaStream_t s);

size_t seq_block_count(size_t nbytes) { return (nbytes + kSeqBlockBytes - 1) / kSeqBlockBytes; }

// Stage 1: positions of the sync entries, in file order.  d_blocks: seq_block_count(nbytes)+1 words;
// d_positions: room for nbytes / 20 + 1 entries.  stats->n_fields <- their number.
cudaError_t launch_seq_find_syncs(const unsigned char *d_bytes, size_t nbytes, const SeqSync &sync, unsigned int *d_blocks,
                                  long long *d_positions, TextStats *d_stats, cudaStream_t s) {
  const TextStats init = {0ull, ~0ull};
  cudaError_t e = cudaMemcpyAsync(d_stats, &init, sizeof init, cudaMemcpyHostToDevice, s);
  if (e != cudaSuccess || nbytes == 0) return e;
  const long long nb = (long long)seq_block_count(nbytes);
  seq_mark_kernel<<<(unsigned)nb, kSeqThreads, 0, s>>>(d_bytes, (long long)nbytes, sync, d_blocks, nullptr, 0);
  e = launch_u32_scan(d_blocks, nb, d_stats, s);
  if (e != cudaSuccess) return e;
  seq_mark_kernel<<<(unsigned)nb, kSeqThreads, 0, s>>>(d_bytes, (long long)nbytes, sync, d_blocks, d_positions, 1);
  return cudaGetLastError();
}

// Stage 2 (pass 0: float counts per segment into d_seg, then scanned in place, total -> stats->n_fields;
// pass 1: the copy).  d_seg: n_sync + 2 words.
cudaError_t launch_seq_walk(const unsigned char *d_bytes, size_t nbytes, const long long *d_positions, long long n_sync,
                            int d, int big_endian, unsigned int *d_seg, float *d_out, size_t out_cap, TextStats *d_stats,
                            int pass, int sm_count, cudaStream_t s) {
  if (nbytes == 0) return cudaSuccess;
  long long blocks = (n_sync + 1 + 7) / 8;  // 8 warps per CTA
  if (blocks > (long long)sm_count * 8) blocks = (long long)sm_count * 8;
  seq_walk_kernel<<<(unsigned)blocks, 256, 0, s>>>(d_bytes, (long long)nbytes, d_positions, n_sync, d, big_endian, d_seg,
                                                   d_out, (long long)out_cap, d_stats, pass);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess || pass) return e;
  return launch_u32_scan(d_seg, n_sync + 1, d_stats, s);
}

}  // namespace hgm
