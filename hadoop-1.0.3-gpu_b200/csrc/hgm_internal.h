// hgm_internal.h -- shared declarations of libhgpumap's translation units (not installed).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/hgpumap.h"

// NVTX ranges on the entry points a trace of a map task should show (the fork's ad-hoc
// wall-clock LOG.info timers, mapred/MapTask.java:443,451, pipes/PipesGPUMapRunner.java:110-112).  Header-only NVTX 3:
// without a tool attached a range is one load and a branch.
#include <nvtx3/nvToolsExt.h>
namespace hgm {
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange &) = delete;
  NvtxRange &operator=(const NvtxRange &) = delete;
};
}  // namespace hgm
#define HGM_RANGE(name) ::hgm::NvtxRange hgm_nvtx_range_(name)

namespace hgm {

// ---- arithmetic contract ------------------------------------
// distance = FP32 sum_j (x_j - c_j)^2 with j ascending and every operation rounded on
// its own.  nvcc contracts a*b+c into FMA by default, so the exact paths spell the
// rounding out with intrinsics instead of relying on a compiler flag.
__device__ __forceinline__ float sqdiff(float x, float c) {
  float t = __fsub_rn(x, c);
  return __fmul_rn(t, t);
}

// ---- d == 2 cell index (kernels_grid2d.cu) -------------------------------------------
// cell coordinate of x:  bits(fma(x, s, B)) - 0x4B000000, where B = b + 2^23 with b an
// integer, i.e. RN(x*s) + b obtained from the mantissa of one FMA.  NaN, +-inf and
// out-of-range x all come out >= G as unsigned and take the brute-force path.
struct Grid2D {
  float sx, Bx, sy, By;
  int32_t bx, by;  // the integer offsets inside Bx, By (builder needs them)
  int32_t G;       // cells per side
  int32_t shape;   // map-kernel shape picked for this job (kernels_grid2d.cu); kShapeFx = the fixed-point form
  float fsx, fsy;  // fixed-point scales of kernels_grid2d_fx.cu: powers of two with |x * fsx| < 2^21 inside the grid; 0 = unusable
};
constexpr int32_t kShapeFx = 6;

constexpr uint32_t kMagicBits = 0x4B000000u;  // bits of 2^23
constexpr uint16_t kCellBrute = 0xFFFFu;      // cell entry: scan all k centroids
constexpr uint16_t kCellMulti = 0x8000u;      // cell entry: 0x8000 | candidate count (2..8)
constexpr int kMaxCand = 8;
// 8-bit copy of the table for k <= kTable8MaxK: cluster id, or kTable8Multi = "look at the 16-bit entry"
constexpr int kTable8MaxK = 254;
constexpr uint8_t kTable8Multi = 0xFF;
// bytes of the table buffer for a G x G grid: 16-bit table, then the 8-bit copy (counting-sort form) or the
// fixed-point form's own 16-bit table
inline size_t grid2d_table_bytes(int G) { return (size_t)G * G * 4 + 32; }

// A second, coarse index over several times the centroids' box: a point outside the fine grid is resolved from its
// cell there instead of by a scan over all k centroids (fixed-point form only; built by the same builder kernel).
struct Grid2DOuter {
  Grid2D g;               // G == 0: none
  const uint16_t *table;  // G*G 16-bit entries
  const uint16_t *cand;   // G*G*kMaxCand
};
constexpr int kOuterCells = 96;  // cells per side of the coarse index

// launchers (each returns the CUDA error of its launch)
cudaError_t launch_grid2d_build(const float *d_cent, int k, const Grid2D &g, uint16_t *d_table,
                                uint16_t *d_cand, cudaStream_t s);
cudaError_t launch_grid2d_map(const float *d_pts, int64_t n, const float *d_cent, int k,
                              const Grid2D &g, const uint16_t *d_table, const uint16_t *d_cand,
                              double *d_sums, unsigned long long *d_counts, int32_t *d_assign,
                              int sm_count, size_t smem_optin, cudaStream_t s, const Grid2DOuter *outer = nullptr);
bool grid2d_plan(int k, size_t smem_optin, int *G_out, int *shape_out);
#ifdef HGM_CHECKED
void grid2d_checked_failures(unsigned int out[2]);
void grid2d_fx_checked_failures(unsigned int out[2]);
#endif
// the fixed-point form (kernels_grid2d_fx.cu): table size it can hold for k clusters (0 = none), and its launcher
int grid2d_fx_table_size(int k, size_t smem_optin, size_t table_cap = 0);
cudaError_t launch_grid2d_fx_map(const float *d_pts, int64_t n, const float *d_cent, int k, const Grid2D &g,
                                 const uint16_t *d_table, const uint16_t *d_cand, double *d_sums,
                                 unsigned long long *d_counts, int32_t *d_assign, int sm_count, size_t smem_optin,
                                 cudaStream_t s, const Grid2DOuter *outer = nullptr);

// scan (assign + per-launch counts in d_meta) followed by the segmented sums; d_assign must not be NULL
cudaError_t launch_exact_map(const float *d_pts, int64_t n, int d, const float *d_cent, int k, int32_t *d_assign,
                             unsigned int *d_meta, int32_t *d_order, double *d_sums, unsigned long long *d_counts,
                             int sm_count, size_t smem_optin, cudaStream_t s);
// per-cluster FP64 sums from assign[n] + the launch's counts in d_meta[0..k) (kernels_tensor.cu)
cudaError_t launch_segmented_sums(const float *d_pts, int64_t n, int d, int k, const int32_t *d_assign,
                                  unsigned int *d_meta, int32_t *d_order, double *d_sums,
                                  unsigned long long *d_counts, int sm_count, cudaStream_t s);

// tcgen05 path (kernels_tensor.cu)
bool tensor_shape_ok(int k, int d, size_t smem_optin);
size_t tensor_caug_floats(int k, int d);
void tensor_set_form(int form);  // HGM_TENSOR_FORM_*
cudaError_t launch_tensor_prep(const float *d_cent, int k, int d, float *d_caug, unsigned int *d_cmax_bits,
                               cudaStream_t s);
size_t tensor_meta_words(int k);  // uint32 words of per-launch cluster bookkeeping
cudaError_t launch_tensor_map(const float *d_pts, int64_t n, int d, const float *d_cent, int k, const float *d_caug,
                              const unsigned int *d_cmax_bits, int32_t *d_assign, int32_t *d_flag_list,
                              unsigned int *d_flag_count, unsigned int *d_meta, int32_t *d_order, double *d_sums,
                              unsigned long long *d_counts, int sm_count, size_t smem_optin, cudaStream_t s,
                              cudaEvent_t *phase /* 4 events or NULL */, int *launches /* kernels launched, or NULL */);

// text decode (kernels_text.cu)
struct TextStats {
  unsigned long long n_fields;     // floats written by the last decode
  unsigned long long first_error;  // (byte position << 4) | HGM_TEXT_*, all ones = none
};
enum { HGM_TEXT_MALFORMED = 1, HGM_TEXT_RAGGED = 2, HGM_TEXT_LONG = 3, HGM_TEXT_INTERNAL = 4 };
size_t text_block_count(size_t nbytes);
cudaError_t launch_text_decode(const unsigned char *d_text, size_t nbytes, int d, float *d_out, size_t out_cap,
                               unsigned int *d_block_counts, TextStats *d_stats, cudaStream_t s);

// SequenceFile decode (kernels_seq.cu)
struct SeqSync { unsigned char marker[16]; };
size_t seq_block_count(size_t nbytes);
cudaError_t launch_seq_find_syncs(const unsigned char *d_bytes, size_t nbytes, const SeqSync &sync, unsigned int *d_blocks,
                                  long long *d_positions, TextStats *d_stats, cudaStream_t s);
cudaError_t launch_seq_walk(const unsigned char *d_bytes, size_t nbytes, const long long *d_positions, long long n_sync,
                            int d, int big_endian, unsigned int *d_seg, float *d_out, size_t out_cap, TextStats *d_stats,
                            int pass, int sm_count, cudaStream_t s);

cudaError_t launch_bswap32(uint32_t *d_words, int64_t n_words, cudaStream_t s);

cudaError_t launch_synth_points(uint64_t seed, int64_t first, int64_t n, int d, int k,
                                const float *d_centres, float *d_out, cudaStream_t s);

// host side of the synthetic generator (synth.cu)
void synth_centres_host(uint64_t seed, int k, int d, float *out);
void synth_perturb_host(uint64_t seed, int k, int d, const float *centres, float *out);
void synth_points_host(uint64_t seed, int64_t first, int64_t n, int d, int k, const float *centres,
                       float *out);

}  // namespace hgm
