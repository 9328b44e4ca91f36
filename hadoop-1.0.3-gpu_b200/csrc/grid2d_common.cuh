// grid2d_common.cuh -- device helpers shared by the two d == 2 map kernels
// (kernels_grid2d.cu: counting-sort combiner; kernels_grid2d_fx.cu: fixed-point combiner).
#pragma once

#include "hgm_internal.h"

// `make CHECKED=1` (-DHGM_CHECKED): every data-dependent shared-memory index of the d == 2 kernels and every
// centroid id read from a candidate list is tested against its bound; a miss is counted in a device word the host
// reads back (hgm_checked_failures) and the access is clamped, so a run ends with a number instead of a fault.
// compute-sanitizer is closed on this pool: this build, small cases and the oracle comparison stand in for memcheck.
#ifdef HGM_CHECKED
#define HGM_BOUND(i, limit) hgm::checked_index((i), (limit), __LINE__)
#else
#define HGM_BOUND(i, limit) (i)
#endif

namespace hgm {
namespace {

#ifdef HGM_CHECKED
__device__ unsigned int g_check_failures[2];  // [0] count, [1] source line of the first one
template <class T, class U>
__device__ __forceinline__ T checked_index(T i, U limit, int line) {
  if ((long long)i >= 0 && (long long)i < (long long)limit) return i;
  if (atomicAdd(&g_check_failures[0], 1u) == 0u) g_check_failures[1] = (unsigned int)line;
  return (T)0;
}
#endif

constexpr int kMaxK = 1024;
constexpr uint16_t kCandEnd = 0xFFFF;  // pads a candidate list; a list of nothing else = scan all k
constexpr uint16_t kCandMore = 0x8000;  // set in a list's first id: the list holds more than two candidates

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier
// (SASS: UBLKCP).  dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes,
                                             uint64_t *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// exact FP32 squared distance for d == 2, exactly as the oracle rounds it:
// acc = (0 + dx*dx) + dy*dy with dx*dx >= +0 so the leading 0 + is the identity.
__device__ __forceinline__ float dist2(float x, float y, float2 c) {
  return __fadd_rn(sqdiff(x, c.x), sqdiff(y, c.y));
}

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

// cell entry of a point: one FMA per axis lands RN(x*s)+b in the mantissa; the unsigned min
// folds every out-of-range / NaN / inf value onto the last row or column, which the builder
// marks "scan everything".
__device__ __forceinline__ uint32_t cell_of(float x, float y, const Grid2D &g) {
  const uint32_t ux = min(__float_as_uint(__fmaf_rn(x, g.sx, g.Bx)) - kMagicBits, (uint32_t)(g.G - 1));
  const uint32_t uy = min(__float_as_uint(__fmaf_rn(y, g.sy, g.By)) - kMagicBits, (uint32_t)(g.G - 1));
  return uy * (uint32_t)g.G + ux;
}

// exact scan of one point over its cell's candidates (or over all k): ascending index, strict '<' from
// (+inf, 0) -- the reference scan restricted to a superset of its winners -- driven by the cell's
// candidate list alone (multi-candidate and scan-all cells only).  The first 32-bit word of a list is the
// whole answer for a two-candidate cell (by far the common case): two ids, kCandMore clear.  All ones =
// scan every centroid; kCandMore set = read the rest of the kCandEnd-padded list.
__device__ __forceinline__ int resolve_cand(float x, float y, uint32_t id01, const uint16_t *__restrict__ list,
                                            const float2 *cent_s, int k) {
  float best = INFINITY;
  int arg = 0;
  if (id01 == 0xFFFFFFFFu) {
    for (int c = 0; c < k; ++c) {
      const float dd = dist2(x, y, cent_s[HGM_BOUND(c, k)]);
      if (dd < best) { best = dd; arg = c; }
    }
  } else {
    const int c0 = id01 & 0x7FFF, c1 = id01 >> 16;
    const float d0 = dist2(x, y, cent_s[HGM_BOUND(c0, k)]), d1 = dist2(x, y, cent_s[HGM_BOUND(c1, k)]);
    if (d0 < best) { best = d0; arg = c0; }
    if (d1 < best) { best = d1; arg = c1; }
    if (id01 & kCandMore) {
      const uint4 ids = __ldg(reinterpret_cast<const uint4 *>(list));
      const unsigned long long lo = ((unsigned long long)ids.y << 32) | ids.x;
      const unsigned long long hi = ((unsigned long long)ids.w << 32) | ids.z;
      for (int i = 2; i < kMaxCand; ++i) {
        const int c = (int)(((i < 4 ? lo : hi) >> (16 * (i & 3))) & 0xFFFF);
        if (c == kCandEnd) break;
        const float dd = dist2(x, y, cent_s[HGM_BOUND(c, k)]);
        if (dd < best) { best = dd; arg = c; }
      }
    }
  }
  return arg;
}

}  // namespace
}  // namespace hgm
