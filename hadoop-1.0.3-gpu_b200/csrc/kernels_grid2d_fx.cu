// kernels_grid2d_fx.cu -- K1, second form: the d == 2 map kernel with a fixed-point combiner.
//
// Same slots as kernels_grid2d.cu (Mapper::map  c++/pipes/api/hadoop/Pipes.hh:158-161 called at
// HadoopPipes.cc:884-885; the combiner's Reducer::reduce  Pipes.hh:166-169 called at
// HadoopPipes.cc:700-702), same cell index, same exact candidate scan for the points of
// multi-candidate cells.  What differs is how a point reaches its cluster's partial sum.
//
// The counting-sort form moves every point through shared memory twice (scatter by cluster, read
// back by the cluster's owner lanes) at addresses decided by the data: 20.7 shared-memory
// wavefronts per 32 points, more than half of them bank-conflict replays, which is what bounded it
// (profiles/r01_v11_grid2d_summary.md).  This form never moves the point.  The only native
// shared-memory atomic on sm_100a is the 32-bit integer ATOMS (FP32/FP64/64-bit adds are CAS
// loops), so the coordinate is turned into integers that ATOMS can add EXACTLY:
//
//   x * S (S a power of two chosen per job so that |x * S| < 2^21 everywhere inside the grid)
//     = hi + lo * 2^-22,   hi = RN(x * S) an integer,  lo = (x * S - hi) * 2^22 an integer,
//   both taken from the mantissa of one FMA each ("magic number" rounding), |hi|, |lo| <= 2^21.
//
// A point of a single-candidate cell is then five ATOMS.ADD: hi and lo of x, hi and lo of y, and 1
// -- into words [cluster][word][lane]: the bank is the LANE, never the data, so every one of them
// is a single conflict-free wavefront (tools/microbench/lsu_rate.cu: table lookup + 5 atomics =
// 7.6 cycles per 32 points against 20.7+ before).  The representation is exact for every x with
// |x * S| >= 2 (or x == 0): a float there has no bits below 2^-22.  Smaller non-zero values
// (|x| below 2^-20 of the grid's extent) can only occur in the cell column that contains x = 0 and
// the row that contains y = 0: the index builder marks those cells "ask the index" in this form's
// table, so the hot path needs no test of its own.  Their points, the points of multi-candidate
// cells, points outside the grid and NaN/inf go to a per-warp queue in shared memory and are settled
// 32 at a time, twice per tile: exact candidate scan, then the same five atomics, or -- out of range /
// too small -- an FP64 CAS add.  The table in shared memory has 16 bits per cell: a two-candidate cell
// (most queued points sit on a single Voronoi edge) names both centroids there and is settled without
// touching global memory; only cells with more candidates, the strips and the outside go to the
// candidate lists in global memory.
//
// A word takes at most kFxMaxAdds adds between two flushes (1000 * (2^21 + 1) < 2^31: the signed
// sum cannot wrap).  The count word of each (cluster, lane) says how many it has had: the add that
// finds it at the threshold raises a CTA-wide flag, every warp finishes the tile it is on, and the
// CTA flushes: lane words -> warp reduction -> int64 totals per cluster.  With evenly loaded
// clusters that happens once, at the end; with every point in one cluster, every ~19 k points.
// At the end the int64 totals become FP64 (hi + lo * 2^-22) / S -- exact up to the final FP64
// rounding -- and leave the SM as k * 3 RED.ADD.F64 / .U64, as in the counting-sort form.
#include "grid2d_common.cuh"
#include "hgm_internal.h"

namespace hgm {

namespace {

constexpr int kFxWords = 5;                     // hi x, lo x, hi y, lo y, count
constexpr uint32_t kFxMagicBits = 0x4B400000u;  // bits of 1.5 * 2^23
constexpr float kFxMagic = 12582912.0f;         // 1.5 * 2^23: RN(v + magic) - magic = RN(v) for |v| < 2^22
constexpr float kFxLoScale = 4194304.0f;        // 2^22
constexpr int kFxMaxAdds = 1000;
// This form's cell table, 16 bits per cell: id (< 256) = the cell's only candidate; id0 | id1 << 8 (id0 < id1, so
// >= 256) = its two candidates; kFxOther = ask the index in global memory.
constexpr uint32_t kFxOther = 0xFFFFu;

template <int kWarpsT, int kPPTT, int kLanesT, int kFlushAtT>
struct FxShape {
  static constexpr int kWarps = kWarpsT;
  static constexpr int kThreads = kWarpsT * 32;
  static constexpr int kPPT = kPPTT;      // points per lane per tile (even)
  static constexpr int kTile = 32 * kPPTT;  // points per warp tile
  static constexpr int kLanes = kLanesT;  // private words per (cluster, word): 32, or 16 when k is large
  static constexpr int kQCap = kTile / 2 + 32;  // a warp's queue: what half a tile can add on top of < 32 left over
  // every add to a count word returns its old value, so some lane sees kFlushAt exactly when a word has
  // had that many adds; after that each warp adds at most what one tile can: kPPT direct adds and kPPT
  // queue batches (+ 1 at the very end) per lane, times the lanes that share a word.  A power of two, so
  // that the test is one OR per add and one compare per tile.
  static constexpr int kFlushAt = kFlushAtT;
  static_assert(kFlushAtT + kWarpsT * (32 / kLanesT) * (2 * kPPTT + 1) + 8 <= kFxMaxAdds,
                "too many warps for the headroom of a 32-bit word");
  static_assert((kFlushAtT & (kFlushAtT - 1)) == 0, "kFlushAt must be a power of two");
};

__host__ __device__ inline size_t fx_head_bytes(int G, int k) {
  // [table G*G u16][centroids k float2][mbarrier 16 B][flag 16 B]
  return align16((size_t)G * G * 2) + align16((size_t)k * 8) + 32;
}
template <class S>
__host__ __device__ inline size_t fx_body_bytes(int k, bool with_assign) {
  // [acc (k+1)*5*lanes u32][tot k*5 i64][late sums k*2 f64][late counts k u64][queues warps*qcap float2][queue index u16]
  return (size_t)(k + 1) * kFxWords * S::kLanes * 4 + (size_t)k * kFxWords * 8 + (size_t)k * 16 + (size_t)k * 8 +
         (size_t)S::kWarps * S::kQCap * 8 + (with_assign ? align16((size_t)S::kWarps * S::kQCap * 2) : 0);
}

struct FxParams {
  float sx, sy;      // fixed-point scales S (powers of two)
  float tinyx, tinyy;  // 2 / S: a non-zero |x| below this has bits the limbs cannot hold
  float limx, limy;  // 2^21 / S: at or beyond this the limbs would overflow
  uint32_t rows;     // accumulator rows: k clusters + the one nobody reads (bound of the checked build)
};

// The five adds of one point into this lane's words of row c, unconditionally: ptxas turns a predicated shared
// atomic into a branch around it, and a branch per atomic costs more issue slots than everything else a point needs.
// A point that must not be added here (multi-candidate cell, past the end of a ragged tile) is given row k instead, a
// row nobody reads, and `one` = 0 so that row's count word never moves.  acc_lane = shared-memory address of this
// lane's word of row 0, word 0.  Returns what the count word held before.
template <int kLanes>
__device__ __forceinline__ uint32_t fx_add_point(uint32_t one, uint32_t acc_lane, const FxParams &fx, float x, float y, uint32_t c) {
  const float tx = __fmaf_rn(x, fx.sx, kFxMagic);         // mantissa = RN(x*S), offset by the magic bits
  const float rx = __fmaf_rn(x, fx.sx, kFxMagic - tx);    // x*S - RN(x*S): exact, |rx| <= 1/2
  const float ux = __fmaf_rn(rx, kFxLoScale, kFxMagic);   // mantissa = rx * 2^22 (an integer for non-tiny x)
  const float ty = __fmaf_rn(y, fx.sy, kFxMagic);
  const float ry = __fmaf_rn(y, fx.sy, kFxMagic - ty);
  const float uy = __fmaf_rn(ry, kFxLoScale, kFxMagic);
  const uint32_t addr = acc_lane + HGM_BOUND(c, fx.rows) * (uint32_t)(kFxWords * kLanes * 4);
  uint32_t old;
  // the magic bits ride along in every add; the flush takes count * kFxMagicBits off again (mod 2^32)
  asm volatile(
      "red.shared.add.u32 [%1], %2;\n\t"
      "red.shared.add.u32 [%1+%7], %3;\n\t"
      "red.shared.add.u32 [%1+%8], %4;\n\t"
      "red.shared.add.u32 [%1+%9], %5;\n\t"
      "atom.shared.add.u32 %0, [%1+%10], %6;"
      : "=r"(old)
      : "r"(addr), "r"(__float_as_uint(tx)), "r"(__float_as_uint(ux)), "r"(__float_as_uint(ty)),
        "r"(__float_as_uint(uy)), "r"(one), "n"(kLanes * 4), "n"(2 * kLanes * 4), "n"(3 * kLanes * 4), "n"(4 * kLanes * 4)
      : "memory");
  return old;
}

// One tile of a warp: table lookups, then per point either the five adds (single-candidate cell) or a push
// onto the warp's queue.  kFull = every point of the tile exists (no bounds tests in the straight-line code).
// `seen` collects the OR of the counts found.
template <class S, bool kAssign, bool kFull, int kFirst>
__device__ __forceinline__ void fx_work_tile(const uint16_t *table_s, const Grid2D &g, const FxParams &fx, uint32_t acc_lane,
                                             uint32_t k, float2 *queue, uint16_t *qindex, int32_t *__restrict__ assign, long long base,
                                             int live, int lane, uint32_t lt_mask, const float (&xs)[S::kPPT],
                                             const float (&ys)[S::kPPT], int &qn, uint32_t &seen) {
  // half a tile per call (points kFirst .. kFirst + kPPT/2 - 1 of every lane): the queue is drained in between, so
  // it needs room for half a tile only and the table can have that much more of the shared memory
  constexpr int kPPT = S::kPPT, kEnd = kFirst + S::kPPT / 2;
  // keep the table loads in a pass of their own, ahead of every shared-memory write of the half tile, or the
  // compiler chains load -> atomics -> load
  uint32_t ent[kPPT];
  uint32_t old[kPPT];  // what each count word held: looked at after the half tile, not behind each atomic's round trip
#pragma unroll
  for (int i = kFirst; i < kEnd; ++i) ent[i] = table_s[cell_of(xs[i], ys[i], g)];
#pragma unroll
  for (int i = kFirst; i < kEnd; ++i) {
    const int local = 2 * ((i >> 1) * 32 + lane) + (i & 1);
    const bool valid = kFull || local < live;
    // single-candidate cells hold their cluster id (< k); every other entry is >= 256 > k: row k, the one nobody reads
    const uint32_t c = kFull ? min(ent[i], k) : (valid ? min(ent[i], k) : k);
    const bool direct = c != k;
    if (kAssign && direct) assign[base + local] = (int)c;
    old[i] = fx_add_point<S::kLanes>(direct ? 1u : 0u, acc_lane, fx, xs[i], ys[i], c);
    const uint32_t queued = __ballot_sync(0xFFFFFFFFu, valid && !direct);
    if (queued) {
      if (valid && !direct) {
        const int pos = qn + __popc(queued & lt_mask);
        queue[HGM_BOUND(pos, S::kQCap)] = make_float2(xs[i], ys[i]);
        if (kAssign) qindex[pos] = (uint16_t)local;
      }
      qn += __popc(queued);
    }
  }
#pragma unroll
  for (int i = kFirst; i < kEnd; ++i) seen |= old[i];
}

template <class S, bool kVec4, bool kAssign>
__global__ void __launch_bounds__(S::kThreads, 1)
    grid2d_fx_kernel(const float *__restrict__ pts, long long n, const float *__restrict__ cent, int k, Grid2D g,
                     FxParams fx, const uint16_t *__restrict__ table, const uint16_t *__restrict__ cand,
                     Grid2DOuter outer, double *__restrict__ g_sums, unsigned long long *__restrict__ g_counts,
                     int32_t *__restrict__ assign) {
  constexpr int kPPT = S::kPPT;
  constexpr int kTile = S::kTile;
  constexpr int kLanes = S::kLanes;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const size_t table_bytes_al = align16((size_t)g.G * g.G * 2);
  uint16_t *const table_s = reinterpret_cast<uint16_t *>(smem_raw);
  float2 *const cent_s = reinterpret_cast<float2 *>(smem_raw + table_bytes_al);
  uint64_t *const bar = reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(cent_s) + align16((size_t)k * 8));
  volatile int *const flag = reinterpret_cast<volatile int *>(bar + 2);
  unsigned char *const body = smem_raw + fx_head_bytes(g.G, k);
  uint32_t *const acc = reinterpret_cast<uint32_t *>(body);
  long long *const tot = reinterpret_cast<long long *>(acc + (size_t)(k + 1) * kFxWords * kLanes);
  double *const late_sum = reinterpret_cast<double *>(tot + (size_t)k * kFxWords);
  unsigned long long *const late_cnt = reinterpret_cast<unsigned long long *>(late_sum + 2 * (size_t)k);
  float2 *const queue = reinterpret_cast<float2 *>(late_cnt + k) + (size_t)warp * S::kQCap;
  uint16_t *const qindex = reinterpret_cast<uint16_t *>(reinterpret_cast<float2 *>(late_cnt + k) + (size_t)S::kWarps * S::kQCap) +
                           (size_t)warp * S::kQCap;

  // stage this form's cell table (it follows the 16-bit index table in the same buffer) and the centroids with TMA
  // bulk copies (SASS: UBLKCP)
  const unsigned char *table_src = reinterpret_cast<const unsigned char *>(table + (size_t)g.G * g.G);
  const uint32_t table_bytes = (uint32_t)table_bytes_al;
  const uint32_t cent_bytes = (uint32_t)align16((size_t)k * 8);
  if (threadIdx.x == 0) mbar_init(bar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, table_bytes + cent_bytes);
    for (uint32_t o = 0; o < table_bytes; o += 32768) {
      const uint32_t sz = table_bytes - o < 32768 ? table_bytes - o : 32768;
      tma_bulk_g2s(smem_raw + o, table_src + o, sz, bar);
    }
    tma_bulk_g2s(cent_s, cent, cent_bytes, bar);
    *flag = 0;
  }
  for (int i = threadIdx.x; i < (k + 1) * kFxWords * kLanes; i += S::kThreads) acc[i] = 0u;
  for (int i = threadIdx.x; i < k * kFxWords; i += S::kThreads) tot[i] = 0;
  for (int i = threadIdx.x; i < k; i += S::kThreads) {
    late_sum[2 * i] = 0.0;
    late_sum[2 * i + 1] = 0.0;
    late_cnt[i] = 0ull;
  }
  mbar_wait(bar, 0);
  __syncthreads();

  const uint32_t acc_lane = smem_u32(acc + (lane & (kLanes - 1)));  // this lane's word of every (cluster, word) row
  const uint32_t lt_mask = (1u << lane) - 1u;

  // a value the two limbs hold exactly: zero, or no bits below 2^-22 after scaling
  auto tiny = [](float v, float lim) { return fabsf(v) < lim && v != 0.0f; };
  // settle up to 32 queued points, one per lane: exact candidate scan, then the limbs or an FP64 add
  auto settle = [&](int first, int count, long long base) -> uint32_t {
    uint32_t full = 0u;
    const bool live = lane < count;
    const float2 v = live ? queue[first + lane] : make_float2(0.f, 0.f);
    const uint32_t last = (uint32_t)(g.G - 1);
    const uint32_t ux = min(__float_as_uint(__fmaf_rn(v.x, g.sx, g.Bx)) - kMagicBits, last);  // cell_of, taken apart
    const uint32_t uy = min(__float_as_uint(__fmaf_rn(v.y, g.sy, g.By)) - kMagicBits, last);
    const uint32_t cell = uy * (uint32_t)g.G + ux;
    const uint32_t ent = table_s[cell];
    uint32_t c = ent;   // < 256: a single-candidate cell (only the tail of a ragged tile queues those)
    bool fits = true;   // an interior cell outside the strips around the axes: the limbs hold its points exactly
    // a two-candidate cell -- by far the most common kind of queued point -- is settled from shared memory alone:
    // the table entry names both centroids; ascending index, strict '<' from (+inf, 0), as the scan does it
    if (ent >= 256u && ent != kFxOther) {
      const uint32_t c0 = ent & 0xFFu, c1 = ent >> 8;
      const float d0 = dist2(v.x, v.y, cent_s[HGM_BOUND(c0, k)]), d1 = dist2(v.x, v.y, cent_s[HGM_BOUND(c1, k)]);
      float best = INFINITY;
      c = 0u;
      if (d0 < best) { best = d0; c = c0; }
      if (d1 < best) { best = d1; c = c1; }
    }
    // everything else asks the index in global memory: more than two candidates, the strips around x = 0 and y = 0,
    // points outside the fine grid (the coarse index), NaN / inf
    const bool far = live && ent == kFxOther;
    if (__any_sync(0xFFFFFFFFu, far)) {
      const uint16_t *list = cand + (size_t)cell * kMaxCand;
      const uint16_t *entry = table + cell;
      // outside the fine grid (its last row and column collect everything out of range): the coarse index over
      // several times the box knows the few centroids that can win out there
      if (outer.g.G != 0 && (ux == last || uy == last)) {
        const uint32_t out_cell = cell_of(v.x, v.y, outer.g);
        list = outer.cand + (size_t)out_cell * kMaxCand;
        entry = outer.table + out_cell;
      }
      // the 16-bit entry tells a single-candidate cell that is only marked for its tiny values from a
      // multi-candidate one; both loads go out together
      const uint32_t e16 = far ? (uint32_t)__ldg(entry) : 0u;
      const uint32_t id01 = far ? __ldg(reinterpret_cast<const uint32_t *>(list)) : 0u;
      if (far) c = e16;
      // "scan every centroid" points (outside both indexes, NaN/inf, a cell with more candidates than a list holds):
      // one at a time with the whole warp on it -- lane l scans centroids l, l + 32, ... in ascending order, the
      // lexicographic (distance, index) minimum over the lanes is the scan's winner.  Left to its own lane such a
      // point kept the other 31 waiting for k distance evaluations.
      bool scan_all = far && e16 >= kCellMulti && id01 == 0xFFFFFFFFu;
      uint32_t todo = __ballot_sync(0xFFFFFFFFu, scan_all);
      if (__popc(todo) > 10) {  // many of them: k evaluations on every lane at once cost less than 32 cooperative rounds
        todo = 0u;
        scan_all = false;
      }
      for (; todo; todo &= todo - 1) {
        const int src = __ffs(todo) - 1;
        const float px = __shfl_sync(0xFFFFFFFFu, v.x, src), py = __shfl_sync(0xFFFFFFFFu, v.y, src);
        float best = INFINITY;
        int arg = 0x7FFFFFFF;
        for (int cc = lane; cc < k; cc += 32) {
          const float dd = dist2(px, py, cent_s[cc]);
          if (dd < best) { best = dd; arg = cc; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
          const int oa = __shfl_xor_sync(0xFFFFFFFFu, arg, o);
          if (ob < best || (ob == best && oa < arg)) { best = ob; arg = oa; }
        }
        if (lane == src) c = best < INFINITY ? (uint32_t)arg : 0u;  // nothing beat +inf: the scan keeps index 0
      }
      if (far) {
        if (e16 >= kCellMulti && !scan_all) c = (uint32_t)resolve_cand(v.x, v.y, id01, list, cent_s, k);
        fits = fabsf(v.x) < fx.limx && fabsf(v.y) < fx.limy && !tiny(v.x, fx.tinyx) && !tiny(v.y, fx.tinyy);
      }
    }
    if (live) {
      if (kAssign) assign[base + qindex[first + lane]] = (int)c;
      full = fx_add_point<kLanes>(fits ? 1u : 0u, acc_lane, fx, v.x, v.y, fits ? c : (uint32_t)k);
      if (!fits) {  // shared FP64 / 64-bit atomics are CAS loops: the rare path only
        atomicAdd(&late_sum[2 * HGM_BOUND(c, k)], (double)v.x);
        atomicAdd(&late_sum[2 * HGM_BOUND(c, k) + 1], (double)v.y);
        atomicAdd(&late_cnt[HGM_BOUND(c, k)], 1ull);
      }
    }
    return full;
  };

  // Tiles are counted in 32 bits (the launcher refuses more than 2^31 of them) and the warp's read pointer is advanced
  // from tile to tile: the 64-bit index arithmetic per tile was a tenth of the loop's instructions.
  const int ntiles = (int)((n + kTile - 1) / kTile), full_tiles = (int)(n / kTile);
  const int tile_stride = (int)gridDim.x * S::kWarps;
  int tile = (int)blockIdx.x * S::kWarps + warp;
  int qn = 0;  // points in this warp's queue

  // lane owns the point pairs (i*32 + lane), i = 0..kPPT/2-1, of its warp's tile
  const float *src = pts + 2 * ((long long)tile * kTile) + 4 * lane;  // this lane's first pair of the warp's current tile
  const long long src_step = 2ll * kTile * tile_stride;
  auto load_tile = [&](int tl, float4 (&buf)[kPPT / 2]) {
    if (tl < full_tiles) {
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        if (kVec4) {
          asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(buf[i].x), "=f"(buf[i].y), "=f"(buf[i].z), "=f"(buf[i].w)
                       : "l"(src + 128 * i));
        } else {
          const float2 a = __ldcs(reinterpret_cast<const float2 *>(src + 128 * i));
          const float2 b = __ldcs(reinterpret_cast<const float2 *>(src + 128 * i + 2));
          buf[i] = make_float4(a.x, a.y, b.x, b.y);
        }
      }
    } else {
      const int live = (int)(n - (long long)tl * kTile);
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        const int local = 2 * (i * 32 + lane);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (local < live) {
          const float2 a = __ldcs(reinterpret_cast<const float2 *>(src + 128 * i));
          v.x = a.x; v.y = a.y;
        }
        if (local + 1 < live) {
          const float2 b = __ldcs(reinterpret_cast<const float2 *>(src + 128 * i + 2));
          v.z = b.x; v.w = b.y;
        }
        buf[i] = v;
      }
    }
  };

  float4 buf[kPPT / 2];
  if (tile < ntiles) load_tile(tile, buf);
  for (;;) {
    // the flag is read by the whole warp at once; the vote makes the loop condition warp-uniform by construction
    while (tile < ntiles && !__any_sync(0xFFFFFFFFu, *flag != 0)) {
      const bool whole = tile < full_tiles;
      const long long base = (long long)tile * kTile;
      const int live = whole ? kTile : (int)(n - base);
      float xs[kPPT], ys[kPPT];
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        xs[2 * i] = buf[i].x; ys[2 * i] = buf[i].y; xs[2 * i + 1] = buf[i].z; ys[2 * i + 1] = buf[i].w;
      }
      tile += tile_stride;
      src += src_step;
      if (tile < ntiles) load_tile(tile, buf);  // the next tile's loads fly while this one is worked on

      uint32_t seen = 0u;  // OR of the counts found: any bit at or above kFlushAt = a word is at the threshold
      if (whole)
        fx_work_tile<S, kAssign, true, 0>(table_s, g, fx, acc_lane, (uint32_t)k, queue, qindex, assign, base, live, lane, lt_mask, xs, ys, qn, seen);
      else
        fx_work_tile<S, kAssign, false, 0>(table_s, g, fx, acc_lane, (uint32_t)k, queue, qindex, assign, base, live, lane, lt_mask, xs, ys, qn, seen);
      __syncwarp();
      while (qn >= 32) {  // whole batches of 32 only: the rest waits for company
        qn -= 32;
        seen |= settle(qn, 32, base);
        __syncwarp();
      }
      if (whole)
        fx_work_tile<S, kAssign, true, kPPT / 2>(table_s, g, fx, acc_lane, (uint32_t)k, queue, qindex, assign, base, live, lane, lt_mask, xs, ys, qn, seen);
      else
        fx_work_tile<S, kAssign, false, kPPT / 2>(table_s, g, fx, acc_lane, (uint32_t)k, queue, qindex, assign, base, live, lane, lt_mask, xs, ys, qn, seen);
      __syncwarp();
      // ... unless the points' tile-relative indices are needed for the assignment output
      while (kAssign ? qn > 0 : qn >= 32) {
        const int take = min(qn, 32);
        qn -= take;
        seen |= settle(qn, take, base);
        __syncwarp();
      }
      if (seen >= (uint32_t)S::kFlushAt) *flag = 1;  // every warp stops after the tile it is on
    }
    const bool done = tile >= ntiles;
    if (done && qn > 0) {  // not reached with kAssign: its queue is empty after every tile
      settle(0, qn, 0);  // one more add per lane at most: inside kFlushAt's allowance, and a flush follows anyway
      qn = 0;
    }
    // every warp has stopped adding: fold the lane words into the per-cluster totals and clear them
    const int all_done = __syncthreads_and(done);
    for (int c = warp; c < k; c += S::kWarps) {
      uint32_t *const row = acc + (size_t)c * (kFxWords * kLanes) + lane;
      const bool on = lane < kLanes;
      const uint32_t cnt = on ? row[4 * kLanes] : 0u;
      if (__any_sync(0xFFFFFFFFu, cnt != 0u)) {
        long long sum[kFxWords];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const int32_t v = on ? (int32_t)(row[w * kLanes] - cnt * kFxMagicBits) : 0;
          const int hi16 = __reduce_add_sync(0xFFFFFFFFu, v >> 16);
          const unsigned lo16 = __reduce_add_sync(0xFFFFFFFFu, (unsigned)v & 0xFFFFu);
          sum[w] = (long long)hi16 * 65536 + (long long)lo16;
          if (on) row[w * kLanes] = 0u;
        }
        sum[4] = (long long)__reduce_add_sync(0xFFFFFFFFu, cnt);
        if (on) row[4 * kLanes] = 0u;
        if (lane < kFxWords) {
          long long mine = sum[0];
#pragma unroll
          for (int w = 1; w < kFxWords; ++w) mine = lane == w ? sum[w] : mine;
          tot[c * kFxWords + lane] += mine;
        }
      }
    }
    if (threadIdx.x == 0) *flag = 0;
    __syncthreads();
    if (all_done) break;
  }

  // ---- totals -> FP64 -> global, once per CTA -------------------------------------------------
  const double inv_sx = 1.0 / (double)fx.sx, inv_sy = 1.0 / (double)fx.sy;
  const double lo_unit = 1.0 / 4194304.0;
  for (int c = threadIdx.x; c < k; c += S::kThreads) {
    const long long *t = tot + c * kFxWords;
    const unsigned long long m = (unsigned long long)t[4] + late_cnt[c];
    if (m != 0ull) {
      const double sx = ((double)t[0] + (double)t[1] * lo_unit) * inv_sx + late_sum[2 * c];
      const double sy = ((double)t[2] + (double)t[3] * lo_unit) * inv_sy + late_sum[2 * c + 1];
      atomicAdd(&g_sums[2 * c], sx);
      atomicAdd(&g_sums[2 * c + 1], sy);
      atomicAdd(&g_counts[c], m);
    }
  }
}

using FxWide = FxShape<24, 8, 32, 512>;    // k <= 128: a private word per lane
using FxNarrow = FxShape<24, 8, 16, 128>;  // k <= 254 (the 8-bit table): two lanes share a word

template <class S>
size_t fx_smem(int G, int k, bool with_assign) {
  return fx_head_bytes(G, k) + fx_body_bytes<S>(k, with_assign);
}

template <class S>
cudaError_t launch_fx(const float *d_pts, int64_t n, const float *d_cent, int k, const Grid2D &g, const FxParams &fx,
                      const uint16_t *d_table, const uint16_t *d_cand, const Grid2DOuter &outer, double *d_sums,
                      unsigned long long *d_counts, int32_t *d_assign, int sm_count, size_t smem_optin, cudaStream_t s) {
  const bool asg = d_assign != nullptr;
  const size_t smem = fx_smem<S>(g.G, k, asg);
  if (smem > smem_optin) return cudaErrorInvalidValue;
  const bool vec4 = (reinterpret_cast<uintptr_t>(d_pts) & 15) == 0;
  void (*kern)(const float *, long long, const float *, int, Grid2D, FxParams, const uint16_t *, const uint16_t *,
               Grid2DOuter, double *, unsigned long long *, int32_t *);
  kern = vec4 ? (asg ? grid2d_fx_kernel<S, true, true> : grid2d_fx_kernel<S, true, false>)
              : (asg ? grid2d_fx_kernel<S, false, true> : grid2d_fx_kernel<S, false, false>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long ntiles = (n + S::kTile - 1) / S::kTile;
  if (ntiles > 0x7FF00000ll) return cudaErrorInvalidValue;  // the kernel counts tiles in 32 bits (5.5e11 points: 4.4 TB)
  const long long want = (ntiles + S::kWarps - 1) / S::kWarps;
  const int grid = (int)(want < sm_count ? want : sm_count);
  kern<<<grid, S::kThreads, smem, s>>>(d_pts, (long long)n, d_cent, k, g, fx, d_table, d_cand, outer, d_sums, d_counts,
                                       d_assign);
  return cudaGetLastError();
}

}  // namespace

// Largest table the fixed-point form can keep beside its accumulators (0 = this k does not fit it).
int grid2d_fx_table_size(int k, size_t smem_optin, size_t table_cap) {
  if (k > kTable8MaxK) return 0;
  for (int G = 240; G >= 96; G -= 8) {
    if (table_cap && grid2d_table_bytes(G) > table_cap) continue;  // a caller with a fixed-size table buffer
    const size_t need = k <= 128 ? fx_smem<FxWide>(G, k, true) : fx_smem<FxNarrow>(G, k, true);
    if (need <= smem_optin) return G;
  }
  return 0;
}

cudaError_t launch_grid2d_fx_map(const float *d_pts, int64_t n, const float *d_cent, int k, const Grid2D &g,
                                 const uint16_t *d_table, const uint16_t *d_cand, double *d_sums,
                                 unsigned long long *d_counts, int32_t *d_assign, int sm_count, size_t smem_optin,
                                 cudaStream_t s, const Grid2DOuter *outer) {
  if (n <= 0) return cudaSuccess;
  if (k > kTable8MaxK || !(g.fsx > 0.f) || !(g.fsy > 0.f)) return cudaErrorInvalidValue;
  FxParams fx;
  fx.sx = g.fsx; fx.sy = g.fsy;
  fx.tinyx = 2.0f / g.fsx; fx.tinyy = 2.0f / g.fsy;
  fx.limx = 2097152.0f / g.fsx; fx.limy = 2097152.0f / g.fsy;
  fx.rows = (uint32_t)k + 1u;
  Grid2DOuter none{};
  const Grid2DOuter &o = outer ? *outer : none;
  if (k <= 128)
    return launch_fx<FxWide>(d_pts, n, d_cent, k, g, fx, d_table, d_cand, o, d_sums, d_counts, d_assign, sm_count,
                             smem_optin, s);
  return launch_fx<FxNarrow>(d_pts, n, d_cent, k, g, fx, d_table, d_cand, o, d_sums, d_counts, d_assign, sm_count,
                             smem_optin, s);
}

#ifdef HGM_CHECKED
void grid2d_fx_checked_failures(unsigned int out[2]) { cudaMemcpyFromSymbol(out, g_check_failures, 8); }
#endif

}  // namespace hgm
