// kernels_grid2d.cu -- K1: the d == 2 K-means map kernel for sm_100a.
//
// Fills the reference's map-body and combiner slots for a whole split at once:
//   Mapper::map   c++/pipes/api/hadoop/Pipes.hh:158-161 (called HadoopPipes.cc:884-885)
//   Reducer::reduce as combiner  Pipes.hh:166-169 (called HadoopPipes.cc:700-702)
//
// Why not a brute-force scan: at d=2, k=128 a scan costs ~8 FP32 instructions per
// (point, centroid) = ~1000 per point, two orders above what an 8-byte point may cost at
// HBM speed.  So the centroid table is turned, once per
// hgm_set_centroids, into a cell index: a G x G grid over the centroids' bounding box
// whose cells list the ONLY centroids that can win the FP32 scan for any point of the
// cell.  Most cells list one centroid (no arithmetic at all); cells on a Voronoi edge
// list 2..8 and are scanned exactly, in ascending index with strict '<', i.e. the
// reference scan restricted to a superset of its possible winners -- bit-identical result.
//
// Exactness of the pruning (builder, FP64): c0 = centroid nearest the cell centre.  c is
// dropped iff for EVERY p of the (slightly inflated) cell box
//     D(c,p) - D(c0,p) > 1e-5 * (maxD(c) + maxD(c0)) + 1e-44,
// D being exact squared distance.  D(c,.) - D(c0,.) is linear in p, so its minimum over the
// box sits at a corner picked per axis by the sign of (c - c0).  The FP32 distance the scan
// computes is within (1 +- 4*2^-24) of D plus at most 3e-45 of underflow, so a dropped c
// has a strictly larger COMPUTED distance than c0 everywhere in the cell and can never be a
// minimiser of the scan, let alone its lowest-index one.
//
// Combiner: per tile (1024 points per thread group at k <= 128) a counting sort by cluster in
// shared memory -- one native integer ATOMS per point gives the rank, the point goes to slot
// [rank][cluster] of a striped buffer (shared FP32/FP64 atomic adds are CAS loops on sm_100) --
// then each cluster's column is summed in FP64 by its owner lanes, which keep their partials in
// registers for the whole kernel.  The FP64 partials leave the SM once, as k*3 RED.ADD.F64/U64.
#include <stdlib.h>

#include <type_traits>

#include "hgm_internal.h"
#include "grid2d_common.cuh"

namespace hgm {

namespace {


// Shape of the map kernel: kGroups independent thread groups of kGroupThreads per CTA, each
// working on tiles of kGroupThreads * kPPT points; kRegPrefetch keeps the next tile's loads
// in registers, otherwise the next tile is pulled into L2 with prefetch instructions.
template <int kGroupsT, int kGroupThreadsT, bool kRegPrefetchT, int kPPTT = 8>
struct Shape {
  static constexpr int kPPT = kPPTT;  // points per thread per tile (even)
  static constexpr int kGroups = kGroupsT;
  static constexpr int kGroupThreads = kGroupThreadsT;
  static constexpr int kThreads = kGroupsT * kGroupThreadsT;
  static constexpr int kTile = kGroupThreadsT * kPPTT;
  static constexpr bool kRegPrefetch = kRegPrefetchT;
};


// ---- index builder: one thread per cell -------------------------------------------------
__global__ void __launch_bounds__(256) grid2d_build_kernel(const float *__restrict__ cent, int k,
                                                           Grid2D g, uint16_t *__restrict__ table,
                                                           uint16_t *__restrict__ cand) {
  // table8 (one byte per cell, for k <= kTable8MaxK) follows the 16-bit table in the same buffer
  uint8_t *const table8 = reinterpret_cast<uint8_t *>(table + (size_t)g.G * g.G);
  extern __shared__ double2 sc[];
  for (int c = threadIdx.x; c < k; c += blockDim.x)
    sc[c] = make_double2((double)cent[2 * c], (double)cent[2 * c + 1]);
  __syncthreads();
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= g.G * g.G) return;
  const int ix = cell % g.G, iy = cell / g.G;

  const double isx = 1.0 / (double)g.sx, isy = 1.0 / (double)g.sy;
  double xlo = ((double)(ix - g.bx) - 0.5) * isx, xhi = ((double)(ix - g.bx) + 0.5) * isx;
  double ylo = ((double)(iy - g.by) - 0.5) * isy, yhi = ((double)(iy - g.by) + 0.5) * isy;
  // inflate: the kernel's FMA rounds x*s + B once, ties land on either side, and the
  // divisions above carry an FP64 rounding of their own.
  const double px = 1e-6 * isx + 1e-9 * fmax(fabs(xlo), fabs(xhi));
  const double py = 1e-6 * isy + 1e-9 * fmax(fabs(ylo), fabs(yhi));
  xlo -= px; xhi += px; ylo -= py; yhi += py;
  const double mx = 0.5 * (xlo + xhi), my = 0.5 * (ylo + yhi);

  int c0 = 0;
  double best = INFINITY;
  for (int c = 0; c < k; ++c) {
    const double dx = mx - sc[c].x, dy = my - sc[c].y;
    const double dd = dx * dx + dy * dy;
    if (dd < best) { best = dd; c0 = c; }
  }
  const double2 p0 = sc[c0];
  const double fx0 = fmax(fabs(xlo - p0.x), fabs(xhi - p0.x));
  const double fy0 = fmax(fabs(ylo - p0.y), fabs(yhi - p0.y));
  const double d0max = fx0 * fx0 + fy0 * fy0;

  int cnt = 0;
  uint16_t list[kMaxCand];
  for (int c = 0; c < k; ++c) {
    bool keep = (c == c0);
    if (!keep) {
      const double2 pc = sc[c];
      const double qx = (pc.x - p0.x) > 0.0 ? xhi : xlo;  // corner minimising D(c)-D(c0)
      const double qy = (pc.y - p0.y) > 0.0 ? yhi : ylo;
      const double dc = (qx - pc.x) * (qx - pc.x) + (qy - pc.y) * (qy - pc.y);
      const double d0 = (qx - p0.x) * (qx - p0.x) + (qy - p0.y) * (qy - p0.y);
      const double fxc = fmax(fabs(xlo - pc.x), fabs(xhi - pc.x));
      const double fyc = fmax(fabs(ylo - pc.y), fabs(yhi - pc.y));
      const double tau = 1e-5 * (fxc * fxc + fyc * fyc + d0max) + 1e-44;
      keep = !((dc - d0) > tau);  // written so that a NaN keeps the candidate
    }
    if (keep) {
      if (cnt < kMaxCand) list[cnt] = (uint16_t)c;
      ++cnt;
    }
  }
  uint16_t entry;
  if (ix == g.G - 1 || iy == g.G - 1) {
    entry = kCellBrute;  // the kernel clamps everything outside the grid onto this border
  } else if (cnt == 1) {
    entry = (uint16_t)c0;
  } else if (cnt <= kMaxCand) {
    entry = (uint16_t)(kCellMulti | cnt);
    for (int i = 0; i < kMaxCand; ++i) cand[(size_t)cell * kMaxCand + i] = i < cnt ? list[i] : kCandEnd;
    if (cnt > 2) cand[(size_t)cell * kMaxCand] = (uint16_t)(list[0] | kCandMore);  // first word = whole answer otherwise
  } else {
    entry = kCellBrute;
  }
  // a "scan everything" cell says so in its candidate list too: the map kernel resolves a queued
  // point from the list alone: one global load instead of two dependent ones (16-bit entry, then list)
  if (entry == kCellBrute)
    for (int i = 0; i < kMaxCand; ++i) cand[(size_t)cell * kMaxCand + i] = kCandEnd;
  table[cell] = entry;
  // The fixed-point combiner (kernels_grid2d_fx.cu) cannot hold a non-zero |x| < 2 / fsx exactly.  Such a
  // value can only lie in a cell whose (inflated) box reaches into that strip around x = 0, or y = 0: those
  // cells are "not direct" in ITS 8-bit table and their points are settled one by one, where the test is made.
  bool fx_slow = false;
  if (g.shape == kShapeFx) {
    const double tx = 2.0 / (double)g.fsx, ty = 2.0 / (double)g.fsy;
    fx_slow = (xlo < tx && xhi > -tx) || (ylo < ty && yhi > -ty);
  }
  if (g.shape == kShapeFx) {
    // that form's own table, 16 bits per cell in place of the 8-bit copy: the only candidate, or both candidates of
    // a two-candidate cell (id0 | id1 << 8, ascending, so >= 256), or 0xFFFF = ask the index in global memory
    uint16_t *const fx_table = table + (size_t)g.G * g.G;
    uint16_t e = 0xFFFFu;
    if (!fx_slow && entry < kCellMulti) e = entry;
    else if (!fx_slow && entry == (kCellMulti | 2)) e = (uint16_t)(list[0] | (list[1] << 8));
    fx_table[cell] = e;
  } else {
    table8[cell] = entry < kCellMulti ? (uint8_t)(entry < kTable8Multi ? entry : kTable8Multi) : kTable8Multi;
  }
}

// ---- the map kernel ---------------------------------------------------------------------
// One persistent CTA per SM made of independent thread groups.  The groups share the cell
// table and the centroids (read-only after the prologue) and nothing else: each owns its
// tiles, its sort buffer, bin counters and FP64 partials, and synchronises on its own named
// barrier, so one group's sum phase overlaps another's load/lookup/scatter phase.
//
// The sort buffer holds, per group:
//   - a [rank][cluster] "striped" array, k clusters wide and R rows high: a point of a
//     single-candidate cell goes to slot rank*k + cluster, its rank coming from one native
//     integer ATOMS on the cluster's counter -- no offsets, no scan;
//   - two queues (they alternate by tile) for the points of multi-candidate cells;
//   - an overflow region at the top for ranks that do not fit (heavy skew).
//
// Per tile (kGroupThreads * 8 points) and group, two phases and two barriers:
//   1a each thread takes 8 points: cell -> (cluster | queue) -> rank -> store.  Straight line.
//   1b the queue the PREVIOUS tile filled is drained, one point per thread: exact candidate
//      scan, then the point is stored into this tile's striped array like any other.
//   2  owner lanes walk their cluster's column (consecutive lanes = consecutive addresses,
//      conflict-free) and add it in FP64 to partials they keep in registers for the whole
//      kernel; overflow points are added with a shared-memory CAS.
// A tile that fills even the overflow region is redone by the packed counting-sort path.

constexpr int kHistExtra = 4;  // queue counters [2], overflow counters [2]

// per-group shared memory:
// [sort buffer slots float2][acc 2k f64][cnt k i64][hist k+4 i32][offs k+1 i32][late k i32]
// [flag 16 B][lidx (slots) u16, assign only]
__host__ __device__ inline size_t group_small_bytes(int k) {
  return (size_t)k * 16 + (size_t)k * 8 + align16((size_t)(k + kHistExtra) * 4) +
         align16((size_t)(k + 1) * 4) + align16((size_t)k * 4) + 16;
}
__host__ __device__ inline size_t group_bytes(int slots, int k, bool with_assign) {
  return (size_t)slots * 8 + group_small_bytes(k) + (with_assign ? align16((size_t)slots * 2) : 0);
}
__host__ __device__ inline bool use_table8(int k) { return k <= kTable8MaxK; }
__host__ __device__ inline size_t shared_head_bytes(int G, int k) {
  return align16((size_t)G * G * (use_table8(k) ? 1 : 2)) + align16((size_t)k * 8) + 16;
}

__device__ __forceinline__ void group_sync(int group, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "r"(nthreads) : "memory");
}


constexpr uint16_t kNoIndex = 0xFFFF;  // lidx value: this entry's assignment is already written

// Threads per cluster in the column-sum phase, and the row stride of the striped buffer that goes
// with it.  A warp's LDS.64 is served in two half-warp phases of 128 bytes; the 16 lanes of a phase
// read (16 / tpc) neighbouring clusters from tpc consecutive rows, so consecutive rows must start
// (16 / tpc) slots further along the 16-slot bank row: stride = that (mod 16).  With stride = k
// (a multiple of 16 at k = 128) the rows of a phase landed on the same banks: the column reads,
// a quarter of all shared-memory wavefronts, ran at 3.7 wavefronts per instruction instead of 2.
__host__ __device__ inline int column_threads(int k, int group_threads) {
  int tpc = 1;
  while (tpc < 32 && tpc * 2 * k <= group_threads) tpc *= 2;
  return tpc;
}
__host__ __device__ inline int row_stride(int k, int tpc) {
  const int off = tpc >= 32 ? 1 : (tpc > 1 ? 16 / tpc : 0);
  int rs = (k / 16) * 16 + off;
  if (rs < k) rs += 16;
  return rs;
}

template <class S, bool kVec4, bool kAssign, bool kT8>
__global__ void __launch_bounds__(S::kThreads, 1)
    grid2d_map_kernel(const float *__restrict__ pts, long long n, const float *__restrict__ cent,
                      int k, Grid2D g, int slots, int rcap, int qcap,
                      const uint16_t *__restrict__ table, const uint16_t *__restrict__ cand,
                      double *__restrict__ g_sums, unsigned long long *__restrict__ g_counts,
                      int32_t *__restrict__ assign) {
  constexpr int kGT = S::kGroupThreads;
  constexpr int kTile = S::kTile;
  constexpr int kPPT = S::kPPT;
  using entry_t = typename std::conditional<kT8, uint8_t, uint16_t>::type;
  constexpr uint32_t kMulti = kT8 ? (uint32_t)kTable8Multi : (uint32_t)kCellMulti;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int group = threadIdx.x / kGT;
  const int t = threadIdx.x % kGT;  // thread within the group
  const int rs = row_stride(k, column_threads(k, kGT));  // slots per row of the striped buffer (>= k)
  const int qoff = rcap * rs;                     // queue 0 starts here, queue 1 at qoff + qcap
  const int tend = slots - qoff - 2 * qcap;       // capacity of the overflow region (slots-1 downwards)

  // carve shared memory (pointers computed, not looked up: no local-memory struct)
  const size_t table_bytes_al = align16((size_t)g.G * g.G * sizeof(entry_t));
  entry_t *const table_s = reinterpret_cast<entry_t *>(smem_raw);
  float2 *const cent_s = reinterpret_cast<float2 *>(smem_raw + table_bytes_al);
  uint64_t *const bar = reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(cent_s) + align16((size_t)k * 8));
  unsigned char *const gb = smem_raw + shared_head_bytes(g.G, k) + (size_t)group * group_bytes(slots, k, kAssign);
  float2 *const sorted = reinterpret_cast<float2 *>(gb);
  double *const acc = reinterpret_cast<double *>(gb + (size_t)slots * 8);
  long long *const cnt = reinterpret_cast<long long *>(acc + 2 * k);
  int *const hist = reinterpret_cast<int *>(cnt + k);  // k cluster counters, 2 queue counters, 2 overflow counters
  int *const offs = reinterpret_cast<int *>(reinterpret_cast<unsigned char *>(hist) + align16((size_t)(k + kHistExtra) * 4));
  int *const late = reinterpret_cast<int *>(reinterpret_cast<unsigned char *>(offs) + align16((size_t)(k + 1) * 4));
  int *const flag = reinterpret_cast<int *>(reinterpret_cast<unsigned char *>(late) + align16((size_t)k * 4));
  uint16_t *const lidx = reinterpret_cast<uint16_t *>(reinterpret_cast<unsigned char *>(flag) + 16);
  // "the sort buffer is free again": one arrival per warp at the end of a tile, waited for just before
  // the next tile's first write to the buffer -- the global loads, the cell arithmetic and the table
  // lookups of the next tile run in the shadow of the slowest warp's column sums
  uint64_t *const free_bar = reinterpret_cast<uint64_t *>(flag + 2);

  // stage the cell table and the centroids with TMA bulk copies (SASS: UBLKCP)
  const unsigned char *table_src = kT8 ? reinterpret_cast<const unsigned char *>(table + (size_t)g.G * g.G)
                                       : reinterpret_cast<const unsigned char *>(table);
  const uint32_t table_bytes = (uint32_t)table_bytes_al;
  const uint32_t cent_bytes = (uint32_t)align16((size_t)k * 8);
  if (threadIdx.x == 0) mbar_init(bar, 1);
  if (t == 0) mbar_init(free_bar, kGT / 32);
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, table_bytes + cent_bytes);
    for (uint32_t o = 0; o < table_bytes; o += 32768) {
      const uint32_t sz = table_bytes - o < 32768 ? table_bytes - o : 32768;
      tma_bulk_g2s(reinterpret_cast<unsigned char *>(table_s) + o, table_src + o, sz, bar);
    }
    tma_bulk_g2s(cent_s, cent, cent_bytes, bar);
  }
  for (int c = t; c < k; c += kGT) {
    late[c] = 0;
    acc[2 * c] = 0.0;
    acc[2 * c + 1] = 0.0;
    cnt[c] = 0;
  }
  for (int c = t; c < k + kHistExtra; c += kGT) hist[c] = 0;
  if (t == 0) *flag = 0;
  mbar_wait(bar, 0);
  __syncthreads();

  // threads per cluster in the column-sum phase (power of two, <= 32); when every cluster has
  // its own fixed lanes (one pass) their partials stay in registers across tiles
  const int tpc = column_threads(k, kGT);
  const int owners = kGT / tpc;  // clusters covered per pass
  const int sub = t & (tpc - 1);
  const bool one_pass = k <= owners;
  const int my_c = t / tpc;
  double reg_x = 0.0, reg_y = 0.0;
  long long reg_n = 0;

  const long long ntiles = (n + kTile - 1) / kTile;
  const long long tile_stride = (long long)gridDim.x * S::kGroups;
  const long long tile0 = (long long)blockIdx.x * S::kGroups + group;

  // raw tile loads: thread t owns the point pairs (i*kGT + t), i = 0..kPPT/2-1
  auto load_tile = [&](long long tile, float4 (&buf)[kPPT / 2]) {
    const long long base = tile * kTile;
    const float *src = pts + 2 * base + 4 * t;
    if (base + kTile <= n) {  // full tile: no bounds checks
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        if (kVec4) {
          // streamed once: no L1 allocation (SASS LDG.E.NA.128), the little L1 left beside 227 KB of shared
          // memory stays with the candidate lists
          asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(buf[i].x), "=f"(buf[i].y), "=f"(buf[i].z), "=f"(buf[i].w)
                       : "l"(src + 4 * i * kGT));
        } else {
          const float2 a = __ldcs(reinterpret_cast<const float2 *>(src + 4 * i * kGT));
          const float2 b = __ldcs(reinterpret_cast<const float2 *>(src + 4 * i * kGT + 2));
          buf[i] = make_float4(a.x, a.y, b.x, b.y);
        }
      }
    } else {
      const int live = (int)(n - base);
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        const int local = 2 * (i * kGT + t);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (local < live) {
          const float2 a = __ldcs(reinterpret_cast<const float2 *>(src + 4 * i * kGT));
          v.x = a.x; v.y = a.y;
        }
        if (local + 1 < live) {
          const float2 b = __ldcs(reinterpret_cast<const float2 *>(src + 4 * i * kGT + 2));
          v.z = b.x; v.w = b.y;
        }
        buf[i] = v;
      }
    }
  };
  // pull the next tile into L2 ahead of use: one 128-byte line per prefetch
  auto prefetch_tile = [&](long long tile) {
    constexpr int kLines = kTile * 8 / 128;
    if (t < kLines && (tile + 1) * kTile <= n)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(pts + 2 * tile * kTile) + t * 128));
  };
  // add a partial to the group's accumulators from any thread (shared FP64 atomic = CAS loop;
  // only overflow points and the packed path come here)
  auto flush = [&](int c, double ax, double ay) {
    atomicAdd(&acc[2 * c], ax);
    atomicAdd(&acc[2 * c + 1], ay);
  };
  // resolve one stored point exactly and add it with CAS (overflow region, packed path)
  auto settle = [&](float2 v, long long base, uint16_t li) {
    const uint32_t cell = cell_of(v.x, v.y, g);
    const uint32_t e = table_s[cell];
    int c = (int)e;
    if (e >= kMulti) {
      const uint16_t *list = cand + (size_t)cell * kMaxCand;
      c = resolve_cand(v.x, v.y, __ldg(reinterpret_cast<const uint32_t *>(list)), list, cent_s, k);
    }
    flush(c, (double)v.x, (double)v.y);
    atomicAdd(&late[c], 1);
    if (kAssign && li != kNoIndex) assign[base + li] = c;
  };
  // rank a resolved point in cluster c and store it; ranks beyond the striped rows go to the
  // overflow region of parity `par`; returns false if even that is full
  auto place = [&](float x, float y, int c, int par, uint16_t li) -> bool {
    const int r = atomicAdd(&hist[HGM_BOUND(c, k)], 1);
    int pos = r * rs + c;
    if (r >= rcap) {
      const int r2 = atomicAdd(&hist[k + 2 + par], 1);
      if (r2 >= tend) return false;
      pos = slots - 1 - r2;
    }
    sorted[HGM_BOUND(pos, slots)] = make_float2(x, y);
    if (kAssign) lidx[pos] = li;
    return true;
  };

  float4 buf[kPPT / 2];
  if (S::kRegPrefetch) {
    if (tile0 < ntiles) load_tile(tile0, buf);
  } else {
    if (tile0 < ntiles) prefetch_tile(tile0);
  }
  int par = 0;
  long long prev_base = 0;
  bool pending = false;  // the previous tile left points in its queue
  uint32_t tiles_done = 0;  // tiles this group has finished = arrivals phases completed or pending on free_bar
  for (long long tile = tile0; tile < ntiles || pending; tile += tile_stride, par ^= 1, ++tiles_done) {
    bool buffer_free = tiles_done == 0;
    auto wait_free = [&]() {
      if (!buffer_free) mbar_wait(free_bar, (tiles_done - 1) & 1);
      buffer_free = true;
    };
    const bool has_tile = tile < ntiles;
    const long long base = tile * kTile;  // first point of the tile (even)
    const int live = has_tile ? (int)min((long long)kTile, n - base) : 0;
    const int my_q = qoff + par * qcap;            // this tile's queue
    const int prev_q = qoff + (par ^ 1) * qcap;    // the previous tile's
    float xs[kPPT], ys[kPPT];

    if (has_tile) {
      if (!S::kRegPrefetch) {
        load_tile(tile, buf);
        if (tile + tile_stride < ntiles) prefetch_tile(tile + tile_stride);
      }
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        xs[2 * i] = buf[i].x; ys[2 * i] = buf[i].y; xs[2 * i + 1] = buf[i].z; ys[2 * i + 1] = buf[i].w;
      }
      if (S::kRegPrefetch && tile + tile_stride < ntiles) load_tile(tile + tile_stride, buf);
    }

    // ---- phase 1a: cell -> (cluster | queue) -> rank -> slot --------------------------------
    bool lost = false;
    uint32_t qid01 = 0;
    int *const my_qcount = &hist[k + par];
    if (live == kTile) {  // full tile: straight-line code, no bounds checks
      // Three passes over the thread's points, not one loop: the table is a byte array, so the
      // compiler must keep every later table load behind every earlier store to the sort buffer, and a
      // single loop became one serial chain LDS -> ATOMS -> STS per point, 8 times in a row (seen in
      // SASS: the same three registers reused).  Split, the 8 loads, then the 8 atomics, then the 8
      // stores each go out back to back and their latencies overlap.
      uint32_t ent[kPPT];
      int rank[kPPT];
#pragma unroll
      for (int i = 0; i < kPPT; ++i) ent[i] = table_s[cell_of(xs[i], ys[i], g)];
      wait_free();
#pragma unroll
      for (int i = 0; i < kPPT; ++i) rank[i] = atomicAdd(ent[i] < kMulti ? &hist[HGM_BOUND(ent[i], k)] : my_qcount, 1);
      // in the shadow of the eight atomics: the first word of the candidate list of this thread's
      // queued point (previous tile), so that phase 1b does not start a global load in front of the
      // group barrier.  One register; 0 = not loaded (two distinct ids never give 0).
      if (pending && t < min(hist[k + (par ^ 1)], qcap)) {
        const float2 qv = sorted[prev_q + t];
        qid01 = __ldg(reinterpret_cast<const uint32_t *>(cand + (size_t)cell_of(qv.x, qv.y, g) * kMaxCand));
      }
#pragma unroll
      for (int i = 0; i < kPPT; ++i) {
        const uint32_t e = ent[i];
        const bool resolved = e < kMulti;
        const int r = rank[i];
        const int local = 2 * ((i >> 1) * kGT + t) + (i & 1);
        if (kAssign && resolved) assign[base + local] = (int)e;
        int pos = resolved ? r * rs + (int)e : my_q + r;
        bool ok = true;
        if (__any_sync(0xFFFFFFFFu, r >= (resolved ? rcap : qcap))) {  // rare: a bin outgrew its room
          if (r >= (resolved ? rcap : qcap)) {
            const int r2 = atomicAdd(&hist[k + 2 + par], 1);
            ok = r2 < tend;
            pos = slots - 1 - r2;
          }
          lost |= !ok;
        }
        if (ok) {
          sorted[HGM_BOUND(pos, slots)] = make_float2(xs[i], ys[i]);
          if (kAssign) lidx[pos] = resolved ? kNoIndex : (uint16_t)local;
        }
      }
    } else if (live > 0) {
      wait_free();
#pragma unroll
      for (int i = 0; i < kPPT; ++i) {
        const int local = 2 * ((i >> 1) * kGT + t) + (i & 1);  // index within the tile
        if (local < live) {
          const uint32_t e = table_s[cell_of(xs[i], ys[i], g)];
          if (e < kMulti) {
            if (kAssign) assign[base + local] = (int)e;
            lost |= !place(xs[i], ys[i], (int)e, par, kNoIndex);
          } else {
            const int r = atomicAdd(my_qcount, 1);
            int pos = my_q + r;
            bool ok = true;
            if (r >= qcap) {
              const int r2 = atomicAdd(&hist[k + 2 + par], 1);
              ok = r2 < tend;
              pos = slots - 1 - r2;
            }
            if (ok) {
              sorted[HGM_BOUND(pos, slots)] = make_float2(xs[i], ys[i]);
              if (kAssign) lidx[pos] = (uint16_t)local;
            }
            lost |= !ok;
          }
        }
      }
    }
    // ---- phase 1b: drain the previous tile's queue into this tile's striped array ------------
    wait_free();
    if (pending) {
      const int nq = min(hist[k + (par ^ 1)], qcap);
      for (int u = t; u < nq; u += kGT) {
        const float2 v = sorted[prev_q + u];
        // the candidate list alone decides (one global load; the 16-bit entry + list were two dependent
        // ones, in front of the group barrier: 0.733 -> 0.693 ms per 200 M points)
        const uint16_t *list = cand + (size_t)cell_of(v.x, v.y, g) * kMaxCand;
        uint32_t id01 = u == t ? qid01 : 0u;
        if (id01 == 0u) id01 = __ldg(reinterpret_cast<const uint32_t *>(list));
        const int c = resolve_cand(v.x, v.y, id01, list, cent_s, k);
        if (kAssign) assign[prev_base + lidx[prev_q + u]] = c;
        lost |= !place(v.x, v.y, c, par, kNoIndex);
      }
    }
    if (lost) *flag = 1;
    group_sync(group, kGT);

    if (*flag == 0) {
      // the other parity's queue and overflow counters were consumed before the barrier
      if (t == 0) { hist[k + (par ^ 1)] = 0; hist[k + 2 + (par ^ 1)] = 0; }
      // ---- phase 2: FP64 column sums by the clusters' owner lanes ----------------------------
      if (one_pass) {
        const bool on = my_c < k;
        const int h = on ? hist[my_c] : 0;
        const int m = min(h, rcap);
        const int step = tpc * rs;
        const float2 *p = sorted + my_c + sub * rs;
        int r = sub;
        for (; r + tpc < m; r += 2 * tpc, p += 2 * step) {  // two rows per trip
          const float2 v0 = p[0];
          const float2 v1 = p[step];
          reg_x += (double)v0.x; reg_y += (double)v0.y;
          reg_x += (double)v1.x; reg_y += (double)v1.y;
        }
        if (r < m) {
          const float2 v = p[0];
          reg_x += (double)v.x; reg_y += (double)v.y;
        }
        __syncwarp();  // all lanes of the cluster have read h
        if (on && sub == 0) { reg_n += m; hist[my_c] = 0; }
      } else {
        for (int c0 = 0; c0 < k; c0 += owners) {  // uniform trip count: shuffles need all lanes
          const int c = c0 + my_c;
          const bool on = c < k;
          const int h = on ? hist[c] : 0;
          const int m = min(h, rcap);
          double ax = 0.0, ay = 0.0;
          const float2 *col = sorted + c;
          for (int r = sub; r < m; r += tpc) {
            const float2 v = col[r * rs];
            ax += (double)v.x;
            ay += (double)v.y;
          }
          for (int o = tpc >> 1; o > 0; o >>= 1) {
            ax += __shfl_xor_sync(0xFFFFFFFFu, ax, o);
            ay += __shfl_xor_sync(0xFFFFFFFFu, ay, o);
          }
          if (on && sub == 0 && h) {
            acc[2 * c] += ax;
            acc[2 * c + 1] += ay;
            cnt[c] += m;
            hist[c] = 0;
          }
        }
      }
      // overflow region: rare, one point per thread, CAS adds.  The owner lanes above update
      // the shared partials without atomics when they are not register-resident: keep the two apart.
      if (!one_pass) group_sync(group, kGT);
      const int nover = min(hist[k + 2 + par], tend);
      for (int u = t; u < nover; u += kGT) {
        const int pos = slots - 1 - u;
        settle(sorted[pos], base, kAssign ? lidx[pos] : kNoIndex);
      }
      pending = has_tile;  // this tile's queue waits for the next phase 1b
      prev_base = base;
    } else {
      // ---- packed path: redo this tile as histogram -> scan -> scatter -> run sums -----------
      // Everything phase 1 stored is discarded; the tile's points are still in xs/ys, the
      // previous tile's queue is still intact (settled here with CAS).
      if (pending) {
        const int nq = min(hist[k + (par ^ 1)], qcap);
        for (int u = t; u < nq; u += kGT) settle(sorted[prev_q + u], prev_base, kAssign ? lidx[prev_q + u] : kNoIndex);
      }
      group_sync(group, kGT);
      for (int c = t; c < k + kHistExtra; c += kGT) hist[c] = 0;
      group_sync(group, kGT);
      int *const ucount = hist + k;  // bin k of the packed layout: every unresolved point
      uint32_t key[kPPT];
#pragma unroll
      for (int i = 0; i < kPPT; ++i) {
        const int local = 2 * ((i >> 1) * kGT + t) + (i & 1);
        key[i] = 0xFFFFFFFFu;
        if (local < live) {
          const uint32_t e = table_s[cell_of(xs[i], ys[i], g)];
          const bool resolved = e < kMulti;
          const int r = atomicAdd(resolved ? &hist[HGM_BOUND(e, k)] : ucount, 1);
          key[i] = (resolved ? e : (uint32_t)k) | ((uint32_t)r << 16);
        }
      }
      group_sync(group, kGT);
      if (t < 32) {  // exclusive scan of the k+1 bins
        const int nb = k + 1;
        const int per = (nb + 31) / 32;
        const int lo = t * per;
        int sum = 0;
        for (int j = 0; j < per; ++j)
          if (lo + j < nb) sum += hist[lo + j];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
          if (t >= o) incl += up;
        }
        int run = incl - sum;
        for (int j = 0; j < per; ++j) {
          const int c = lo + j;
          if (c < nb) { offs[c] = run; run += hist[c]; }
        }
      }
      group_sync(group, kGT);
#pragma unroll
      for (int i = 0; i < kPPT; ++i) {
        if (key[i] != 0xFFFFFFFFu) {
          const int pos = offs[key[i] & 0xFFFF] + (int)(key[i] >> 16);
          sorted[HGM_BOUND(pos, slots)] = make_float2(xs[i], ys[i]);
          if (kAssign) lidx[pos] = (uint16_t)(2 * ((i >> 1) * kGT + t) + (i & 1));
        }
      }
      group_sync(group, kGT);
      for (int c0 = 0; c0 < k; c0 += owners) {
        const int c = c0 + my_c;
        const bool on = c < k;
        const int start = on ? offs[c] : 0, m = on ? hist[c] : 0;
        double ax = 0.0, ay = 0.0;
        for (int j = sub; j < m; j += tpc) {
          const float2 v = sorted[start + j];
          ax += (double)v.x;
          ay += (double)v.y;
        }
        for (int o = tpc >> 1; o > 0; o >>= 1) {
          ax += __shfl_xor_sync(0xFFFFFFFFu, ax, o);
          ay += __shfl_xor_sync(0xFFFFFFFFu, ay, o);
        }
        if (on && sub == 0 && m) {
          acc[2 * c] += ax;
          acc[2 * c + 1] += ay;
          cnt[c] += m;
        }
      }
      group_sync(group, kGT);  // the run sums above write the partials without atomics
      for (int j = offs[k] + t; j < live; j += kGT) settle(sorted[j], base, kAssign ? lidx[j] : kNoIndex);
      group_sync(group, kGT);
      for (int c = t; c < k + kHistExtra; c += kGT) hist[c] = 0;
      if (t == 0) *flag = 0;
      pending = false;
    }
    __syncwarp();  // sort buffer, counters and flag are free for the next tile once every warp has said so
    if ((t & 31) == 0) mbar_arrive(free_bar);
  }
  if (tiles_done) mbar_wait(free_bar, (tiles_done - 1) & 1);  // everybody's late points are in

  // ---- partials -> global: registers first (one_pass), then the group's shared partials ----
  if (one_pass) {
    for (int o = tpc >> 1; o > 0; o >>= 1) {  // every lane takes part in the shuffles
      reg_x += __shfl_xor_sync(0xFFFFFFFFu, reg_x, o);
      reg_y += __shfl_xor_sync(0xFFFFFFFFu, reg_y, o);
    }
    if (my_c < k && sub == 0 && reg_n) {
      atomicAdd(&g_sums[2 * my_c], reg_x);
      atomicAdd(&g_sums[2 * my_c + 1], reg_y);
      atomicAdd(&g_counts[my_c], (unsigned long long)reg_n);
    }
  }
  for (int c = t; c < k; c += kGT) {
    const long long m = cnt[c] + late[c];
    if (m != 0) {
      atomicAdd(&g_sums[2 * c], acc[2 * c]);
      atomicAdd(&g_sums[2 * c + 1], acc[2 * c + 1]);
      atomicAdd(&g_counts[c], (unsigned long long)m);
    }
  }
}

using ShapeA = Shape<2, 512, false>;  // 1024 threads, two groups, L2 prefetch
using ShapeB = Shape<2, 512, true>;   // 1024 threads, register prefetch
using ShapeC = Shape<4, 256, false>;  // 1024 threads, four groups, L2 prefetch
using ShapeD = Shape<3, 256, true>;   // 768 threads, register prefetch
using ShapeE = Shape<8, 128, false>;  // 1024 threads, eight groups of four warps: shorter barriers, smaller tiles
using ShapeF = Shape<6, 128, false>;  // 768 threads, six groups

// Builds made with -DHGM_EXPERIMENTS let HGM_GRID2D_VARIANT force a shape; the product picks it per job.
int forced_variant() {
#ifdef HGM_EXPERIMENTS
  static const int variant = [] {
    const char *v = getenv("HGM_GRID2D_VARIANT");
    return v ? atoi(v) : -1;
  }();
  return variant;
#else
  return -1;
#endif
}

// How the shared memory left after the table is split: slots of the sort buffer per group and
// rows of the striped layout.  The overflow region keeps at least a quarter tile.
struct Plan {
  int slots, rcap, qcap;
  size_t smem;
  bool ok;
};
template <class S>
Plan make_plan(int G, int k, bool with_assign, size_t smem_optin) {
  Plan p{0, 0, 0, 0, false};
  const size_t head = shared_head_bytes(G, k);
  if (head >= smem_optin) return p;
  const size_t per_group = (smem_optin - head) / S::kGroups;
  const size_t fixed = group_small_bytes(k) + 32;
  if (per_group <= fixed) return p;
  size_t slots = (per_group - fixed) / (with_assign ? 10 : 8);
  slots &= ~size_t(7);
  if (slots > 65536) slots = 65536;  // lidx and slot arithmetic are 16-bit
  if ((long long)slots < S::kTile) return p;  // the packed path needs a whole tile
  // queues: 3/16 of a tile each (the benchmark queues ~5 %); overflow region: >= 1/4 tile
  const int qcap = S::kTile * 3 / 16;
  int rcap = ((int)slots - 2 * qcap - S::kTile / 4) / row_stride(k, column_threads(k, S::kGroupThreads));
  if (rcap < 1) rcap = 0;  // no striped rows: every tile takes the packed path
  if (rcap > 1023) rcap = 1023;
  p.slots = (int)slots;
  p.rcap = rcap;
  p.qcap = qcap;
  p.smem = head + S::kGroups * group_bytes((int)slots, k, with_assign);
  p.ok = p.smem <= smem_optin;
  return p;
}

template <class S>
cudaError_t launch_shape(const float *d_pts, int64_t n, const float *d_cent, int k, const Grid2D &g,
                         const uint16_t *d_table, const uint16_t *d_cand, double *d_sums,
                         unsigned long long *d_counts, int32_t *d_assign, int sm_count,
                         size_t smem_optin, cudaStream_t s) {
  const Plan plan = make_plan<S>(g.G, k, d_assign != nullptr, smem_optin);
  if (!plan.ok) return cudaErrorInvalidValue;
  const bool vec4 = (reinterpret_cast<uintptr_t>(d_pts) & 15) == 0;
  const bool asg = d_assign != nullptr;
  const bool t8 = use_table8(k);
  void (*kern)(const float *, long long, const float *, int, Grid2D, int, int, int, const uint16_t *,
               const uint16_t *, double *, unsigned long long *, int32_t *);
  if (t8) {
    kern = vec4 ? (asg ? grid2d_map_kernel<S, true, true, true> : grid2d_map_kernel<S, true, false, true>)
                : (asg ? grid2d_map_kernel<S, false, true, true> : grid2d_map_kernel<S, false, false, true>);
  } else {
    kern = vec4 ? (asg ? grid2d_map_kernel<S, true, true, false> : grid2d_map_kernel<S, true, false, false>)
                : (asg ? grid2d_map_kernel<S, false, true, false> : grid2d_map_kernel<S, false, false, false>);
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem);
  if (e != cudaSuccess) return e;
  const long long ntiles = (n + S::kTile - 1) / S::kTile;
  const long long want = (ntiles + S::kGroups - 1) / S::kGroups;
  const int grid = (int)(want < sm_count ? want : sm_count);
  kern<<<grid, S::kThreads, plan.smem, s>>>(d_pts, (long long)n, d_cent, k, g, plan.slots, plan.rcap,
                                            plan.qcap, d_table, d_cand, d_sums, d_counts, d_assign);
  return cudaGetLastError();
}

template <class S>
bool shape_fits(int G, int k, size_t smem_optin) {
  // a table is usable if the assign variant fits and the plain variant gets >= 2 tiles of slots
  const Plan a = make_plan<S>(G, k, true, smem_optin);
  const Plan b = make_plan<S>(G, k, false, smem_optin);
  return a.ok && b.ok && b.slots >= 2 * S::kTile;
}

}  // namespace

// Pick the table size and kernel shape for k clusters: six groups of 128 threads at k <= 128
// (measured best at k = 128: 0.765 ms against 0.84 for four groups of 256 -- a third of all warp time
// was spent waiting at the two group barriers of a tile, and four warps meet faster than eight),
// four groups of 256 above that while the table can stay at 160+ cells per side, otherwise two
// groups of 512, whose per-group arrays cost half as much shared memory.
bool grid2d_plan(int k, size_t smem_optin, int *G_out, int *shape_out) {
  const int forced = forced_variant();
  auto fits = [&](int shape, int G) {
    switch (shape) {
      case 0: return shape_fits<ShapeA>(G, k, smem_optin);
      case 1: return shape_fits<ShapeB>(G, k, smem_optin);
      case 3: return shape_fits<ShapeD>(G, k, smem_optin);
      case 4: return shape_fits<ShapeE>(G, k, smem_optin);
      case 5: return shape_fits<ShapeF>(G, k, smem_optin);
      default: return shape_fits<ShapeC>(G, k, smem_optin);
    }
  };
  // six groups of 128 need one owner lane per cluster to keep the partials in registers: k <= 128 only
  const int order[3] = {k <= 128 ? 5 : 2, 2, 0};
  for (int pass = 0; pass < 3; ++pass) {
    const int shape = forced >= 0 ? forced : order[pass];
    const int gmin = (forced >= 0 || pass == 2) ? 64 : 160;
    for (int G = 256; G >= gmin; G -= 32)
      if (fits(shape, G)) { *G_out = G; *shape_out = shape; return true; }
    if (forced >= 0) break;
  }
  return false;
}

cudaError_t launch_grid2d_build(const float *d_cent, int k, const Grid2D &g, uint16_t *d_table,
                                uint16_t *d_cand, cudaStream_t s) {
  const int cells = g.G * g.G;
  grid2d_build_kernel<<<(cells + 255) / 256, 256, (size_t)k * sizeof(double2), s>>>(d_cent, k, g,
                                                                                  d_table, d_cand);
  return cudaGetLastError();
}

cudaError_t launch_grid2d_map(const float *d_pts, int64_t n, const float *d_cent, int k,
                              const Grid2D &g, const uint16_t *d_table, const uint16_t *d_cand,
                              double *d_sums, unsigned long long *d_counts, int32_t *d_assign,
                              int sm_count, size_t smem_optin, cudaStream_t s, const Grid2DOuter *outer) {
  if (n <= 0) return cudaSuccess;
  if (k > kMaxK) return cudaErrorInvalidValue;
#define HGM_ARGS d_pts, n, d_cent, k, g, d_table, d_cand, d_sums, d_counts, d_assign, sm_count, smem_optin, s
  switch (g.shape) {
    case 0: return launch_shape<ShapeA>(HGM_ARGS);
    case 1: return launch_shape<ShapeB>(HGM_ARGS);
    case 3: return launch_shape<ShapeD>(HGM_ARGS);
    case 4: return launch_shape<ShapeE>(HGM_ARGS);
    case 5: return launch_shape<ShapeF>(HGM_ARGS);
    case kShapeFx: return launch_grid2d_fx_map(HGM_ARGS, outer);
    default: return launch_shape<ShapeC>(HGM_ARGS);
  }
#undef HGM_ARGS
}

#ifdef HGM_CHECKED
void grid2d_checked_failures(unsigned int out[2]) { cudaMemcpyFromSymbol(out, g_check_failures, 8); }
#endif

}  // namespace hgm
