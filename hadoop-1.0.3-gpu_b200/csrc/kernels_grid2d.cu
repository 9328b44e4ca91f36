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
// Combiner: per tile of 8192 points a counting sort by cluster in shared memory (one native
// integer ATOMS per point gives the rank; shared FP32/FP64 atomic adds are CAS loops on
// sm_100), then each cluster's contiguous run is summed in FP64 by its owner threads.  The
// per-CTA FP64 partials leave the SM once, as k*3 RED.ADD.F64/U64 to global.
#include "hgm_internal.h"

namespace hgm {

namespace {

constexpr int kThreads = 512;
constexpr int kPPT = 16;                    // points per thread per tile
constexpr int kTile = kThreads * kPPT;      // 8192 points = 64 KB of float2
constexpr int kMaxK = 1024;

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

// ---- index builder: one thread per cell -------------------------------------------------
__global__ void __launch_bounds__(256) grid2d_build_kernel(const float *__restrict__ cent, int k,
                                                           Grid2D g, uint16_t *__restrict__ table,
                                                           uint16_t *__restrict__ cand) {
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
  if (cnt == 1) {
    table[cell] = (uint16_t)c0;
  } else if (cnt <= kMaxCand) {
    table[cell] = (uint16_t)(kCellMulti | cnt);
    for (int i = 0; i < kMaxCand; ++i) cand[(size_t)cell * kMaxCand + i] = i < cnt ? list[i] : 0;
  } else {
    table[cell] = kCellBrute;
  }
}

// ---- the map kernel ---------------------------------------------------------------------
struct Smem {
  uint16_t *table;     // G*G
  float2 *sorted;      // kTile
  float2 *cent;        // k (padded to even)
  int *hist;           // k
  int *offs;           // k
  double *acc;         // 2k
  long long *cnt;      // k
  uint64_t *bar;
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

__device__ __forceinline__ Smem carve(unsigned char *base, int G, int k) {
  Smem s;
  size_t o = 0;
  s.table = reinterpret_cast<uint16_t *>(base + o); o += align16((size_t)G * G * 2);
  s.sorted = reinterpret_cast<float2 *>(base + o);  o += (size_t)kTile * 8;
  s.cent = reinterpret_cast<float2 *>(base + o);    o += align16((size_t)k * 8);
  s.acc = reinterpret_cast<double *>(base + o);     o += (size_t)k * 16;
  s.cnt = reinterpret_cast<long long *>(base + o);  o += (size_t)k * 8;
  s.hist = reinterpret_cast<int *>(base + o);       o += align16((size_t)k * 4);
  s.offs = reinterpret_cast<int *>(base + o);       o += align16((size_t)k * 4);
  s.bar = reinterpret_cast<uint64_t *>(base + o);
  return s;
}

// nearest centroid of one point given its cell entry; exact scan over the candidates
__device__ __forceinline__ int resolve_slow(float x, float y, uint32_t entry, uint32_t cell,
                                            const float2 *cent_s, int k,
                                            const uint16_t *__restrict__ cand) {
  float best = INFINITY;
  int arg = 0;
  if (entry == kCellBrute) {
    for (int c = 0; c < k; ++c) {
      const float dd = dist2(x, y, cent_s[c]);
      if (dd < best) { best = dd; arg = c; }
    }
  } else {
    const int cnt = entry & 0x7FFF;
    const uint4 ids = __ldg(reinterpret_cast<const uint4 *>(cand + (size_t)cell * kMaxCand));
    const uint32_t w[4] = {ids.x, ids.y, ids.z, ids.w};
#pragma unroll
    for (int i = 0; i < kMaxCand; ++i) {
      if (i < cnt) {
        const int c = (w[i >> 1] >> ((i & 1) * 16)) & 0xFFFF;
        const float dd = dist2(x, y, cent_s[c]);
        if (dd < best) { best = dd; arg = c; }
      }
    }
  }
  return arg;
}

template <bool kVec4, bool kAssign>
__global__ void __launch_bounds__(kThreads, 1)
    grid2d_map_kernel(const float *__restrict__ pts, long long n, const float *__restrict__ cent,
                      int k, Grid2D g, const uint16_t *__restrict__ table,
                      const uint16_t *__restrict__ cand, double *__restrict__ g_sums,
                      unsigned long long *__restrict__ g_counts, int32_t *__restrict__ assign) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const Smem s = carve(smem_raw, g.G, k);
  const int t = threadIdx.x;
  const int G = g.G;

  // stage the cell table and the centroids with TMA bulk copies
  const uint32_t table_bytes = (uint32_t)align16((size_t)G * G * 2);
  const uint32_t cent_bytes = (uint32_t)align16((size_t)k * 8);
  if (t == 0) mbar_init(s.bar, 1);
  __syncthreads();
  if (t == 0) {
    mbar_expect_tx(s.bar, table_bytes + cent_bytes);
    for (uint32_t o = 0; o < table_bytes; o += 32768) {
      const uint32_t sz = table_bytes - o < 32768 ? table_bytes - o : 32768;
      tma_bulk_g2s(reinterpret_cast<unsigned char *>(s.table) + o,
                   reinterpret_cast<const unsigned char *>(table) + o, sz, s.bar);
    }
    tma_bulk_g2s(s.cent, cent, cent_bytes, s.bar);
  }
  for (int c = t; c < k; c += kThreads) {
    s.hist[c] = 0;
    s.acc[2 * c] = 0.0;
    s.acc[2 * c + 1] = 0.0;
    s.cnt[c] = 0;
  }
  mbar_wait(s.bar, 0);
  __syncthreads();

  // threads per cluster in the segment-sum phase (power of two, <= 32)
  int tpc = 1;
  while (tpc < 32 && tpc * 2 * k <= kThreads) tpc *= 2;
  const int groups = kThreads / tpc;

  const long long ntiles = (n + kTile - 1) / kTile;

  // raw tile loads: thread t owns the point pairs (i*kThreads + t), i = 0..kPPT/2-1
  auto load_tile = [&](long long tile, float4 (&buf)[kPPT / 2]) {
    const long long base = tile * kTile;
#pragma unroll
    for (int i = 0; i < kPPT / 2; ++i) {
      const long long p = base + 2ll * (i * kThreads + t);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (kVec4) {
        if (p + 1 < n) {
          v = __ldcs(reinterpret_cast<const float4 *>(pts + 2 * p));
        } else if (p < n) {
          const float2 h = __ldcs(reinterpret_cast<const float2 *>(pts + 2 * p));
          v.x = h.x; v.y = h.y;
        }
      } else {
        if (p < n) {
          const float2 h = __ldcs(reinterpret_cast<const float2 *>(pts + 2 * p));
          v.x = h.x; v.y = h.y;
        }
        if (p + 1 < n) {
          const float2 h = __ldcs(reinterpret_cast<const float2 *>(pts + 2 * p + 2));
          v.z = h.x; v.w = h.y;
        }
      }
      buf[i] = v;
    }
  };

  float4 nxt[kPPT / 2];
  if ((long long)blockIdx.x < ntiles) load_tile(blockIdx.x, nxt);
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long base = tile * kTile;  // first point of the tile (even)
    float xs[kPPT], ys[kPPT];
    uint32_t key[kPPT];  // cluster | rank << 16, or 0xFFFFFFFF for "no point"

    // ---- phase A: cell lookup, exact resolve, rank (loads were issued a tile ahead) ------
#pragma unroll
    for (int i = 0; i < kPPT / 2; ++i) {
      xs[2 * i] = nxt[i].x; ys[2 * i] = nxt[i].y; xs[2 * i + 1] = nxt[i].z; ys[2 * i + 1] = nxt[i].w;
    }
    if (tile + gridDim.x < ntiles) load_tile(tile + gridDim.x, nxt);
    uint32_t entry[kPPT];
    uint32_t cellv[kPPT];
#pragma unroll
    for (int i = 0; i < kPPT; ++i) {
      const uint32_t ux = __float_as_uint(__fmaf_rn(xs[i], g.sx, g.Bx)) - kMagicBits;
      const uint32_t uy = __float_as_uint(__fmaf_rn(ys[i], g.sy, g.By)) - kMagicBits;
      const bool inside = (ux < (uint32_t)G) && (uy < (uint32_t)G);
      const uint32_t cell = inside ? uy * G + ux : 0u;
      cellv[i] = cell;
      const uint32_t e = s.table[cell];
      entry[i] = inside ? e : (uint32_t)kCellBrute;
    }
#pragma unroll
    for (int i = 0; i < kPPT; ++i) {
      const long long p = base + 2ll * ((i >> 1) * kThreads + t) + (i & 1);
      int c = (int)entry[i];
      if (entry[i] >= kCellMulti) c = resolve_slow(xs[i], ys[i], entry[i], cellv[i], s.cent, k, cand);
      if (p < n) {
        const int r = atomicAdd(&s.hist[c], 1);
        key[i] = (uint32_t)c | ((uint32_t)r << 16);
      } else {
        key[i] = 0xFFFFFFFFu;
      }
    }
    if (kAssign) {
      const bool assign8 = (reinterpret_cast<uintptr_t>(assign) & 7) == 0;
#pragma unroll
      for (int i = 0; i < kPPT / 2; ++i) {
        const long long p = base + 2ll * (i * kThreads + t);
        if (p + 1 < n && assign8) {
          *reinterpret_cast<int2 *>(assign + p) =
              make_int2((int)(key[2 * i] & 0xFFFF), (int)(key[2 * i + 1] & 0xFFFF));
        } else {
          if (p < n) assign[p] = (int)(key[2 * i] & 0xFFFF);
          if (p + 1 < n) assign[p + 1] = (int)(key[2 * i + 1] & 0xFFFF);
        }
      }
    }
    __syncthreads();

    // ---- phase B: exclusive scan of the tile histogram (warp 0) ------------------------
    if (t < 32) {
      const int per = (k + 31) / 32;
      const int lo = t * per;
      int sum = 0;
      for (int j = 0; j < per; ++j)
        if (lo + j < k) sum += s.hist[lo + j];
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (t >= o) incl += up;
      }
      int run = incl - sum;
      for (int j = 0; j < per; ++j)
        if (lo + j < k) { s.offs[lo + j] = run; run += s.hist[lo + j]; }
    }
    __syncthreads();

    // ---- phase C: scatter the points into cluster order --------------------------------
#pragma unroll
    for (int i = 0; i < kPPT; ++i) {
      if (key[i] != 0xFFFFFFFFu) {
        const int c = key[i] & 0xFFFF;
        s.sorted[s.offs[c] + (int)(key[i] >> 16)] = make_float2(xs[i], ys[i]);
      }
    }
    __syncthreads();

    // ---- phase D: FP64 run sums, owner threads fold them into the CTA partials ---------
    {
      const int sub = t & (tpc - 1);
      for (int c0 = 0; c0 < k; c0 += groups) {  // uniform trip count: the shuffles need all lanes
        const int c = c0 + t / tpc;
        const bool live = c < k;
        const int start = live ? s.offs[c] : 0, m = live ? s.hist[c] : 0;
        double ax = 0.0, ay = 0.0;
        for (int j = sub; j < m; j += tpc) {
          const float2 v = s.sorted[start + j];
          ax += (double)v.x;
          ay += (double)v.y;
        }
        for (int o = tpc >> 1; o > 0; o >>= 1) {
          ax += __shfl_xor_sync(0xFFFFFFFFu, ax, o);
          ay += __shfl_xor_sync(0xFFFFFFFFu, ay, o);
        }
        if (live && sub == 0) {
          s.acc[2 * c] += ax;
          s.acc[2 * c + 1] += ay;
          s.cnt[c] += m;
        }
      }
    }
    __syncthreads();
    for (int c = t; c < k; c += kThreads) s.hist[c] = 0;
    // the next tile's phase A only touches hist after its loads; the barrier that follows
    // phase A orders this zeroing against phase B.  It must not race the atomics though:
    __syncthreads();
  }

  // ---- CTA partials -> global (one RED per value per CTA) ------------------------------
  for (int c = t; c < k; c += kThreads) {
    if (s.cnt[c] != 0) {
      atomicAdd(&g_sums[2 * c], s.acc[2 * c]);
      atomicAdd(&g_sums[2 * c + 1], s.acc[2 * c + 1]);
      atomicAdd(&g_counts[c], (unsigned long long)s.cnt[c]);
    }
  }
}

}  // namespace

size_t grid2d_map_smem_bytes(int G, int k) {
  return align16((size_t)G * G * 2) + (size_t)kTile * 8 + align16((size_t)k * 8) + (size_t)k * 16 +
         (size_t)k * 8 + 2 * align16((size_t)k * 4) + 16;
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
                              int sm_count, size_t smem_optin, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  if (k > kMaxK) return cudaErrorInvalidValue;
  const size_t smem = grid2d_map_smem_bytes(g.G, k);
  if (smem > smem_optin) return cudaErrorInvalidValue;
  const bool vec4 = (reinterpret_cast<uintptr_t>(d_pts) & 15) == 0;
  const bool asg = d_assign != nullptr;
  auto kern = vec4 ? (asg ? grid2d_map_kernel<true, true> : grid2d_map_kernel<true, false>)
                   : (asg ? grid2d_map_kernel<false, true> : grid2d_map_kernel<false, false>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long ntiles = (n + kTile - 1) / kTile;
  const int grid = (int)(ntiles < sm_count ? ntiles : sm_count);
  kern<<<grid, kThreads, smem, s>>>(d_pts, (long long)n, d_cent, k, g, d_table, d_cand, d_sums,
                                    d_counts, d_assign);
  return cudaGetLastError();
}

}  // namespace hgm
