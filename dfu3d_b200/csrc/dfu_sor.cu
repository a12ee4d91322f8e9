// Global statistical outlier removal (S4): Open3D PointCloud::RemoveStatisticalOutliers semantics as restated in
// oracle/fakes.py (statistical_outlier_mask).  The call is commented out in the shipped reference
// (tools/PENet/dataloaders/my_loader0.py:735, my_loader_KittiPandasetWaymo.py:767) and named by the north star
// (README.md:22 "global statistical filtering"; BASELINE.json config 3: k = 20, ratio = 2.0).
//
//   mean_i  = (1/k') * sum of the k' = min(k, n) smallest distances from point i (itself, 0.0, included), the
//             square roots added in ascending order
//   mu      = sum over mean_i > 0 of mean_i / n,   sigma = sqrt(sum over mean_i > 0 of (mean_i - mu)^2 / (n - 1))
//   keep_i  = 0 < mean_i < mu + ratio * sigma
//
// Exact k-NN, staged through shared memory with TMA bulk copies (round 2 rebuild; the round-1 kernel walked an implicit
// octree with one thread per point and kept its top-k list in local memory: 7 of 32 lanes active, 25 ms per 4 M points):
//   * ONE radix sort by a 48-bit Hilbert key per segment puts spatial neighbours next to each other; every segment is
//     padded to a multiple of 32 points.  Leaves ("groups") are 8 consecutive sorted points: an exact float64 record
//     (SorGroup) and a float32 screening record (SorGroupF: coordinates relative to the segment's box minimum + their
//     box); an 8-ary pyramid of boxes over the group array is the search structure.
//   * k_sor_sweep: one WARP per tile of 32 consecutive sorted queries.
//       1. a proven bound of the k-th neighbour distance from the +-32 curve neighbours (float32 insertion network in
//          registers over the bulk-copied window, turned into a bound by the error model below);
//       2. the tile's search box (union of the lanes' balls) is intersected with the pyramid level by level; the
//          surviving level-1 nodes are visited from the tile's own node outwards;
//       3. the group boxes of 8 nodes are tested per step (the next step's boxes in flight), the groups that pass are
//          staged 64 at a time with cp.async.bulk + mbarrier (one copy per run of consecutive groups) and SWEPT:
//          every lane evaluates every staged point in float32 (broadcast reads, no atomics, no work list; a group
//          whose box no lane can reach is skipped by one vote) and appends the points inside its ball to its own
//          shared-memory column.  A column about to fill triggers a tightening of all balls (guess a smaller radius
//          from the density seen so far, verify by counting, compact), so one sweep over the region suffices;
//       4. the exact float64 distances of the few dozen survivors are computed in place, the k smallest extracted in
//          ascending order on 32-bit order keys (with an exact fallback for ties) and their square roots added
//          sequentially (bit-identical to std::accumulate over the sorted distances).
//     float32 decides which points are looked at, never a value.  No local memory.
//   * Queries whose bound is far larger than their tile's (salt outliers: their curve neighbours are far away), tiles
//     whose search box meets too many nodes (a jump of the curve inside the tile), and lanes whose column holds more
//     equal distances than it has room for are DEFERRED to k_sor_single: one warp per query, nearest-first descent of
//     the same pyramid with pruning against the current k-th distance, the k best kept sorted across the lanes of the
//     warp (registers + shuffles).  k > 32 runs entirely on that kernel.
// The reductions for mu and sigma are two-stage and order-fixed (chunk partials, then one block per segment), so the
// result does not depend on scheduling.
#include <cub/device/device_radix_sort.cuh>
#include <stdlib.h>
#include "dfu_ops.cuh"

#define SOR_KMAX 64
#define SOR_KFAST 32        /* k_sor_sweep handles k <= 32 */
#define SOR_CHUNK 4096      /* points per partial sum of the mu / sigma reductions */
#ifndef SOR_WARPS
#define SOR_WARPS 2         /* warps (tiles in flight) per CTA of k_sor_sweep */
#endif
#ifndef SOR_FRONT
#define SOR_FRONT 384       /* level-1 frontier capacity per tile */
#endif
#define SOR_MAXLEV 10
#ifndef SOR_DEFER
#define SOR_DEFER 9.0f      /* lanes whose squared bound exceeds 9 x the tile median go to k_sor_single */
#endif
enum { SO_SEGOFF = 170, SO_POFF, SO_BOX, SO_KEY, SO_IDX, SO_SKEY, SO_SIDX, SO_REC, SO_TMP, SO_SPI,
       SO_MEAN, SO_PART, SO_SEGSTAT, SO_SEGCH, SO_PYR, SO_TODO, SO_SCAL, SO_FREC };

struct SorBox { float4 lo, hi; };   // 32 bytes
// 8 consecutive points of the sorted, padded array + their bounding box: ONE contiguous 224-byte record, so that a run
// of consecutive groups is one bulk copy
struct __align__(16) SorGroup {
  double x[8], y[8], z[8];          // pads: +inf
  SorBox box;
};
static_assert(sizeof(SorGroup) == 224, "SorGroup must be 224 bytes");
// float32 copy of a group for the screening sweep: coordinates relative to the segment's box minimum (pads: +inf)
// and the box of those eight float triples (empty for an all-pad group)
struct __align__(16) SorGroupF { float x[8], y[8], z[8]; float lo[3], hi[3], pad[2]; };
static_assert(sizeof(SorGroupF) == 128, "SorGroupF must be 128 bytes");

struct SorTree {
  const SorGroup* rec;               // [NP / 8]
  const SorGroupF* frec;             // [NP / 8] float32 screening copy
  const unsigned long long* box;     // [nseg][6] ordered-bit min xyz, max xyz of the segments
  const int32_t* spi;                // [NP] original index, -1 for pads
  const int64_t* seg_off;            // [nseg+1]
  const int64_t* poff;               // [nseg+1] padded offsets (multiples of 32)
  const SorBox* lv[SOR_MAXLEV];      // lv[0]: compact copy of the group boxes; lv[L]: boxes of 8^L groups
  int nlv[SOR_MAXLEV];
  int nlev, nseg, ntiles, k;
};

// ---------------------------------------------------------------- keys
__device__ __forceinline__ unsigned long long sor_spread3(unsigned v) {
  unsigned long long x = v & 0xffffu;
  x = (x | (x << 32)) & 0x001f00000000ffffull;
  x = (x | (x << 16)) & 0x001f0000ff0000ffull;
  x = (x | (x << 8)) & 0x100f00f00f00f00full;
  x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
  x = (x | (x << 2)) & 0x1249249249249249ull;
  return x;
}
// 48-bit Hilbert index of a 16-bit cell (Skilling's transpose algorithm).  Only the ORDER matters (locality of the
// groups), never the result: any key gives the same neighbours.
__device__ __forceinline__ unsigned long long sor_hilbert(unsigned x, unsigned y, unsigned z) {
  unsigned X0 = x, X1 = y, X2 = z;
  for (unsigned Q = 1u << 15; Q > 1; Q >>= 1) {
    const unsigned P = Q - 1;
    if (X0 & Q) X0 ^= P;
    if (X1 & Q) X0 ^= P; else { const unsigned t = (X0 ^ X1) & P; X0 ^= t; X1 ^= t; }
    if (X2 & Q) X0 ^= P; else { const unsigned t = (X0 ^ X2) & P; X0 ^= t; X2 ^= t; }
  }
  X1 ^= X0; X2 ^= X1;
  unsigned t = 0;
  for (unsigned Q = 1u << 15; Q > 1; Q >>= 1)
    if (X2 & Q) t ^= Q - 1;
  X0 ^= t; X1 ^= t; X2 ^= t;
  return (sor_spread3(X0) << 2) | (sor_spread3(X1) << 1) | sor_spread3(X2);
}

__device__ __forceinline__ int sor_seg_of(const int64_t* __restrict__ off, int nseg, int64_t i) {
  int lo = 0, hi = nseg;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (off[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

// per-segment bounding box (ordered-bit atomics; box[s] = min xyz, max xyz)
__global__ void __launch_bounds__(256) k_sor_bbox(const double* __restrict__ rows, int64_t n, const int64_t* __restrict__ seg_off,
                                                  int nseg, unsigned long long* __restrict__ box, int32_t* __restrict__ errflag) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  int s = -1;
  unsigned long long b[3] = {0, 0, 0};
  if (i < n) {
    s = sor_seg_of(seg_off, nseg, i);
    const double px = rows[3 * i], py = rows[3 * i + 1], pz = rows[3 * i + 2];
    if (!(px == px && py == py && pz == pz) || isinf(px) || isinf(py) || isinf(pz)) atomicOr(errflag, 1);
    b[0] = dfu_order_f64(px); b[1] = dfu_order_f64(py); b[2] = dfu_order_f64(pz);
  }
  // warp reduction when the warp sits in one segment, then ONE atomic per block and axis when the block does too
  // (a single segment of millions of points would otherwise send 6 atomics per warp to the same 6 words)
  __shared__ unsigned long long sh[8][6];
  __shared__ int shs[8];
  const int s0 = __shfl_sync(0xffffffffu, s, 0);
  const bool uniform = __all_sync(0xffffffffu, s == s0) && s0 >= 0;
  if (uniform) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      unsigned long long mn = b[a], mx = b[a];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long on = __shfl_xor_sync(0xffffffffu, mn, o), ox = __shfl_xor_sync(0xffffffffu, mx, o);
        mn = on < mn ? on : mn; mx = ox > mx ? ox : mx;
      }
      if ((threadIdx.x & 31) == 0) { sh[threadIdx.x >> 5][a] = mn; sh[threadIdx.x >> 5][3 + a] = mx; }
    }
  } else if (s >= 0) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      atomicMin(box + 6 * s + a, b[a]);
      atomicMax(box + 6 * s + 3 + a, b[a]);
    }
  }
  if ((threadIdx.x & 31) == 0) shs[threadIdx.x >> 5] = uniform ? s0 : -1;
  __syncthreads();
  if (threadIdx.x < 6) {
    const int a = threadIdx.x;
    bool same = true;
    for (int w = 0; w < 8; w++) same = same && shs[w] == shs[0];
    if (same && shs[0] >= 0) {
      unsigned long long v = sh[0][a];
      for (int w = 1; w < 8; w++) v = a < 3 ? (sh[w][a] < v ? sh[w][a] : v) : (sh[w][a] > v ? sh[w][a] : v);
      if (a < 3) atomicMin(box + 6 * shs[0] + a, v); else atomicMax(box + 6 * shs[0] + a, v);
    } else {
      for (int w = 0; w < 8; w++)
        if (shs[w] >= 0) {
          if (a < 3) atomicMin(box + 6 * shs[w] + a, sh[w][a]); else atomicMax(box + 6 * shs[w] + a, sh[w][a]);
        }
    }
  }
}

__global__ void __launch_bounds__(256) k_sor_keys(const double* __restrict__ rows, int64_t n, const int64_t* __restrict__ seg_off,
                                                  int nseg, const unsigned long long* __restrict__ box,
                                                  unsigned long long* __restrict__ key, int32_t* __restrict__ idx) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int s = sor_seg_of(seg_off, nseg, i);
  double mn[3], ext = 0.0;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    mn[a] = dfu_unorder_f64(box[6 * s + a]);
    const double e = dfu_unorder_f64(box[6 * s + 3 + a]) - mn[a];
    ext = e > ext ? e : ext;
  }
  const double c0 = ext > 0.0 ? ext * (1.0 / 65536.0) * (1.0 + 1e-6) : 1.0;
  unsigned c[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const double v = floor((rows[3 * i + a] - mn[a]) / c0);
    c[a] = v < 0.0 ? 0u : (v > 65535.0 ? 65535u : (unsigned)v);      // NaN -> 0 (the call fails with DFU_ERR_DATA anyway)
  }
  key[i] = ((unsigned long long)(unsigned)s << 48) | sor_hilbert(c[0], c[1], c[2]);
  idx[i] = (int32_t)i;
}

// float32 screening coordinate: the same expression in k_sor_gather (points) and k_sor_sweep (queries)
__device__ __forceinline__ float sor_rel_f32(double v, double origin) { return __double2float_rn(__dsub_rn(v, origin)); }

// one warp per tile of the PADDED sorted domain: group records in sorted order (+inf pads), original index (-1 pads),
// and the bounding box of every group of 8 points, rounded outwards to float (in the record and in lv[0])
__global__ void __launch_bounds__(256) k_sor_gather(const int32_t* __restrict__ sidx, const double* __restrict__ rows,
                                                    const int64_t* __restrict__ seg_off, const int64_t* __restrict__ poff, int nseg,
                                                    int ntiles, SorGroup* __restrict__ rec, int32_t* __restrict__ spi,
                                                    SorBox* __restrict__ gbox, const unsigned long long* __restrict__ box,
                                                    SorGroupF* __restrict__ frec) {
  const int tile = (int)(((int64_t)blockIdx.x * 256 + threadIdx.x) >> 5);
  if (tile >= ntiles) return;
  const int lane = threadIdx.x & 31;
  const int64_t t = (int64_t)tile * 32 + lane;
  const int s = sor_seg_of(poff, nseg, (int64_t)tile * 32);
  const int64_t r = t - poff[s];
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  double px = inf, py = inf, pz = inf;
  int i = -1;
  if (r < seg_off[s + 1] - seg_off[s]) {
    i = sidx[seg_off[s] + r];
    px = rows[3 * (int64_t)i]; py = rows[3 * (int64_t)i + 1]; pz = rows[3 * (int64_t)i + 2];
  }
  SorGroup& G = rec[t >> 3];
  G.x[lane & 7] = px; G.y[lane & 7] = py; G.z[lane & 7] = pz;
  spi[t] = i;
  {
    SorGroupF& F = frec[t >> 3];
    const float rx = sor_rel_f32(px, dfu_unorder_f64(box[6 * s])), ry = sor_rel_f32(py, dfu_unorder_f64(box[6 * s + 1])),
                rz = sor_rel_f32(pz, dfu_unorder_f64(box[6 * s + 2]));
    F.x[lane & 7] = rx; F.y[lane & 7] = ry; F.z[lane & 7] = rz;
    float ax = rx, ay = ry, az = rz, bx = i >= 0 ? rx : -INFINITY, by = i >= 0 ? ry : -INFINITY, bz = i >= 0 ? rz : -INFINITY;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      ax = fminf(ax, __shfl_xor_sync(0xffffffffu, ax, o)); ay = fminf(ay, __shfl_xor_sync(0xffffffffu, ay, o));
      az = fminf(az, __shfl_xor_sync(0xffffffffu, az, o)); bx = fmaxf(bx, __shfl_xor_sync(0xffffffffu, bx, o));
      by = fmaxf(by, __shfl_xor_sync(0xffffffffu, by, o)); bz = fmaxf(bz, __shfl_xor_sync(0xffffffffu, bz, o));
    }
    if ((lane & 7) == 0) { F.lo[0] = ax; F.lo[1] = ay; F.lo[2] = az; F.hi[0] = bx; F.hi[1] = by; F.hi[2] = bz; F.pad[0] = F.pad[1] = 0.f; }
  }
  float lx = i >= 0 ? __double2float_rd(px) : INFINITY, ly = i >= 0 ? __double2float_rd(py) : INFINITY,
        lz = i >= 0 ? __double2float_rd(pz) : INFINITY;
  float hx = i >= 0 ? __double2float_ru(px) : -INFINITY, hy = i >= 0 ? __double2float_ru(py) : -INFINITY,
        hz = i >= 0 ? __double2float_ru(pz) : -INFINITY;
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) {
    lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, o)); ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, o));
    lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, o)); hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, o));
    hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, o));
  }
  if ((lane & 7) == 0) {
    SorBox b;
    b.lo = make_float4(lx, ly, lz, 0.f);
    b.hi = make_float4(hx, hy, hz, 0.f);
    G.box = b;
    gbox[t >> 3] = b;
  }
}

// next pyramid level: union of 8 children (empty boxes have lo = +inf, hi = -inf and never intersect anything)
__global__ void __launch_bounds__(256) k_sor_pyr(const SorBox* __restrict__ child, int nchild, SorBox* __restrict__ parent, int nparent) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= nparent) return;
  float4 lo = make_float4(INFINITY, INFINITY, INFINITY, 0.f), hi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
  for (int c = 8 * j; c < 8 * j + 8 && c < nchild; c++) {
    const SorBox b = child[c];
    lo.x = fminf(lo.x, b.lo.x); lo.y = fminf(lo.y, b.lo.y); lo.z = fminf(lo.z, b.lo.z);
    hi.x = fmaxf(hi.x, b.hi.x); hi.y = fmaxf(hi.y, b.hi.y); hi.z = fmaxf(hi.z, b.hi.z);
  }
  SorBox o;
  o.lo = lo; o.hi = hi;
  parent[j] = o;
}

__device__ __forceinline__ double sor_d2(double ax, double ay, double az, double px, double py, double pz) {
  // ((a - b) ** 2).sum(-1): separately rounded products, (dx^2 + dy^2) + dz^2
  const double dx = __dsub_rn(ax, px), dy = __dsub_rn(ay, py), dz = __dsub_rn(az, pz);
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

#ifdef SOR_STATS
__device__ unsigned long long g_sor_stats[8];   // tiles, groups tested, groups swept, tightenings, final entries, final lanes, lanes hit, appends
#endif
__device__ __forceinline__ bool sor_box_hits(const SorBox& b, const float4 qlo, const float4 qhi) {
  return b.lo.x <= qhi.x && b.hi.x >= qlo.x && b.lo.y <= qhi.y && b.hi.y >= qlo.y && b.lo.z <= qhi.z && b.hi.z >= qlo.z;
}

// cause: 0 far bound, 1 search box meets too many nodes, 2 more equal distances than the column holds
__device__ __forceinline__ void sor_defer(bool want, int64_t t, int32_t* __restrict__ todo, int32_t* __restrict__ ntodo, int cause, int lane) {
  const unsigned m = __ballot_sync(0xffffffffu, want);
  if (!m) return;
  int base = 0;
  if (lane == 0) {
    base = atomicAdd(ntodo, __popc(m));
    atomicAdd(ntodo + 3 + cause, __popc(m));
  }
  base = __shfl_sync(0xffffffffu, base, 0);
  if (want) todo[base + __popc(m & ((1u << lane) - 1u))] = (int32_t)t;
}

// ---------------------------------------------------------------- the sweep kernel (float32 screen, exact finish)
// One warp per tile of 32 consecutive sorted queries; the candidates are SCREENED in float32:
// every lane sweeps every staged point (the point is a shared-memory broadcast, the arithmetic is 7 float
// instructions, all 32 lanes busy, no atomics, no work list) and appends the points inside its own ball to its own
// column as (float d2, sorted position).  When a column is about to fill, the warp tightens: each lane guesses a
// smaller radius from the density it has seen, verifies by counting that at least k entries lie inside, compacts the
// column and shrinks its ball, so one sweep over the tile's search region suffices.  Only the few dozen survivors
// are evaluated exactly (float64, from the group records) and the k smallest of those are summed in ascending
// order: the float32 stage decides WHICH points are looked at, never a value.
//
// Error model of the screen.  Coordinates are stored as float32 relative to the segment's box minimum, E = largest
// extent of the box, delta = 2^-22 E.  Per axis |float difference - true difference| <= 2^-23 E + 2^-24 |dx| (+ the
// float64 subtraction, 2^-53 relative), hence for a pair at true distance d
//     |d2f - d2| <= 4 delta d + 1e-6 d^2 + 4 delta^2      (twice what the rounding analysis gives)
//   * a point with d2 <= B passes the screen  d2f <= S(B) = B + 4 delta sqrt(B) + 1e-6 B + 4 delta^2;
//   * a point with d2f <= X has d <= dmax(X), the positive root of (1 - 1e-6) d^2 - 4 delta d - (X + 4 delta^2) = 0,
//     so k entries with d2f <= X prove that the k-th neighbour lies within B' = dmax(X)^2.
#ifndef SOR_SCG
#define SOR_SCG 64          /* float groups staged per chunk of the sweep (32: one more CTA per SM, measured no faster) */
#endif
#ifndef SOR_SLCAP20
#define SOR_SLCAP20 64      /* column entries per lane (k <= 20) */
#endif

template <int LCAP>
struct __align__(128) SorSweepSmem {
  union {
    struct {
      SorGroupF fst[SOR_SCG];                               // staged float groups (step 1: the curve window)
      int front[SOR_FRONT];                                 // level-1 frontier (the descent's second buffer is fst)
      int grp[SOR_SCG + 64];                                // group ids that passed the box test (a step adds up to 64)
    };
    unsigned key[LCAP][32];                                 // step 4: order keys of the columns
  };
  uint2 ent[LCAP][32];                                      // (float d2 bits, sorted position); step 4: the exact d2 (double)
  unsigned long long mbar;
};

struct SorErr { double A, D; };                             // 4 delta, 4 delta^2
#define SOR_ERR_C 1e-6

__device__ __forceinline__ float sor_screen(double B, const SorErr& e) {
  const double s = __fma_rn(e.A, __dsqrt_ru(B), B) + SOR_ERR_C * B + e.D;
  return __double2float_ru(s * (1.0 + 1e-12));
}
__device__ __forceinline__ double sor_proven(float X, const SorErr& e) {
  const double c = 1.0 - SOR_ERR_C, x = (double)X + e.D;
  const double d = (e.A + __dsqrt_ru(__fma_rn(e.A, e.A, 4.0 * c * x) * (1.0 + 1e-12))) / (2.0 * c);
  return d * d * (1.0 + 1e-12);
}

struct SorSweepLane {
  double px, py, pz;      // the query
  double B;               // proven squared bound of the k-th neighbour distance
  float fx, fy, fz;       // float32 screening coordinates of the query
  float thr;              // screen threshold S(B); < 0: the lane does not sweep
  bool active;
};

__device__ __forceinline__ void sor_sweep_box(const SorSweepLane& L, float4& qlo, float4& qhi) {
  const unsigned full = 0xffffffffu;
  qlo = make_float4(INFINITY, INFINITY, INFINITY, 0.f);
  qhi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
  if (L.thr >= 0.f) {
    const double r = __dsqrt_ru(L.B) * (1.0 + 1e-12);
    qlo = make_float4(__double2float_rd(L.px - r), __double2float_rd(L.py - r), __double2float_rd(L.pz - r), 0.f);
    qhi = make_float4(__double2float_ru(L.px + r), __double2float_ru(L.py + r), __double2float_ru(L.pz + r), 0.f);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    qlo.x = fminf(qlo.x, __shfl_xor_sync(full, qlo.x, o)); qlo.y = fminf(qlo.y, __shfl_xor_sync(full, qlo.y, o));
    qlo.z = fminf(qlo.z, __shfl_xor_sync(full, qlo.z, o)); qhi.x = fmaxf(qhi.x, __shfl_xor_sync(full, qhi.x, o));
    qhi.y = fmaxf(qhi.y, __shfl_xor_sync(full, qhi.y, o)); qhi.z = fmaxf(qhi.z, __shfl_xor_sync(full, qhi.z, o));
  }
}

// entries of the lane's column with d2f <= X (the column holds cnt entries; cmax = the warp's largest cnt, rounded up)
template <int LCAP>
__device__ __forceinline__ int sor_count_le(const SorSweepSmem<LCAP>& w, int cnt, int cmax, float X, int lane) {
  int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
  for (int j = 0; j < cmax; j += 4) {
    c0 += (j < cnt && __uint_as_float(w.ent[j][lane].x) <= X) ? 1 : 0;
    c1 += (j + 1 < cnt && __uint_as_float(w.ent[j + 1][lane].x) <= X) ? 1 : 0;
    c2 += (j + 2 < cnt && __uint_as_float(w.ent[j + 2][lane].x) <= X) ? 1 : 0;
    c3 += (j + 3 < cnt && __uint_as_float(w.ent[j + 3][lane].x) <= X) ? 1 : 0;
  }
  return (c0 + c1) + (c2 + c3);
}

// Shrink the balls of the lanes that hold at least kk + 8 entries.  Returns true in the lanes that MUST shrink
// (column nearly full) and could not (more than LCAP - 8 entries share one float distance): they are deferred.
template <int LCAP>
__device__ __forceinline__ bool sor_tighten(SorSweepSmem<LCAP>& w, SorSweepLane& L, int& cnt, int kk, const SorErr& er, int lane, int slack = 8) {
  const unsigned full = 0xffffffffu;
  const bool want = L.thr >= 0.f && cnt >= kk + slack, need = cnt > LCAP - 8;
  int cmax = cnt;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cmax = max(cmax, __shfl_xor_sync(full, cmax, o));
  cmax = (cmax + 3) & ~3;
  // count(<= lo) < kk <= count(<= hi); first guess: the entries grow like the area of the ball (points on a surface)
  float lo = -1.f, hi = L.thr, X = want ? L.thr * ((float)(kk + 4) / (float)cnt) : 0.f;
  int chi = cnt;
  bool exhausted = false;
  for (int a = 0; a < 48; a++) {
    const bool open = want && !exhausted && chi > kk + slack + 4 && a < 3;
    const bool forced = want && !exhausted && need && chi > (LCAP - 24 > kk + 4 ? LCAP - 24 : kk + 4);
    if (!__any_sync(full, open || forced)) break;
    const int c = sor_count_le<LCAP>(w, cnt, cmax, X, lane);
    if (open || forced) {
      if (c >= kk) { hi = X; chi = c; } else lo = X;
      // next guess: proportional to the area while nothing below is known, then bisection
      const float nx = lo < 0.f ? hi * ((float)(kk + 4) / (float)chi) : 0.5f * (lo + hi);
      if (nx > lo && nx < hi) X = nx; else exhausted = true;      // adjacent floats: nothing left to try
    }
  }
  bool stuck = false;
  if (want && hi < L.thr) {
    const double Bn = sor_proven(hi, er);
    if (Bn < L.B) {
      L.B = Bn;
      const float tn = sor_screen(Bn, er);
      L.thr = tn < L.thr ? tn : L.thr;
    }
  }
  // compaction (lane-private, in place)
  int wj = 0;
  for (int j = 0; j < cmax; j += 4) {                       // wj <= j: a store never reaches an entry not yet loaded
    const uint2 e0 = w.ent[j][lane], e1 = w.ent[j + 1][lane], e2 = w.ent[j + 2][lane], e3 = w.ent[j + 3][lane];
    if (j < cnt && __uint_as_float(e0.x) <= L.thr) { w.ent[wj][lane] = e0; wj++; }
    if (j + 1 < cnt && __uint_as_float(e1.x) <= L.thr) { w.ent[wj][lane] = e1; wj++; }
    if (j + 2 < cnt && __uint_as_float(e2.x) <= L.thr) { w.ent[wj][lane] = e2; wj++; }
    if (j + 3 < cnt && __uint_as_float(e3.x) <= L.thr) { w.ent[wj][lane] = e3; wj++; }
  }
  cnt = wj;
  stuck = need && cnt > LCAP - 8;
  return stuck;
}

// sweep the ng staged groups: every lane against every point
template <int LCAP>
__device__ __forceinline__ bool sor_sweep_chunk(SorSweepSmem<LCAP>& w, const int* __restrict__ grp, int ng, SorSweepLane& L, int& cnt, int kk, const SorErr& er,
                                                int64_t t, int32_t* __restrict__ todo, int32_t* __restrict__ ntodo, int lane) {
  const unsigned full = 0xffffffffu;
  bool tightened = false;
  for (int g = 0; g < ng; g++) {
    const SorGroupF& G = w.fst[g];
    // squared distance from the query to the group's float box, by the same monotone operations as the point distances
    // below: it is <= the screened distance of every point of the group, so a group no lane can reach is skipped
    const float4 ba = *reinterpret_cast<const float4*>(&G.lo[0]), bb = *reinterpret_cast<const float4*>(&G.hi[1]);   // lo xyz hi x | hi y z
    const float ex = fmaxf(fmaxf(ba.x - L.fx, L.fx - ba.w), 0.f), ey = fmaxf(fmaxf(ba.y - L.fy, L.fy - bb.x), 0.f),
                ez = fmaxf(fmaxf(ba.z - L.fz, L.fz - bb.y), 0.f);
    const float bd = __fmaf_rn(ez, ez, __fmaf_rn(ey, ey, __fmul_rn(ex, ex)));
#ifdef SOR_STATS
    if (lane == 0) atomicAdd(&g_sor_stats[1], 1ull);
#endif
    if (!__any_sync(full, bd <= L.thr)) continue;
    const float4 xa = *reinterpret_cast<const float4*>(&G.x[0]), xb = *reinterpret_cast<const float4*>(&G.x[4]);
    const float4 ya = *reinterpret_cast<const float4*>(&G.y[0]), yb = *reinterpret_cast<const float4*>(&G.y[4]);
    const float4 za = *reinterpret_cast<const float4*>(&G.z[0]), zb = *reinterpret_cast<const float4*>(&G.z[4]);
    const float gx[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
    const float gy[8] = {ya.x, ya.y, ya.z, ya.w, yb.x, yb.y, yb.z, yb.w};
    const float gz[8] = {za.x, za.y, za.z, za.w, zb.x, zb.y, zb.z, zb.w};
    if (__any_sync(full, cnt > LCAP - 8)) {
#ifdef SOR_STATS
      if (lane == 0) atomicAdd(&g_sor_stats[3], 1ull);
#endif
      const bool stuck = sor_tighten<LCAP>(w, L, cnt, kk, er, lane);
      sor_defer(stuck, t, todo, ntodo, 2, lane);
      if (stuck) { L.active = false; L.thr = -1.f; cnt = 0; }
      tightened = true;
    }
    const unsigned pos0 = (unsigned)grp[g] * 8u;
#ifdef SOR_STATS
    const int cnt0 = cnt;
#endif
#pragma unroll
    for (int q = 0; q < 8; q++) {
      const float dx = gx[q] - L.fx, dy = gy[q] - L.fy, dz = gz[q] - L.fz;
      const float d = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
      if (d <= L.thr) {                                     // at most 8 appends between two fill checks
        w.ent[cnt][lane] = make_uint2(__float_as_uint(d), pos0 + q);
        cnt++;
      }
    }
#ifdef SOR_STATS
    {
      const unsigned hm = __ballot_sync(full, cnt > cnt0);
      if (lane == 0) { atomicAdd(&g_sor_stats[2], 1ull); atomicAdd(&g_sor_stats[6], (unsigned long long)__popc(hm)); }
      atomicAdd(&g_sor_stats[7], (unsigned long long)(cnt - cnt0));
    }
#endif
  }
  return tightened;
}

template <int KCAP, int LCAP>
__global__ void __launch_bounds__(SOR_WARPS * 32) k_sor_sweep(SorTree T, double* __restrict__ mean_out, int32_t* __restrict__ todo,
                                                              int32_t* __restrict__ ntodo, int32_t* __restrict__ tile_ctr) {
  extern __shared__ __align__(128) unsigned char sor_smem_raw[];
  SorSweepSmem<LCAP>& w = reinterpret_cast<SorSweepSmem<LCAP>*>(sor_smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const unsigned full = 0xffffffffu;
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  if (lane == 0) dfu_mbar_init(&w.mbar, 1);
  __syncwarp();
  unsigned phase = 0;
  const int W = T.k > 32 ? T.k : 32;
  for (;;) {
    int tile = 0;
    if (lane == 0) tile = atomicAdd(tile_ctr, 1);
    tile = __shfl_sync(full, tile, 0);
    if (tile >= T.ntiles) break;
    const int64_t t0 = (int64_t)tile * 32, t = t0 + lane;
    const int s = sor_seg_of(T.poff, T.nseg, t0);
    const int64_t ps = T.poff[s], pe = T.poff[s + 1];
    const int ns = (int)(T.seg_off[s + 1] - T.seg_off[s]);
    const int kk = ns < T.k ? ns : T.k;
    SorSweepLane L;
    const int orig = T.spi[t];
    L.active = orig >= 0;
    // segment origin and error constants
    SorErr er;
    {
      const double lx = dfu_unorder_f64(T.box[6 * s]), ly = dfu_unorder_f64(T.box[6 * s + 1]), lz = dfu_unorder_f64(T.box[6 * s + 2]);
      const double ex = dfu_unorder_f64(T.box[6 * s + 3]) - lx, ey = dfu_unorder_f64(T.box[6 * s + 4]) - ly,
                   ez = dfu_unorder_f64(T.box[6 * s + 5]) - lz;
      const double E = fmax(fmax(ex, ey), fmax(ez, 1e-30));
      const double delta = E * (1.0 / 4194304.0);
      er.A = 4.0 * delta; er.D = 4.0 * delta * delta;
    }
    // ---- 1. bound from the curve neighbours, float32: positions t-32 .. t+32 of the segment, staged from
    //         [t0-32, t0+64); the k-th smallest screened distance X proves the bound B = sor_proven(X)
    {
      const int64_t a = t0 - 32 > ps ? t0 - 32 : ps, b = t0 + 64 < pe ? t0 + 64 : pe;
      const int off = (int)(a - (t0 - 32));
      if (lane == 0) {
        const uint32_t bytes = (uint32_t)((b - a) >> 3) * (uint32_t)sizeof(SorGroupF);
        dfu_mbar_arrive_expect_tx(&w.mbar, bytes);
        dfu_bulk_g2s(w.fst + (off >> 3), T.frec + (a >> 3), bytes, &w.mbar);
      }
      {
        const SorGroup& G = T.rec[t >> 3];
        L.px = G.x[lane & 7]; L.py = G.y[lane & 7]; L.pz = G.z[lane & 7];
      }
      dfu_mbar_wait(&w.mbar, phase);
      phase ^= 1u;
      {
        const int p = 32 + lane;
        L.fx = w.fst[p >> 3].x[p & 7]; L.fy = w.fst[p >> 3].y[p & 7]; L.fz = w.fst[p >> 3].z[p & 7];
      }
      float best[KCAP];
#pragma unroll
      for (int i = 0; i < KCAP; i++) best[i] = INFINITY;
      const int last = (int)(b - (t0 - 32)) - 1;
      const int jlo = off - (32 + lane) > -W ? off - (32 + lane) : -W;
      const int jhi = last - (32 + lane) < W ? last - (32 + lane) : W;
#pragma unroll 4
      for (int j = jlo; j <= jhi; j++) {
        const int p = 32 + lane + j;
        const SorGroupF& G = w.fst[p >> 3];
        const float dx = G.x[p & 7] - L.fx, dy = G.y[p & 7] - L.fy, dz = G.z[p & 7] - L.fz;
        float v = dx * dx + dy * dy + dz * dz;              // pads: +inf
#pragma unroll
        for (int i = 0; i < KCAP; i++) {
          const float lo = fminf(best[i], v);
          v = fmaxf(best[i], v);
          best[i] = lo;
        }
      }
      float kth = 0.f;
#pragma unroll
      for (int i = 0; i < KCAP; i++) kth = fmaxf(kth, i < kk ? best[i] : 0.f);
      L.B = L.active && kth < INFINITY ? sor_proven(kth, er) : inf;
      __syncwarp();
    }
    // ---- 1b. lanes whose bound is far above the tile's median are handled one by one
    {
      float v = L.active ? (float)L.B : INFINITY;
#pragma unroll
      for (int k2 = 2; k2 <= 32; k2 <<= 1)
#pragma unroll
        for (int j = k2 >> 1; j > 0; j >>= 1) {
          const float o = __shfl_xor_sync(full, v, j);
          const bool up = (lane & k2) == 0 || k2 == 32, lower = (lane & j) == 0;
          v = (lower == up) ? fminf(v, o) : fmaxf(v, o);
        }
      const int nv = __popc(__ballot_sync(full, L.active));
      const float med = __shfl_sync(full, v, nv > 0 ? nv >> 1 : 0);
      const bool far = L.active && (!(L.B < (double)INFINITY) || (float)L.B > SOR_DEFER * med);
      sor_defer(far, t, todo, ntodo, 0, lane);
      if (far) L.active = false;
    }
    if (!__any_sync(full, L.active)) continue;
    L.thr = L.active ? sor_screen(L.B, er) : -1.f;
    // ---- 2. pyramid descent to the level-1 nodes that meet the tile's search box
    const int gfirst = (int)(ps >> 3), glast = (int)(pe >> 3) - 1;
    int nf = 0;
    float4 qlo, qhi;
    sor_sweep_box(L, qlo, qhi);
    {
      int lev = 1;
      while (lev < T.nlev - 1 && ((glast >> (3 * lev)) - (gfirst >> (3 * lev)) + 1) > 32) lev++;
      // ping-pong between the frontier array and the (still idle) staging buffer, arranged to END in the frontier array
      int* const fb0 = w.front; int* const fb1 = reinterpret_cast<int*>(w.fst);
      static_assert(sizeof(w.fst) >= sizeof(w.front), "the staging buffer doubles as the second frontier");
      int fsel = (lev - 1) & 1;
      const int nfirst = gfirst >> (3 * lev), nlast = glast >> (3 * lev);
      bool overflow = false;
      for (int b0 = nfirst; b0 <= nlast; b0 += 32) {
        const int node = b0 + lane;
        const bool hit = node <= nlast && sor_box_hits(T.lv[lev][node], qlo, qhi);
        const unsigned bm = __ballot_sync(full, hit);
        if (nf + __popc(bm) > SOR_FRONT) { overflow = true; break; }
        if (hit) (fsel ? fb1 : fb0)[nf + __popc(bm & ((1u << lane) - 1u))] = node;
        nf += __popc(bm);
      }
      __syncwarp();
      while (!overflow && lev > 1) {
        const int cl = lev - 1, cfirst = gfirst >> (3 * cl), clast = glast >> (3 * cl);
        int nn = 0;
        for (int f0 = 0; f0 < nf; f0 += 4) {
          const int fi = f0 + (lane >> 3);
          bool hit = false;
          int c = -1;
          if (fi < nf) {
            c = (fsel ? fb1 : fb0)[fi] * 8 + (lane & 7);
            if (c >= cfirst && c <= clast) hit = sor_box_hits(T.lv[cl][c], qlo, qhi);
          }
          const unsigned bm = __ballot_sync(full, hit);
          if (nn + __popc(bm) > SOR_FRONT) { overflow = true; break; }
          if (hit) (fsel ? fb0 : fb1)[nn + __popc(bm & ((1u << lane) - 1u))] = c;
          nn += __popc(bm);
        }
        __syncwarp();
        fsel ^= 1; nf = nn; lev = cl;
      }
      if (overflow) {
        sor_defer(L.active, t, todo, ntodo, 1, lane);
        continue;
      }
    }
    // ---- 3. one sweep over the groups that meet the (shrinking) search box.  The level-1 nodes are visited from the
    //         tile's own node outwards along the curve (the near groups hold most neighbours, so the first tightening
    //         already brings the balls close to their final size and the far groups fail the box test), 8 nodes =
    //         64 group boxes per step, the boxes of the next step in flight while this one is handled.
    int cnt = 0;
    {
      const int own = (int)(t0 >> 6);
      int m = 0;                                            // frontier nodes below the tile's own
      for (int f0 = 0; f0 < nf; f0 += 32)
        m += __popc(__ballot_sync(full, f0 + lane < nf && w.front[f0 + lane] < own));
      const int nu = nf - m, nd = m, nalt = 2 * (nu < nd ? nu : nd);
      const int nsteps = (nf + 7) >> 3;
      int ng = 0;
      bool first = true;
      SorBox bxa, bxb;
      int ga = -1, gb = -1;
      auto fetch = [&](int step) {
        ga = -1; gb = -1;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int i = step * 8 + h * 4 + (lane >> 3);     // position in the outward visiting order
          int g = -1;
          if (i < nf) {
            const int fi = i < nalt ? ((i & 1) ? m - 1 - (i >> 1) : m + (i >> 1)) : (nu > nd ? m + (i - nd) : m - 1 - (i - nu));
            g = w.front[fi] * 8 + (lane & 7);
            if (g < gfirst || g > glast) g = -1;
          }
          if (h == 0) { ga = g; if (g >= 0) bxa = T.lv[0][g]; } else { gb = g; if (g >= 0) bxb = T.lv[0][g]; }
        }
      };
      fetch(0);
      for (int step = 0; step < nsteps; step++) {
        const bool hita = ga >= 0 && sor_box_hits(bxa, qlo, qhi), hitb = gb >= 0 && sor_box_hits(bxb, qlo, qhi);
        const int cga = ga, cgb = gb;
        if (step + 1 < nsteps) fetch(step + 1);
        const unsigned ma = __ballot_sync(full, hita), mb = __ballot_sync(full, hitb);
        const unsigned below = (1u << lane) - 1u;
        if (hita) w.grp[ng + __popc(ma & below)] = cga;
        if (hitb) w.grp[ng + __popc(ma) + __popc(mb & below)] = cgb;
        ng += __popc(ma) + __popc(mb);
        __syncwarp();
        if (ng >= SOR_SCG || (step + 1 == nsteps && ng > 0)) {
          bool tg = false;
          for (int c0 = 0; c0 < ng; c0 += SOR_SCG) {
            const int filled = ng - c0 < SOR_SCG ? ng - c0 : SOR_SCG;
            // runs of consecutive groups -> one bulk copy each
            if (lane == 0) dfu_mbar_arrive_expect_tx(&w.mbar, (uint32_t)filled * (uint32_t)sizeof(SorGroupF));
            __syncwarp();
#pragma unroll
            for (int h = 0; h < SOR_SCG; h += 32) {
              const int i = h + lane;
              const int gi = i < filled ? w.grp[c0 + i] : -2;
              const bool st = i < filled && (lane == 0 || w.grp[c0 + i - 1] + 1 != gi);
              const unsigned starts = __ballot_sync(full, st);
              if (st) {
                const unsigned later = lane < 31 ? (starts >> (lane + 1)) : 0u;
                const int lim = filled - h < 32 ? filled - h : 32;
                const int len = later ? __ffs((int)later) : lim - lane;
                dfu_bulk_g2s(w.fst + i, T.frec + gi, (uint32_t)len * (uint32_t)sizeof(SorGroupF), &w.mbar);
              }
            }
            dfu_mbar_wait(&w.mbar, phase);
            phase ^= 1u;
            tg |= sor_sweep_chunk<LCAP>(w, w.grp + c0, filled, L, cnt, kk, er, t, todo, ntodo, lane);
            __syncwarp();
          }
          if (first) {                                      // the near groups are in: shrink every ball that can be shrunk
            first = false;
            const bool stuck = sor_tighten<LCAP>(w, L, cnt, kk, er, lane, 3);
            sor_defer(stuck, t, todo, ntodo, 2, lane);
            if (stuck) { L.active = false; L.thr = -1.f; cnt = 0; }
            tg = true;
          }
          if (tg) sor_sweep_box(L, qlo, qhi);
          ng = 0;
          if (!__any_sync(full, L.active)) break;
        }
      }
    }
    // ---- 4. exact distances of the survivors (in place), the k' smallest in ascending order, sequential sum of roots
    {
      const bool fin = L.active;
      const int nlf = fin ? cnt : 0;
#ifdef SOR_STATS
      if (lane == 0) atomicAdd(&g_sor_stats[0], 1ull);
      atomicAdd(&g_sor_stats[4], (unsigned long long)nlf);
      atomicAdd(&g_sor_stats[5], (unsigned long long)(fin ? 1 : 0));
#endif
      int nmax = nlf;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) nmax = max(nmax, __shfl_xor_sync(full, nmax, o));
      nmax = (nmax + 3) & ~3;
      double (*list)[32] = reinterpret_cast<double (*)[32]>(w.ent);
      for (int j = 0; j < nmax; j += 4) {                   // four entries' loads in flight
        double cx[4], cy[4], cz[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          cx[u] = cy[u] = cz[u] = 0.0;
          if (j + u < nlf) {
            const unsigned pos = w.ent[j + u][lane].y;
            const SorGroup& G = T.rec[pos >> 3];
            cx[u] = G.x[pos & 7]; cy[u] = G.y[pos & 7]; cz[u] = G.z[pos & 7];
          }
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (j + u < nlf) list[j + u][lane] = sor_d2(cx[u], cy[u], cz[u], L.px, L.py, L.pz);
      }
      __syncwarp();
      // order keys: T = trunc(d2 * (2^26 - 2) / top) with top >= every entry, monotone non-decreasing in d2
      const double top = fin ? sor_proven(L.thr, er) : 1.0;
      const double sc = __ddiv_rn(67108862.0, top);
      for (int j = 0; j < nmax; j++) {
        unsigned key = 0xffffffffu;
        if (j < nlf) key = (__double2uint_rz(__dmul_rn(list[j][lane], sc)) << 6) | (unsigned)j;
        w.key[j][lane] = key;
      }
      double acc = 0.0, prev = 0.0;
      unsigned mlast = 0u;
      bool redo = false;
      for (int r = 0; r < kk; r++) {
        unsigned m0 = 0xffffffffu, m1 = 0xffffffffu, m2 = 0xffffffffu, m3 = 0xffffffffu;
        for (int j = 0; j < nmax; j += 4) {
          m0 = min(m0, w.key[j][lane]); m1 = min(m1, w.key[j + 1][lane]);
          m2 = min(m2, w.key[j + 2][lane]); m3 = min(m3, w.key[j + 3][lane]);
        }
        const unsigned m = min(min(m0, m1), min(m2, m3));
        if (fin) {
          const int slot = (int)(m & 63u);
          const double d = list[slot][lane];
          redo |= d < prev;
          prev = d;
          acc = __dadd_rn(acc, __dsqrt_rn(d));
          w.key[slot][lane] = 0xffffffffu;
          mlast = m;
        }
      }
      bool chk = fin && !redo;
      while (__any_sync(full, chk)) {
        unsigned m0 = 0xffffffffu, m1 = 0xffffffffu, m2 = 0xffffffffu, m3 = 0xffffffffu;
        for (int j = 0; j < nmax; j += 4) {
          m0 = min(m0, w.key[j][lane]); m1 = min(m1, w.key[j + 1][lane]);
          m2 = min(m2, w.key[j + 2][lane]); m3 = min(m3, w.key[j + 3][lane]);
        }
        const unsigned m = min(min(m0, m1), min(m2, m3));
        if (chk) {
          if ((m >> 6) != (mlast >> 6)) chk = false;
          else if (list[m & 63u][lane] < prev) { redo = true; chk = false; }
          else w.key[m & 63u][lane] = 0xffffffffu;
        }
      }
      if (__any_sync(full, redo)) {
        const int nr = redo ? nlf : 0;
        double acc2 = 0.0;
        for (int r = 0; r < kk; r++) {
          double b0 = inf;
          int i0 = 0;
          for (int j = 0; j < nr; j++) {
            const double x = list[j][lane];
            if (x < b0) { b0 = x; i0 = j; }
          }
          acc2 = __dadd_rn(acc2, __dsqrt_rn(b0));
          if (redo) list[i0][lane] = inf;
        }
        if (redo) acc = acc2;
      }
      if (fin) mean_out[orig] = __ddiv_rn(acc, (double)kk);
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------- one warp per query (deferred queries, k > 32)
#define SOR_STACK 128
struct SorSingleSmem {
  double sd[SOR_STACK];
  int snode[SOR_STACK];
  signed char slev[SOR_STACK];
};

__device__ __forceinline__ double sor_box_d2(const SorBox& b, double px, double py, double pz) {
  const double dx = fmax(fmax((double)b.lo.x - px, px - (double)b.hi.x), 0.0), dy = fmax(fmax((double)b.lo.y - py, py - (double)b.hi.y), 0.0),
               dz = fmax(fmax((double)b.lo.z - pz, pz - (double)b.hi.z), 0.0);
  return (dx * dx + dy * dy + dz * dz) * (1.0 - 1e-12);      // lower bound of the distance to any point of the box
}

// insert v into the ascending list held one element per lane in b0 (ranks 0..31) and b1 (ranks 32..63)
__device__ __forceinline__ void sor_insert(double v, double& b0, double& b1, bool two, int lane) {
  const unsigned full = 0xffffffffu;
  const int p0 = __popc(__ballot_sync(full, b0 <= v));
  const double up0 = __shfl_up_sync(full, b0, 1);
  const double carry = __shfl_sync(full, b0, 31);
  if (p0 < 32) {
    b0 = lane < p0 ? b0 : (lane == p0 ? v : up0);
    if (two) {
      const double up1 = __shfl_up_sync(full, b1, 1);
      b1 = lane == 0 ? carry : up1;
    }
  } else if (two) {
    const int p1 = __popc(__ballot_sync(full, b1 <= v));
    const double up1 = __shfl_up_sync(full, b1, 1);
    b1 = lane < p1 ? b1 : (lane == p1 ? v : up1);
  }
}

__global__ void __launch_bounds__(128) k_sor_single(SorTree T, const int32_t* __restrict__ todo, const int32_t* __restrict__ ntodo,
                                                    int all, double* __restrict__ mean_out) {
  __shared__ SorSingleSmem sm[4];
  SorSingleSmem& S = sm[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const unsigned full = 0xffffffffu;
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  const int64_t nq = all ? (int64_t)T.ntiles * 32 : (int64_t)*ntodo;
  const int64_t nw = (int64_t)gridDim.x * 4;
  for (int64_t e = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5); e < nq; e += nw) {
    const int64_t t = all ? e : (int64_t)todo[e];
    const int orig = T.spi[t];
    if (orig < 0) continue;
    const int s = sor_seg_of(T.poff, T.nseg, t);
    const int64_t ps = T.poff[s], pe = T.poff[s + 1];
    const int ns = (int)(T.seg_off[s + 1] - T.seg_off[s]);
    const int kk = ns < T.k ? ns : T.k;
    const bool two = kk > 32;
    const double px = T.rec[t >> 3].x[t & 7], py = T.rec[t >> 3].y[t & 7], pz = T.rec[t >> 3].z[t & 7];
    const int gfirst = (int)(ps >> 3), glast = (int)(pe >> 3) - 1;
    double b0 = inf, b1 = inf, kth = inf;
    int sp = 0;
    // start level: at most 32 nodes of the segment (or the top level); push them far-to-near
    int lev = 1;
    while (lev < T.nlev - 1 && ((glast >> (3 * lev)) - (gfirst >> (3 * lev)) + 1) > 32) lev++;
    {
      const int nfirst = gfirst >> (3 * lev), nlast = glast >> (3 * lev);
      for (int b0i = nfirst; b0i <= nlast; b0i += 32) {
        const int node = b0i + lane;
        double d = inf;
        if (node <= nlast) d = sor_box_d2(T.lv[lev][node], px, py, pz);
        // rank by descending distance among the lanes of this batch (ties by lane)
        int rank = 0;
        for (int o = 0; o < 32; o++) {
          const double od = __shfl_sync(full, d, o);
          rank += (od > d || (od == d && o < lane)) ? 1 : 0;
        }
        const int nvalid = __popc(__ballot_sync(full, d < inf));
        // lanes with d = inf rank first (they are the largest); valid lanes get ranks (32 - nvalid) .. 31
        if (d < inf && sp + nvalid <= SOR_STACK) {
          const int slot = sp + rank - (32 - nvalid);
          S.sd[slot] = d; S.snode[slot] = node; S.slev[slot] = (signed char)lev;
        }
        sp += nvalid <= SOR_STACK - sp ? nvalid : 0;
        __syncwarp();
      }
    }
    while (sp > 0) {
      sp--;
      const double nd = S.sd[sp];
      const int node = S.snode[sp], nl = S.slev[sp];
      __syncwarp();
      if (!(nd < kth)) continue;
      if (nl == 1) {
        // leaf: 64 consecutive sorted points (pads are +inf; positions outside the segment are skipped)
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int64_t j = (int64_t)node * 64 + h * 32 + lane;
          double c = inf;
          if (j >= ps && j < pe) {
            const SorGroup& G = T.rec[j >> 3];
            c = sor_d2(G.x[j & 7], G.y[j & 7], G.z[j & 7], px, py, pz);
          }
          unsigned m = __ballot_sync(full, c < kth);
          while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const double v = __shfl_sync(full, c, src);
            if (v < kth) {
              sor_insert(v, b0, b1, two, lane);
              kth = kk <= 32 ? __shfl_sync(full, b0, kk - 1) : __shfl_sync(full, b1, kk - 33);
            }
          }
        }
      } else {
        // inner node: its 8 children, nearest pushed last
        const int cl = nl - 1, cfirst = gfirst >> (3 * cl), clast = glast >> (3 * cl);
        const int c = node * 8 + (lane & 7);
        double d = inf;
        if (lane < 8 && c >= cfirst && c <= clast) {
          d = sor_box_d2(T.lv[cl][c], px, py, pz);
          if (!(d < kth)) d = inf;
        }
        int rank = 0;
#pragma unroll
        for (int o = 0; o < 8; o++) {
          const double od = __shfl_sync(full, d, o);
          rank += (od > d || (od == d && o < lane)) ? 1 : 0;
        }
        const int nvalid = __popc(__ballot_sync(full, d < inf));
        if (d < inf && sp + nvalid <= SOR_STACK) {
          const int slot = sp + rank - (8 - nvalid);
          S.sd[slot] = d; S.snode[slot] = c; S.slev[slot] = (signed char)cl;
        }
        // the stack holds at most 7 siblings per level above the current path: SOR_STACK = 96 > 7 * SOR_MAXLEV + 32
        sp += nvalid;
        __syncwarp();
      }
    }
    // ascending sum of square roots
    const double q0 = __dsqrt_rn(b0), q1 = __dsqrt_rn(b1);
    double acc = 0.0;
    for (int r = 0; r < kk; r++) acc = __dadd_rn(acc, r < 32 ? __shfl_sync(full, q0, r) : __shfl_sync(full, q1, r - 32));
    if (lane == 0) mean_out[orig] = __ddiv_rn(acc, (double)kk);
  }
}

// ---- order-fixed two-stage reductions ------------------------------------------------------------------
__device__ __forceinline__ double sor_block_sum(double v) {
  __shared__ double sh[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < 8; w++) r += sh[w];
  __syncthreads();
  return r;   // valid in thread 0
}

// chunk c of segment s = points [seg_off[s] + (c - seg_ch[s]) * SOR_CHUNK, ...): chunks never straddle segments
// PASS 0: sum of mean_i > 0;  PASS 1: sum of (mean_i - mu)^2 over mean_i > 0
template <int PASS>
__global__ void __launch_bounds__(256) k_sor_chunk_sum(const double* __restrict__ mean, const int32_t* __restrict__ seg_ch, int nseg,
                                                       const int64_t* __restrict__ seg_off, const double* __restrict__ segstat,
                                                       double* __restrict__ part) {
  const int c = blockIdx.x;
  int lo = 0, hi = nseg;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (seg_ch[mid] <= c) lo = mid; else hi = mid;
  }
  // empty segments own no chunk: step over them
  int s = lo;
  while (s + 1 < nseg && seg_ch[s + 1] <= c) s++;
  const int64_t a = seg_off[s] + (int64_t)(c - seg_ch[s]) * SOR_CHUNK;
  int64_t b = a + SOR_CHUNK;
  if (b > seg_off[s + 1]) b = seg_off[s + 1];
  const double mu = PASS ? segstat[4 * s] : 0.0;
  double acc = 0.0;
  for (int64_t i = a + threadIdx.x; i < b; i += 256) {
    const double m = mean[i];
    if (m > 0.0) acc += PASS ? (m - mu) * (m - mu) : m;
  }
  const double r = sor_block_sum(acc);
  if (threadIdx.x == 0) part[c] = r;
}

// one block per segment: segstat[s] = mu, sigma, threshold, n
template <int PASS>
__global__ void __launch_bounds__(256) k_sor_seg_fin(const double* __restrict__ part, const int32_t* __restrict__ seg_ch,
                                                     const int64_t* __restrict__ seg_off, double ratio, double* __restrict__ segstat) {
  const int s = blockIdx.x;
  double acc = 0.0;
  for (int c = seg_ch[s] + threadIdx.x; c < seg_ch[s + 1]; c += 256) acc += part[c];
  const double r = sor_block_sum(acc);
  if (threadIdx.x == 0) {
    const double n = (double)(seg_off[s + 1] - seg_off[s]);
    if (PASS == 0) {
      segstat[4 * s] = n > 0 ? r / n : 0.0;
      segstat[4 * s + 3] = n;
    } else {
      const double sd = n > 1 ? sqrt(r / (n - 1.0)) : 0.0;
      segstat[4 * s + 1] = sd;
      segstat[4 * s + 2] = segstat[4 * s] + ratio * sd;
    }
  }
}

__global__ void __launch_bounds__(256) k_sor_mask(const double* __restrict__ mean, const int64_t* __restrict__ seg_off, int nseg,
                                                  int64_t n, const double* __restrict__ segstat, uint8_t* __restrict__ keep) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const double m = mean[i];
  keep[i] = (m > 0.0 && m < segstat[4 * sor_seg_of(seg_off, nseg, i) + 2]) ? 1 : 0;
}

extern "C" int dfu_sor_filter(dfu_ctx* ctx, void* stream, const double* xyz, int nseg, const int64_t* seg_off_host, int k,
                              double std_ratio, uint8_t* out_keep, double* out_mean, double* out_stats_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || nseg <= 0 || !seg_off_host || !out_keep) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sor_filter: bad argument");
  if (k < 1 || k > SOR_KMAX) return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "dfu_sor_filter: k = %d (supported: 1..%d)", k, SOR_KMAX);
  if (nseg > 65535) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_sor_filter: %d segments (max 65535)", nseg);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = seg_off_host[nseg];
  for (int s = 0; s < nseg; s++)
    if (seg_off_host[s + 1] < seg_off_host[s]) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sor_filter: seg_off must ascend");
  if (out_stats_host) memset(out_stats_host, 0, (size_t)nseg * 16);
  if (n == 0) return DFU_OK;
  if (!xyz) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sor_filter: xyz is NULL");
  // per-segment tables (host, O(nseg)): padded offsets (multiples of 32) and first chunk of the reductions.  They are
  // uploaded from pageable memory: cudaMemcpyAsync returns once the bytes are staged, so no stream sync is needed.
  int64_t* tab = (int64_t*)malloc((size_t)(nseg + 1) * 12);
  if (!tab) return dfu_set_err(ctx, DFU_ERR_CUDA, "dfu_sor_filter: out of host memory");
  int64_t* poff_h = tab;
  int32_t* segch_h = (int32_t*)(tab + nseg + 1);
  int64_t np = 0, nch64 = 0;
  for (int s = 0; s < nseg; s++) {
    const int64_t c = seg_off_host[s + 1] - seg_off_host[s];
    poff_h[s] = np; segch_h[s] = (int32_t)nch64;
    np += (c + 31) / 32 * 32;
    nch64 += (c + SOR_CHUNK - 1) / SOR_CHUNK;
  }
  poff_h[nseg] = np; segch_h[nseg] = (int32_t)nch64;
  if (np >= 0x7fffffffLL) { free(tab); return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_sor_filter: too many points"); }
  const int nch = (int)nch64, ntiles = (int)(np / 32);
  // pyramid geometry
  SorTree T;
  memset(&T, 0, sizeof(T));
  int64_t nodes = 0;
  {
    int c = (int)(np / 8), L = 0;
    while (true) {
      T.nlv[L] = c; nodes += c; L++;
      if (c <= 8 || L == SOR_MAXLEV) break;
      c = (c + 7) / 8;
    }
    T.nlev = L;
    if (T.nlev < 2) { T.nlv[1] = 1; nodes += 1; T.nlev = 2; }
  }
#define SOR_SLOT(var, type, id, bytes)                      \
  type var = (type)dfu_slot(ctx, id, bytes);                \
  if (!var) { free(tab); return DFU_ERR_CUDA; }
  SOR_SLOT(seg_off, int64_t*, SO_SEGOFF, (size_t)(nseg + 1) * 8);
  SOR_SLOT(poff, int64_t*, SO_POFF, (size_t)(nseg + 1) * 8);
  SOR_SLOT(segch, int32_t*, SO_SEGCH, (size_t)(nseg + 1) * 4);
  SOR_SLOT(box, unsigned long long*, SO_BOX, (size_t)nseg * 48);
  SOR_SLOT(key, unsigned long long*, SO_KEY, n * 8);
  SOR_SLOT(idx, int32_t*, SO_IDX, n * 4);
  SOR_SLOT(skey, unsigned long long*, SO_SKEY, n * 8);
  SOR_SLOT(sidx, int32_t*, SO_SIDX, n * 4);
  SOR_SLOT(rec, SorGroup*, SO_REC, (size_t)(np / 8) * sizeof(SorGroup));
  SOR_SLOT(frec, SorGroupF*, SO_FREC, (size_t)(np / 8) * sizeof(SorGroupF));
  SOR_SLOT(spi, int32_t*, SO_SPI, np * 4);
  SOR_SLOT(pyr, SorBox*, SO_PYR, (size_t)nodes * sizeof(SorBox));
  SOR_SLOT(todo, int32_t*, SO_TODO, np * 4);
  SOR_SLOT(scal, int32_t*, SO_SCAL, 64);          // [0] error flag, [1] deferred count, [2] tile counter
  SOR_SLOT(part, double*, SO_PART, (size_t)(nch > 0 ? nch : 1) * 8);
  SOR_SLOT(segstat, double*, SO_SEGSTAT, (size_t)nseg * 32);
  double* mean = out_mean;
  if (!mean) {
    mean = (double*)dfu_slot(ctx, SO_MEAN, n * 8);
    if (!mean) { free(tab); return DFU_ERR_CUDA; }
  }
  cudaError_t ce = cudaMemcpyAsync(seg_off, seg_off_host, (size_t)(nseg + 1) * 8, cudaMemcpyHostToDevice, st);
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(poff, poff_h, (size_t)(nseg + 1) * 8, cudaMemcpyHostToDevice, st);
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(segch, segch_h, (size_t)(nseg + 1) * 4, cudaMemcpyHostToDevice, st);
  free(tab);
  if (ce != cudaSuccess) return dfu_set_err(ctx, DFU_ERR_CUDA, "dfu_sor_filter: upload: %s", cudaGetErrorString(ce));
  {
    SorBox* p = pyr;
    for (int L = 0; L < T.nlev; L++) { T.lv[L] = p; p += T.nlv[L]; }
  }
  T.rec = rec; T.frec = frec; T.box = box; T.spi = spi; T.seg_off = seg_off; T.poff = poff; T.nseg = nseg; T.ntiles = ntiles; T.k = k;
  const unsigned nb = (unsigned)((n + 255) / 256);
  {
    DFU_PROF(ctx, st, PT_GRID_BUILD);
    // box minima start at all-ones, maxima at zero; scalars at zero
    DFU_CUDA(ctx, cudaMemsetAsync(box, 0, (size_t)nseg * 48, st));
    DFU_CUDA(ctx, cudaMemset2DAsync(box, 48, 0xFF, 24, (size_t)nseg, st));
    DFU_CUDA(ctx, cudaMemsetAsync(scal, 0, 64, st));
    k_sor_bbox<<<nb, 256, 0, st>>>(xyz, n, seg_off, nseg, box, scal);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_keys<<<nb, 256, 0, st>>>(xyz, n, seg_off, nseg, box, key, idx);
    DFU_LAUNCH_CHECK(ctx);
    int segbits = 0;
    while ((1 << segbits) < nseg) segbits++;
    size_t tb = 0;
    DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, skey, idx, sidx, (int)n, 0, 48 + segbits, st));
    DFU_SLOT(ctx, tmp, void*, SO_TMP, tb);
    DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, skey, idx, sidx, (int)n, 0, 48 + segbits, st));
    ctx->launches += 7;
    k_sor_gather<<<(unsigned)(((int64_t)ntiles * 32 + 255) / 256), 256, 0, st>>>(sidx, xyz, seg_off, poff, nseg, ntiles, rec, spi,
                                                                                  (SorBox*)T.lv[0], box, frec);
    DFU_LAUNCH_CHECK(ctx);
    for (int L = 1; L < T.nlev; L++) {
      k_sor_pyr<<<(T.nlv[L] + 255) / 256, 256, 0, st>>>(T.lv[L - 1], T.nlv[L - 1], (SorBox*)T.lv[L], T.nlv[L]);
      DFU_LAUNCH_CHECK(ctx);
    }
  }
  {
    DFU_PROF(ctx, st, PT_SOR_KNN);
    const int grid_single = 148 * 8;
    if (k <= SOR_KFAST) {
      int grid = 148 * 16;
      if (grid > (ntiles + SOR_WARPS - 1) / SOR_WARPS) grid = (ntiles + SOR_WARPS - 1) / SOR_WARPS;
      grid = 148 * 8;
      if (grid > (ntiles + SOR_WARPS - 1) / SOR_WARPS) grid = (ntiles + SOR_WARPS - 1) / SOR_WARPS;
      if (k <= 20) {
        const size_t smem = sizeof(SorSweepSmem<SOR_SLCAP20>) * SOR_WARPS;
        DFU_CUDA(ctx, cudaFuncSetAttribute(k_sor_sweep<20, SOR_SLCAP20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DFU_CUDA(ctx, cudaFuncSetAttribute(k_sor_sweep<20, SOR_SLCAP20>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        k_sor_sweep<20, SOR_SLCAP20><<<grid, SOR_WARPS * 32, smem, st>>>(T, mean, todo, scal + 1, scal + 2);
      } else {
        const size_t smem = sizeof(SorSweepSmem<64>) * SOR_WARPS;
        DFU_CUDA(ctx, cudaFuncSetAttribute(k_sor_sweep<32, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DFU_CUDA(ctx, cudaFuncSetAttribute(k_sor_sweep<32, 64>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        k_sor_sweep<32, 64><<<grid, SOR_WARPS * 32, smem, st>>>(T, mean, todo, scal + 1, scal + 2);
      }
      DFU_LAUNCH_CHECK(ctx);
      k_sor_single<<<grid_single, 128, 0, st>>>(T, todo, scal + 1, 0, mean);
      DFU_LAUNCH_CHECK(ctx);
    } else {
      k_sor_single<<<grid_single, 128, 0, st>>>(T, todo, scal + 1, 1, mean);
      DFU_LAUNCH_CHECK(ctx);
    }
  }
  {
    DFU_PROF(ctx, st, PT_SOR_MASK);
    k_sor_chunk_sum<0><<<nch, 256, 0, st>>>(mean, segch, nseg, seg_off, segstat, part);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_seg_fin<0><<<nseg, 256, 0, st>>>(part, segch, seg_off, std_ratio, segstat);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_chunk_sum<1><<<nch, 256, 0, st>>>(mean, segch, nseg, seg_off, segstat, part);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_seg_fin<1><<<nseg, 256, 0, st>>>(part, segch, seg_off, std_ratio, segstat);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_mask<<<nb, 256, 0, st>>>(mean, seg_off, nseg, n, segstat, out_keep);
    DFU_LAUNCH_CHECK(ctx);
  }
  if (out_stats_host) {
    double* hs = (double*)malloc((size_t)nseg * 32 + 64);
    if (!hs) return dfu_set_err(ctx, DFU_ERR_CUDA, "dfu_sor_filter: out of host memory");
    int32_t e[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int32_t* e4 = e + 4;
    cudaError_t c2 = cudaMemcpyAsync(hs, segstat, (size_t)nseg * 32, cudaMemcpyDeviceToHost, st);
    if (c2 == cudaSuccess) c2 = cudaMemcpyAsync(e, scal, 32, cudaMemcpyDeviceToHost, st);
    if (c2 == cudaSuccess) c2 = cudaStreamSynchronize(st);
    if (c2 == cudaSuccess)
      for (int s = 0; s < nseg; s++) { out_stats_host[2 * s] = hs[4 * s]; out_stats_host[2 * s + 1] = hs[4 * s + 1]; }
    free(hs);
    if (c2 != cudaSuccess) return dfu_set_err(ctx, DFU_ERR_CUDA, "dfu_sor_filter: read-back: %s", cudaGetErrorString(c2));
    if (e[0]) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_sor_filter: a coordinate is NaN or infinite");
    ctx->sor_deferred = e[1];
    for (int c = 0; c < 3; c++) ctx->sor_cause[c] = e4[c];
    ctx->sor_points = n;
  }
  return DFU_OK;
}

extern "C" int64_t dfu_sor_counter(const dfu_ctx* ctx, int which) {
  if (!ctx) return -1;
#ifdef SOR_STATS
  if (which >= 10 && which < 18) {
    unsigned long long h[8];
    cudaMemcpyFromSymbol(h, g_sor_stats, sizeof(h));
    return (int64_t)h[which - 10];
  }
#endif
  if (which >= 2 && which <= 4) return ctx->sor_cause[which - 2];
  return which == 0 ? ctx->sor_deferred : ctx->sor_points;
}
