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
// Exact k-NN on a MULTI-RESOLUTION grid obtained from ONE radix sort: points are sorted by a 48-bit Morton code of
// their 16-bit cell coordinates (finest cell = bounding-box extent / 65536), so every cell of every coarser level
// (edge 2^M finest cells) is one contiguous range of the sorted array.  Per query (one thread per SORTED point, so
// that neighbouring threads walk the same ranges):
//   1. the k nearest among the 2k Morton neighbours give an upper bound r_ub of the k-th neighbour distance;
//   2. the smallest level M whose 2x2x2 block of cells around the query's nearest cell corner contains the ball of
//      radius r_ub is scanned exhaustively (8 ranges, two binary searches each).  Dense regions (4 mm point spacing
//      next to the sensor) and sparse ones (salt outliers metres from anything) therefore cost the same few hundred
//      candidates, which a single-resolution grid cannot offer.
// The reductions for mu and sigma are two-stage and order-fixed (chunk partials, then one block per segment), so the
// result does not depend on scheduling.
#include <cub/device/device_radix_sort.cuh>
#include <vector>
#include "dfu_ops.cuh"

#define SOR_THREADS 128
#define SOR_KMAX 64
#define SOR_CHUNK 4096
enum { SO_X = 170, SO_Y, SO_Z, SO_SEG, SO_SEGOFF, SO_BOX, SO_KEY, SO_IDX, SO_SKEY, SO_SIDX, SO_SX, SO_SY, SO_SZ, SO_TMP,
       SO_MEAN, SO_CHSEG, SO_CHSTART, SO_PART, SO_SEGSTAT, SO_SEGCH };

__device__ __forceinline__ unsigned long long sor_spread3(unsigned v) {
  unsigned long long x = v & 0xffffu;
  x = (x | (x << 32)) & 0x001f00000000ffffull;
  x = (x | (x << 16)) & 0x001f0000ff0000ffull;
  x = (x | (x << 8)) & 0x100f00f00f00f00full;
  x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
  x = (x | (x << 2)) & 0x1249249249249249ull;
  return x;
}
__device__ __forceinline__ unsigned long long sor_morton(unsigned cx, unsigned cy, unsigned cz) {
  return (sor_spread3(cx) << 2) | (sor_spread3(cy) << 1) | sor_spread3(cz);
}

// rows [n][3] -> SoA + segment id + per-segment bounding box (ordered-bit atomics; box[s] = min xyz, max xyz)
__global__ void __launch_bounds__(256) k_sor_prepare(const double* __restrict__ rows, int64_t n, const int64_t* __restrict__ seg_off,
                                                     int nseg, double* __restrict__ x, double* __restrict__ y, double* __restrict__ z,
                                                     int32_t* __restrict__ seg, unsigned long long* __restrict__ box) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  int s = -1;
  unsigned long long b[3] = {0, 0, 0};
  if (i < n) {
    int lo = 0, hi = nseg;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (seg_off[mid] <= i) lo = mid; else hi = mid;
    }
    s = lo;
    const double px = rows[3 * i], py = rows[3 * i + 1], pz = rows[3 * i + 2];
    x[i] = px; y[i] = py; z[i] = pz; seg[i] = s;
    b[0] = dfu_order_f64(px); b[1] = dfu_order_f64(py); b[2] = dfu_order_f64(pz);
  }
  const int s0 = __shfl_sync(0xffffffffu, s, 0);
  const bool uniform = __all_sync(0xffffffffu, s == s0);
  if (uniform && s0 >= 0) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      unsigned long long mn = b[a], mx = b[a];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long on = __shfl_xor_sync(0xffffffffu, mn, o), ox = __shfl_xor_sync(0xffffffffu, mx, o);
        mn = on < mn ? on : mn; mx = ox > mx ? ox : mx;
      }
      if ((threadIdx.x & 31) == 0) {
        atomicMin(box + 6 * s0 + a, mn);
        atomicMax(box + 6 * s0 + 3 + a, mx);
      }
    }
  } else if (s >= 0) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
      atomicMin(box + 6 * s + a, b[a]);
      atomicMax(box + 6 * s + 3 + a, b[a]);
    }
  }
}

struct SorSeg { double mn[3]; double c0; };   // finest cell edge and origin of a segment

__device__ __forceinline__ SorSeg sor_seg(const unsigned long long* __restrict__ box, int s) {
  SorSeg g;
  double ext = 0.0;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    g.mn[a] = dfu_unorder_f64(box[6 * s + a]);
    const double e = dfu_unorder_f64(box[6 * s + 3 + a]) - g.mn[a];
    ext = e > ext ? e : ext;
  }
  g.c0 = ext > 0.0 ? ext * (1.0 / 65536.0) * (1.0 + 1e-6) : 1.0;
  return g;
}
__device__ __forceinline__ unsigned sor_coord(double v, double mn, double c0) {
  const double c = floor((v - mn) / c0);
  return c < 0.0 ? 0u : (c > 65535.0 ? 65535u : (unsigned)c);
}

__global__ void __launch_bounds__(256) k_sor_keys(const double* __restrict__ x, const double* __restrict__ y,
                                                  const double* __restrict__ z, const int32_t* __restrict__ seg, int64_t n,
                                                  const unsigned long long* __restrict__ box, unsigned long long* __restrict__ key,
                                                  int32_t* __restrict__ idx, int32_t* __restrict__ errflag) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int s = seg[i];
  const SorSeg g = sor_seg(box, s);
  const double px = x[i], py = y[i], pz = z[i];
  if (!(px == px && py == py && pz == pz) || isinf(px) || isinf(py) || isinf(pz)) atomicOr(errflag, 1);
  key[i] = ((unsigned long long)(unsigned)s << 48) |
           sor_morton(sor_coord(px, g.mn[0], g.c0), sor_coord(py, g.mn[1], g.c0), sor_coord(pz, g.mn[2], g.c0));
  idx[i] = (int32_t)i;
}

__global__ void __launch_bounds__(256) k_sor_gather(const int32_t* __restrict__ sidx, const double* __restrict__ x,
                                                    const double* __restrict__ y, const double* __restrict__ z, int64_t n,
                                                    double* __restrict__ sx, double* __restrict__ sy, double* __restrict__ sz) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= n) return;
  const int i = sidx[t];
  sx[t] = x[i]; sy[t] = y[i]; sz[t] = z[i];
}

__device__ __forceinline__ int sor_lower_bound(const unsigned long long* __restrict__ a, int lo, int hi, unsigned long long v) {
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// bounded list of the k smallest squared distances seen so far
struct SorTopK {
  double v[SOR_KMAX];
  int k, cnt, imax;
  double lim;   // candidates with d2 >= lim cannot enter once the list is full
  __device__ __forceinline__ void init(int kk) { k = kk; cnt = 0; imax = 0; lim = 1.7976931348623157e308; }
  __device__ __forceinline__ void push(double d2) {
    if (cnt < k) {
      v[cnt] = d2;
      if (cnt == 0 || d2 > v[imax]) imax = cnt;
      cnt++;
      if (cnt == k) lim = v[imax];
    } else if (d2 < lim) {
      v[imax] = d2;
      int im = 0;
      double m = v[0];
      for (int j = 1; j < k; j++)
        if (v[j] > m) { m = v[j]; im = j; }
      imax = im; lim = m;
    }
  }
};

__device__ __forceinline__ double sor_d2(double ax, double ay, double az, double px, double py, double pz) {
  // ((a - b) ** 2).sum(-1): separately rounded products, (dx^2 + dy^2) + dz^2
  const double dx = __dsub_rn(ax, px), dy = __dsub_rn(ay, py), dz = __dsub_rn(az, pz);
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// squared distance from q to the level-M cell with level-M coordinates e[], enlarged by `slack` (conservative)
__device__ __forceinline__ double sor_cell_d2(const SorSeg& g, const double* q, int ex, int ey, int ez, int M, double slack) {
  const int e[3] = {ex, ey, ez};
  double acc = 0.0;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const double blo = g.mn[a] + ldexp((double)e[a], M) * g.c0 - slack, bhi = g.mn[a] + ldexp((double)(e[a] + 1), M) * g.c0 + slack;
    const double d = q[a] < blo ? blo - q[a] : (q[a] > bhi ? q[a] - bhi : 0.0);
    acc += d * d;
  }
  return acc * (1.0 - 1e-12);
}
__device__ __forceinline__ unsigned sor_compact3(unsigned long long x) {
  x &= 0x1249249249249249ull;
  x = (x | (x >> 2)) & 0x10c30c30c30c30c3ull;
  x = (x | (x >> 4)) & 0x100f00f00f00f00full;
  x = (x | (x >> 8)) & 0x001f0000ff0000ffull;
  x = (x | (x >> 16)) & 0x001f00000000ffffull;
  x = (x | (x >> 32)) & 0xffffull;
  return (unsigned)x;
}

#define SOR_LEAF 96   /* measured on the config-3 cloud: 4: 50.6 ms, 16: 35.7, 32: 30.0, 64: 27.0, 128: 26.7, 256: 28.1, 2048: 48.6 */
#define SOR_STACK 136

__global__ void __launch_bounds__(SOR_THREADS) k_sor_knn(const unsigned long long* __restrict__ skey, const int32_t* __restrict__ sidx,
                                                         const double* __restrict__ sx, const double* __restrict__ sy,
                                                         const double* __restrict__ sz, int64_t n,
                                                         const int64_t* __restrict__ seg_off,
                                                         const unsigned long long* __restrict__ box, int k,
                                                         double* __restrict__ mean_out) {
  const int64_t t64 = (int64_t)blockIdx.x * SOR_THREADS + threadIdx.x;
  if (t64 >= n) return;
  const int t = (int)t64;
  const unsigned long long key = skey[t];
  const int s = (int)(key >> 48);
  const int lo = (int)seg_off[s], hi = (int)seg_off[s + 1];     // the segment's range is the same in sorted order
  const int kk = (hi - lo) < k ? (hi - lo) : k;
  const double px = sx[t], py = sy[t], pz = sz[t];
  SorTopK top;
  top.init(kk);
  // 1. Morton window
  const int w0 = t - k < lo ? lo : t - k, w1 = t + k + 1 > hi ? hi : t + k + 1;
  for (int j = w0; j < w1; j++) top.push(sor_d2(__ldg(sx + j), __ldg(sy + j), __ldg(sz + j), px, py, pz));
  if (w1 - w0 < hi - lo) {
    const SorSeg g = sor_seg(box, s);
    const double rub = __dsqrt_ru(top.lim);
    const double slack = g.c0 * 1e-3;
    const unsigned cq[3] = {sor_coord(px, g.mn[0], g.c0), sor_coord(py, g.mn[1], g.c0), sor_coord(pz, g.mn[2], g.c0)};
    const double q[3] = {px, py, pz};
    const unsigned long long segbits = (unsigned long long)(unsigned)s << 48;
    // node stack: sorted range [r0, r1) that is (part of) one level-M cell
    int st_r0[SOR_STACK], st_r1[SOR_STACK];
    signed char st_M[SOR_STACK];
    int sp = 0;
    // 2. smallest level whose 2x2x2 block around the nearest cell corner contains the ball (q, rub), domain faces
    //    excepted (nothing lies beyond the bounding box)
    int M = 1;
    while (M < 16 && ldexp(g.c0, M) < 2.0 * rub) M++;
    int b0[3] = {0, 0, 0};
    for (; M < 16; M++) {
      bool ok = true;
      const int top_cell = (int)(65535u >> M);
#pragma unroll
      for (int a = 0; a < 3; a++) {
        const int c = (int)(cq[a] >> M), half = (int)((cq[a] >> (M - 1)) & 1u);
        b0[a] = c - 1 + half;
        const double blo = g.mn[a] + ldexp((double)b0[a], M) * g.c0, bhi = g.mn[a] + ldexp((double)(b0[a] + 2), M) * g.c0;
        if (b0[a] > 0 && !(q[a] - rub >= blo + slack)) ok = false;
        if (b0[a] + 1 < top_cell && !(q[a] + rub <= bhi - slack)) ok = false;
      }
      if (ok) break;
    }
    if (M >= 16) {
      st_r0[0] = lo; st_r1[0] = hi; st_M[0] = 16; sp = 1;
    } else {
      const int top_cell = (int)(65535u >> M);
      for (int o = 0; o < 8; o++) {
        const int ex = b0[0] + (o >> 2), ey = b0[1] + ((o >> 1) & 1), ez = b0[2] + (o & 1);
        if (ex < 0 || ey < 0 || ez < 0 || ex > top_cell || ey > top_cell || ez > top_cell) continue;
        if (sor_cell_d2(g, q, ex, ey, ez, M, slack) >= top.lim) continue;
        const unsigned long long klo = segbits | (sor_morton((unsigned)ex, (unsigned)ey, (unsigned)ez) << (3 * M));
        const unsigned long long khi = klo + (1ull << (3 * M));
        const int r0 = sor_lower_bound(skey, lo, hi, klo);
        if (r0 >= hi || __ldg(skey + r0) >= khi) continue;
        const int r1 = sor_lower_bound(skey, r0, hi, khi);
        st_r0[sp] = r0; st_r1[sp] = r1; st_M[sp] = (signed char)M; sp++;
      }
    }
    // 3. depth-first descent with pruning against the current k-th distance; children are pushed far-to-near
    while (sp > 0) {
      sp--;
      const int r0 = st_r0[sp], r1 = st_r1[sp], Mn = st_M[sp];
      if (Mn < 16) {
        const unsigned long long pre = (__ldg(skey + r0) & 0xffffffffffffull) >> (3 * Mn);
        if (sor_cell_d2(g, q, (int)sor_compact3(pre >> 2), (int)sor_compact3(pre >> 1), (int)sor_compact3(pre), Mn, slack) >= top.lim)
          continue;
      }
      if (r1 - r0 <= SOR_LEAF || Mn == 0) {
        for (int j = r0; j < r1; j++) {
          if (j >= w0 && j < w1) continue;
          top.push(sor_d2(__ldg(sx + j), __ldg(sy + j), __ldg(sz + j), px, py, pz));
        }
        continue;
      }
      // split into the (up to 8) children at level Mn - 1
      const int Mc = Mn - 1;
      const unsigned long long pre = (Mn < 16) ? ((__ldg(skey + r0) & 0xffffffffffffull) >> (3 * Mn)) : 0ull;
      int cr[9];
      cr[0] = r0; cr[8] = r1;
      for (int c = 1; c < 8; c++)
        cr[c] = sor_lower_bound(skey, cr[c - 1], r1, segbits | (((pre << 3) | (unsigned long long)c) << (3 * Mc)));
      double cd[8];
      int ord[8], nc = 0;
      for (int c = 0; c < 8; c++) {
        if (cr[c + 1] <= cr[c]) continue;
        const unsigned long long cp = (pre << 3) | (unsigned long long)c;
        const double d = sor_cell_d2(g, q, (int)sor_compact3(cp >> 2), (int)sor_compact3(cp >> 1), (int)sor_compact3(cp), Mc, slack);
        if (d >= top.lim) continue;
        int b = nc++;
        while (b > 0 && cd[ord[b - 1]] < d) { ord[b] = ord[b - 1]; b--; }   // descending distance
        ord[b] = c; cd[c] = d;
      }
      for (int a = 0; a < nc && sp < SOR_STACK; a++) {
        const int c = ord[a];
        st_r0[sp] = cr[c]; st_r1[sp] = cr[c + 1]; st_M[sp] = (signed char)Mc; sp++;
      }
    }
  }
  // 4. ascending order, sequential sum of square roots (std::accumulate over the sorted k-NN distances)
  for (int a = 1; a < kk; a++) {
    const double v = top.v[a];
    int b = a - 1;
    while (b >= 0 && top.v[b] > v) { top.v[b + 1] = top.v[b]; b--; }
    top.v[b + 1] = v;
  }
  double acc = 0.0;
  for (int a = 0; a < kk; a++) acc = __dadd_rn(acc, __dsqrt_rn(top.v[a]));
  mean_out[sidx[t]] = __ddiv_rn(acc, (double)kk);
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

// PASS 0: sum of mean_i > 0;  PASS 1: sum of (mean_i - mu)^2 over mean_i > 0
template <int PASS>
__global__ void __launch_bounds__(256) k_sor_chunk_sum(const double* __restrict__ mean, const int32_t* __restrict__ ch_seg,
                                                       const int64_t* __restrict__ ch_start, const int64_t* __restrict__ seg_off,
                                                       const double* __restrict__ segstat, double* __restrict__ part) {
  const int c = blockIdx.x;
  const int s = ch_seg[c];
  const int64_t a = ch_start[c];
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

__global__ void __launch_bounds__(256) k_sor_mask(const double* __restrict__ mean, const int32_t* __restrict__ seg, int64_t n,
                                                  const double* __restrict__ segstat, uint8_t* __restrict__ keep) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const double m = mean[i];
  keep[i] = (m > 0.0 && m < segstat[4 * seg[i] + 2]) ? 1 : 0;
}

extern "C" int dfu_sor_filter(dfu_ctx* ctx, void* stream, const double* xyz, int nseg, const int64_t* seg_off_host, int k,
                              double std_ratio, uint8_t* out_keep, double* out_mean, double* out_stats_host) {
  if (!ctx || nseg <= 0 || !seg_off_host || !out_keep) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sor_filter: bad argument");
  if (k < 1 || k > SOR_KMAX) return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "dfu_sor_filter: k = %d (supported: 1..%d)", k, SOR_KMAX);
  if (nseg > 65535) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_sor_filter: %d segments (max 65535)", nseg);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = seg_off_host[nseg];
  if (n >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_sor_filter: too many points");
  for (int s = 0; s < nseg; s++)
    if (seg_off_host[s + 1] < seg_off_host[s]) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sor_filter: seg_off must ascend");
  if (out_stats_host) memset(out_stats_host, 0, (size_t)nseg * 16);
  if (n == 0) return DFU_OK;
  if (!xyz) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sor_filter: xyz is NULL");
  // chunk table (host): chunks never straddle segments
  std::vector<int32_t> ch_seg, seg_ch(nseg + 1);
  std::vector<int64_t> ch_start;
  for (int s = 0; s < nseg; s++) {
    seg_ch[s] = (int32_t)ch_seg.size();
    for (int64_t a = seg_off_host[s]; a < seg_off_host[s + 1]; a += SOR_CHUNK) { ch_seg.push_back(s); ch_start.push_back(a); }
  }
  seg_ch[nseg] = (int32_t)ch_seg.size();
  const int nch = (int)ch_seg.size();
  DFU_SLOT(ctx, x, double*, SO_X, n * 8);
  DFU_SLOT(ctx, y, double*, SO_Y, n * 8);
  DFU_SLOT(ctx, z, double*, SO_Z, n * 8);
  DFU_SLOT(ctx, seg, int32_t*, SO_SEG, n * 4);
  DFU_SLOT(ctx, seg_off, int64_t*, SO_SEGOFF, (size_t)(nseg + 1) * 8);
  DFU_SLOT(ctx, box, unsigned long long*, SO_BOX, (size_t)nseg * 48 + 64);
  DFU_SLOT(ctx, key, unsigned long long*, SO_KEY, n * 8);
  DFU_SLOT(ctx, idx, int32_t*, SO_IDX, n * 4);
  DFU_SLOT(ctx, skey, unsigned long long*, SO_SKEY, n * 8);
  DFU_SLOT(ctx, sidx, int32_t*, SO_SIDX, n * 4);
  DFU_SLOT(ctx, sx, double*, SO_SX, n * 8);
  DFU_SLOT(ctx, sy, double*, SO_SY, n * 8);
  DFU_SLOT(ctx, sz, double*, SO_SZ, n * 8);
  DFU_SLOT(ctx, chs, int32_t*, SO_CHSEG, (size_t)nch * 4);
  DFU_SLOT(ctx, cha, int64_t*, SO_CHSTART, (size_t)nch * 8);
  DFU_SLOT(ctx, part, double*, SO_PART, (size_t)nch * 8);
  DFU_SLOT(ctx, segstat, double*, SO_SEGSTAT, (size_t)nseg * 32);
  DFU_SLOT(ctx, segch, int32_t*, SO_SEGCH, (size_t)(nseg + 1) * 4);
  double* mean = out_mean;
  if (!mean) {
    mean = (double*)dfu_slot(ctx, SO_MEAN, n * 8);
    if (!mean) return DFU_ERR_CUDA;
  }
  int32_t* errflag = (int32_t*)(box + (size_t)nseg * 6);
  DFU_CUDA(ctx, cudaMemcpyAsync(seg_off, seg_off_host, (size_t)(nseg + 1) * 8, cudaMemcpyHostToDevice, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(chs, ch_seg.data(), (size_t)nch * 4, cudaMemcpyHostToDevice, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(cha, ch_start.data(), (size_t)nch * 8, cudaMemcpyHostToDevice, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(segch, seg_ch.data(), (size_t)(nseg + 1) * 4, cudaMemcpyHostToDevice, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));   // the host vectors above go out of scope
  const unsigned nb = (unsigned)((n + 255) / 256);
  {
    DFU_PROF(ctx, st, PT_GRID_BUILD);
    // box minima start at all-ones, maxima (and the error flag behind them) at zero
    DFU_CUDA(ctx, cudaMemsetAsync(box, 0, (size_t)nseg * 48 + 64, st));
    DFU_CUDA(ctx, cudaMemset2DAsync(box, 48, 0xFF, 24, (size_t)nseg, st));
    k_sor_prepare<<<nb, 256, 0, st>>>(xyz, n, seg_off, nseg, x, y, z, seg, box);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_keys<<<nb, 256, 0, st>>>(x, y, z, seg, n, box, key, idx, errflag);
    DFU_LAUNCH_CHECK(ctx);
    size_t tb = 0;
    DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, skey, idx, sidx, (int)n, 0, 64, st));
    DFU_SLOT(ctx, tmp, void*, SO_TMP, tb);
    DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, skey, idx, sidx, (int)n, 0, 64, st));
    ctx->launches += 8;
    k_sor_gather<<<nb, 256, 0, st>>>(sidx, x, y, z, n, sx, sy, sz);
    DFU_LAUNCH_CHECK(ctx);
  }
  {
    DFU_PROF(ctx, st, PT_SOR_KNN);
    k_sor_knn<<<(unsigned)((n + SOR_THREADS - 1) / SOR_THREADS), SOR_THREADS, 0, st>>>(skey, sidx, sx, sy, sz, n, seg_off, box, k, mean);
    DFU_LAUNCH_CHECK(ctx);
  }
  {
    DFU_PROF(ctx, st, PT_SOR_MASK);
    k_sor_chunk_sum<0><<<nch, 256, 0, st>>>(mean, chs, cha, seg_off, segstat, part);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_seg_fin<0><<<nseg, 256, 0, st>>>(part, segch, seg_off, std_ratio, segstat);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_chunk_sum<1><<<nch, 256, 0, st>>>(mean, chs, cha, seg_off, segstat, part);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_seg_fin<1><<<nseg, 256, 0, st>>>(part, segch, seg_off, std_ratio, segstat);
    DFU_LAUNCH_CHECK(ctx);
    k_sor_mask<<<nb, 256, 0, st>>>(mean, seg, n, segstat, out_keep);
    DFU_LAUNCH_CHECK(ctx);
  }
  if (out_stats_host) {
    std::vector<double> hs((size_t)nseg * 4);
    int32_t e = 0;
    DFU_CUDA(ctx, cudaMemcpyAsync(hs.data(), segstat, (size_t)nseg * 32, cudaMemcpyDeviceToHost, st));
    DFU_CUDA(ctx, cudaMemcpyAsync(&e, errflag, 4, cudaMemcpyDeviceToHost, st));
    DFU_CUDA(ctx, cudaStreamSynchronize(st));
    if (e) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_sor_filter: a coordinate is NaN or infinite");
    for (int s = 0; s < nseg; s++) { out_stats_host[2 * s] = hs[4 * s]; out_stats_host[2 * s + 1] = hs[4 * s + 1]; }
  }
  return DFU_OK;
}
