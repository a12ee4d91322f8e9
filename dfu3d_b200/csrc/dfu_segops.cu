// Uniform hash grid (radix-sorted cell keys) + the two neighbourhood operators built on it:
//   * radius filter  — Open3D PointCloud::RemoveRadiusOutliers semantics as restated in oracle/fakes.py
//                      (call sites tools/PENet/dataloaders/my_loader.py:569-599)
//   * NN gate        — BallQuery (tools/PENet/dataloaders/my_loader.py:489-494, 602-604)
// Cell edge = radius * (1 + 2^-30): every neighbour closer than the radius lies in the 3x3x3 block around the query's
// cell.  Keys put z in the lowest bits, so the three z-cells of one (x, y) column are ONE contiguous range of the
// sorted array: 9 range look-ups per query instead of 27.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "dfu_ops.cuh"

#define G_THREADS 256

__global__ void __launch_bounds__(G_THREADS) k_grid_bbox(const double* __restrict__ x, const double* __restrict__ y,
                                                         const double* __restrict__ z, const int32_t* __restrict__ seg,
                                                         const uint8_t* __restrict__ active, int64_t n,
                                                         unsigned long long* __restrict__ segmin) {
  const int64_t i = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  bool ok = i < n && (active == nullptr || active[i]);
  int s = ok ? seg[i] : -1;
  unsigned long long bx = ~0ull, by = ~0ull, bz = ~0ull;
  if (ok) {
    bx = dfu_order_f64(x[i]);
    by = dfu_order_f64(y[i]);
    bz = z ? dfu_order_f64(z[i]) : dfu_order_f64(0.0);
  }
  // warp-aggregate when the whole warp sits in one segment (the common case)
  const int s0 = __shfl_sync(0xffffffffu, s, 0);
  const bool uniform = __all_sync(0xffffffffu, s == s0 || s == -1);
  if (uniform) {
    int sv = s;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      unsigned long long ox = __shfl_xor_sync(0xffffffffu, bx, o), oy = __shfl_xor_sync(0xffffffffu, by, o),
                         oz = __shfl_xor_sync(0xffffffffu, bz, o);
      int os = __shfl_xor_sync(0xffffffffu, sv, o);
      bx = ox < bx ? ox : bx; by = oy < by ? oy : by; bz = oz < bz ? oz : bz;
      sv = os > sv ? os : sv;
    }
    s = sv;
  } else if (ok) {
    atomicMin(segmin + 3 * s + 0, bx);
    atomicMin(segmin + 3 * s + 1, by);
    atomicMin(segmin + 3 * s + 2, bz);
  }
  // one atomic per axis and BLOCK when all its warps sit in one segment (a single segment of millions of points would
  // otherwise send three atomics per warp to the same three words: 258 us per 4 M points, 8 % issue utilisation)
  __shared__ unsigned long long shv[G_THREADS / 32][3];
  __shared__ int shs[G_THREADS / 32];
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    shs[w] = uniform ? s : -2;                               // -1: nothing to add, -2: handled above
    shv[w][0] = bx; shv[w][1] = by; shv[w][2] = bz;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int a = threadIdx.x;
    int sb = -1;
    bool same = true;
    for (int k = 0; k < G_THREADS / 32; k++) {
      if (shs[k] == -1) continue;
      if (shs[k] == -2 || (sb >= 0 && shs[k] != sb)) same = false;
      sb = shs[k] >= 0 ? shs[k] : sb;
    }
    if (same && sb >= 0) {
      unsigned long long v = ~0ull;
      for (int k = 0; k < G_THREADS / 32; k++)
        if (shs[k] >= 0 && shv[k][a] < v) v = shv[k][a];
      atomicMin(segmin + 3 * sb + a, v);
    } else {
      for (int k = 0; k < G_THREADS / 32; k++)
        if (shs[k] >= 0) atomicMin(segmin + 3 * shs[k] + a, shv[k][a]);
    }
  }
}

__global__ void __launch_bounds__(G_THREADS) k_grid_keys(const double* __restrict__ x, const double* __restrict__ y,
                                                         const double* __restrict__ z, const int32_t* __restrict__ seg,
                                                         const uint8_t* __restrict__ active, int64_t n,
                                                         const unsigned long long* __restrict__ segmin,
                                                         const double* __restrict__ cell, unsigned long long* __restrict__ key,
                                                         int32_t* __restrict__ idx, int32_t* __restrict__ errflag) {
  const int64_t i = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (i >= n) return;
  idx[i] = (int32_t)i;
  unsigned long long k = ~0ull;
  if (active == nullptr || active[i]) {
    const int s = seg[i];
    const double c = cell[s];
    int cx = dfu_cell_coord(x[i], dfu_unorder_f64(segmin[3 * s + 0]), c);
    int cy = dfu_cell_coord(y[i], dfu_unorder_f64(segmin[3 * s + 1]), c);
    int cz = z ? dfu_cell_coord(z[i], dfu_unorder_f64(segmin[3 * s + 2]), c) : 1;
    if (cx < 1 || cy < 1 || cz < 1) atomicOr(errflag, 1);   // NaN / overflow of the 16-bit cell range
    else k = dfu_grid_key(s, cx, cy, cz);
  }
  key[i] = k;
}

__global__ void __launch_bounds__(G_THREADS) k_grid_heads(const unsigned long long* __restrict__ skey,
                                                          const int32_t* __restrict__ sidx, const double* __restrict__ x,
                                                          const double* __restrict__ y, const double* __restrict__ z,
                                                          int64_t n, double* __restrict__ sx, double* __restrict__ sy,
                                                          double* __restrict__ sz, int32_t* __restrict__ head) {
  const int64_t t = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (t >= n) return;
  const int i = sidx[t];
  sx[t] = x[i];
  sy[t] = y[i];
  sz[t] = z ? z[i] : 0.0;
  head[t] = (t == 0 || skey[t] != skey[t - 1]) ? 1 : 0;
}

// pcell (inclusive scan of head) -> cell ordinal; scatter the cell table
__global__ void __launch_bounds__(G_THREADS) k_grid_cells(const unsigned long long* __restrict__ skey, int64_t n,
                                                          int32_t* __restrict__ pcell, unsigned long long* __restrict__ ukey,
                                                          int32_t* __restrict__ ustart, int32_t* __restrict__ ncell,
                                                          int32_t* __restrict__ segcell, int nseg,
                                                          unsigned long long* __restrict__ hkey, int32_t* __restrict__ hval,
                                                          unsigned hmask) {
  const int64_t t = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (t >= n) return;
  const int c = pcell[t] - 1;
  pcell[t] = c;
  if (t == 0 || skey[t] != skey[t - 1]) {
    ukey[c] = skey[t];
    ustart[c] = (int32_t)t;
    const unsigned long long k = skey[t];
    if (k != ~0ull) {                        // inactive points (key ~0 = the empty marker) own no cell
      unsigned h = dfu_hash_u64(k) & hmask;
      while (atomicCAS(hkey + h, ~0ull, k) != ~0ull) h = (h + 1) & hmask;      // keys are unique: never finds itself
      hval[h] = c;
    }
    // first cell of every segment that starts here (inactive points carry segment 0xffff and sort last)
    int s1 = (int)(skey[t] >> 48);
    if (s1 > nseg) s1 = nseg;
    const int s0 = t == 0 ? -1 : (int)(skey[t - 1] >> 48);
    for (int s = s0 + 1; s <= s1; s++) segcell[s] = c;
  }
  if (t == n - 1) {
    ustart[c + 1] = (int32_t)n;
    *ncell = c + 1;
    int s1 = (int)(skey[t] >> 48);
    if (s1 > nseg) s1 = nseg;
    for (int s = s1 + 1; s <= nseg; s++) segcell[s] = c + 1;
  }
}

int dfu_grid_build(dfu_ctx* ctx, cudaStream_t st, int sb, const double* x, const double* y, const double* z,
                   const int32_t* seg, const uint8_t* active, int64_t n, int nseg, const double* cell_dev, DfuGrid* g) {
  if (nseg > 65534) return dfu_set_err(ctx, DFU_ERR_DATA, "grid: %d segments (max 65534)", nseg);
  if (n >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_DATA, "grid: too many points");
  const int64_t m = n > 0 ? n : 1;
  DFU_SLOT(ctx, segmin, unsigned long long*, sb + 0, (size_t)(nseg > 0 ? nseg : 1) * 3 * 8);
  DFU_SLOT(ctx, key, unsigned long long*, sb + 1, m * 8);
  DFU_SLOT(ctx, idx, int32_t*, sb + 2, m * 4);
  DFU_SLOT(ctx, skey, unsigned long long*, sb + 3, m * 8);
  DFU_SLOT(ctx, sidx, int32_t*, sb + 4, m * 4);
  DFU_SLOT(ctx, sx, double*, sb + 5, m * 8);
  DFU_SLOT(ctx, sy, double*, sb + 6, m * 8);
  DFU_SLOT(ctx, sz, double*, sb + 7, m * 8);
  DFU_SLOT(ctx, pcell, int32_t*, sb + 8, m * 4);
  DFU_SLOT(ctx, ukey, unsigned long long*, sb + 9, m * 8);
  DFU_SLOT(ctx, ustart, int32_t*, sb + 10, (m + 1) * 4);
  DFU_SLOT(ctx, scal, int32_t*, sb + 11, 64);
  DFU_SLOT(ctx, segcell, int32_t*, sb + 13, (size_t)(nseg + 2) * 4);
  g->segcell = segcell;
  size_t hcap = 64;
  while (hcap < (size_t)m * 2) hcap <<= 1;
  DFU_SLOT(ctx, htab, unsigned long long*, sb + 14, hcap * 12);
  g->hkey = htab; g->hval = (int32_t*)(htab + hcap); g->hmask = (unsigned)(hcap - 1); g->dims = z ? 3 : 2;
  g->n = n; g->nseg = nseg; g->cell = cell_dev; g->segmin = segmin; g->skey = skey; g->sidx = sidx;
  g->sx = sx; g->sy = sy; g->sz = sz; g->pcell = pcell; g->ukey = ukey; g->ustart = ustart;
  g->ncell = scal; g->errflag = scal + 1;
  DFU_CUDA(ctx, cudaMemsetAsync(scal, 0, 4, st));   // ncell only; errflag (scal[1]) is reset by dfu_grid_err_reset
  DFU_CUDA(ctx, cudaMemsetAsync(ustart, 0, 8, st));
  DFU_CUDA(ctx, cudaMemsetAsync(segcell, 0, (size_t)(nseg + 2) * 4, st));
  DFU_CUDA(ctx, cudaMemsetAsync(htab, 0xFF, hcap * 8, st));
  if (n == 0) return DFU_OK;
  DFU_CUDA(ctx, cudaMemsetAsync(segmin, 0xFF, (size_t)nseg * 3 * 8, st));
  const unsigned nb = (unsigned)((n + G_THREADS - 1) / G_THREADS);
  k_grid_bbox<<<nb, G_THREADS, 0, st>>>(x, y, z, seg, active, n, segmin);
  DFU_LAUNCH_CHECK(ctx);
  k_grid_keys<<<nb, G_THREADS, 0, st>>>(x, y, z, seg, active, n, segmin, cell_dev, key, idx, g->errflag);
  DFU_LAUNCH_CHECK(ctx);
  size_t tb = 0, tb2 = 0;
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, skey, idx, sidx, (int)n, 0, 64, st));
  DFU_CUDA(ctx, cub::DeviceScan::InclusiveSum(nullptr, tb2, pcell, pcell, (int)n, st));
  if (tb2 > tb) tb = tb2;
  DFU_SLOT(ctx, tmp, void*, sb + 12, tb);
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, skey, idx, sidx, (int)n, 0, 64, st));
  ctx->launches += 8;
  k_grid_heads<<<nb, G_THREADS, 0, st>>>(skey, sidx, x, y, z, n, sx, sy, sz, pcell);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cub::DeviceScan::InclusiveSum(tmp, tb, pcell, pcell, (int)n, st));
  ctx->launches += 2;
  k_grid_cells<<<nb, G_THREADS, 0, st>>>(skey, n, pcell, ukey, ustart, g->ncell, segcell, nseg, g->hkey, g->hval, g->hmask);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

// ---------------------------------------------------------------- neighbourhood walks
// MODE 0: count points with d2 < r2 until the count exceeds nb (self included)     -> keep = count > nb
// MODE 1: any reference point with sqrt(d2) < maxdist                             -> keep
// The squared distance is ((dx*dx + dy*dy) + dz*dz) with separately rounded products, as numpy evaluates
// ((a - b) ** 2).sum(-1) (oracle/fakes.py: radius_outlier_mask; oracle/restate.py: ball_gate).
// first index in a[lo..hi) with a[i] >= v
__device__ __forceinline__ int grid_lb(const unsigned long long* __restrict__ a, int lo, int hi, unsigned long long v) {
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

template <int MODE>
__device__ __forceinline__ bool grid_walk(const DfuGrid& g, int ncell, int s, double px, double py, double pz, double lim,
                                          int nb) {
  (void)ncell;
  const double c = g.cell[s];
  const int cx = dfu_cell_coord(px, dfu_unorder_f64(g.segmin[3 * s + 0]), c);
  const int cy = dfu_cell_coord(py, dfu_unorder_f64(g.segmin[3 * s + 1]), c);
  const int cz = dfu_cell_coord(pz, dfu_unorder_f64(g.segmin[3 * s + 2]), c);
  if (cx < 0 || cy < 0 || cz < 0) return false;   // far outside the reference bounding box (or an empty segment)
  int count = 0;
  // 27 (9 for a 2-D grid) hashed cell look-ups, the query's own cell first: with a small nb the walk usually ends there
  const int nz = g.dims == 3 ? 3 : 1;
  for (int o = 0; o < 9 * nz; o++) {
    const int oo = o / nz, zz = o % nz;
    const int dx = (oo == 0) ? 0 : ((oo <= 4 ? oo - 1 : oo) / 3 - 1), dy = (oo == 0) ? 0 : ((oo <= 4 ? oo - 1 : oo) % 3 - 1);
    const int dz = zz == 0 ? 0 : (zz == 1 ? -1 : 1);
    const int ex = cx + dx, ey = cy + dy, ez = cz + dz;
    if (ex < 1 || ey < 1 || ez < 1) continue;
    const int ci = dfu_grid_find(g, dfu_grid_key(s, ex, ey, ez));
    if (ci < 0) continue;
    const int lo = __ldg(g.ustart + ci), hi = __ldg(g.ustart + ci + 1);
    for (int j = lo; j < hi; j++) {
      const double ax = __dsub_rn(__ldg(g.sx + j), px), ay = __dsub_rn(__ldg(g.sy + j), py),
                   az = __dsub_rn(__ldg(g.sz + j), pz);
      const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az));
      if (MODE == 0) {
        if (d2 < lim) {
          if (++count > nb) return true;
        }
      } else {
        if (__dsqrt_rn(d2) < lim) return true;
      }
    }
  }
  return false;
}

// one thread per SORTED position: neighbouring threads walk the same cells (broadcast loads)
__global__ void __launch_bounds__(G_THREADS) k_radius_count(DfuGrid g, const SegPrm* __restrict__ prm,
                                                            uint8_t* __restrict__ keep, int32_t* __restrict__ todo,
                                                            int32_t* __restrict__ ntodo) {
  const int64_t t = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (t >= g.n) return;
  const unsigned long long k = g.skey[t];
  if (k == ~0ull) return;                       // inactive: keep stays 0
  const int s = (int)(k >> 48);
  const SegPrm p = prm[s];
  const double r2 = __dmul_rn(p.r, p.r);
  const double px = g.sx[t], py = g.sy[t], pz = g.sz[t];
  // Fast accept: the sorted array puts points of the same / adjacent cells next to each other, so in a dense cloud
  // the nb + 1 points needed are almost always among the nb + 2 entries on each side of t (coalesced loads, no cell look-up).
  // Only a "keep" can be decided here; everything else takes the exact walk below, which counts from scratch.
  {
    int count = 1;   // the point itself (d2 = 0 < r2)
    const int reach = p.nb + 2 < 4 ? 4 : (p.nb + 2 > 48 ? 48 : p.nb + 2);   // entries examined on each side
    for (int o = 1; o <= reach && count <= p.nb; o++) {
#pragma unroll
      for (int sgn = -1; sgn <= 1; sgn += 2) {
        const int64_t u = t + sgn * o;
        if (u < 0 || u >= g.n) continue;
        if ((g.skey[u] >> 48) != (unsigned long long)s) continue;     // other segment or inactive (~0)
        const double ax = __dsub_rn(g.sx[u], px), ay = __dsub_rn(g.sy[u], py), az = __dsub_rn(g.sz[u], pz);
        if (__dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az)) < r2) count++;
      }
    }
    if (count > p.nb) { keep[g.sidx[t]] = 1; return; }
  }
  // undecided: queue the sorted position for the warp-per-point walk (k_radius_walk).  A per-thread walk here would
  // make nearly every warp wait for its one outlier lane to scan cells that hold thousands of points.
  todo[atomicAdd(ntodo, 1)] = (int32_t)t;
}

// Phase 2 of the radius filter: one WARP per undecided point.  The lanes stride over the points of each of the 9
// (x, y) columns around the query (coalesced loads of the sorted SoA) and stop as soon as the count exceeds nb.
__global__ void __launch_bounds__(G_THREADS) k_radius_walk(DfuGrid g, const SegPrm* __restrict__ prm,
                                                           const int32_t* __restrict__ todo, const int32_t* __restrict__ ntodo,
                                                           uint8_t* __restrict__ keep) {
  const int lane = threadIdx.x & 31;
  const int nw = (gridDim.x * G_THREADS) >> 5;
  const int nt = *ntodo;
  for (int e = (blockIdx.x * G_THREADS + threadIdx.x) >> 5; e < nt; e += nw) {
    const int t = todo[e];
    const int s = (int)(g.skey[t] >> 48);
    const SegPrm p = prm[s];
    const double r2 = __dmul_rn(p.r, p.r);
    const double px = g.sx[t], py = g.sy[t], pz = g.sz[t];
    const double c = g.cell[s];
    const int cx = dfu_cell_coord(px, dfu_unorder_f64(g.segmin[3 * s + 0]), c);
    const int cy = dfu_cell_coord(py, dfu_unorder_f64(g.segmin[3 * s + 1]), c);
    const int cz = dfu_cell_coord(pz, dfu_unorder_f64(g.segmin[3 * s + 2]), c);
    bool kp = false;
    if (cx >= 0 && cy >= 0 && cz >= 0) {
      // lane l < 27 looks up cell l of the 3x3x3 block (9 cells of a 2-D grid): one or two hash probes each, in parallel
      const int nz = g.dims == 3 ? 3 : 1;
      int lo = 0, hi = 0;
      if (lane < 9 * nz) {
        const int ex = cx + (lane / nz) / 3 - 1, ey = cy + (lane / nz) % 3 - 1, ez = cz + (nz == 3 ? lane % 3 - 1 : 0);
        if (ex >= 1 && ey >= 1 && ez >= 1) {
          const int ci = dfu_grid_find(g, dfu_grid_key(s, ex, ey, ez));
          if (ci >= 0) { lo = __ldg(g.ustart + ci); hi = __ldg(g.ustart + ci + 1); }
        }
      }
      unsigned have = __ballot_sync(0xffffffffu, hi > lo);
      int count = 0;
      while (have && !kp) {
        const int src = __ffs(have) - 1;
        have &= have - 1;
        const int clo = __shfl_sync(0xffffffffu, lo, src), chi = __shfl_sync(0xffffffffu, hi, src);
        for (int j0 = clo; j0 < chi && !kp; j0 += 32) {
          const int j = j0 + lane;
          bool in = false;
          if (j < chi) {
            const double ax = __dsub_rn(__ldg(g.sx + j), px), ay = __dsub_rn(__ldg(g.sy + j), py),
                         az = __dsub_rn(__ldg(g.sz + j), pz);
            in = __dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az)) < r2;
          }
          count += __popc(__ballot_sync(0xffffffffu, in));
          kp = count > p.nb;
        }
      }
    }
    if (lane == 0) keep[g.sidx[t]] = kp ? 1 : 0;
  }
}

// count-dependent radius / stride (my_loader0.py:739-766): one thread per segment, in place
__global__ void __launch_bounds__(G_THREADS) k_seg_resolve(SegPrm* __restrict__ prm, const int64_t* __restrict__ seg_off, int nseg) {
  const int s = blockIdx.x * G_THREADS + threadIdx.x;
  if (s >= nseg) return;
  SegPrm p = prm[s];
  const int64_t n = seg_off[s + 1] - seg_off[s];
  if (p.n0 >= 0 && !(n > p.n0)) p.r = p.r_small;
  if (p.n_hi >= 0) p.every_k = n > p.n_hi ? p.k_hi : (n > p.n_mid ? p.k_mid : p.every_k);
  prm[s] = p;
}

__global__ void __launch_bounds__(G_THREADS) k_seg_active(const int32_t* __restrict__ seg, const int64_t* __restrict__ seg_off,
                                                          int64_t n, const SegPrm* __restrict__ prm, uint8_t* __restrict__ active,
                                                          double* __restrict__ cell, int nseg) {
  const int64_t i = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (i < nseg) cell[i] = prm[i].r * (1.0 + 9.313225746154785e-10);
  if (i >= n) return;
  const int s = seg[i];
  const SegPrm p = prm[s];
  const int64_t local = i - seg_off[s];
  active[i] = (p.enabled && p.r > 0.0 && (p.every_k <= 1 || (local % p.every_k) == 0)) ? 1 : 0;
}

int dfu_radius_keep(dfu_ctx* ctx, cudaStream_t st, int sb, const double* x, const double* y, const double* z,
                    const int32_t* seg, const int64_t* seg_off_dev, int64_t n, int nseg, SegPrm* prm_dev,
                    uint8_t* keep) {
  if (n == 0) return DFU_OK;
  k_seg_resolve<<<(nseg + G_THREADS - 1) / G_THREADS, G_THREADS, 0, st>>>(prm_dev, seg_off_dev, nseg);
  DFU_LAUNCH_CHECK(ctx);
  DFU_SLOT(ctx, active, uint8_t*, sb + GRID_SLOTS - 1, n);
  DFU_SLOT(ctx, cellv, double*, sb + GRID_SLOTS, (size_t)nseg * 8);
  const int64_t span = n > nseg ? n : nseg;
  const unsigned nb = (unsigned)((span + G_THREADS - 1) / G_THREADS);
  k_seg_active<<<nb, G_THREADS, 0, st>>>(seg, seg_off_dev, n, prm_dev, active, cellv, nseg);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cudaMemsetAsync(keep, 0, n, st));
  DfuGrid g;
  {
    DFU_PROF(ctx, st, PT_GRID_BUILD);
    int rc = dfu_grid_build(ctx, st, sb, x, y, z, seg, active, n, nseg, cellv, &g);
    if (rc) return rc;
  }
  {
    DFU_PROF(ctx, st, PT_RADIUS);
    // the unsorted-key buffer of the grid build is free after the sort: it holds the list of undecided points
    int32_t* todo = (int32_t*)ctx->slot[sb + 1];
    int32_t* ntodo = g.ncell + 2;
    DFU_CUDA(ctx, cudaMemsetAsync(ntodo, 0, 4, st));
    k_radius_count<<<(unsigned)((n + G_THREADS - 1) / G_THREADS), G_THREADS, 0, st>>>(g, prm_dev, keep, todo, ntodo);
    DFU_LAUNCH_CHECK(ctx);
    k_radius_walk<<<148 * 8, G_THREADS, 0, st>>>(g, prm_dev, todo, ntodo, keep);
    DFU_LAUNCH_CHECK(ctx);
  }
  return DFU_OK;
}

// ---------------------------------------------------------------- NN gate on a hashed cell list (no sort)
// The gate's reference points (the kept LiDAR points of an instance) are sparse: a 0.1 m cell holds one to three of them.
// So the cells are not sorted into ranges at all: every active reference point is pushed onto the singly linked list of its
// cell (open-addressing hash keyed by segment + absolute cell coordinates, atomicExch on the list head), and a query
// probes its 27 cells and walks their short lists.  Round 1 built a second sorted grid for this (8 radix passes over
// every (point, instance) pair of the batch).
__device__ __forceinline__ bool gate_cell(double v, double cell, int& c) {
  const double f = floor(v / cell);
  if (!(f >= -32760.0 && f <= 32760.0)) return false;      // NaN / a cloud more than 32 k cells from the origin
  c = (int)f + 32768;
  return true;
}

// number of active reference points -> scal[3]; the hash then uses the smallest power of two >= 2 x that count (<= the
// allocated capacity), so that the probes of ~2 M queries stay inside a few MB of L2 instead of a table sized for every pair
__global__ void __launch_bounds__(G_THREADS) k_gate_count(const uint8_t* __restrict__ ractive, int64_t nr, int32_t* __restrict__ nact) {
  const int64_t i = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  const bool a = i < nr && (!ractive || ractive[i]);
  const unsigned m = __ballot_sync(0xffffffffu, a);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(nact, __popc(m));
}
__device__ __forceinline__ unsigned gate_mask(int nact, unsigned cap_mask) {
  unsigned want = nact > 16 ? (unsigned)nact * 2u : 32u;
  unsigned m = 0xffffffffu >> __clz(want - 1u);          // pow2ceil(want) - 1
  return m < cap_mask ? m : cap_mask;
}

__global__ void __launch_bounds__(G_THREADS) k_gate_insert(const double* __restrict__ rx, const double* __restrict__ ry,
                                                           const double* __restrict__ rz, const int32_t* __restrict__ rseg,
                                                           const uint8_t* __restrict__ ractive, int64_t nr,
                                                           const double* __restrict__ maxdist, unsigned long long* __restrict__ hkey,
                                                           int32_t* __restrict__ head, int32_t* __restrict__ next, unsigned cap_mask,
                                                           int32_t* __restrict__ errflag, const int32_t* __restrict__ nact) {
  const int64_t i = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (i >= nr || (ractive && !ractive[i])) return;
  const unsigned hmask = gate_mask(*nact, cap_mask);
  const int s = rseg[i];
  const double cell = maxdist[s] * (1.0 + 9.313225746154785e-10);
  int cx, cy, cz;
  if (!(cell > 0.0) || !gate_cell(rx[i], cell, cx) || !gate_cell(ry[i], cell, cy) || !gate_cell(rz[i], cell, cz)) {
    atomicOr(errflag, 1);
    return;
  }
  const unsigned long long k = dfu_grid_key(s, cx, cy, cz);
  unsigned h = dfu_hash_u64(k) & hmask;
  while (true) {
    const unsigned long long prev = atomicCAS(hkey + h, ~0ull, k);
    if (prev == ~0ull || prev == k) break;
    h = (h + 1) & hmask;
  }
  next[i] = atomicExch(head + h, (int32_t)i);
}

__global__ void __launch_bounds__(G_THREADS) k_gate(const double* __restrict__ qx, const double* __restrict__ qy,
                                                    const double* __restrict__ qz, const int32_t* __restrict__ qseg, int64_t nq,
                                                    const double* __restrict__ rx, const double* __restrict__ ry,
                                                    const double* __restrict__ rz, const double* __restrict__ maxdist,
                                                    const unsigned long long* __restrict__ hkey, const int32_t* __restrict__ head,
                                                    const int32_t* __restrict__ next, unsigned cap_mask,
                                                    const uint8_t* __restrict__ qactive, uint8_t* __restrict__ keep,
                                                    const int32_t* __restrict__ nact) {
  const int64_t i = (int64_t)blockIdx.x * G_THREADS + threadIdx.x;
  if (i >= nq) return;
  const unsigned hmask = gate_mask(*nact, cap_mask);
  if (qactive && !qactive[i]) { keep[i] = 0; return; }      // the caller does not need this query's answer
  const int s = qseg[i];
  const double lim = maxdist[s], cell = lim * (1.0 + 9.313225746154785e-10);
  const double px = qx[i], py = qy[i], pz = qz[i];
  int cx, cy, cz;
  bool kp = false;
  if (cell > 0.0 && gate_cell(px, cell, cx) && gate_cell(py, cell, cy) && gate_cell(pz, cell, cz)) {
    for (int o = 0; o < 27 && !kp; o++) {
      // the query's own cell first
      const int oo = o == 0 ? 13 : (o <= 13 ? o - 1 : o);
      const unsigned long long k = dfu_grid_key(s, cx + oo / 9 - 1, cy + (oo / 3) % 3 - 1, cz + oo % 3 - 1);
      unsigned h = dfu_hash_u64(k) & hmask;
      int j = -1;
      while (true) {
        const unsigned long long hk = __ldg(hkey + h);
        if (hk == k) { j = __ldg(head + h); break; }
        if (hk == ~0ull) break;
        h = (h + 1) & hmask;
      }
      for (; j >= 0 && !kp; j = __ldg(next + j)) {
        const double ax = __dsub_rn(__ldg(rx + j), px), ay = __dsub_rn(__ldg(ry + j), py), az = __dsub_rn(__ldg(rz + j), pz);
        const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(ax, ax), __dmul_rn(ay, ay)), __dmul_rn(az, az));
        kp = __dsqrt_rn(d2) < lim;       // BallQuery: sqrt of the sum of squares, strict (my_loader.py:489-494)
      }
    }
  }
  keep[i] = kp ? 1 : 0;
}

int dfu_gate_keep(dfu_ctx* ctx, cudaStream_t st, int sb, const double* qx, const double* qy, const double* qz,
                  const int32_t* qseg, int64_t nq, const double* rx, const double* ry, const double* rz,
                  const int32_t* rseg, const uint8_t* ractive, int64_t nr, int nseg, const double* maxdist_dev,
                  uint8_t* keep, const uint8_t* qactive) {
  (void)nseg;
  if (nq == 0) return DFU_OK;
  if (nr == 0) {
    DFU_CUDA(ctx, cudaMemsetAsync(keep, 0, nq, st));
    return DFU_OK;
  }
  if (nr >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_DATA, "gate: too many points");
  size_t hcap = 64;
  while (hcap < (size_t)nr * 2) hcap <<= 1;
  DFU_SLOT(ctx, htab, unsigned long long*, sb + 14, hcap * 12);
  DFU_SLOT(ctx, next, int32_t*, sb + 1, (size_t)nr * 8);
  DFU_SLOT(ctx, scal, int32_t*, sb + 11, 64);
  int32_t* head = (int32_t*)(htab + hcap);
  {
    DFU_PROF(ctx, st, PT_GRID_BUILD);
    DFU_CUDA(ctx, cudaMemsetAsync(htab, 0xFF, hcap * 12, st));      // keys: empty, heads: -1
    DFU_CUDA(ctx, cudaMemsetAsync(scal + 3, 0, 4, st));
    k_gate_count<<<(unsigned)((nr + G_THREADS - 1) / G_THREADS), G_THREADS, 0, st>>>(ractive, nr, scal + 3);
    DFU_LAUNCH_CHECK(ctx);
    k_gate_insert<<<(unsigned)((nr + G_THREADS - 1) / G_THREADS), G_THREADS, 0, st>>>(rx, ry, rz, rseg, ractive, nr, maxdist_dev, htab,
                                                                                       head, next, (unsigned)(hcap - 1), scal + 1, scal + 3);
    DFU_LAUNCH_CHECK(ctx);
  }
  DFU_PROF(ctx, st, PT_GATE);
  k_gate<<<(unsigned)((nq + G_THREADS - 1) / G_THREADS), G_THREADS, 0, st>>>(qx, qy, qz, qseg, nq, rx, ry, rz, maxdist_dev, htab, head, next,
                                                                             (unsigned)(hcap - 1), qactive, keep, scal + 3);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

int dfu_grid_err_reset(dfu_ctx* ctx, cudaStream_t st, int sb) {
  DFU_SLOT(ctx, scal, int32_t*, sb + 11, 64);
  DFU_CUDA(ctx, cudaMemsetAsync(scal, 0, 64, st));
  return DFU_OK;
}
const int32_t* dfu_grid_err_ptr(dfu_ctx* ctx, int sb) { return (const int32_t*)ctx->slot[sb + 11] + 1; }
