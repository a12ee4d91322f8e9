// S5 adaptive-range clustering and S6 rectangle search.
// Reference: tools/PENet/rectangle_fitting/rectangle_fitting.py:161-191 (_adoptive_range_segmentation),
// :113-159 (_rectangle_search), :83-111 (_calc_variance_criterion).
//
// The reference builds S_i = { j : d(i,j) <= R0 + Rd*|p_i| } for every point and merges intersecting sets to a
// fixpoint in O(n^3) Python with a 15 s wall-clock abort.  The fixpoint is the set of connected components of the
// graph i~j <=> d(i,j) <= max(R_i, R_j), listed by minimum point index.  Here: a 2-D grid of
// edge R0/sqrt(2) (any two points of one cell are closer than R0 <= R_i, so a cell is connected by construction),
// union-find over CELLS with atomic hooking, and an exact "exists a connecting pair" test per pair of nearby cells,
// pruned by the cells' bounding boxes and skipped when the two cells already share a root.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "dfu_ops.cuh"

#define C_THREADS 128

struct CellAux {
  unsigned long long minx, maxx, miny, maxy, maxR;   // orderable bits
};

__global__ void k_cluster_cellsize(const double* __restrict__ R0, double* __restrict__ cell, int ninst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ninst) cell[i] = R0[i] / sqrt(2.0) * (1.0 - 1e-9);
}

__global__ void __launch_bounds__(256) k_cluster_aux_init(CellAux* __restrict__ aux, int64_t n) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= n) return;
  CellAux a;
  a.minx = ~0ull; a.maxx = 0ull; a.miny = ~0ull; a.maxy = 0ull; a.maxR = 0ull;
  aux[t] = a;
}

__global__ void __launch_bounds__(256) k_cluster_aux(DfuGrid g, const double* __restrict__ R0, const double* __restrict__ Rd,
                                                     double* __restrict__ sR, CellAux* __restrict__ aux,
                                                     int32_t* __restrict__ parent, int32_t* __restrict__ rootmin) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const bool ok = t < g.n;
  int c = -1;
  unsigned long long bx = 0, by = 0, bR = 0;
  if (ok) {
    const int s = (int)(g.skey[t] >> 48);
    const double x = g.sx[t], y = g.sy[t];
    // R = R0 + Rd * np.linalg.norm([x, y])   (rectangle_fitting.py:167)
    const double R = __dadd_rn(R0[s], __dmul_rn(Rd[s], __dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)))));
    sR[t] = R;
    // The pair test visits the 5x5 cell neighbourhood (cell edge R0 / sqrt(2)): exact only while R <= 2 cells.  A
    // caller-supplied R0 <= 0, Rd < 0 or an R beyond that is an error (sticky flag, bit 1), never silently wrong clusters.
    if (!(R0[s] > 0.0) || Rd[s] < 0.0 || !(R <= 2.0 * g.cell[s])) atomicOr(g.errflag, 2);
    c = g.pcell[t];
    bx = dfu_order_f64(x); by = dfu_order_f64(y); bR = dfu_order_f64(R);
    if (t == g.ustart[c]) {
      parent[c] = c;
      rootmin[c] = 0x7fffffff;
    }
  }
  const int c0 = __shfl_sync(0xffffffffu, c, 0);
  if (__all_sync(0xffffffffu, c == c0)) {
    if (c0 < 0) return;
    // the whole warp sits in one cell (cells hold hundreds of points): shuffle reduction, 5 atomics per warp
    unsigned long long mnx = bx, mxx = bx, mny = by, mxy = by, mxR = bR;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long a0 = __shfl_xor_sync(0xffffffffu, mnx, o), a1 = __shfl_xor_sync(0xffffffffu, mxx, o),
                               a2 = __shfl_xor_sync(0xffffffffu, mny, o), a3 = __shfl_xor_sync(0xffffffffu, mxy, o),
                               a4 = __shfl_xor_sync(0xffffffffu, mxR, o);
      mnx = a0 < mnx ? a0 : mnx; mxx = a1 > mxx ? a1 : mxx; mny = a2 < mny ? a2 : mny; mxy = a3 > mxy ? a3 : mxy;
      mxR = a4 > mxR ? a4 : mxR;
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&aux[c0].minx, mnx); atomicMax(&aux[c0].maxx, mxx); atomicMin(&aux[c0].miny, mny); atomicMax(&aux[c0].maxy, mxy);
      atomicMax(&aux[c0].maxR, mxR);
    }
  } else if (ok) {
    atomicMin(&aux[c].minx, bx);
    atomicMax(&aux[c].maxx, bx);
    atomicMin(&aux[c].miny, by);
    atomicMax(&aux[c].maxy, by);
    atomicMax(&aux[c].maxR, bR);
  }
}

__device__ __forceinline__ int uf_find(volatile int32_t* parent, int c) {
  int p = parent[c];
  while (p != c) {
    const int gp = parent[p];
    if (gp != p) parent[c] = gp;   // path halving (benign race: only ever points further up)
    c = p;
    p = gp;
  }
  return c;
}

__device__ __forceinline__ void uf_union(int32_t* parent, int a, int b) {
  while (true) {
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a == b) return;
    if (a < b) { int t = a; a = b; b = t; }            // hook the larger root under the smaller
    if (atomicCAS(parent + a, a, b) == a) return;
  }
}

// the 12 "forward" offsets of the 5x5 neighbourhood (each unordered cell pair is visited once)
__constant__ int c_off[12][2] = {{0, 1}, {0, 2}, {1, -2}, {1, -1}, {1, 0}, {1, 1}, {1, 2}, {2, -2}, {2, -1}, {2, 0}, {2, 1}, {2, 2}};

__global__ void __launch_bounds__(C_THREADS) k_cluster_pairs(DfuGrid g, const double* __restrict__ sR,
                                                             const CellAux* __restrict__ aux, int32_t* __restrict__ parent,
                                                             int pass) {
  __shared__ int s_c2, s_found;
  const int ncell = *g.ncell;
  const long long ntask = (long long)ncell * 12;
  for (long long task = blockIdx.x; task < ntask; task += gridDim.x) {
    const int c = (int)(task / 12), o = (int)(task % 12);
    // pass 0: the 4 edge/corner-adjacent offsets only; pass 1: the 8 remaining (distance-2) offsets
    const bool near = (c_off[o][0] <= 1 && c_off[o][1] >= -1 && c_off[o][1] <= 1);
    if ((pass == 0) != near) continue;
    __syncthreads();
    if (threadIdx.x == 0) {
      s_c2 = -1;
      s_found = 0;
      const unsigned long long k = g.ukey[c];
      if (k != ~0ull) {
        const int cx = (int)((k >> 32) & 0xffff) + c_off[o][0], cy = (int)((k >> 16) & 0xffff) + c_off[o][1];
        if (cx >= 1 && cy >= 1 && cx < 65535 && cy < 65535) {
          const int sg = (int)(k >> 48);
          const unsigned long long k2 = dfu_grid_key(sg, cx, cy, 1);
          const int sc0 = g.segcell[sg], sc1 = g.segcell[sg + 1];
          const int c2 = sc0 + dfu_lower_bound_u64(g.ukey + sc0, sc1 - sc0, k2);
          if (c2 < sc1 && g.ukey[c2] == k2 && uf_find(parent, c) != uf_find(parent, c2)) s_c2 = c2;
        }
      }
    }
    __syncthreads();
    const int c2 = s_c2;
    if (c2 < 0) continue;
    const int a0 = g.ustart[c], a1 = g.ustart[c + 1], b0 = g.ustart[c2], b1 = g.ustart[c2 + 1];
    const double bminx = dfu_unorder_f64(aux[c2].minx), bmaxx = dfu_unorder_f64(aux[c2].maxx),
                 bminy = dfu_unorder_f64(aux[c2].miny), bmaxy = dfu_unorder_f64(aux[c2].maxy),
                 bR = dfu_unorder_f64(aux[c2].maxR);
    volatile int* vfound = &s_found;
    for (int i = a0 + threadIdx.x; i < a1 && !*vfound; i += C_THREADS) {
      const double xi = g.sx[i], yi = g.sy[i], Ri = sR[i];
      // conservative prune: distance from p_i to the bounding box of the other cell
      const double ex = fmax(fmax(bminx - xi, xi - bmaxx), 0.0), ey = fmax(fmax(bminy - yi, yi - bmaxy), 0.0);
      const double lim = fmax(Ri, bR) * (1.0 + 1e-12);
      if (ex * ex + ey * ey > lim * lim) continue;
      for (int j = b0; j < b1; j++) {
        const double dx = __dsub_rn(xi, __ldg(g.sx + j)), dy = __dsub_rn(yi, __ldg(g.sy + j));
        const double d = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));   // rectangle_fitting.py:169
        if (d <= Ri || d <= __ldg(sR + j)) {
          *vfound = 1;
          break;
        }
        if ((j & 31) == 31 && *vfound) break;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_found) uf_union(parent, c, c2);
  }
}

__global__ void __launch_bounds__(256) k_cluster_rootmin(DfuGrid g, int32_t* __restrict__ parent, int32_t* __restrict__ rootmin) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= *g.ncell) return;
  if (g.ukey[c] == ~0ull) return;
  // radix sort is stable and idx ascends, so the first sorted point of a cell has the cell's minimum index
  atomicMin(rootmin + uf_find(parent, c), g.sidx[g.ustart[c]]);
}

__global__ void __launch_bounds__(256) k_cluster_minidx(DfuGrid g, int32_t* __restrict__ parent, const int32_t* __restrict__ rootmin,
                                                        int32_t* __restrict__ minidx, int32_t* __restrict__ head) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= g.n) return;
  const int i = g.sidx[t];
  const int m = (g.skey[t] == ~0ull) ? i : rootmin[uf_find(parent, g.pcell[t])];
  minidx[i] = m;
  head[i] = (m == i) ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_cluster_number(const int32_t* __restrict__ minidx, const int32_t* __restrict__ cnum,
                                                        int64_t n, int32_t* __restrict__ cluster_of_point,
                                                        int32_t* __restrict__ first_point, int32_t* __restrict__ ncluster) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int m = minidx[i];
  const int k = cnum[m];             // exclusive scan of head
  cluster_of_point[i] = k;
  if (m == i) first_point[k] = (int32_t)i;
  if (i == n - 1) *ncluster = cnum[i] + ((m == i) ? 1 : 0);
}

int dfu_cluster_points(dfu_ctx* ctx, cudaStream_t st, int sb, const double* x, const double* y, const int32_t* inst,
                       const int64_t* inst_off_dev, int64_t n, int ninst, const double* R0_dev, const double* Rd_dev,
                       ClusterOut* out) {
  (void)inst_off_dev;
  const int64_t m = n > 0 ? n : 1;
  const int ob = sb + GRID_SLOTS + 1;
  DFU_SLOT(ctx, cellv, double*, sb + GRID_SLOTS, (size_t)(ninst > 0 ? ninst : 1) * 8);
  DFU_SLOT(ctx, sR, double*, ob + 0, m * 8);
  DFU_SLOT(ctx, aux, CellAux*, ob + 1, m * sizeof(CellAux));
  DFU_SLOT(ctx, parent, int32_t*, ob + 2, m * 4);
  DFU_SLOT(ctx, rootmin, int32_t*, ob + 3, m * 4);
  DFU_SLOT(ctx, minidx, int32_t*, ob + 4, m * 4);
  DFU_SLOT(ctx, head, int32_t*, ob + 5, m * 4);
  DFU_SLOT(ctx, cop, int32_t*, ob + 6, m * 4);
  DFU_SLOT(ctx, first, int32_t*, ob + 7, m * 4);
  DFU_SLOT(ctx, ncl, int32_t*, ob + 8, 16);
  out->n = n; out->cluster_of_point = cop; out->ncluster = ncl; out->first_point = first;
  DFU_CUDA(ctx, cudaMemsetAsync(ncl, 0, 16, st));
  if (n == 0) return DFU_OK;
  DFU_PROF(ctx, st, PT_CLUSTER);
  k_cluster_cellsize<<<(ninst + 127) / 128, 128, 0, st>>>(R0_dev, cellv, ninst);
  DFU_LAUNCH_CHECK(ctx);
  DfuGrid g;
  int rc = dfu_grid_build(ctx, st, sb, x, y, nullptr, inst, nullptr, n, ninst, cellv, &g);
  if (rc) return rc;
  // aux: min fields start at ~0, max fields at 0 (a kernel: strided cudaMemset2D over 2.5 M rows costs milliseconds)
  const unsigned nb = (unsigned)((n + 255) / 256);
  k_cluster_aux_init<<<nb, 256, 0, st>>>(aux, n);
  DFU_LAUNCH_CHECK(ctx);
  k_cluster_aux<<<nb, 256, 0, st>>>(g, R0_dev, Rd_dev, sR, aux, parent, rootmin);
  DFU_LAUNCH_CHECK(ctx);
  const int grid = 148 * 8;
  k_cluster_pairs<<<grid, C_THREADS, 0, st>>>(g, sR, aux, parent, 0);
  DFU_LAUNCH_CHECK(ctx);
  k_cluster_pairs<<<grid, C_THREADS, 0, st>>>(g, sR, aux, parent, 1);
  DFU_LAUNCH_CHECK(ctx);
  k_cluster_rootmin<<<nb, 256, 0, st>>>(g, parent, rootmin);
  DFU_LAUNCH_CHECK(ctx);
  k_cluster_minidx<<<nb, 256, 0, st>>>(g, parent, rootmin, minidx, head);
  DFU_LAUNCH_CHECK(ctx);
  size_t tb = 0;
  DFU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb, head, head, (int)n, st));
  DFU_SLOT(ctx, tmp, void*, ob + 9, tb);
  DFU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tb, head, head, (int)n, st));
  ctx->launches += 2;
  k_cluster_number<<<nb, 256, 0, st>>>(minidx, head, n, cop, first, ncl);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

// ---------------------------------------------------------------- S6: rectangle search
// rect record (16 doubles): a[4] b[4] c[4] theta best_cost runner_up_cost n_points
__global__ void __launch_bounds__(256) k_fit_iota(int32_t* __restrict__ idx, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < n) idx[i] = (int32_t)i;
}

__global__ void __launch_bounds__(256) k_fit_gather(const int32_t* __restrict__ scop, const int32_t* __restrict__ order,
                                                    const double* __restrict__ x, const double* __restrict__ y, int64_t n,
                                                    int K, double* __restrict__ cx, double* __restrict__ cy,
                                                    int32_t* __restrict__ coff) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= n) return;
  const int i = order[t];
  cx[t] = x[i];
  cy[t] = y[i];
  if (t == 0 || scop[t] != scop[t - 1]) coff[scop[t]] = (int32_t)t;
  if (t == n - 1) coff[K] = (int32_t)n;
}

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void fit_project(double x, double y, double cs, double sn, double& c1, double& c2) {
  c1 = __dadd_rn(__dmul_rn(x, cs), __dmul_rn(y, sn));      // X @ [cos, sin]    (rectangle_fitting.py:124)
  c2 = __dadd_rn(__dmul_rn(x, -sn), __dmul_rn(y, cs));     // X @ [-sin, cos]   (:125)
}

// One CTA per cluster.  The cluster's points (contiguous in the gathered arrays) are staged ONCE in shared memory with two
// TMA bulk copies (x and y; cp.async.bulk + mbarrier) and the 89 angles loop over the staged copy: round 1 re-read the
// points from global memory 89 x 3 times per cluster.  Clusters of up to FIT_WARP_MAX points give every warp its own angles
// (shuffle reductions only); larger ones work on one angle at a time with all 256 threads.  Clusters that do not fit
// the staging buffer (> FIT_STAGE points) read global memory as before.
#define FIT_STAGE 1024      /* points staged per cluster (16 KB: 13 CTAs per SM) */
#define FIT_WARP_MAX 512

__device__ __forceinline__ double fit_block_reduce(double v, int op, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? fmin(v, w) : (op == 1 ? fmax(v, w) : v + w);
  }
  __syncthreads();                       // previous use of sh is over
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = sh[0];
#pragma unroll
  for (int i = 1; i < 8; i++) r = op == 0 ? fmin(r, sh[i]) : (op == 1 ? fmax(r, sh[i]) : r + sh[i]);
  return r;
}
__device__ __forceinline__ double fit_warp_reduce(double v, int op) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double w = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? fmin(v, w) : (op == 1 ? fmax(v, w) : v + w);
  }
  return v;
}

// variance criterion of one angle over points px/py [n], lanes lane, lane + G, ...  (rectangle_fitting.py:83-111;
// np.var = mean((d - mean(d))^2)).  BLOCK: reductions over the 256 threads, else over the warp.
template <bool BLOCK>
__device__ __forceinline__ double fit_angle_cost(const double* __restrict__ px, const double* __restrict__ py, int n, double cs, double sn,
                                                 int lane, double* sh, double* ext) {
  const int G = BLOCK ? 256 : 32;
  double mn1 = INFINITY, mx1 = -INFINITY, mn2 = INFINITY, mx2 = -INFINITY;
  for (int p = lane; p < n; p += G) {
    double c1, c2;
    fit_project(px[p], py[p], cs, sn, c1, c2);
    mn1 = fmin(mn1, c1); mx1 = fmax(mx1, c1); mn2 = fmin(mn2, c2); mx2 = fmax(mx2, c2);
  }
  if (BLOCK) {
    mn1 = fit_block_reduce(mn1, 0, sh); mx1 = fit_block_reduce(mx1, 1, sh); mn2 = fit_block_reduce(mn2, 0, sh); mx2 = fit_block_reduce(mx2, 1, sh);
  } else {
    mn1 = fit_warp_reduce(mn1, 0); mx1 = fit_warp_reduce(mx1, 1); mn2 = fit_warp_reduce(mn2, 0); mx2 = fit_warp_reduce(mx2, 1);
  }
  if (ext) { ext[0] = mn1; ext[1] = mn2; ext[2] = mx1; ext[3] = mx2; }
  // One pass for both moments of the two distance sets, on data shifted by the middle of their range (D1 lies in
  // [0, (mx1 - mn1) / 2]): var = (sum e^2 - (sum e)^2 / n) / n with e = D - pivot.  The shift keeps |e| below a quarter
  // of the extent, so the cancellation costs ~n ulps (1e-12 relative for 10^4 points), far below the 1e-9 gap under
  // which a maximum is reported as a near-tie; np.var's own two-pass value differs from the exact one by as much.
  const double p1 = 0.25 * (mx1 - mn1), p2 = 0.25 * (mx2 - mn2);
  double s1 = 0, s2 = 0, n1 = 0, n2 = 0, q1 = 0, q2 = 0;
  for (int p = lane; p < n; p += G) {
    double c1, c2;
    fit_project(px[p], py[p], cs, sn, c1, c2);
    const double D1 = fmin(fabs(mx1 - c1), fabs(c1 - mn1)), D2 = fmin(fabs(mx2 - c2), fabs(c2 - mn2));
    if (D1 < D2) { const double e = D1 - p1; s1 += e; q1 = fma(e, e, q1); n1 += 1; }
    else { const double e = D2 - p2; s2 += e; q2 = fma(e, e, q2); n2 += 1; }
  }
  if (BLOCK) {
    s1 = fit_block_reduce(s1, 2, sh); s2 = fit_block_reduce(s2, 2, sh); n1 = fit_block_reduce(n1, 2, sh); n2 = fit_block_reduce(n2, 2, sh);
    q1 = fit_block_reduce(q1, 2, sh); q2 = fit_block_reduce(q2, 2, sh);
  } else {
    s1 = fit_warp_reduce(s1, 2); s2 = fit_warp_reduce(s2, 2); n1 = fit_warp_reduce(n1, 2); n2 = fit_warp_reduce(n2, 2);
    q1 = fit_warp_reduce(q1, 2); q2 = fit_warp_reduce(q2, 2);
  }
  const double V1 = n1 > 0 ? -((q1 - s1 * s1 / n1) / n1) : 0.0, V2 = n2 > 0 ? -((q2 - s2 * s2 / n2) / n2) : 0.0;
  return V1 + V2;
}

#define FIT_SLICES 8        /* CTAs per cluster: each takes every 8th angle */
__global__ void __launch_bounds__(256) k_fit_cost2(const double* __restrict__ cx, const double* __restrict__ cy,
                                                   const int32_t* __restrict__ coff, int K, const double* __restrict__ angles,
                                                   int na, double* __restrict__ cost) {
  extern __shared__ __align__(128) unsigned char fit_smem[];
  double* sx = reinterpret_cast<double*>(fit_smem);                    // [FIT_STAGE]
  double* sy = sx + FIT_STAGE;                                         // [FIT_STAGE]
  double* sh = sy + FIT_STAGE;                                         // [8]
  unsigned long long* mbar = reinterpret_cast<unsigned long long*>(sh + 8);
  const int k = blockIdx.x, sl = blockIdx.y;
  const int p0 = coff[k], n = coff[k + 1] - p0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double *px = cx + p0, *py = cy + p0;
  if (n <= FIT_STAGE - 2 && n > 64) {
    // the bulk copy needs 16-byte alignment: start at the even element at or below p0
    const int a0 = p0 & ~1, skew = p0 - a0;
    const uint32_t bytes = (uint32_t)(((n + skew + 1) & ~1) * 8);
    if (threadIdx.x == 0) {
      dfu_mbar_init(mbar, 1);
      dfu_mbar_arrive_expect_tx(mbar, 2u * bytes);
      dfu_bulk_g2s(sx, cx + a0, bytes, mbar);
      dfu_bulk_g2s(sy, cy + a0, bytes, mbar);
    }
    __syncthreads();
    dfu_mbar_wait(mbar, 0);
    px = sx + skew; py = sy + skew;
  }
  if (n <= FIT_WARP_MAX) {
    for (int a = sl + FIT_SLICES * warp; a < na; a += FIT_SLICES * 8) {
      const double c = fit_angle_cost<false>(px, py, n, angles[2 * a], angles[2 * a + 1], lane, nullptr, nullptr);
      if (lane == 0) cost[(long long)k * na + a] = c;
    }
  } else {
    for (int a = sl; a < na; a += FIT_SLICES) {
      const double c = fit_angle_cost<true>(px, py, n, angles[2 * a], angles[2 * a + 1], threadIdx.x, sh, nullptr);
      if (threadIdx.x == 0) cost[(long long)k * na + a] = c;
    }
  }
}

// one CTA per cluster: first maximum (rectangle_fitting.py:135-136), runner-up, rectangle at the best angle (:138-157);
// every warp scans the costs redundantly (identical result in each)
__global__ void __launch_bounds__(256) k_fit_finish2(const double* __restrict__ cx, const double* __restrict__ cy,
                                                     const int32_t* __restrict__ coff, int K, const double* __restrict__ angles,
                                                     int na, const double* __restrict__ cost, double* __restrict__ rect) {
  __shared__ double sh[8];
  const int k = blockIdx.x;
  const int p0 = coff[k], n = coff[k + 1] - p0;
  const int lane = threadIdx.x & 31;
  const double *px = cx + p0, *py = cy + p0;
  const double* scost = cost + (long long)k * na;
  double best = -INFINITY;
  int bi = 0x7fffffff;
  for (int a = lane; a < na; a += 32) {
    const double c = scost[a];
    if (c > best) { best = c; bi = a; }        // ascending a per lane: keeps the first maximum of the lane
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  double run = -INFINITY;
  for (int a = lane; a < na; a += 32)
    if (a != bi) run = fmax(run, scost[a]);
  run = warp_max(run);
  if (bi == 0x7fffffff) bi = 0;               // all costs NaN: the reference keeps theta = None; flagged by best = -inf
  const double cs = angles[2 * bi], sn = angles[2 * bi + 1];
  double mn1 = INFINITY, mx1 = -INFINITY, mn2 = INFINITY, mx2 = -INFINITY;
  for (int p = threadIdx.x; p < n; p += 256) {
    double c1, c2;
    fit_project(px[p], py[p], cs, sn, c1, c2);
    mn1 = fmin(mn1, c1); mx1 = fmax(mx1, c1); mn2 = fmin(mn2, c2); mx2 = fmax(mx2, c2);
  }
  mn1 = fit_block_reduce(mn1, 0, sh); mx1 = fit_block_reduce(mx1, 1, sh); mn2 = fit_block_reduce(mn2, 0, sh); mx2 = fit_block_reduce(mx2, 1, sh);
  if (threadIdx.x == 0) {
    double* r = rect + (long long)k * 16;
    r[0] = cs; r[1] = -sn; r[2] = cs; r[3] = -sn;
    r[4] = sn; r[5] = cs; r[6] = sn; r[7] = cs;
    r[8] = mn1; r[9] = mn2; r[10] = mx1; r[11] = mx2;
    r[12] = atan2(sn, cs); r[13] = best; r[14] = run; r[15] = (double)n;
  }
}

// rect_dev [K][16].  K must be the host copy of *ClusterOut.ncluster.
int dfu_fit_clusters(dfu_ctx* ctx, cudaStream_t st, int sb, const double* x, const double* y, const int32_t* cop, int64_t n,
                     int K, const double* angles_dev, int na, double* rect_dev) {
  if (K == 0 || n == 0) return DFU_OK;
  DFU_PROF(ctx, st, PT_FIT);
  DFU_SLOT(ctx, idx, int32_t*, sb + 0, n * 4);
  DFU_SLOT(ctx, scop, int32_t*, sb + 1, n * 4);
  DFU_SLOT(ctx, order, int32_t*, sb + 2, n * 4);
  DFU_SLOT(ctx, cx, double*, sb + 3, n * 8);
  DFU_SLOT(ctx, cy, double*, sb + 4, n * 8);
  DFU_SLOT(ctx, coff, int32_t*, sb + 5, ((size_t)K + 1) * 4);
  const unsigned nb = (unsigned)((n + 255) / 256);
  k_fit_iota<<<nb, 256, 0, st>>>(idx, n);
  DFU_LAUNCH_CHECK(ctx);
  int bits = 1;
  while ((1ll << bits) < K) bits++;
  size_t tb = 0;
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, cop, scop, idx, order, (int)n, 0, bits, st));
  DFU_SLOT(ctx, tmp, void*, sb + 7, tb);
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, cop, scop, idx, order, (int)n, 0, bits, st));
  ctx->launches += 4;
  k_fit_gather<<<nb, 256, 0, st>>>(scop, order, x, y, n, K, cx, cy, coff);
  DFU_LAUNCH_CHECK(ctx);
  DFU_SLOT(ctx, cost, double*, sb + 6, (size_t)K * na * 8);
  const size_t smem = (size_t)(2 * FIT_STAGE + 8) * 8 + 16;
  DFU_CUDA(ctx, cudaFuncSetAttribute(k_fit_cost2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_fit_cost2<<<dim3((unsigned)K, FIT_SLICES), 256, smem, st>>>(cx, cy, coff, K, angles_dev, na, cost);
  DFU_LAUNCH_CHECK(ctx);
  k_fit_finish2<<<K, 256, 0, st>>>(cx, cy, coff, K, angles_dev, na, cost, rect_dev);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}
