// S2: pixel -> 3-D back-projection, height gate, angular-voxel dedup (evaluated once per view), cloud emit.
// Reference arithmetic: tools/PENet/dataloaders/my_loader.py:507-509,532-542,554-557 (depth2pointsrgbpm),
// :166-180 (to_sphere_coords), :247-275 (la_sampling2 / la_sampling20), :69-83 (voxel generators),
// tools/PENet/dataloaders/calibration_kitti.py:89-102,134-144.
//
// The reference materialises new_p [M,8] float64, compacts it twice and re-voxelises the whole cloud once per
// instance.  Here the pixel's linear index IS the point's rank (np.nonzero is row-major), so nothing is
// materialised: three passes over the depth map resolve, per angular voxel, the opening pixel, the point count, the
// min-z representative (la_sampling2) and the min-y representative (la_sampling20, quirk Q2) in a direct-address
// table, and a fourth kernel gathers the representatives.
#include <mutex>
#include <cuda_fp16.h>
#include <cub/cub.cuh>
#include "dfu_common.cuh"

#define S2_THREADS 256
// theta in (1.5, pi] -> bins 3249..4071 ; atan in [-pi/2, pi/2] -> bins 1714..3285 (voxel 0.002f, origin -5)
#define CT0 3249
#define NT 823
#define CF0 1714
#define NF 1572
#define NCELL (NT * NF)
#define KEY_INVALID 0xFFFFFFFFu
#define KEY_AMBIG 0xFFFFFFFEu   /* pass A only: resolved by k_densify_amb before pass B */
#define S2_NQ 64                /* sub-queues of ambiguous pixels (spreads the counter atomics) */
#define S2_EPS 2e-6f   /* rad: twice the error bound of the float32 angle evaluation in s2_key */

struct __align__(32) VoxCell {
  unsigned long long minz;  // orderable bits of the smallest z
  unsigned long long miny;
  unsigned int first;       // opening pixel (smallest linear index)
  unsigned int cnt;         // count - 1 (starts at 0xFFFFFFFF)
  unsigned int argz;        // smallest pixel index attaining minz
  unsigned int argy;
};

struct S2Prm {
  int H, W, Hm, Wm, image_is_i32, max_per_voxel;
  float min_depth;
  double max_z, min_theta;
  double rW;   // RN(1 / W): row = floor((pix + 0.5) * rW) is exact for pix < 2^31, W < 2^20 (error 2^-21 << 0.5 / W)
};

static S2Prm make_s2(const dfu_fusion_params* p) {
  S2Prm s;
  s.H = p->H; s.W = p->W; s.Hm = p->Hm; s.Wm = p->Wm; s.image_is_i32 = p->image_is_i32;
  s.max_per_voxel = p->max_per_voxel; s.min_depth = p->min_depth; s.max_z = p->max_z; s.min_theta = p->min_theta;
  s.rW = 1.0 / (double)p->W;
  return s;
}

// Per-view float64 constants of the back-projection for passes A and B, in constant memory: the operands reach the
// FP64 pipe straight from the constant bank (uniform per CTA), so no thread widens 18 float32 constants per pixel and
// the CTA does not stage + barrier the view block.  Derived on the host from the float32 dfu_view (widening is exact;
// the two reciprocals are IEEE divisions = correctly rounded, which dfu_div_by requires).  Batches with more views than
// fit in the 64 KB bank run in chunks of S2_CVIEWS.
#define S2_CVIEWS 400
struct S2View { double cu, fu, tx, cv, fv, ty, rfu, rfv, m[3][4]; };   // m[j][k] = inv44[k][j]
static_assert(sizeof(S2View) * S2_CVIEWS <= 64000, "constant bank");
__constant__ S2View c_view[S2_CVIEWS];

// c_view is one per device and process: users on different streams (several contexts) are ordered through an event
// recorded behind the last kernel that read it, under a host mutex.
static std::mutex g_cview_mu;
static cudaEvent_t g_cview_free[64];

static void s2_view_consts(const dfu_view& w, S2View* o) {
  o->cu = (double)w.cu; o->fu = (double)w.fu; o->tx = (double)w.tx;
  o->cv = (double)w.cv; o->fv = (double)w.fv; o->ty = (double)w.ty;
  o->rfu = 1.0 / (double)w.fu; o->rfv = 1.0 / (double)w.fv;
  for (int j = 0; j < 3; j++)
    for (int k = 0; k < 4; k++) o->m[j][k] = (double)w.inv44[k * 4 + j];
}

// dfu_backproject (calibration_kitti.py:134-144, :89-102) on the constant-bank view: identical operation order
// pixel index -> (row, col) as float64 values (the back-projection needs them as float64 anyway): one conversion, one
// FMA, one floor, one FMA instead of an integer division by a run-time W and two conversions
__device__ __forceinline__ void s2_rowcol(const S2Prm& q, int pix, double& row, double& col) {
  const double p = (double)pix;
  row = floor(__fma_rn(p, q.rW, 0.5 * q.rW));
  col = __fma_rn(-row, (double)q.W, p);       // exact: integers below 2^53
}

__device__ __forceinline__ void s2_backproject(const S2View& c, double col, double row, float d, double& X, double& Y, double& Z) {
  const double dd = (double)d;
  const double xr = __dadd_rn(dfu_div_by(__dmul_rn(__dsub_rn(col, c.cu), dd), c.fu, c.rfu), c.tx);
  const double yr = __dadd_rn(dfu_div_by(__dmul_rn(__dsub_rn(row, c.cv), dd), c.fv, c.rfv), c.ty);
  X = __dadd_rn(__fma_rn(dd, c.m[0][2], __fma_rn(yr, c.m[0][1], __dmul_rn(xr, c.m[0][0]))), c.m[0][3]);
  Y = __dadd_rn(__fma_rn(dd, c.m[1][2], __fma_rn(yr, c.m[1][1], __dmul_rn(xr, c.m[1][0]))), c.m[1][3]);
  Z = __dadd_rn(__fma_rn(dd, c.m[2][2], __fma_rn(yr, c.m[2][1], __dmul_rn(xr, c.m[2][0]))), c.m[2][3]);
}

// to_sphere_coords (:166-180) + spconv v1 voxel coordinates with float32 constants (oracle/fakes.py), float64 as numpy
// evaluates them.  Returns the table key or KEY_INVALID; *miss: inside the voxel range but outside the table window.
__device__ __noinline__ unsigned s2_key_exact(const S2Prm& q, double X, double Y, double Z, bool& miss) {
  miss = false;
  double r = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(X, X), __dmul_rn(Y, Y)), __dmul_rn(Z, Z)));
  double theta = acos(__ddiv_rn(Z, r));
  double fan = atan(__ddiv_rn(Y, X));
  if (!(theta > q.min_theta)) return KEY_INVALID;
  const double vs = (double)0.002f;
  double cr = floor(__ddiv_rn(__dadd_rn(r, 100.0), 200.0));
  double ct = floor(__ddiv_rn(__dadd_rn(theta, 5.0), vs));
  double cf = floor(__ddiv_rn(__dadd_rn(fan, 5.0), vs));
  if (!(cr >= 0.0 && cr < 1.0 && ct >= 0.0 && ct < 5000.0 && cf >= 0.0 && cf < 5000.0)) return KEY_INVALID;
  int it = (int)ct - CT0, jf = (int)cf - CF0;
  if (it < 0 || it >= NT || jf < 0 || jf >= NF) {
    miss = true;
    return KEY_INVALID;
  }
  return (unsigned)(it * NF + jf);
}

// Pass A's key: the table key of the pixel's point, KEY_INVALID, or KEY_AMBIG when the float32 angles lie within
// S2_EPS of a bin edge (or of the theta gate) and the float64 evaluation has to decide (~0.4 % of the points: those go
// to a queue and are resolved densely by k_densify_amb instead of stalling 31 idle lanes of this warp).
// bit flags in st: 1 valid pixel, 2 passes height gate, 4 out-of-table (should not happen).
// atan2f(rho, z) equals acos(z / r) on [0, pi] and is well conditioned everywhere; CUDA's atan2f / atanf are accurate to
// 3 / 1 ulp, the float32 inputs carry <= 2 ulp: |error| < 1e-6 rad, half of S2_EPS.
__device__ __forceinline__ unsigned s2_key(const S2View& vw, const S2Prm& q, int pix, float d, double& X, double& Y,
                                           double& Z, unsigned& st) {
  st = 0;
  // depth[depth<0.001]=0; nonzero(); depth>0   (my_loader.py:507-509)
  if (!(d >= q.min_depth)) return KEY_INVALID;
  st |= 1u;
  double row, col;
  s2_rowcol(q, pix, row, col);
  s2_backproject(vw, col, row, d, X, Y, Z);
  if (!(Z < q.max_z)) return KEY_INVALID;  // new_p[:, 2] < 1.  (:540)
  st |= 2u;
  const float Xf = (float)X, Yf = (float)Y, Zf = (float)Z;
  const float rho = sqrtf(__fmaf_rn(Xf, Xf, Yf * Yf));
  const float rf = sqrtf(__fmaf_rn(rho, rho, Zf * Zf));
  const float th = atan2f(rho, Zf);
  const float fa = atanf(__fdividef(Yf, Xf));
  const double inv_vs = 1.0 / (double)0.002f;
  const double tt = ((double)th + 5.0) * inv_vs, tf = ((double)fa + 5.0) * inv_vs;
  const double ft = tt - floor(tt), ff = tf - floor(tf);
  const double D = (double)S2_EPS * inv_vs;
  const bool sure = (rf < 99.0f) && (rf > 1e-3f) && (fabsf(Xf) > 1e-6f) && (fabsf(th - (float)q.min_theta) > S2_EPS) &&
                    (ft > D) && (ft < 1.0 - D) && (ff > D) && (ff < 1.0 - D);
  if (!sure) return KEY_AMBIG;
  if (!(th > (float)q.min_theta)) return KEY_INVALID;
  const int it = (int)tt - CT0, jf = (int)tf - CF0;
  if (it < 0 || it >= NT || jf < 0 || jf >= NF) {
    st |= 4u;
    return KEY_INVALID;
  }
  return (unsigned)(it * NF + jf);
}

// stages the view constants in shared memory, with the correctly rounded reciprocals of fu and fv (dfu_div_by)
#define LOAD_VIEW(vw, views, v)                                            \
  __shared__ dfu_view vw;                                                  \
  __shared__ double vw##_rcp[2];                                           \
  for (int i_ = threadIdx.x; i_ < (int)(sizeof(dfu_view) / 4); i_ += blockDim.x) \
    ((uint32_t*)&vw)[i_] = ((const uint32_t*)((views) + (v)))[i_];       \
  if (threadIdx.x == 64) vw##_rcp[0] = __drcp_rn((double)__ldg(&(views)[v].fu)); \
  if (threadIdx.x == 96) vw##_rcp[1] = __drcp_rn((double)__ldg(&(views)[v].fv)); \
  __syncthreads();

// pass A: keys + table atomics.  No barrier: the view constants come from the constant bank and the per-warp counts
// (valid pixels | gated points << 10 | table misses << 20, accumulated over the CTA's tiles) go to warpstat, one word per
// warp and CTA, summed by k_blk_reduce.
// blockIdx.y = view inside the chunk that c_view currently holds; v0 = first view of the chunk.
#ifndef S2_TILES
#define S2_TILES 8
#endif
#ifndef S2_MINB
#define S2_MINB 6
#endif
static_assert(S2_TILES >= 1 && S2_TILES <= 31, "10-bit per-warp count fields");
/* S2_TILES 256-pixel tiles per CTA: the next tile's depth is in flight while this one is computed */
__global__ void __launch_bounds__(S2_THREADS, S2_MINB) k_densify_a(int v0, int bpv, S2Prm q, const float* __restrict__ depth,
                                                          unsigned* __restrict__ key, VoxCell* __restrict__ table,
                                                          unsigned* __restrict__ warpstat, uint2* __restrict__ ambq,
                                                          unsigned* __restrict__ ambcnt, unsigned capq) {
  const S2View& vw = c_view[blockIdx.y];
  const int v = v0 + blockIdx.y;
  const int npix = q.H * q.W;
  const int lane = threadIdx.x & 31;
  const float* dv = depth + (size_t)v * npix;
  int pix = blockIdx.x * (S2_TILES * S2_THREADS) + threadIdx.x;
  float dn = (pix < npix) ? __ldg(dv + pix) : 0.0f;
  unsigned acc = 0;
#pragma unroll 1
  for (int j = 0; j < S2_TILES; j++, pix += S2_THREADS) {
    const int blk = blockIdx.x * S2_TILES + j;
    if (blk >= bpv) break;
    const float d = dn;
    if (j + 1 < S2_TILES) dn = (pix + S2_THREADS < npix) ? __ldg(dv + pix + S2_THREADS) : 0.0f;
    unsigned st = 0, k = KEY_INVALID;
    unsigned long long oz = ~0ull, oy = ~0ull;
    if (pix < npix) {
      double X, Y, Z;
      k = s2_key(vw, q, pix, d, X, Y, Z, st);
      key[(size_t)v * npix + pix] = k;
      if (k < KEY_AMBIG) { oz = dfu_order_f64(Z); oy = dfu_order_f64(Y); }
    }
    {
      const unsigned amb = __ballot_sync(0xffffffffu, k == KEY_AMBIG);
      if (amb) {
        const int qi = (blockIdx.x + blockIdx.y) & (S2_NQ - 1);
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(ambcnt + qi, (unsigned)__popc(amb));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (k == KEY_AMBIG) {
          const unsigned slot = base + dfu_warp_excl(amb, lane);
          if (slot < capq) ambq[(size_t)qi * capq + slot] = make_uint2((unsigned)v, (unsigned)pix);
          k = KEY_INVALID;      // not part of any run below
        }
      }
    }
    // Neighbouring pixels of a row mostly fall into the same angular voxel (a voxel spans ~2.5 pixels), so each RUN of
    // equal keys inside the warp is reduced with shuffles and only its first lane touches the table.
    {
      const unsigned kprev = __shfl_up_sync(0xffffffffu, k, 1);
      const bool head = (lane == 0) || (k != kprev);
      const unsigned heads = __ballot_sync(0xffffffffu, head);
      const unsigned above = heads & ~((2u << lane) - 1u);
      const int end = above ? (__ffs(above) - 2) : 31;          // last lane of this lane's run
      // runs are short: stop as soon as no lane has a partner left
      for (int o = 1; o < 32 && __any_sync(0xffffffffu, lane + o <= end && k != KEY_INVALID); o <<= 1) {
        const unsigned long long z2 = __shfl_down_sync(0xffffffffu, oz, o), y2 = __shfl_down_sync(0xffffffffu, oy, o);
        if (lane + o <= end) { oz = z2 < oz ? z2 : oz; oy = y2 < oy ? y2 : oy; }
      }
      if (head && k != KEY_INVALID) {
        VoxCell* c = table + (size_t)v * NCELL + k;
        atomicMin(&c->first, (unsigned)pix);                    // pixels ascend inside a run: the head is its minimum
        atomicAdd(&c->cnt, (unsigned)(end - lane + 1));
        atomicMin(&c->minz, oz);
        atomicMin(&c->miny, oy);
      }
    }
    acc += (st & 1u) | ((st & 2u) << 9) | ((st & 4u) << 18);      // per-lane counts, 10-bit fields (<= S2_TILES each)
  }
  // one word per warp and CTA, stored in the slot of the CTA's first tile: valid | gated << 10 | misses << 20
  const unsigned a1 = __reduce_add_sync(0xffffffffu, acc & 1023u), a2 = __reduce_add_sync(0xffffffffu, (acc >> 10) & 1023u),
                 a4 = __reduce_add_sync(0xffffffffu, (acc >> 20) & 1023u);
  if (lane == 0)
    warpstat[((size_t)v * bpv + (size_t)blockIdx.x * S2_TILES) * (S2_THREADS / 32) + (threadIdx.x >> 5)] = a1 | (a2 << 10) | (a4 << 20);
}

// The ambiguous pixels of pass A, one thread each: exact float64 key, key array patched, table atomics.  Normally they
// come from the sub-queues; if any sub-queue overflowed (pathological input), the key array is scanned for KEY_AMBIG
// instead.  grid (S2_NQ, S2_AMB_Y).
#define S2_AMB_Y 16
__device__ __forceinline__ void s2_resolve_amb(const dfu_view* __restrict__ views, const S2Prm& q, int v, int pix, int npix,
                                               const float* __restrict__ depth, unsigned* __restrict__ key,
                                               VoxCell* __restrict__ table, unsigned long long* __restrict__ stats) {
  const int row = pix / q.W, col = pix - row * q.W;
  double X, Y, Z;
  dfu_backproject(views[v], col, row, __ldg(depth + (size_t)v * npix + pix), X, Y, Z);
  bool miss;
  const unsigned k = s2_key_exact(q, X, Y, Z, miss);
  key[(size_t)v * npix + pix] = k;
  if (miss) atomicAdd(stats + v * 8 + 4, 1ull);
  if (k != KEY_INVALID) {
    VoxCell* c = table + (size_t)v * NCELL + k;
    atomicMin(&c->first, (unsigned)pix);
    atomicAdd(&c->cnt, 1u);
    atomicMin(&c->minz, dfu_order_f64(Z));
    atomicMin(&c->miny, dfu_order_f64(Y));
  }
}

__global__ void __launch_bounds__(256) k_densify_amb(const dfu_view* __restrict__ views, S2Prm q, int V,
                                                     const float* __restrict__ depth, unsigned* __restrict__ key,
                                                     VoxCell* __restrict__ table, const uint2* __restrict__ ambq,
                                                     const unsigned* __restrict__ ambcnt, unsigned capq,
                                                     unsigned long long* __restrict__ stats) {
  const int npix = q.H * q.W;
  const int over = __syncthreads_or(threadIdx.x < S2_NQ && ambcnt[threadIdx.x] > capq);
  if (!over) {
    const unsigned n = ambcnt[blockIdx.x];
    for (unsigned i = blockIdx.y * 256 + threadIdx.x; i < n; i += S2_AMB_Y * 256) {
      const uint2 e = ambq[(size_t)blockIdx.x * capq + i];
      s2_resolve_amb(views, q, (int)e.x, (int)e.y, npix, depth, key, table, stats);
    }
  } else {
    const int64_t total = (int64_t)V * npix;
    for (int64_t g = ((int64_t)blockIdx.y * S2_NQ + blockIdx.x) * 256 + threadIdx.x; g < total; g += (int64_t)S2_NQ * S2_AMB_Y * 256)
      if (key[g] == KEY_AMBIG) s2_resolve_amb(views, q, (int)(g / npix), (int)(g % npix), npix, depth, key, table, stats);
  }
}

// pass B: arg-min resolution and the bitmap of voxel-opening pixels (one ballot per warp).  The voxel ordinal is the
// rank of the opening pixel in (view, pixel) order = exclusive scan of the block counts (k_blk_reduce) + the rank
// inside the bitmap, so no further pass over the pixels is needed: pass C runs one thread per VOXEL.  No barrier.
__global__ void __launch_bounds__(S2_THREADS) k_densify_b(int v0, int bpv, S2Prm q, const float* __restrict__ depth,
                                                          const unsigned* __restrict__ key, VoxCell* __restrict__ table,
                                                          unsigned* __restrict__ blkbits,
                                                          unsigned long long* __restrict__ stats, uint2* __restrict__ ovf,
                                                          unsigned ovf_cap, unsigned long long* __restrict__ ovf_cnt) {
  const S2View& vw = c_view[blockIdx.y];
  const int v = v0 + blockIdx.y;
  const int npix = q.H * q.W;
  const unsigned* kv = key + (size_t)v * npix;
  int pix = blockIdx.x * (S2_TILES * S2_THREADS) + threadIdx.x;
  unsigned kn = (pix < npix) ? __ldg(kv + pix) : KEY_INVALID;
#pragma unroll 1
  for (int j = 0; j < S2_TILES; j++, pix += S2_THREADS) {
    const int blk = blockIdx.x * S2_TILES + j;
    if (blk >= bpv) break;
    const unsigned k = kn;
    if (j + 1 < S2_TILES) kn = (pix + S2_THREADS < npix) ? __ldg(kv + pix + S2_THREADS) : KEY_INVALID;
    bool opener = false;
    if (k != KEY_INVALID) {
      VoxCell* c = table + (size_t)v * NCELL + k;
      // minz / miny / first / cnt are final after pass A: the whole cell is fetched up front (two 128-bit loads, in
      // flight together with the depth load) instead of field by field between the atomics
      const ulonglong2 mm = *reinterpret_cast<const ulonglong2*>(c);            // minz, miny
      const uint2 fc = *reinterpret_cast<const uint2*>(&c->first);              // first, cnt
      const float d = __ldg(depth + (size_t)v * npix + pix);
      double row, col, X, Y, Z;
      s2_rowcol(q, pix, row, col);
      s2_backproject(vw, col, row, d, X, Y, Z);
      if (dfu_order_f64(Z) == mm.x) atomicMin(&c->argz, (unsigned)pix);
      if (dfu_order_f64(Y) == mm.y) atomicMin(&c->argy, (unsigned)pix);
      opener = (fc.x == (unsigned)pix);
      if (opener && (fc.y + 1u > (unsigned)q.max_per_voxel)) {
        // more points than the voxeliser keeps (my_loader.py:73): listed for k_vox_overflow, which redoes the two
        // arg-mins over the FIRST max_per_voxel points in input order
        atomicAdd(stats + v * 8 + 3, 1ull);
        const unsigned long long slot = atomicAdd(ovf_cnt, 1ull);
        if (slot < ovf_cap) ovf[slot] = make_uint2((unsigned)v, k);
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, opener);
    if ((threadIdx.x & 31) == 0) blkbits[((size_t)v * bpv + blk) * (S2_THREADS / 32) + (threadIdx.x >> 5)] = bal;
  }
}

// Voxels with more than max_per_voxel points (rare: surfaces seen edge-on from the LiDAR origin but face-on from the
// camera).  The reference's voxeliser keeps the first max_per_voxel points of a voxel in INPUT order and la_sampling2 / 20
// take the arg-min over those only (my_loader.py:69-83,253-257,268-272).  Input order = pixel order, so: one CTA per listed
// voxel scans the view's keys forward from the voxel's opening pixel, collects its first max_per_voxel pixels in
// order, back-projects them exactly and replaces the cell's two arg-mins (first minimum in that order).
#define S2_OVF_MAX 256
__global__ void __launch_bounds__(256) k_vox_overflow(const dfu_view* __restrict__ views, S2Prm q, const float* __restrict__ depth,
                                                      const unsigned* __restrict__ key, VoxCell* __restrict__ table,
                                                      const uint2* __restrict__ ovf, unsigned ovf_cap,
                                                      const unsigned long long* __restrict__ ovf_cnt) {
  __shared__ unsigned s_pix[S2_OVF_MAX];
  __shared__ unsigned long long s_val[2][S2_OVF_MAX];
  __shared__ int s_wsum[8], s_found;
  const int npix = q.H * q.W;
  const unsigned long long total = *ovf_cnt;
  const unsigned n = total < ovf_cap ? (unsigned)total : ovf_cap;
  const int keep = q.max_per_voxel < S2_OVF_MAX ? q.max_per_voxel : S2_OVF_MAX;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (unsigned e = blockIdx.x; e < n; e += gridDim.x) {
    const int v = (int)ovf[e].x;
    const unsigned k = ovf[e].y;
    VoxCell* c = table + (size_t)v * NCELL + k;
    const unsigned* kv = key + (size_t)v * npix;
    __syncthreads();
    if (threadIdx.x == 0) s_found = 0;
    __syncthreads();
    for (int p0 = (int)c->first; p0 < npix; p0 += 256) {
      const int pix = p0 + threadIdx.x;
      const bool m = pix < npix && __ldg(kv + pix) == k;
      const unsigned bal = __ballot_sync(0xffffffffu, m);
      if (lane == 0) s_wsum[warp] = __popc(bal);
      __syncthreads();
      int base = s_found, tot = 0;
      for (int w = 0; w < 8; w++) { if (w < warp) base += s_wsum[w]; tot += s_wsum[w]; }
      const int pos = base + __popc(bal & ((1u << lane) - 1u));
      if (m && pos < keep) s_pix[pos] = (unsigned)pix;
      __syncthreads();
      if (threadIdx.x == 0) s_found += tot;
      __syncthreads();
      if (s_found >= keep) break;
    }
    const int nf = s_found < keep ? s_found : keep;
    // exact back-projection of the kept pixels; first minimum in input order = smallest (value, position)
    if ((int)threadIdx.x < S2_OVF_MAX) {
      unsigned long long vz = ~0ull, vy = ~0ull;
      if ((int)threadIdx.x < nf) {
        const int pix = (int)s_pix[threadIdx.x], row = pix / q.W, col = pix - row * q.W;
        double X, Y, Z;
        dfu_backproject(views[v], col, row, __ldg(depth + (size_t)v * npix + pix), X, Y, Z);
        vz = dfu_order_f64(Z); vy = dfu_order_f64(Y);
      }
      s_val[0][threadIdx.x] = vz; s_val[1][threadIdx.x] = vy;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
      const unsigned long long* a = s_val[threadIdx.x];
      int best = 0;
      for (int i = 1; i < nf; i++)
        if (a[i] < a[best]) best = i;                 // strict: the first minimum stays
      if (threadIdx.x == 0) c->argz = s_pix[best]; else c->argy = s_pix[best];
    }
  }
}

// one thread per pixel block: openers of the block (popcount of its bitmap words) for the scan, and the per-view sums
// of pass A's per-warp counts.  stats int64 [V][8]: 0 valid pixels, 1 gated points, 3 overflowing voxels, 4 table misses
__global__ void __launch_bounds__(256) k_blk_reduce(int bpv, const unsigned* __restrict__ blkbits,
                                                    const unsigned* __restrict__ warpstat, int32_t* __restrict__ blkcnt,
                                                    unsigned long long* __restrict__ stats) {
  const int v = blockIdx.y;
  const int b = blockIdx.x * 256 + threadIdx.x;
  unsigned n1 = 0, n2 = 0, n4 = 0;
  if (b < bpv) {
    const uint4* bb = reinterpret_cast<const uint4*>(blkbits + ((size_t)v * bpv + b) * (S2_THREADS / 32));
    const uint4* ws = reinterpret_cast<const uint4*>(warpstat + ((size_t)v * bpv + b) * (S2_THREADS / 32));
    int n = 0;
#pragma unroll
    for (int h = 0; h < S2_THREADS / 128; h++) {
      const uint4 m = __ldg(bb + h);
      n += __popc(m.x) + __popc(m.y) + __popc(m.z) + __popc(m.w);
      if (b % S2_TILES == 0) {      // pass A stores its per-warp counts in the slot of each CTA's first tile
        const uint4 w = __ldg(ws + h);
        const unsigned ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int j = 0; j < 4; j++) { n1 += ww[j] & 1023u; n2 += (ww[j] >> 10) & 1023u; n4 += (ww[j] >> 20) & 1023u; }
      }
    }
    blkcnt[(size_t)v * bpv + b] = n;
  }
  __shared__ unsigned sacc[3];
  if (threadIdx.x < 3) sacc[threadIdx.x] = 0;
  __syncthreads();
  n1 = __reduce_add_sync(0xffffffffu, n1); n2 = __reduce_add_sync(0xffffffffu, n2); n4 = __reduce_add_sync(0xffffffffu, n4);
  if ((threadIdx.x & 31) == 0) { atomicAdd(&sacc[0], n1); atomicAdd(&sacc[1], n2); atomicAdd(&sacc[2], n4); }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (sacc[0]) atomicAdd(stats + v * 8 + 0, (unsigned long long)sacc[0]);
    if (sacc[1]) atomicAdd(stats + v * 8 + 1, (unsigned long long)sacc[1]);
    if (sacc[2]) atomicAdd(stats + v * 8 + 4, (unsigned long long)sacc[2]);
  }
}

// pass C, one thread per VOXEL: find the opening pixel (block by binary search in the scanned counts, lane by rank in
// the bitmap), read the representatives out of its cell, put the cell back to all-ones (so that the next batch
// starts from a clean table without a 41 MB-per-view memset), and gather the min-y representative's xyz + label word
// for the instance stage.
// owner[first voxel of pixel block b] = b for every block that opens a voxel; an inclusive max-scan then gives every
// voxel ordinal its pixel block (blkoff is monotone), replacing two dependent binary searches per voxel
__global__ void __launch_bounds__(256) k_vox_owner(int64_t nblocks, const int32_t* __restrict__ blkcnt,
                                                   const int64_t* __restrict__ blkoff, int32_t* __restrict__ owner) {
  const int64_t b = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (b < nblocks && blkcnt[b] > 0) owner[blkoff[b]] = (int32_t)b;
}
struct S2MaxOp { __device__ __forceinline__ int32_t operator()(int32_t a, int32_t b) const { return a > b ? a : b; } };

__global__ void __launch_bounds__(256) k_vox_finish(const dfu_view* __restrict__ views, S2Prm q, int V, int bpv,
                                                    const int32_t* __restrict__ owner, const int64_t* __restrict__ blkoff,
                                                    const unsigned* __restrict__ blkbits, const unsigned* __restrict__ key,
                                                    const float* __restrict__ depth, const uint32_t* __restrict__ labelmap,
                                                    VoxCell* __restrict__ table, int32_t* __restrict__ repz,
                                                    int32_t* __restrict__ repy, double* __restrict__ repy_xyz,
                                                    uint32_t* __restrict__ repy_label, int64_t total) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= total) return;
  // the pixel block that opens voxel t (k_vox_owner + max-scan): no search
  const int64_t blo = __ldg(owner + t);
  const int v = (int)(blo / bpv);
  int r = (int)(t - __ldg(blkoff + blo));
  const unsigned* bits = blkbits + (size_t)blo * (S2_THREADS / 32);
  int w = 0;
  unsigned m = __ldg(bits);
  while (r >= __popc(m)) { r -= __popc(m); w++; m = __ldg(bits + w); }
  const int lane = (int)__fns(m, 0, r + 1);
  const size_t npix = (size_t)q.H * q.W;
  const int opix = (int)(blo - (int64_t)v * bpv) * S2_THREADS + w * 32 + lane;
  VoxCell* c = table + (size_t)v * NCELL + __ldg(key + v * npix + opix);
  const int pz = (int)c->argz, py = (int)c->argy;
  uint4* w4 = reinterpret_cast<uint4*>(c);
  const uint4 ones = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
  w4[0] = ones; w4[1] = ones;
  repz[t] = pz;
  repy[t] = py;
  const dfu_view& vw = views[v];
  const int row = py / q.W, col = py - row * q.W;
  double X, Y, Z;
  dfu_backproject(vw, col, row, __ldg(depth + v * npix + py), X, Y, Z);
  repy_xyz[t] = X;
  repy_xyz[total + t] = Y;
  repy_xyz[2 * total + t] = Z;
  repy_label[t] = labelmap ? __ldg(labelmap + ((size_t)v * q.Hm + row) * q.Wm + col) : 0u;
}

// The voxeliser's max_num_voxels (my_loader.py:74): a view keeps its first `maxv` angular voxels in first-appearance order
// (= the order of the voxel lists), the points of later voxels are dropped.  Rare (the table has 1.29 M cells against
// the default limit of 10^6), so the lists are simply re-packed: thread per KEPT voxel, out of place.
__global__ void __launch_bounds__(256) k_vox_cut(int V, const int64_t* __restrict__ old_off, const int64_t* __restrict__ new_off,
                                                  int64_t old_total, int64_t new_total, const int32_t* __restrict__ repz,
                                                  const int32_t* __restrict__ repy, const double* __restrict__ repy_xyz,
                                                  const uint32_t* __restrict__ repy_label, int32_t* __restrict__ oz,
                                                  int32_t* __restrict__ oy, double* __restrict__ oxyz, uint32_t* __restrict__ ol) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= new_total) return;
  int lo = 0, hi = V;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (new_off[mid] <= t) lo = mid; else hi = mid;
  }
  const int64_t src = old_off[lo] + (t - new_off[lo]);
  oz[t] = repz[src];
  oy[t] = repy[src];
  ol[t] = repy_label[src];
  oxyz[t] = repy_xyz[src];
  oxyz[new_total + t] = repy_xyz[old_total + src];
  oxyz[2 * new_total + t] = repy_xyz[2 * old_total + src];
}

__global__ void k_vox_view_offsets(const int64_t* __restrict__ blkoff, int V, int bpv, int64_t* out) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v <= V) out[v] = blkoff[(int64_t)v * bpv];
}

// gathers per voxel: min-y representative xyz (float64 SoA: x[n] y[n] z[n]) + label word, optional min-z xyz
__global__ void __launch_bounds__(256) k_vox_gather(const dfu_view* __restrict__ views, S2Prm q, int V,
                                                    const int64_t* __restrict__ voxoff, const float* __restrict__ depth,
                                                    const uint32_t* __restrict__ labelmap,
                                                    const int32_t* __restrict__ repz, const int32_t* __restrict__ repy,
                                                    double* __restrict__ repy_xyz, uint32_t* __restrict__ repy_label,
                                                    double* __restrict__ repz_xyz, int64_t total) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= total) return;
  // binary search the view
  int lo = 0, hi = V;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (voxoff[mid] <= t) lo = mid; else hi = mid;
  }
  const int v = lo;
  const dfu_view& vw = views[v];
  const size_t npix = (size_t)q.H * q.W;
  {
    int pix = repy[t];
    int row = pix / q.W, col = pix - row * q.W;
    double X, Y, Z;
    dfu_backproject(vw, col, row, __ldg(depth + v * npix + pix), X, Y, Z);
    repy_xyz[t] = X;
    repy_xyz[total + t] = Y;
    repy_xyz[2 * total + t] = Z;
    repy_label[t] = labelmap ? __ldg(labelmap + ((size_t)v * q.Hm + row) * q.Wm + col) : 0u;
  }
  if (repz_xyz) {
    int pix = repz[t];
    int row = pix / q.W, col = pix - row * q.W;
    double X, Y, Z;
    dfu_backproject(vw, col, row, __ldg(depth + v * npix + pix), X, Y, Z);
    repz_xyz[t] = X;
    repz_xyz[total + t] = Y;
    repz_xyz[2 * total + t] = Z;
  }
}

// ---------------------------------------------------------------- cloud emit (my_loader.py:611-617, vis_utils.py:164)
template <typename T> struct Cvt;
template <> struct Cvt<double> { static __device__ __forceinline__ double f(double v) { return v; } };
template <> struct Cvt<float>  { static __device__ __forceinline__ float f(double v) { return (float)v; } };
template <> struct Cvt<__half> { static __device__ __forceinline__ __half f(double v) { return __double2half(v); } };

// one row = 8 values: written with 128-bit stores (cudaMalloc'd / torch buffers are 256 B aligned, rows are 16 / 32 / 64 B)
__device__ __forceinline__ void emit_row(__half* dst, const double* o) {
  __half2 h[4];
#pragma unroll
  for (int j = 0; j < 4; j++) h[j] = __halves2half2(__double2half(o[2 * j]), __double2half(o[2 * j + 1]));
  *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(h);
}
__device__ __forceinline__ void emit_row(float* dst, const double* o) {
  reinterpret_cast<float4*>(dst)[0] = make_float4((float)o[0], (float)o[1], (float)o[2], (float)o[3]);
  reinterpret_cast<float4*>(dst)[1] = make_float4((float)o[4], (float)o[5], (float)o[6], (float)o[7]);
}
__device__ __forceinline__ void emit_row(double* dst, const double* o) {
#pragma unroll
  for (int j = 0; j < 4; j++) reinterpret_cast<double2*>(dst)[j] = make_double2(o[2 * j], o[2 * j + 1]);
}

template <typename T>
__global__ void __launch_bounds__(256) k_emit_cloud(const dfu_view* __restrict__ views, S2Prm q, int V,
                                                    const int64_t* __restrict__ rowoff /*[V+1]*/,
                                                    const int64_t* __restrict__ foff /*[V+1]*/,
                                                    const int64_t* __restrict__ voxoff /*[V+1]*/,
                                                    const float4* __restrict__ lidar, const int32_t* __restrict__ flist,
                                                    const float* __restrict__ depth, const void* __restrict__ image,
                                                    const int32_t* __restrict__ repz, T* __restrict__ out, int64_t total) {
  const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= total) return;
  int lo = 0, hi = V;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (rowoff[mid] <= t) lo = mid; else hi = mid;
  }
  const int v = lo;
  const dfu_view& vw = views[v];
  const int64_t r = t - rowoff[v];
  const int64_t nF = foff[v + 1] - foff[v];
  double o[8];
  if (r < nF) {
    // new_lidar (:611-614): float32 values widened to float64, intensity * 10 in float64
    float4 p = __ldg(lidar + vw.lidar_off + flist[foff[v] + r]);
    o[0] = (double)p.x; o[1] = (double)p.y; o[2] = (double)p.z; o[3] = __dmul_rn((double)p.w, 10.0);
    o[4] = 0; o[5] = 0; o[6] = 0; o[7] = 2.0;
  } else {
    const int pix = repz[voxoff[v] + (r - nF)];
    const int row = pix / q.W, col = pix - row * q.W;
    const size_t npix = (size_t)q.H * q.W;
    double X, Y, Z;
    dfu_backproject(vw, col, row, __ldg(depth + v * npix + pix), X, Y, Z);
    o[0] = X; o[1] = Y; o[2] = Z; o[3] = 0;
    const size_t ip = (((size_t)v * q.Hm + row) * q.Wm + col) * 3;
    int R, G, B;
    if (image == nullptr) { R = G = B = 0; }
    else if (q.image_is_i32) { const int* im = (const int*)image; R = im[ip]; G = im[ip + 1]; B = im[ip + 2]; }
    else { const uint8_t* im = (const uint8_t*)image; R = im[ip]; G = im[ip + 1]; B = im[ip + 2]; }
    const double third = 0x1.5555555555555p-2;   // RN(1/3)
    o[4] = dfu_div_by((double)R, 3.0, third); o[5] = dfu_div_by((double)G, 3.0, third);   // image/3 (:539)
    o[6] = dfu_div_by((double)B, 3.0, third);
    o[7] = 1.0;
  }
  emit_row(out + t * 8, o);
}

// ---------------------------------------------------------------- C ABI
extern "C" int dfu_fuse_densify_sync(dfu_ctx* ctx, int64_t* counts_host, int allow_overflow) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !(ctx->stage & 1)) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_densify_sync: no batch bound");
  const int V = ctx->V;
  cudaStream_t st = ctx->stream;
  S2Prm q = make_s2(&ctx->prm);
  const int64_t npix = (int64_t)q.H * q.W;
  if (npix >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_ARG, "depth map too large");
  const int bpv = (int)((npix + S2_THREADS - 1) / S2_THREADS);
  ctx->bpv_pix = bpv;
  dfu_view* vd = (dfu_view*)ctx->slot[SL_VIEWS];
  DFU_SLOT(ctx, key, unsigned*, SL_KEY, V * npix * sizeof(unsigned));
  DFU_SLOT(ctx, table, VoxCell*, SL_TABLE, (size_t)V * NCELL * sizeof(VoxCell));
  DFU_SLOT(ctx, stats, unsigned long long*, SL_VOXCNT, (V * 8 + 8) * sizeof(unsigned long long));   // [V][8] + the overflow-list counter
  DFU_SLOT(ctx, blkcnt, int32_t*, SL_BLKCNT, (int64_t)V * bpv * sizeof(int32_t));
  DFU_SLOT(ctx, blkoff, int64_t*, SL_BLKOFF, ((int64_t)V * bpv + 1) * sizeof(int64_t));
  DFU_SLOT(ctx, blkbits, unsigned*, SL_VOXCELL, (size_t)V * bpv * (S2_THREADS / 32) * sizeof(unsigned));
  DFU_SLOT(ctx, voxoff, int64_t*, SL_VOXOFF, (V + 1) * sizeof(int64_t));
  DFU_SLOT(ctx, warpstat, unsigned*, SL_WARPSTAT, (size_t)V * bpv * (S2_THREADS / 32) * sizeof(unsigned));
  // queue of pixels whose float32 angles are too close to a bin edge: 1/64 of the pixels (8x the expected 0.2 %)
  unsigned capq = (unsigned)(((int64_t)V * npix / 64 + S2_NQ - 1) / S2_NQ + 64);
  if (const char* e = getenv("DFU3D_S2_CAPQ")) capq = (unsigned)atoi(e);   // test hook: forces the overflow (scan) path
  DFU_SLOT(ctx, ambq, uint2*, SL_AMBQ, (size_t)S2_NQ * (capq + 1) * sizeof(uint2) + S2_NQ * sizeof(unsigned));
  unsigned* ambcnt = (unsigned*)(ambq + (size_t)S2_NQ * capq);
  const unsigned ovf_cap = (unsigned)S2_NQ * capq;      // pass B re-uses the (by then consumed) queue as the list of overflowing voxels
  {
    DFU_PROF(ctx, st, PT_MISC);
    const size_t tbytes = (size_t)V * NCELL * sizeof(VoxCell);
    if (ctx->table_clean_ptr != (void*)table || ctx->table_clean_bytes < tbytes) {
      // first use, or the slot was re-allocated: the whole allocation becomes the clean region
      DFU_CUDA(ctx, cudaMemsetAsync(table, 0xFF, ctx->slot_bytes[SL_TABLE], st));
    }
    ctx->table_clean_ptr = nullptr;   // dirty until pass C has been queued
    ctx->table_clean_bytes = 0;
    DFU_CUDA(ctx, cudaMemsetAsync(stats, 0, (V * 8 + 8) * sizeof(unsigned long long), st));
    DFU_CUDA(ctx, cudaMemsetAsync(ambcnt, 0, S2_NQ * sizeof(unsigned), st));
  }
  // float64 view constants (host copy owned by the context: it must outlive the asynchronous symbol copies)
  ctx->s2_consts = realloc(ctx->s2_consts, sizeof(S2View) * (size_t)V);
  if (!ctx->s2_consts) return dfu_set_err(ctx, DFU_ERR_CUDA, "out of host memory");
  S2View* hc = (S2View*)ctx->s2_consts;
  for (int v = 0; v < V; v++) s2_view_consts(ctx->views_host[v], hc + v);
  for (int v0 = 0; v0 < V; v0 += S2_CVIEWS) {
    const int nv = V - v0 < S2_CVIEWS ? V - v0 : S2_CVIEWS;
    dim3 grid((bpv + S2_TILES - 1) / S2_TILES, nv);
    std::lock_guard<std::mutex> lock(g_cview_mu);
    cudaEvent_t& ev = g_cview_free[ctx->device & 63];
    if (!ev) DFU_CUDA(ctx, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    DFU_CUDA(ctx, cudaStreamWaitEvent(st, ev, 0));
    DFU_CUDA(ctx, cudaMemcpyToSymbolAsync(c_view, hc + v0, sizeof(S2View) * (size_t)nv, 0, cudaMemcpyHostToDevice, st));
    {
      DFU_PROF(ctx, st, PT_DENSIFY_A);
      k_densify_a<<<grid, S2_THREADS, 0, st>>>(v0, bpv, q, ctx->depth, key, table, warpstat, ambq, ambcnt, capq);
      DFU_LAUNCH_CHECK(ctx);
    }
    if (v0 + nv >= V) {   // all keys written: resolve the queued pixels before the first pass B
      DFU_PROF(ctx, st, PT_DENSIFY_A);
      k_densify_amb<<<dim3(S2_NQ, S2_AMB_Y), 256, 0, st>>>(vd, q, V, ctx->depth, key, table, ambq, ambcnt, capq, stats);
      DFU_LAUNCH_CHECK(ctx);
    }
    if (V <= S2_CVIEWS) {
      DFU_PROF(ctx, st, PT_DENSIFY_B);
      k_densify_b<<<grid, S2_THREADS, 0, st>>>(v0, bpv, q, ctx->depth, key, table, blkbits, stats, ambq, ovf_cap, stats + V * 8);
      DFU_LAUNCH_CHECK(ctx);
    }
    DFU_CUDA(ctx, cudaEventRecord(ev, st));
  }
  for (int v0 = 0; V > S2_CVIEWS && v0 < V; v0 += S2_CVIEWS) {   // several chunks: pass B needs its constants again
    const int nv = V - v0 < S2_CVIEWS ? V - v0 : S2_CVIEWS;
    dim3 grid((bpv + S2_TILES - 1) / S2_TILES, nv);
    std::lock_guard<std::mutex> lock(g_cview_mu);
    cudaEvent_t& ev = g_cview_free[ctx->device & 63];
    DFU_CUDA(ctx, cudaStreamWaitEvent(st, ev, 0));
    DFU_CUDA(ctx, cudaMemcpyToSymbolAsync(c_view, hc + v0, sizeof(S2View) * (size_t)nv, 0, cudaMemcpyHostToDevice, st));
    DFU_PROF(ctx, st, PT_DENSIFY_B);
    k_densify_b<<<grid, S2_THREADS, 0, st>>>(v0, bpv, q, ctx->depth, key, table, blkbits, stats, ambq, ovf_cap, stats + V * 8);
    DFU_LAUNCH_CHECK(ctx);
    DFU_CUDA(ctx, cudaEventRecord(ev, st));
  }
  {
    DFU_PROF(ctx, st, PT_DENSIFY_B);
    k_vox_overflow<<<148 * 4, 256, 0, st>>>(vd, q, ctx->depth, key, table, ambq, ovf_cap, stats + V * 8);
    DFU_LAUNCH_CHECK(ctx);
  }
  {
    DFU_PROF(ctx, st, PT_MISC);
    k_blk_reduce<<<dim3((bpv + 255) / 256, V), 256, 0, st>>>(bpv, blkbits, warpstat, blkcnt, stats);
    DFU_LAUNCH_CHECK(ctx);
    int rc = dfu_exclusive_sum_i32(ctx, st, blkcnt, blkoff, (int64_t)V * bpv);
    if (rc) return rc;
    k_vox_view_offsets<<<(V + 1 + 127) / 128, 128, 0, st>>>(blkoff, V, bpv, voxoff);
    DFU_LAUNCH_CHECK(ctx);
  }
  // read back offsets + stats (one sync)
  size_t nb = (V + 1) * sizeof(int64_t) + (V * 8 + 8) * sizeof(unsigned long long);
  char* pin = (char*)dfu_pinned(ctx, nb);
  if (!pin) return DFU_ERR_CUDA;
  DFU_CUDA(ctx, cudaMemcpyAsync(pin, voxoff, (V + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(pin + (V + 1) * sizeof(int64_t), stats, (V * 8 + 8) * sizeof(unsigned long long),
                                cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  memcpy(ctx->off_vox, pin, (V + 1) * sizeof(int64_t));
  const unsigned long long* hs = (const unsigned long long*)(pin + (V + 1) * sizeof(int64_t));
  int64_t total_over = 0, total_miss = 0;
  for (int v = 0; v < V; v++) {
    ctx->cnt_vox[v] = ctx->off_vox[v + 1] - ctx->off_vox[v];
    if (counts_host) {
      counts_host[v * 4 + 0] = (int64_t)hs[v * 8 + 0];
      counts_host[v * 4 + 1] = (int64_t)hs[v * 8 + 1];
      counts_host[v * 4 + 2] = ctx->cnt_vox[v];
      counts_host[v * 4 + 3] = (int64_t)hs[v * 8 + 3];
    }
    total_over += (int64_t)hs[v * 8 + 3];
    total_miss += (int64_t)hs[v * 8 + 4];
  }
  const int64_t nvox = ctx->off_vox[V];
  const int64_t cap = nvox > 0 ? nvox : 1;
  DFU_SLOT(ctx, repz, int32_t*, SL_REPZ, cap * sizeof(int32_t));
  DFU_SLOT(ctx, repy, int32_t*, SL_REPY, cap * sizeof(int32_t));
  DFU_SLOT(ctx, repy_xyz, double*, SL_REPY_XYZ, cap * 3 * sizeof(double));
  DFU_SLOT(ctx, repy_label, uint32_t*, SL_REPY_LABEL, cap * sizeof(uint32_t));
  if (nvox > 0) {
    DFU_PROF(ctx, st, PT_DENSIFY_C);
    const int64_t nblocks = (int64_t)V * bpv;
    DFU_SLOT(ctx, owner, int32_t*, SL_WARPSTAT, (size_t)nvox * sizeof(int32_t));   // pass A's per-warp counts are consumed
    DFU_CUDA(ctx, cudaMemsetAsync(owner, 0, (size_t)nvox * sizeof(int32_t), st));
    k_vox_owner<<<(unsigned)((nblocks + 255) / 256), 256, 0, st>>>(nblocks, blkcnt, blkoff, owner);
    DFU_LAUNCH_CHECK(ctx);
    size_t tb = 0;
    DFU_CUDA(ctx, cub::DeviceScan::InclusiveScan(nullptr, tb, owner, owner, S2MaxOp(), nvox, st));
    DFU_SLOT(ctx, tmp, void*, SL_CUBTMP, tb);
    DFU_CUDA(ctx, cub::DeviceScan::InclusiveScan(tmp, tb, owner, owner, S2MaxOp(), nvox, st));
    ctx->launches += 2;
    k_vox_finish<<<(unsigned)((nvox + 255) / 256), 256, 0, st>>>(vd, q, V, bpv, owner, blkoff, blkbits, key, ctx->depth,
                                                                 ctx->labelmap, table, repz, repy, repy_xyz, repy_label, nvox);
    DFU_LAUNCH_CHECK(ctx);
  }
  // ---- max_num_voxels: keep the first max_voxels voxels of every view
  {
    const int64_t maxv = ctx->prm.max_voxels > 0 ? (int64_t)ctx->prm.max_voxels : INT64_MAX;
    bool cut = false;
    for (int v = 0; v < V; v++) cut = cut || ctx->cnt_vox[v] > maxv;
    if (cut) {
      int64_t* noff_h = (int64_t*)malloc((size_t)(V + 1) * 8);
      if (!noff_h) return dfu_set_err(ctx, DFU_ERR_CUDA, "out of host memory");
      int64_t tot = 0;
      for (int v = 0; v < V; v++) {
        noff_h[v] = tot;
        tot += ctx->cnt_vox[v] < maxv ? ctx->cnt_vox[v] : maxv;
      }
      noff_h[V] = tot;
      // scratch: new offsets | repz | repy | label | xyz   (the scans' temporary storage is free by now)
      const size_t offb = ((size_t)(V + 1) * 8 + 255) & ~(size_t)255;
      char* tmpc = (char*)dfu_slot(ctx, SL_CUBTMP, offb + (size_t)tot * 36 + 1024);
      if (!tmpc) { free(noff_h); return DFU_ERR_CUDA; }
      int64_t* noff = (int64_t*)tmpc;
      double* oxyz = (double*)(tmpc + offb);
      int32_t* oz = (int32_t*)(oxyz + 3 * tot);
      int32_t* oy = oz + tot;
      uint32_t* ol = (uint32_t*)(oy + tot);
      cudaError_t ce = cudaMemcpyAsync(noff, noff_h, (size_t)(V + 1) * 8, cudaMemcpyHostToDevice, st);   // pageable: staged before return
      if (ce == cudaSuccess && tot > 0) {
        k_vox_cut<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(V, voxoff, noff, nvox, tot, repz, repy, repy_xyz, repy_label, oz, oy, oxyz, ol);
        ctx->launches++;
        ce = cudaGetLastError();
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(repz, oz, (size_t)tot * 4, cudaMemcpyDeviceToDevice, st);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(repy, oy, (size_t)tot * 4, cudaMemcpyDeviceToDevice, st);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(repy_label, ol, (size_t)tot * 4, cudaMemcpyDeviceToDevice, st);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(repy_xyz, oxyz, (size_t)tot * 24, cudaMemcpyDeviceToDevice, st);
      }
      if (ce == cudaSuccess) ce = cudaMemcpyAsync(voxoff, noff, (size_t)(V + 1) * 8, cudaMemcpyDeviceToDevice, st);
      for (int v = 0; v <= V; v++) ctx->off_vox[v] = noff_h[v];
      for (int v = 0; v < V; v++) {
        ctx->cnt_vox[v] = noff_h[v + 1] - noff_h[v];
        if (counts_host) counts_host[v * 4 + 2] = ctx->cnt_vox[v];
      }
      free(noff_h);
      if (ce != cudaSuccess) return dfu_set_err(ctx, DFU_ERR_CUDA, "voxel cut: %s", cudaGetErrorString(ce));
    }
  }
  ctx->table_clean_ptr = (void*)table;
  ctx->table_clean_bytes = ctx->slot_bytes[SL_TABLE];
  ctx->stage |= 4;
  if (total_miss > 0)
    return dfu_set_err(ctx, DFU_ERR_DATA, "%lld points fell outside the angular voxel table", (long long)total_miss);
  // voxels with more than max_per_voxel points were re-resolved over their first max_per_voxel points in input order
  // (k_vox_overflow), as the reference's voxeliser does; only a list that did not fit is an error
  (void)allow_overflow;
  if (total_over > 0 && ctx->prm.max_per_voxel > S2_OVF_MAX)
    return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "max_per_voxel = %d with overflowing voxels (supported: up to %d)",
                       ctx->prm.max_per_voxel, S2_OVF_MAX);
  if ((int64_t)hs[V * 8] > (int64_t)ovf_cap)
    return dfu_set_err(ctx, DFU_ERR_DATA, "%lld angular voxels hold more than max_per_voxel points (list capacity %u)",
                       (long long)hs[V * 8], ovf_cap);
  return DFU_OK;
}

extern "C" int dfu_fuse_emit_cloud(dfu_ctx* ctx, void* out, int dtype, int64_t capacity_rows, int64_t* row_off_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || (ctx->stage & 6) != 6)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_emit_cloud: project and densify stages must have run");
  const int V = ctx->V;
  cudaStream_t st = ctx->stream;
  int64_t* rowoff_h = (int64_t*)dfu_pinned(ctx, (V + 1) * sizeof(int64_t));
  if (!rowoff_h) return DFU_ERR_CUDA;
  int64_t tot = 0;
  for (int v = 0; v < V; v++) {
    rowoff_h[v] = tot;
    tot += ctx->cnt_project[v * 3 + 0] + ctx->cnt_vox[v];
  }
  rowoff_h[V] = tot;
  if (row_off_host) memcpy(row_off_host, rowoff_h, (V + 1) * sizeof(int64_t));
  if (!out) return DFU_OK;  // size query
  if ((uintptr_t)out & 15) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_emit_cloud: out must be 16-byte aligned");
  if (capacity_rows < tot) return dfu_set_err(ctx, DFU_ERR_CAPACITY, "cloud needs %lld rows, capacity %lld", (long long)tot, (long long)capacity_rows);
  if (tot == 0) return DFU_OK;
  DFU_SLOT(ctx, rowoff, int64_t*, SL_SEGOUT, (V + 1) * sizeof(int64_t));
  DFU_CUDA(ctx, cudaMemcpyAsync(rowoff, rowoff_h, (V + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  const int64_t* foff = (const int64_t*)ctx->slot[SL_STATS];  // [3][V+1], list 0 = FOV
  const int64_t* voxoff = (const int64_t*)ctx->slot[SL_VOXOFF];
  S2Prm q = make_s2(&ctx->prm);
  dfu_view* vd = (dfu_view*)ctx->slot[SL_VIEWS];
  unsigned nblk = (unsigned)((tot + 255) / 256);
  {
  DFU_PROF(ctx, st, PT_EMIT_CLOUD);
#define EMIT(T)                                                                                                   \
  k_emit_cloud<T><<<nblk, 256, 0, st>>>(vd, q, V, rowoff, foff, voxoff, (const float4*)ctx->lidar,                 \
                                        (const int32_t*)ctx->slot[SL_FLIST], ctx->depth, ctx->image,               \
                                        (const int32_t*)ctx->slot[SL_REPZ], (T*)out, tot)
  if (dtype == 0) EMIT(double);
  else if (dtype == 1) EMIT(float);
  else if (dtype == 2) EMIT(__half);
  else return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_emit_cloud: dtype must be 0, 1 or 2");
#undef EMIT
  DFU_LAUNCH_CHECK(ctx);
  }
  // rowoff_h (pinned) is re-used by later calls: make sure the H2D copy is done
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  return DFU_OK;
}

__global__ void k_soa_to_rows(const double* __restrict__ soa, double* __restrict__ rows, int64_t n) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  rows[t * 3 + 0] = soa[t];
  rows[t * 3 + 1] = soa[n + t];
  rows[t * 3 + 2] = soa[2 * n + t];
}

extern "C" int dfu_fuse_emit_voxels(dfu_ctx* ctx, int32_t* repz_pix, int32_t* repy_pix, double* repz_xyz,
                                    double* repy_xyz, int64_t capacity) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !(ctx->stage & 4)) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_emit_voxels: densify stage not run");
  const int64_t n = ctx->off_vox[ctx->V];
  if (capacity < n) return dfu_set_err(ctx, DFU_ERR_CAPACITY, "voxel outputs need %lld entries", (long long)n);
  if (n == 0) return DFU_OK;
  cudaStream_t st = ctx->stream;
  if (repz_pix) DFU_CUDA(ctx, cudaMemcpyAsync(repz_pix, ctx->slot[SL_REPZ], n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  if (repy_pix) DFU_CUDA(ctx, cudaMemcpyAsync(repy_pix, ctx->slot[SL_REPY], n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  if (repy_xyz) {
    k_soa_to_rows<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const double*)ctx->slot[SL_REPY_XYZ], repy_xyz, n);
    DFU_LAUNCH_CHECK(ctx);
  }
  if (repz_xyz) {
    DFU_SLOT(ctx, tmp, double*, SL_REPZ_XYZ, n * 3 * sizeof(double));
    S2Prm q = make_s2(&ctx->prm);
    k_vox_gather<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((dfu_view*)ctx->slot[SL_VIEWS], q, ctx->V,
                                                              (const int64_t*)ctx->slot[SL_VOXOFF], ctx->depth, ctx->labelmap,
                                                              (const int32_t*)ctx->slot[SL_REPZ], (const int32_t*)ctx->slot[SL_REPY],
                                                              (double*)ctx->slot[SL_REPY_XYZ], (uint32_t*)ctx->slot[SL_REPY_LABEL], tmp, n);
    DFU_LAUNCH_CHECK(ctx);
    k_soa_to_rows<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(tmp, repz_xyz, n);
    DFU_LAUNCH_CHECK(ctx);
  }
  return DFU_OK;
}

// ---------------------------------------------------------------- stand-alone back-projection
// my_loader.py:507-509 (depth threshold, row-major nonzero), calibration_kitti.py:134-144 + 89-102 (float64), and
// optionally the height gate of my_loader.py:540.  Ordered compaction: block counts -> scan -> ballot-prefix scatter.
__global__ void __launch_bounds__(S2_THREADS) k_bp_count(const dfu_view* __restrict__ views, S2Prm q, int gate,
                                                         const float* __restrict__ depth, int32_t* __restrict__ blkcnt) {
  const int v = blockIdx.y;
  LOAD_VIEW(vw, views, v);
  const int npix = q.H * q.W;
  const int pix = blockIdx.x * S2_THREADS + threadIdx.x;
  bool ok = false;
  if (pix < npix) {
    const float d = __ldg(depth + (size_t)v * npix + pix);
    if (d >= q.min_depth) {
      ok = true;
      if (gate) {
        const int row = pix / q.W, col = pix - row * q.W;
        double X, Y, Z;
        dfu_backproject_r(vw, vw_rcp[0], vw_rcp[1], col, row, d, X, Y, Z);
        ok = Z < q.max_z;
      }
    }
  }
  const int n = __syncthreads_count(ok);
  if (threadIdx.x == 0) blkcnt[v * gridDim.x + blockIdx.x] = n;
}

__global__ void __launch_bounds__(S2_THREADS) k_bp_scatter(const dfu_view* __restrict__ views, S2Prm q, int gate,
                                                           const float* __restrict__ depth, const int64_t* __restrict__ blkoff,
                                                           double* __restrict__ xyz, int32_t* __restrict__ pixout) {
  __shared__ int wsum[S2_THREADS / 32];
  const int v = blockIdx.y;
  LOAD_VIEW(vw, views, v);
  const int npix = q.H * q.W;
  const int pix = blockIdx.x * S2_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  bool ok = false;
  double X = 0, Y = 0, Z = 0;
  if (pix < npix) {
    const float d = __ldg(depth + (size_t)v * npix + pix);
    if (d >= q.min_depth) {
      const int row = pix / q.W, col = pix - row * q.W;
      dfu_backproject_r(vw, vw_rcp[0], vw_rcp[1], col, row, d, X, Y, Z);
      ok = !gate || Z < q.max_z;
    }
  }
  const unsigned bal = __ballot_sync(0xffffffffu, ok);
  if (lane == 0) wsum[w] = __popc(bal);
  __syncthreads();
  if (ok) {
    int base = 0;
    for (int i = 0; i < w; i++) base += wsum[i];
    const int64_t pos = blkoff[v * gridDim.x + blockIdx.x] + base + dfu_warp_excl(bal, lane);
    xyz[3 * pos] = X; xyz[3 * pos + 1] = Y; xyz[3 * pos + 2] = Z;
    if (pixout) pixout[pos] = pix;
  }
}

extern "C" int dfu_backproject_sync(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_host,
                                    const dfu_fusion_params* prm_host, const float* depth, int height_gate,
                                    double* out_xyz, int32_t* out_pix, int64_t capacity, int64_t* counts_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || V <= 0 || !views_host || !prm_host || !depth || !counts_host)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_backproject_sync: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  S2Prm q = make_s2(prm_host);
  const int64_t npix = (int64_t)q.H * q.W;
  if (npix <= 0 || npix >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_backproject_sync: bad depth map size");
  const int bpv = (int)((npix + S2_THREADS - 1) / S2_THREADS);
  DFU_SLOT(ctx, vd, dfu_view*, SL_VIEWS, sizeof(dfu_view) * V);
  DFU_SLOT(ctx, blkcnt, int32_t*, SL_BLKCNT, (int64_t)V * bpv * sizeof(int32_t));
  DFU_SLOT(ctx, blkoff, int64_t*, SL_BLKOFF, ((int64_t)V * bpv + 1) * sizeof(int64_t));
  DFU_SLOT(ctx, voxoff, int64_t*, SL_VOXOFF, (V + 1) * sizeof(int64_t));
  ctx->stage = 0;   // SL_VIEWS / SL_BLK* are shared with the batched pipeline: any bound batch is gone
  DFU_CUDA(ctx, cudaMemcpyAsync(vd, views_host, sizeof(dfu_view) * V, cudaMemcpyHostToDevice, st));
  dim3 grid(bpv, V);
  k_bp_count<<<grid, S2_THREADS, 0, st>>>(vd, q, height_gate, depth, blkcnt);
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_exclusive_sum_i32(ctx, st, blkcnt, blkoff, (int64_t)V * bpv);
  if (rc) return rc;
  k_vox_view_offsets<<<(V + 1 + 127) / 128, 128, 0, st>>>(blkoff, V, bpv, voxoff);
  DFU_LAUNCH_CHECK(ctx);
  int64_t* pin = (int64_t*)dfu_pinned(ctx, (V + 1) * sizeof(int64_t));
  if (!pin) return DFU_ERR_CUDA;
  DFU_CUDA(ctx, cudaMemcpyAsync(pin, voxoff, (V + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));   // also covers the pageable views_host copy
  for (int v = 0; v < V; v++) counts_host[v] = pin[v + 1] - pin[v];
  const int64_t total = pin[V];
  if (!out_xyz) return DFU_OK;   // size query
  if (capacity < total)
    return dfu_set_err(ctx, DFU_ERR_CAPACITY, "back-projection needs %lld rows, capacity %lld", (long long)total, (long long)capacity);
  if (total == 0) return DFU_OK;
  DFU_PROF(ctx, st, PT_MISC);
  k_bp_scatter<<<grid, S2_THREADS, 0, st>>>(vd, q, height_gate, depth, blkoff, out_xyz, out_pix);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}
