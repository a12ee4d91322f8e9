// Self-evolution glue: score-thresholded detector predictions -> refreshed pseudo-boxes.
// What the reference runs for it: pcdet/models/model_utils/model_nms_utils.py:6-27 (class_agnostic_nms: score mask,
// top-k by score, rotated NMS, post-max cut) over pcdet/ops/iou3d_nms (iou3d_nms_utils.py:31-42,120-136;
// src/iou3d_nms_kernel.cu:128-234 rotated-rectangle overlap, :295 64x64 bitmask, src/iou3d_nms.cpp:137-183 greedy
// reduction on the HOST after a D2H copy of the whole mask).
//
// Here: every box is prepared ONCE (double-routine cos/sin rounded to float, oriented corners, margin half extents)
// instead of once per pair; the suppression matrix is computed one PAIR per lane (a 64x64 tile = 8 warps x 8 rows x
// 2 ballots) and only for tiles on or above the diagonal; the greedy reduction runs on the device in one block, so
// the only D2H traffic is the number of kept boxes.
//
// Arithmetic: float32 with every product / sum rounded separately (`__fmul_rn` ...: no FMA contraction), in the
// operation order of the reference's CPU build of the same routine (src/iou3d_cpu.cpp:59-230), which is what
// oracle/_ref compiles and the parity tests compare against.  The reference's CUDA build lets nvcc contract into FMAs,
// so its own two builds already differ in the last ulp; the CPU one is the reproducible definition.
#include <cub/cub.cuh>
#include "dfu_common.cuh"

enum { EV_A = 224, EV_B, EV_KEY, EV_IDX, EV_SKEY, EV_SIDX, EV_TMP, EV_CNT, EV_MASK, EV_KEEP };
static_assert(EV_A >= 224 && EV_KEEP < DFU_NSLOTS, "EV_* ids: own range 224..233 (see dfu_common.cuh)");

#define EV_EPS 1e-8f                   // iou3d_cpu.cpp:38
#define EV_MAXPTS 24                   // the reference reserves 16 (iou3d_cpu.cpp:170) and can overrun it; 16 edge pairs + 8 corners

struct __align__(16) BevBox {
  float cx, cy, cosr, sinr;            // centre, cos / sin of +heading
  float hxm, hym, area, rad;           // dx / 2 + 1e-2, dy / 2 + 1e-2 (check_in_box2d, :75-85), dx * dy, reach (below)
  float px[4], py[4];                  // oriented corners (:135-163)
};

__device__ __forceinline__ float ev_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float ev_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float ev_sub(float a, float b) { return __fsub_rn(a, b); }

__device__ __forceinline__ BevBox ev_prepare(const float* __restrict__ b) {
  BevBox g;
  const float cx = b[0], cy = b[1], dx = b[3], dy = b[4], ang = b[6];
  g.cx = cx; g.cy = cy;
  // host libm cosf / sinf are correctly rounded in practice; CUDA's cosf is up to 2 ulp off -> double routine, one rounding
  g.cosr = (float)cos((double)ang);
  g.sinr = (float)sin((double)ang);
  const float hx = __fdiv_rn(dx, 2.f), hy = __fdiv_rn(dy, 2.f);
  g.hxm = ev_add(hx, 1e-2f);
  g.hym = ev_add(hy, 1e-2f);
  g.area = ev_mul(dx, dy);
  // every point the overlap routine can collect for this box (edge crossings, own corners, the other box's corners inside
  // the 1e-2-inflated rectangle) lies within half a diagonal + 1e-2 * sqrt(2) of the centre; 0.03 leaves room for rounding
  g.rad = sqrtf(hx * hx + hy * hy) + 0.03f;
  const float x1 = ev_sub(cx, hx), y1 = ev_sub(cy, hy), x2 = ev_add(cx, hx), y2 = ev_add(cy, hy);
  const float qx[4] = {x1, x2, x2, x1}, qy[4] = {y1, y1, y2, y2};
#pragma unroll
  for (int k = 0; k < 4; k++) {        // rotate_around_center (:118-122)
    const float ux = ev_sub(qx[k], cx), uy = ev_sub(qy[k], cy);
    g.px[k] = ev_add(ev_add(ev_mul(ux, g.cosr), ev_mul(uy, -g.sinr)), cx);
    g.py[k] = ev_add(ev_add(ev_mul(ux, g.sinr), ev_mul(uy, g.cosr)), cy);
  }
  return g;
}

// (p1 - p0) x (p2 - p0)  (:63-65)
__device__ __forceinline__ float ev_cross3(float p1x, float p1y, float p2x, float p2y, float p0x, float p0y) {
  return ev_sub(ev_mul(ev_sub(p1x, p0x), ev_sub(p2y, p0y)), ev_mul(ev_sub(p2x, p0x), ev_sub(p1y, p0y)));
}

// segment p0->p1 against q0->q1 (:87-116); true + the crossing point when they properly cross
__device__ __forceinline__ bool ev_segments(float p1x, float p1y, float p0x, float p0y, float q1x, float q1y, float q0x,
                                            float q0y, float& ox, float& oy) {
  const bool boxes_touch = fminf(p0x, p1x) <= fmaxf(q0x, q1x) && fminf(q0x, q1x) <= fmaxf(p0x, p1x) &&
                           fminf(p0y, p1y) <= fmaxf(q0y, q1y) && fminf(q0y, q1y) <= fmaxf(p0y, p1y);
  if (!boxes_touch) return false;
  const float s1 = ev_cross3(q0x, q0y, p1x, p1y, p0x, p0y);
  const float s2 = ev_cross3(p1x, p1y, q1x, q1y, p0x, p0y);
  const float s3 = ev_cross3(p0x, p0y, q1x, q1y, q0x, q0y);
  const float s4 = ev_cross3(q1x, q1y, p1x, p1y, q0x, q0y);
  if (!(ev_mul(s1, s2) > 0.f && ev_mul(s3, s4) > 0.f)) return false;
  const float s5 = ev_cross3(q1x, q1y, p1x, p1y, p0x, p0y);
  const float den = ev_sub(s5, s1);
  if (fabsf(den) > EV_EPS) {
    ox = __fdiv_rn(ev_sub(ev_mul(s5, q0x), ev_mul(s1, q1x)), den);
    oy = __fdiv_rn(ev_sub(ev_mul(s5, q0y), ev_mul(s1, q1y)), den);
  } else {
    const float a0 = ev_sub(p0y, p1y), b0 = ev_sub(p1x, p0x), c0 = ev_sub(ev_mul(p0x, p1y), ev_mul(p1x, p0y));
    const float a1 = ev_sub(q0y, q1y), b1 = ev_sub(q1x, q0x), c1 = ev_sub(ev_mul(q0x, q1y), ev_mul(q1x, q0y));
    const float D = ev_sub(ev_mul(a0, b1), ev_mul(a1, b0));
    ox = __fdiv_rn(ev_sub(ev_mul(b0, c1), ev_mul(b1, c0)), D);
    oy = __fdiv_rn(ev_sub(ev_mul(a1, c0), ev_mul(a0, c1)), D);
  }
  return true;
}

// point inside the margin-inflated box (:75-85; the rotation by -heading uses cos(-a) = cos a, sin(-a) = -sin a, exact)
__device__ __forceinline__ bool ev_inside(const BevBox& g, float x, float y) {
  const float ux = ev_sub(x, g.cx), uy = ev_sub(y, g.cy);
  const float rx = ev_add(ev_mul(ux, g.cosr), ev_mul(uy, g.sinr));
  const float ry = ev_add(ev_mul(ux, -g.sinr), ev_mul(uy, g.cosr));
  return fabsf(rx) < g.hxm && fabsf(ry) < g.hym;
}

// area of the intersection polygon (:128-220): crossing points of the 16 edge pairs, then the corners of one box inside
// the other (b's corner k, then a's corner k), ordered by atan2 about their mean, fan-summed from the first point
__device__ float ev_overlap(const BevBox& A, const BevBox& B) {
  {  // centres further apart than the two reaches: the reference finds no crossing and no contained corner -> area 0
     // (NaN / inf boxes fail the comparison and take the full routine)
    const float ddx = A.cx - B.cx, ddy = A.cy - B.cy, reach = A.rad + B.rad;
    if (ddx * ddx + ddy * ddy > reach * reach) return 0.f;
  }
  float vx[EV_MAXPTS], vy[EV_MAXPTS];
  int cnt = 0;
  float sx = 0.f, sy = 0.f;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int i1 = (i + 1) & 3;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int j1 = (j + 1) & 3;
      float ox, oy;
      if (ev_segments(A.px[i1], A.py[i1], A.px[i], A.py[i], B.px[j1], B.py[j1], B.px[j], B.py[j], ox, oy)) {
        sx = ev_add(sx, ox); sy = ev_add(sy, oy);
        vx[cnt] = ox; vy[cnt] = oy; cnt++;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; k++) {
    if (ev_inside(A, B.px[k], B.py[k])) { sx = ev_add(sx, B.px[k]); sy = ev_add(sy, B.py[k]); vx[cnt] = B.px[k]; vy[cnt] = B.py[k]; cnt++; }
    if (ev_inside(B, A.px[k], A.py[k])) { sx = ev_add(sx, A.px[k]); sy = ev_add(sy, A.py[k]); vx[cnt] = A.px[k]; vy[cnt] = A.py[k]; cnt++; }
  }
  if (cnt < 3) return 0.f;             // the reference's fan sum over fewer than 3 points is empty or a zero cross product
  const float mx = __fdiv_rn(sx, (float)cnt), my = __fdiv_rn(sy, (float)cnt);
  float ang[EV_MAXPTS];
  for (int k = 0; k < cnt; k++) ang[k] = atan2f(ev_sub(vy[k], my), ev_sub(vx[k], mx));
  // the reference bubble-sorts ascending with a strict `>` swap test (:193-202): a stable sort; so is this insertion
  for (int k = 1; k < cnt; k++) {
    const float a = ang[k], x = vx[k], y = vy[k];
    int m = k - 1;
    while (m >= 0 && ang[m] > a) { ang[m + 1] = ang[m]; vx[m + 1] = vx[m]; vy[m + 1] = vy[m]; m--; }
    ang[m + 1] = a; vx[m + 1] = x; vy[m + 1] = y;
  }
  float area = 0.f;
  for (int k = 0; k < cnt - 1; k++) {  // cross(v[k] - v[0], v[k+1] - v[0])  (:214-216, :59-61)
    const float ax = ev_sub(vx[k], vx[0]), ay = ev_sub(vy[k], vy[0]);
    const float bx = ev_sub(vx[k + 1], vx[0]), by = ev_sub(vy[k + 1], vy[0]);
    area = ev_add(area, ev_sub(ev_mul(ax, by), ev_mul(ay, bx)));
  }
  return fabsf(area) * 0.5f;           // `fabs(area) / 2.0` in double, rounded back: exact halving
}

__device__ __forceinline__ float ev_iou_bev(const BevBox& A, const BevBox& B) {   // :222-230
  const float ov = ev_overlap(A, B);
  return __fdiv_rn(ov, fmaxf(ev_sub(ev_add(A.area, B.area), ov), EV_EPS));
}

// ---------------------------------------------------------------- kernels
// one thread per box; order (nullable) gathers rows of the score-sorted candidates
// ndev (nullable): device-side count, the effective n is min(*ndev, n) — lets the NMS run without reading the number of
// score-passing candidates back to the host
__global__ void k_bev_prepare(const float* __restrict__ boxes, int stride, const int32_t* __restrict__ order, int n,
                              const int32_t* __restrict__ ndev, BevBox* __restrict__ out) {
  if (ndev) n = min(n, *ndev);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = ev_prepare(boxes + (int64_t)(order ? order[i] : i) * stride);
}

// mode 0: BEV IoU (boxes_iou_bev, iou3d_nms_utils.py:31-42); 1: BEV overlap area (boxes_overlap_bev_gpu);
// 2: 3-D IoU (boxes_iou3d_gpu, iou3d_nms_utils.py:45-77: float32 torch elementwise arithmetic around the BEV overlap)
__global__ void __launch_bounds__(128) k_bev_pairs(const BevBox* __restrict__ A, const float* __restrict__ rawA, int sa, int na,
                                                   const BevBox* __restrict__ B, const float* __restrict__ rawB, int sb, int nb,
                                                   int mode, int paired, float* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = paired ? j : blockIdx.y;                 // paired: row i against row i only (paired_boxes_iou3d_gpu, :155-188)
  if (j >= nb || i >= na) return;
  const BevBox a = A[i], b = B[j];
  float r;
  if (mode == 0) r = ev_iou_bev(a, b);
  else {
    r = ev_overlap(a, b);
    if (mode == 2) {
      const float* pa = rawA + (int64_t)i * sa;
      const float* pb = rawB + (int64_t)j * sb;
      const float ha = __fdiv_rn(pa[5], 2.f), hb = __fdiv_rn(pb[5], 2.f);
      const float top = fminf(ev_add(pa[2], ha), ev_add(pb[2], hb)), bot = fmaxf(ev_sub(pa[2], ha), ev_sub(pb[2], hb));
      const float oh = fmaxf(ev_sub(top, bot), 0.f);
      const float o3 = ev_mul(r, oh);
      const float va = ev_mul(ev_mul(pa[3], pa[4]), pa[5]), vb = ev_mul(ev_mul(pb[3], pb[4]), pb[5]);
      r = __fdiv_rn(o3, fmaxf(ev_sub(ev_add(va, vb), o3), 1e-6f));
    }
  }
  out[paired ? (int64_t)j : (int64_t)i * nb + j] = r;
}

// order-preserving key of a score, inverted: an ascending stable radix sort yields descending scores, ties by index.
// rows failing `score >= thresh` (NaN included) get the last key and are not counted
__global__ void k_nms_keys(const float* __restrict__ scores, int n, int use_thresh, float thresh, uint32_t* __restrict__ key,
                           int32_t* __restrict__ idx, int32_t* __restrict__ npass) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  bool pass = false;
  if (i < n) {
    const float s = scores[i];
    pass = use_thresh ? (s >= thresh) : (s == s);
    const uint32_t u = __float_as_uint(s);
    const uint32_t asc = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    key[i] = pass ? ~asc : 0xffffffffu;
    idx[i] = i;
  }
  const unsigned bal = __ballot_sync(0xffffffffu, pass);
  if ((threadIdx.x & 31) == 0 && bal) atomicAdd(npass, __popc(bal));
}

// suppression words: bit c of mask[r * nw + cw] = IoU(box r, box 64 * cw + c) > thresh, for tiles with cw >= r / 64 only
// (a box is only ever suppressed by a better-scored one; the reference computes the lower triangle too and never reads it).
// grid (col tile, row tile), 256 threads: warp w owns rows 8w..8w+7 of the tile, lane l the columns l and l + 32.
__global__ void __launch_bounds__(256) k_nms_mask(const BevBox* __restrict__ P, int n, const int32_t* __restrict__ ndev,
                                                  float thresh, int nw, unsigned long long* __restrict__ mask) {
  const int ct = blockIdx.x, rt = blockIdx.y;
  if (ct < rt) return;
  if (ndev) n = min(n, *ndev);
  if (ct * 64 >= n) return;
  __shared__ BevBox cols[64];
  const int c0 = ct * 64, r0 = rt * 64;
  {
    const float4* src = reinterpret_cast<const float4*>(P + c0);
    float4* dst = reinterpret_cast<float4*>(cols);
    const int nvec = min(64, n - c0) * (int)(sizeof(BevBox) / 16);
    for (int t = threadIdx.x; t < nvec; t += 256) dst[t] = src[t];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int q = 0; q < 8; q++) {
    const int r = r0 + 8 * w + q;
    if (r >= n) break;                 // warp-uniform
    const BevBox a = P[r];
    bool h0 = false, h1 = false;
    const int ca = c0 + lane, cb = c0 + lane + 32;
    if (ca < n && ca > r) h0 = ev_iou_bev(a, cols[lane]) > thresh;          // iou3d_nms_kernel.cu:327-335 (start = row + 1)
    if (cb < n && cb > r) h1 = ev_iou_bev(a, cols[lane + 32]) > thresh;
    const unsigned lo = __ballot_sync(0xffffffffu, h0), hi = __ballot_sync(0xffffffffu, h1);
    if (lane == 0) mask[(int64_t)r * nw + ct] = ((unsigned long long)hi << 32) | lo;
  }
}

// nms_normal_gpu (iou3d_nms_kernel.cu:341-400): the same suppression words from the axis-aligned overlap of the boxes'
// x / y extents, heading ignored.  The only products are dx * dy and width * height (halving is exact), so the value does
// not depend on FMA contraction and equals the reference's CUDA arithmetic.
__device__ __forceinline__ float ev_iou_normal(const float4 a, const float4 b) {      // x y dx dy
  const float ahx = a.z * 0.5f, ahy = a.w * 0.5f, bhx = b.z * 0.5f, bhy = b.w * 0.5f;
  const float left = fmaxf(ev_sub(a.x, ahx), ev_sub(b.x, bhx)), right = fminf(ev_add(a.x, ahx), ev_add(b.x, bhx));
  const float top = fmaxf(ev_sub(a.y, ahy), ev_sub(b.y, bhy)), bottom = fminf(ev_add(a.y, ahy), ev_add(b.y, bhy));
  const float inter = ev_mul(fmaxf(ev_sub(right, left), 0.f), fmaxf(ev_sub(bottom, top), 0.f));
  return __fdiv_rn(inter, fmaxf(ev_sub(ev_add(ev_mul(a.z, a.w), ev_mul(b.z, b.w)), inter), EV_EPS));
}

__global__ void __launch_bounds__(256) k_nms_mask_normal(const float* __restrict__ boxes, int stride, const int32_t* __restrict__ order,
                                                         int n, const int32_t* __restrict__ ndev, float thresh, int nw,
                                                         unsigned long long* __restrict__ mask) {
  const int ct = blockIdx.x, rt = blockIdx.y;
  if (ct < rt) return;
  if (ndev) n = min(n, *ndev);
  if (ct * 64 >= n) return;
  __shared__ float4 cols[64], rows[64];
  const int c0 = ct * 64, r0 = rt * 64;
  if (threadIdx.x < 128) {
    const int k = threadIdx.x & 63, g = (threadIdx.x < 64 ? c0 : r0) + k;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g < n) { const float* p = boxes + (int64_t)order[g] * stride; v = make_float4(p[0], p[1], p[3], p[4]); }
    (threadIdx.x < 64 ? cols : rows)[k] = v;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int q = 0; q < 8; q++) {
    const int r = r0 + 8 * w + q;
    if (r >= n) break;
    const float4 a = rows[8 * w + q];
    const int ca = c0 + lane, cb = c0 + lane + 32;
    const bool h0 = ca < n && ca > r && ev_iou_normal(a, cols[lane]) > thresh;
    const bool h1 = cb < n && cb > r && ev_iou_normal(a, cols[lane + 32]) > thresh;
    const unsigned lo = __ballot_sync(0xffffffffu, h0), hi = __ballot_sync(0xffffffffu, h1);
    if (lane == 0) mask[(int64_t)r * nw + ct] = ((unsigned long long)hi << 32) | lo;
  }
}

// greedy pass of iou3d_nms.cpp:166-181 on the device, 64 boxes (one mask word) at a time: thread 0 resolves the chunk
// against its diagonal words (a 64-step register chain, the shared loads issued up front), then the rows of the
// chunk's survivors are ORed into the later removed-words, one (word, box) pair per thread.  Stops once post_max boxes are kept.
__global__ void __launch_bounds__(1024) k_nms_reduce(const unsigned long long* __restrict__ mask, int n,
                                                     const int32_t* __restrict__ ndev, int nw, int post_max,
                                                     const int32_t* __restrict__ order, int64_t* __restrict__ keep,
                                                     int32_t* __restrict__ nkeep) {
  __shared__ unsigned long long removed[1024];
  __shared__ unsigned long long diag[64];
  __shared__ unsigned long long survivors;
  const int t = threadIdx.x;
  if (ndev) n = min(n, *ndev);
  removed[t] = 0ull;
  int cnt = 0;
  for (int c = 0; c < nw; c++) {
    const int base = c * 64;
    if (base >= n) break;
    if (t < 64) diag[t] = (base + t < n) ? mask[(int64_t)(base + t) * nw + c] : ~0ull;
    __syncthreads();                                         // diag ready; removed[c] final (written in chunk c - 1's phase 2)
    if (t == 0) {
      unsigned long long rem = removed[c], kb = 0ull;
      const int lim = min(64, n - base);
#pragma unroll
      for (int b = 0; b < 64; b++) {
        const bool alive = b < lim && !((rem >> b) & 1ull);
        kb |= (unsigned long long)alive << b;
        rem |= alive ? diag[b] : 0ull;
      }
      survivors = kb;
    }
    __syncthreads();
    const unsigned long long kb = survivors;
    if (t < 64 && ((kb >> t) & 1ull)) {
      const int pos = cnt + __popcll(kb & ((1ull << t) - 1ull));
      if (pos < post_max) keep[pos] = order[base + t];
    }
    cnt += __popcll(kb);
    if (cnt >= post_max) break;                              // uniform: the remaining boxes cannot enter keep[:post_max]
    // one (later word, box of the chunk) pair per thread: every load is independent; the 64 boxes of a word are two warps,
    // OR-reduced by the warp and merged with one shared atomic each
    const int npairs = ((n + 63) / 64 - c - 1) * 64;
    for (int p = t; p < npairs; p += 1024) {                 // npairs is a multiple of 64: warps enter and leave whole
      const int w = c + 1 + (p >> 6), b = p & 63;
      const unsigned long long v = ((kb >> b) & 1ull) ? mask[(int64_t)(base + b) * nw + w] : 0ull;
      const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v), hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
      if ((t & 31) == 0 && (lo | hi)) atomicOr(&removed[w], ((unsigned long long)hi << 32) | lo);
    }
  }
  if (t == 0) *nkeep = min(cnt, post_max);
}

// ---------------------------------------------------------------- C ABI
extern "C" int dfu_boxes_iou_sync(dfu_ctx* ctx, void* stream, const float* boxes_a, int stride_a, int na, const float* boxes_b,
                                  int stride_b, int nb, int mode, float* out) {
  DFU_ON_DEVICE(ctx);
  const int paired = (mode & 4) != 0;
  mode &= ~4;
  if (!ctx || na < 0 || nb < 0 || stride_a < 7 || stride_b < 7 || mode < 0 || mode > 2 || (paired && na != nb) ||
      (na > 0 && nb > 0 && (!boxes_a || !boxes_b || !out)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_boxes_iou_sync: bad argument");
  if (na == 0 || nb == 0) return DFU_OK;
  if (na > 65535 && !paired) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_boxes_iou_sync: more than 65535 rows");
  cudaStream_t st = (cudaStream_t)stream;
  DFU_SLOT(ctx, A, BevBox*, EV_A, (size_t)na * sizeof(BevBox));
  DFU_SLOT(ctx, B, BevBox*, EV_B, (size_t)nb * sizeof(BevBox));
  DFU_PROF(ctx, st, PT_MISC);
  k_bev_prepare<<<(na + 127) / 128, 128, 0, st>>>(boxes_a, stride_a, nullptr, na, nullptr, A);
  DFU_LAUNCH_CHECK(ctx);
  k_bev_prepare<<<(nb + 127) / 128, 128, 0, st>>>(boxes_b, stride_b, nullptr, nb, nullptr, B);
  DFU_LAUNCH_CHECK(ctx);
  dim3 grid((nb + 127) / 128, paired ? 1 : na);
  k_bev_pairs<<<grid, 128, 0, st>>>(A, boxes_a, stride_a, na, B, boxes_b, stride_b, nb, mode, paired, out);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  return DFU_OK;
}

static int ev_nms(dfu_ctx* ctx, void* stream, const float* boxes, int stride, const float* scores, int n, int use_score_thresh,
                  float score_thresh, float iou_thresh, int pre_maxsize, int post_maxsize, int64_t* keep_out,
                  int64_t* n_keep_host, int normal) {
  if (!ctx || !n_keep_host || n < 0 || stride < 7 || (n > 0 && (!boxes || !scores || !keep_out)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_nms_sync: bad argument");
  *n_keep_host = 0;
  if (n == 0 || post_maxsize == 0 || pre_maxsize == 0) return DFU_OK;
  cudaStream_t st = (cudaStream_t)stream;
  DFU_SLOT(ctx, key, uint32_t*, EV_KEY, (size_t)n * 4);
  DFU_SLOT(ctx, idx, int32_t*, EV_IDX, (size_t)n * 4);
  DFU_SLOT(ctx, skey, uint32_t*, EV_SKEY, (size_t)n * 4);
  DFU_SLOT(ctx, sidx, int32_t*, EV_SIDX, (size_t)n * 4);
  DFU_SLOT(ctx, cnt, int32_t*, EV_CNT, 2 * sizeof(int32_t));
  DFU_PROF(ctx, st, PT_MISC);
  DFU_CUDA(ctx, cudaMemsetAsync(cnt, 0, 2 * sizeof(int32_t), st));
  k_nms_keys<<<(n + 255) / 256, 256, 0, st>>>(scores, n, use_score_thresh, score_thresh, key, idx, cnt);
  DFU_LAUNCH_CHECK(ctx);
  size_t tb = 0;
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, skey, idx, sidx, n, 0, 32, st));
  DFU_SLOT(ctx, tmp, void*, EV_TMP, tb);
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, skey, idx, sidx, n, 0, 32, st));   // stable
  ctx->launches += 3;
  int32_t* pin = (int32_t*)dfu_pinned(ctx, 2 * sizeof(int32_t));
  if (!pin) return DFU_ERR_CUDA;
  // m = the candidates: score-passing rows, cut at pre_maxsize (model_nms_utils.py:15 / iou3d_nms_utils.py:129-130).  Its
  // upper bound min(n, pre_maxsize) sizes the grids and the mask; the kernels clamp to the device-side count, so the count
  // is read back only when the bound does not fit the single-block greedy pass
  int m = (pre_maxsize > 0 && n > pre_maxsize) ? pre_maxsize : n;
  const int32_t* mdev = cnt;
  if (m > 65536) {
    DFU_CUDA(ctx, cudaMemcpyAsync(pin, cnt, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    DFU_CUDA(ctx, cudaStreamSynchronize(st));
    if (pin[0] < m) m = pin[0];
    if (m == 0) return DFU_OK;
  }
  const int nw = (m + 63) / 64;
  if (nw > 1024) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_nms_sync: %d candidates, at most 65536 (set pre_maxsize)", m);
  DFU_SLOT(ctx, mask, unsigned long long*, EV_MASK, (size_t)m * nw * sizeof(unsigned long long));
  if (normal) {
    k_nms_mask_normal<<<dim3(nw, nw), 256, 0, st>>>(boxes, stride, sidx, m, mdev, iou_thresh, nw, mask);
    DFU_LAUNCH_CHECK(ctx);
  } else {
    DFU_SLOT(ctx, P, BevBox*, EV_A, (size_t)m * sizeof(BevBox));
    k_bev_prepare<<<(m + 127) / 128, 128, 0, st>>>(boxes, stride, sidx, m, mdev, P);
    DFU_LAUNCH_CHECK(ctx);
    k_nms_mask<<<dim3(nw, nw), 256, 0, st>>>(P, m, mdev, iou_thresh, nw, mask);
    DFU_LAUNCH_CHECK(ctx);
  }
  const int post = post_maxsize < 0 ? m : post_maxsize;
  k_nms_reduce<<<1, 1024, 0, st>>>(mask, m, mdev, nw, post, sidx, keep_out, cnt + 1);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cudaMemcpyAsync(pin + 1, cnt + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  *n_keep_host = pin[1];
  return DFU_OK;
}

extern "C" int dfu_nms_sync(dfu_ctx* ctx, void* stream, const float* boxes, int stride, const float* scores, int n,
                            int use_score_thresh, float score_thresh, float iou_thresh, int pre_maxsize, int post_maxsize,
                            int64_t* keep_out, int64_t* n_keep_host) {
  DFU_ON_DEVICE(ctx);
  return ev_nms(ctx, stream, boxes, stride, scores, n, use_score_thresh, score_thresh, iou_thresh, pre_maxsize, post_maxsize,
                keep_out, n_keep_host, 0);
}

extern "C" int dfu_nms_normal_sync(dfu_ctx* ctx, void* stream, const float* boxes, int stride, const float* scores, int n,
                                   int use_score_thresh, float score_thresh, float iou_thresh, int pre_maxsize,
                                   int post_maxsize, int64_t* keep_out, int64_t* n_keep_host) {
  DFU_ON_DEVICE(ctx);
  return ev_nms(ctx, stream, boxes, stride, scores, n, use_score_thresh, score_thresh, iou_thresh, pre_maxsize, post_maxsize,
                keep_out, n_keep_host, 1);
}
