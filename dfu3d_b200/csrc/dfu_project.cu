// S1: LiDAR -> image projection, frustum / z / plane tests, label inheritance, ordered compaction.
// Reference arithmetic: tools/PENet/dataloaders/calibration_kitti.py:104-132, tools/PENet/vis_utils.py:108-123,
// tools/PENet/dataloaders/my_loader.py:471-484,518-530,102-122.
#include "dfu_common.cuh"

#define S1_THREADS 256

struct S1Prm {
  int H, W, Hm, Wm, bound_x, bound_y, round_mode, no_fov;
  double plane_offset, xlo, xhi, ylo, yhi;
};

static S1Prm make_s1(const dfu_fusion_params* p) {
  S1Prm s;
  s.H = p->H; s.W = p->W; s.Hm = p->Hm; s.Wm = p->Wm;
  s.bound_x = p->bound_x; s.bound_y = p->bound_y; s.round_mode = p->round_mode; s.no_fov = p->no_fov_cull;
  s.plane_offset = p->plane_offset;
  s.xlo = p->plane_range[0]; s.xhi = p->plane_range[1]; s.ylo = p->plane_range[2]; s.yhi = p->plane_range[3];
  return s;
}

struct S1Out {
  float u, v, depth;
  int px, py;
  unsigned flags;
  uint32_t label;
};

// norm of the plane's normal, the same expression for every point of a view (np.linalg.norm: sqrt of the sum of squares)
__device__ __forceinline__ double s1_plane_norm(const dfu_view& vw) {
  return __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(vw.plane[0], vw.plane[0]), __dmul_rn(vw.plane[1], vw.plane[1])),
                              __dmul_rn(vw.plane[2], vw.plane[2])));
}

// fov_only: the fused pipeline reads the plane / bounds flags and the label of FOV points only (the three lists nest),
// so a point outside the FOV (5 of 6 for a six-camera frame) skips the float64 plane distance and the label gather.
__device__ __forceinline__ S1Out s1_point(const dfu_view& vw, const S1Prm& q, float4 p, const uint32_t* __restrict__ lm,
                                          double nrm, bool fov_only = false) {
  S1Out o;
  // lidar_to_rect (calibration_kitti.py:104-112)
  float rx = dfu_row4_f32(p.x, p.y, p.z, 1.0f, vw.M43, 0, 3);
  float ry = dfu_row4_f32(p.x, p.y, p.z, 1.0f, vw.M43, 1, 3);
  float rz = dfu_row4_f32(p.x, p.y, p.z, 1.0f, vw.M43, 2, 3);
  // rect_to_img (:114-123): divide by the rect Z, not by the homogeneous w
  float h0 = dfu_row4_f32(rx, ry, rz, 1.0f, vw.P2T, 0, 3);
  float h1 = dfu_row4_f32(rx, ry, rz, 1.0f, vw.P2T, 1, 3);
  float h2 = dfu_row4_f32(rx, ry, rz, 1.0f, vw.P2T, 2, 3);
  o.u = __fdiv_rn(h0, rz);
  o.v = __fdiv_rn(h1, rz);
  o.depth = __fsub_rn(h2, vw.P2T[3 * 3 + 2]);
  // get_fov_flag (vis_utils.py:108-123): un-rounded coordinates against the depth-map shape
  bool fov = q.no_fov || ((o.u >= 0.0f) && (o.u < (float)q.W) && (o.v >= 0.0f) && (o.v < (float)q.H) && (o.depth >= 0.0f));
  if (fov_only && !fov) {
    o.px = o.py = 0;
    o.label = 0;
    o.flags = 0;
    return o;
  }
  // above_plane (my_loader.py:471-484); dgemv order measured on the reference container: y*b first
  double acc = __dmul_rn((double)p.y, vw.plane[1]);
  acc = __fma_rn((double)p.x, vw.plane[0], acc);
  acc = __fma_rn((double)p.z, vw.plane[2], acc);
  double dist = __dadd_rn(acc, vw.plane[3]);
  dist = __ddiv_rn(dist, nrm);
  bool in_rng = ((double)p.x < q.xhi) && ((double)p.x > q.xlo) && ((double)p.y < q.yhi) && ((double)p.y > q.ylo);
  bool above = !((dist < q.plane_offset) && in_rng);
  // pixel index: np.round half-to-even (my_loader.py:520) or astype(int32) truncation (:113)
  float fx = q.round_mode == 0 ? rintf(o.u) : truncf(o.u);
  float fy = q.round_mode == 0 ? rintf(o.v) : truncf(o.v);
  bool inb = (fx >= 0.0f) && (fx < (float)q.bound_x) && (fy >= 0.0f) && (fy < (float)q.bound_y);
  // clamp before the int conversion so that NaN / huge values stay defined
  o.px = (fx >= -2147483000.0f && fx <= 2147483000.0f) ? (int)fx : (fx > 0 ? 2147483647 : -2147483647 - 1);
  o.py = (fy >= -2147483000.0f && fy <= 2147483000.0f) ? (int)fy : (fy > 0 ? 2147483647 : -2147483647 - 1);
  o.label = 0;
  if (lm != nullptr && inb && o.px < q.Wm && o.py < q.Hm) o.label = __ldg(lm + (size_t)o.py * q.Wm + o.px);
  o.flags = (fov ? 1u : 0u) | (above ? 2u : 0u) | (inb ? 4u : 0u);
  return o;
}

// grid (bpv, V).  blkcnt (nullable): int32 [3][V*bpv] block counts of the three ordered lists.
__global__ void __launch_bounds__(S1_THREADS) k_project(const dfu_view* __restrict__ views, S1Prm q,
                                                        const float4* __restrict__ lidar,
                                                        const uint32_t* __restrict__ labelmap, float2* out_uv,
                                                        float* out_depth, int2* out_pix, uint8_t* out_flags,
                                                        uint32_t* out_label, int32_t* blkcnt) {
  __shared__ dfu_view vw;
  __shared__ double s_nrm;
  const int v = blockIdx.y;
  for (int i = threadIdx.x; i < (int)(sizeof(dfu_view) / 4); i += blockDim.x)
    ((uint32_t*)&vw)[i] = ((const uint32_t*)(views + v))[i];
  __syncthreads();
  if (threadIdx.x == 0) s_nrm = s1_plane_norm(vw);
  __syncthreads();
  const int i = blockIdx.x * S1_THREADS + threadIdx.x;
  unsigned fl = 0;
  if (i < vw.lidar_cnt) {
    float4 p = __ldg(lidar + vw.lidar_off + i);
    const uint32_t* lm = labelmap ? labelmap + (size_t)v * q.Hm * q.Wm : nullptr;
    S1Out o = s1_point(vw, q, p, lm, s_nrm, blkcnt != nullptr && !q.no_fov);
    const int64_t s = vw.slot_off + i;
    if (out_uv) out_uv[s] = make_float2(o.u, o.v);
    if (out_depth) out_depth[s] = o.depth;
    if (out_pix) out_pix[s] = make_int2(o.px, o.py);
    if (out_flags) out_flags[s] = (uint8_t)o.flags;
    if (out_label) out_label[s] = o.label;
    fl = o.flags;
  }
  if (blkcnt) {
    const int VB = gridDim.x * gridDim.y;
    const int b = v * gridDim.x + blockIdx.x;
    int c0 = __syncthreads_count((fl & 1u) != 0);
    int c1 = __syncthreads_count((fl & 3u) == 3u);
    int c2 = __syncthreads_count((fl & 7u) == 7u);
    if (threadIdx.x == 0) {
      blkcnt[0 * VB + b] = c0;
      blkcnt[1 * VB + b] = c1;
      blkcnt[2 * VB + b] = c2;
    }
  }
}

// ordered scatter of the three lists; blkoff int64 [3*VB+1] is the exclusive sum of blkcnt.
__global__ void __launch_bounds__(S1_THREADS) k_lists_scatter(const dfu_view* __restrict__ views,
                                                              const uint8_t* __restrict__ flags,
                                                              const uint32_t* __restrict__ label,
                                                              const int64_t* __restrict__ blkoff, int32_t* f_list,
                                                              int32_t* a_list, uint32_t* b_label) {
  __shared__ int wsum[3][S1_THREADS / 32];
  const int v = blockIdx.y;
  const int cnt = views[v].lidar_cnt;
  const int64_t slot_off = views[v].slot_off;
  const int i = blockIdx.x * S1_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned fl = (i < cnt) ? flags[slot_off + i] : 0u;
  // a camera sees about a sixth of the sweep and the rows are in azimuth order: most blocks hold no FOV point at all
  if (!__syncthreads_or((fl & 1u) != 0)) return;
  bool p[3] = {(fl & 1u) != 0, (fl & 3u) == 3u, (fl & 7u) == 7u};
  int pre[3];
#pragma unroll
  for (int j = 0; j < 3; j++) {
    unsigned bal = __ballot_sync(0xffffffffu, p[j]);
    pre[j] = dfu_warp_excl(bal, lane);
    if (lane == 0) wsum[j][w] = __popc(bal);
  }
  __syncthreads();
  const int VB = gridDim.x * gridDim.y;
  const int b = v * gridDim.x + blockIdx.x;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    int base = 0;
    for (int k = 0; k < w; k++) base += wsum[j][k];
    if (p[j]) {
      int64_t pos = blkoff[(int64_t)j * VB + b] - blkoff[(int64_t)j * VB] + base + pre[j];
      if (j == 0) f_list[pos] = i;
      if (j == 1) a_list[pos] = i;
      if (j == 2) b_label[pos] = label[slot_off + i];
    }
  }
}

// per-view counts / offsets out of the block scan: out [3][V+1]
__global__ void k_lists_view_offsets(const int64_t* __restrict__ blkoff, int V, int bpv, int64_t* out) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 3 * (V + 1)) return;
  int j = t / (V + 1), v = t % (V + 1);
  int64_t VB = (int64_t)V * bpv;
  out[t] = blkoff[j * VB + (int64_t)v * bpv] - blkoff[j * VB];
}

// ---------------------------------------------------------------- mask packing
// grid (blocks over Hm*Wm, V); masks of a view are contiguous: [inst_off .. inst_off+n_inst)
template <typename T>
__global__ void __launch_bounds__(256) k_pack_masks(const dfu_view* __restrict__ views, const T* __restrict__ masks,
                                                    int HW, uint32_t* __restrict__ labelmap) {
  const int v = blockIdx.y;
  const int n = views[v].n_inst;
  const int64_t moff = (int64_t)views[v].inst_off * HW;
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p >= HW) return;
  uint32_t w = 0;
  for (int j = 0; j < n; j++) {
    T m = masks[moff + (int64_t)j * HW + p];
    if (m != (T)0) w |= (1u << j);
  }
  labelmap[(int64_t)v * HW + p] = w;
}

// uint8 masks, H*W a multiple of 16: one thread per 16 pixels, 128-bit loads of every mask plane, four 128-bit stores
__global__ void __launch_bounds__(256) k_pack_masks_v16(const dfu_view* __restrict__ views, const uint8_t* __restrict__ masks,
                                                        int HW, uint32_t* __restrict__ labelmap) {
  const int v = blockIdx.y;
  const int n = views[v].n_inst;
  const int64_t moff = (int64_t)views[v].inst_off * HW;
  const int p = (blockIdx.x * 256 + threadIdx.x) * 16;
  if (p >= HW) return;
  uint32_t w[16];
#pragma unroll
  for (int i = 0; i < 16; i++) w[i] = 0;
  for (int j = 0; j < n; j++) {
    const uint4 m = __ldg(reinterpret_cast<const uint4*>(masks + moff + (int64_t)j * HW + p));
    const uint32_t q[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] |= (((q[i >> 2] >> (8 * (i & 3))) & 0xffu) ? 1u : 0u) << j;
  }
  uint4* out = reinterpret_cast<uint4*>(labelmap + (int64_t)v * HW + p);
#pragma unroll
  for (int i = 0; i < 4; i++) out[i] = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
}

// ---------------------------------------------------------------- sparse depth (load_depth_input)
// last point in input order wins on duplicate pixels (numpy fancy assignment, my_loader.py:117):
// pass 1 atomicMax of the row index per pixel, pass 2 writes the depth of the winning row.
__global__ void __launch_bounds__(S1_THREADS) k_sparse_winner(const dfu_view* __restrict__ views, S1Prm q,
                                                              const float4* __restrict__ lidar, int* __restrict__ winner) {
  __shared__ dfu_view vw;
  const int v = blockIdx.y;
  for (int i = threadIdx.x; i < (int)(sizeof(dfu_view) / 4); i += blockDim.x)
    ((uint32_t*)&vw)[i] = ((const uint32_t*)(views + v))[i];
  __syncthreads();
  const int i = blockIdx.x * S1_THREADS + threadIdx.x;
  if (i >= vw.lidar_cnt) return;
  S1Out o = s1_point(vw, q, __ldg(lidar + vw.lidar_off + i), nullptr, 1.0, true);
  // fov implies 0 <= u < W, 0 <= v < H, so the second bounds test of my_loader.py:110-111 is implied
  if (o.flags & 1u) atomicMax(winner + ((int64_t)v * q.H + o.py) * q.W + o.px, i);
}

__global__ void __launch_bounds__(256) k_sparse_write(const dfu_view* __restrict__ views, S1Prm q,
                                                      const float4* __restrict__ lidar, const int* __restrict__ winner,
                                                      double* __restrict__ out) {
  __shared__ dfu_view vw;
  const int v = blockIdx.y;
  for (int i = threadIdx.x; i < (int)(sizeof(dfu_view) / 4); i += blockDim.x)
    ((uint32_t*)&vw)[i] = ((const uint32_t*)(views + v))[i];
  __syncthreads();
  const int p = blockIdx.x * 256 + threadIdx.x;
  if (p >= q.H * q.W) return;
  const int64_t g = (int64_t)v * q.H * q.W + p;
  int wi = winner[g];
  double d = 0.0;
  if (wi >= 0) {
    S1Out o = s1_point(vw, q, __ldg(lidar + vw.lidar_off + wi), nullptr, s1_plane_norm(vw));
    d = (double)o.depth;
  }
  out[g] = d;
}

// ---------------------------------------------------------------- C ABI
extern "C" int dfu_project_label(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_dev, int max_lidar_cnt,
                                 const dfu_fusion_params* prm, const void* lidar_xyzi, const uint32_t* labelmap,
                                 float* out_uv, float* out_depth, int32_t* out_pix, uint8_t* out_flags,
                                 uint32_t* out_label) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !views_dev || !prm || !lidar_xyzi || V <= 0 || max_lidar_cnt < 0)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_project_label: bad argument");
  if (max_lidar_cnt == 0) return DFU_OK;
  dim3 grid((max_lidar_cnt + S1_THREADS - 1) / S1_THREADS, V);
  DFU_PROF(ctx, (cudaStream_t)stream, PT_PROJECT);
  k_project<<<grid, S1_THREADS, 0, (cudaStream_t)stream>>>(views_dev, make_s1(prm), (const float4*)lidar_xyzi, labelmap,
                                                            (float2*)out_uv, out_depth, (int2*)out_pix, out_flags,
                                                            out_label, nullptr);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

extern "C" int dfu_pack_masks(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_dev, const void* masks,
                              int is_f32, int Hm, int Wm, uint32_t* labelmap) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !views_dev || !labelmap || V <= 0 || Hm <= 0 || Wm <= 0)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_pack_masks: bad argument");
  const int HW = Hm * Wm;
  dim3 grid((HW + 255) / 256, V);
  DFU_PROF(ctx, (cudaStream_t)stream, PT_PACK_MASKS);
  if (is_f32)
    k_pack_masks<float><<<grid, 256, 0, (cudaStream_t)stream>>>(views_dev, (const float*)masks, HW, labelmap);
  else if ((HW & 15) == 0 && (((uintptr_t)masks | (uintptr_t)labelmap) & 15) == 0)
    k_pack_masks_v16<<<dim3((HW / 16 + 255) / 256, V), 256, 0, (cudaStream_t)stream>>>(views_dev, (const uint8_t*)masks, HW, labelmap);
  else
    k_pack_masks<uint8_t><<<grid, 256, 0, (cudaStream_t)stream>>>(views_dev, (const uint8_t*)masks, HW, labelmap);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

// compact label maps (1 or 2 bytes per pixel, bit j = instance j) -> the uint32 label words the kernels read;
// 16 pixels per thread, 128-bit loads and stores
template <typename T>
__global__ void __launch_bounds__(256) k_widen_labelmap(const T* __restrict__ src, int64_t n, uint32_t* __restrict__ dst) {
  const int64_t i0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 16;
  if (i0 + 16 <= n && ((((uintptr_t)src + i0 * sizeof(T)) | (uintptr_t)(dst + i0)) & 15) == 0) {
    T v[16];
    const uint4* s4 = reinterpret_cast<const uint4*>(src + i0);
    uint4* vv = reinterpret_cast<uint4*>(v);
#pragma unroll
    for (int j = 0; j < (int)sizeof(T); j++) vv[j] = __ldg(s4 + j);
    uint4* d4 = reinterpret_cast<uint4*>(dst + i0);
#pragma unroll
    for (int j = 0; j < 4; j++) d4[j] = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  } else {
    for (int64_t i = i0; i < i0 + 16 && i < n; i++) dst[i] = (uint32_t)src[i];
  }
}

extern "C" int dfu_widen_labelmap(dfu_ctx* ctx, void* stream, const void* src, int bytes_per_pixel, int64_t n, uint32_t* labelmap) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || n < 0 || (n > 0 && (!src || !labelmap)) || (bytes_per_pixel != 1 && bytes_per_pixel != 2))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_widen_labelmap: bad argument (bytes_per_pixel must be 1 or 2)");
  if (n == 0) return DFU_OK;
  DFU_PROF(ctx, (cudaStream_t)stream, PT_PACK_MASKS);
  const unsigned nb = (unsigned)((n + 4095) / 4096);
  if (bytes_per_pixel == 1) k_widen_labelmap<uint8_t><<<nb, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)src, n, labelmap);
  else k_widen_labelmap<uint16_t><<<nb, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)src, n, labelmap);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

extern "C" int dfu_sparse_depth(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_dev, int max_lidar_cnt,
                                const dfu_fusion_params* prm, const void* lidar_xyzi, double* out_depth) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !views_dev || !prm || !out_depth || V <= 0)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_sparse_depth: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  S1Prm q = make_s1(prm);
  q.round_mode = 1;
  q.no_fov = 0;
  const int64_t npix = (int64_t)V * q.H * q.W;
  DFU_SLOT(ctx, winner, int*, SL_OP0, npix * sizeof(int));
  DFU_PROF(ctx, st, PT_SPARSE_DEPTH);
  DFU_CUDA(ctx, cudaMemsetAsync(winner, 0xFF, npix * sizeof(int), st));
  if (max_lidar_cnt > 0) {
    dim3 g1((max_lidar_cnt + S1_THREADS - 1) / S1_THREADS, V);
    k_sparse_winner<<<g1, S1_THREADS, 0, st>>>(views_dev, q, (const float4*)lidar_xyzi, winner);
    DFU_LAUNCH_CHECK(ctx);
  }
  dim3 g2((q.H * q.W + 255) / 256, V);
  k_sparse_write<<<g2, 256, 0, st>>>(views_dev, q, (const float4*)lidar_xyzi, winner, out_depth);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

// ---------------------------------------------------------------- Calibration methods on caller arrays
// op 0 lidar_to_rect f32 [n][3] -> f32 [n][3]            (calibration_kitti.py:104-112)
// op 1 lidar_to_rect f64 [n][3] -> f64 [n][3]            (same call with a float64 argument, my_loader.py:679)
// op 2 rect_to_img   f32 [n][3] -> f32 [n][2], f32 [n]   (:114-123)
// op 3 img_to_rect   u i64 [n], v i64 [n], d f32 [n] -> f64 [n][3]   (:134-144)
// op 4 rect_to_lidar f64 [n][3] -> f64 [n][3]            (:89-102)
// op 5 lidar_to_img  f32 [n][3] -> f32 [n][2], f32 [n]   (:125-132)
// op 6 get_fov_flag  f32 rect [n][3], in1 = int32 {H, W} -> u8 [n]   (vis_utils.py:108-123 == my_loader.py:85-100)
// op 7 corners3d_to_img_boxes  f64 corners [n][3] -> f64 [n][2] image x, y   (:146-164; min / max over the 8 corners on the host)
__global__ void __launch_bounds__(256) k_calib_apply(dfu_view vw, int op, const void* __restrict__ in0,
                                                     const void* __restrict__ in1, const void* __restrict__ in2, int64_t n,
                                                     void* __restrict__ out0, void* __restrict__ out1) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  if (op == 0 || op == 2 || op == 5 || op == 6) {
    const float* a = (const float*)in0 + 3 * i;
    float x = a[0], y = a[1], z = a[2];
    if (op != 2 && op != 6) {
      const float rx = dfu_row4_f32(x, y, z, 1.0f, vw.M43, 0, 3), ry = dfu_row4_f32(x, y, z, 1.0f, vw.M43, 1, 3),
                  rz = dfu_row4_f32(x, y, z, 1.0f, vw.M43, 2, 3);
      x = rx; y = ry; z = rz;
    }
    if (op == 0) {
      float* o = (float*)out0 + 3 * i;
      o[0] = x; o[1] = y; o[2] = z;
      return;
    }
    const float h0 = dfu_row4_f32(x, y, z, 1.0f, vw.P2T, 0, 3), h1 = dfu_row4_f32(x, y, z, 1.0f, vw.P2T, 1, 3),
                h2 = dfu_row4_f32(x, y, z, 1.0f, vw.P2T, 2, 3);
    const float u = __fdiv_rn(h0, z), v = __fdiv_rn(h1, z), dep = __fsub_rn(h2, vw.P2T[3 * 3 + 2]);
    if (op == 6) {
      const int H = ((const int*)in1)[0], W = ((const int*)in1)[1];
      ((uint8_t*)out0)[i] = (u >= 0.0f && u < (float)W && v >= 0.0f && v < (float)H && dep >= 0.0f) ? 1 : 0;
      return;
    }
    float* o = (float*)out0 + 2 * i;
    o[0] = u;
    o[1] = v;
    ((float*)out1)[i] = dep;
  } else if (op == 1) {
    const double* a = (const double*)in0 + 3 * i;
    double* o = (double*)out0 + 3 * i;
    for (int j = 0; j < 3; j++) o[j] = dfu_row4_f64(a[0], a[1], a[2], 1.0, vw.M43, j, 3);
  } else if (op == 3) {
    const long long u = ((const long long*)in0)[i], v = ((const long long*)in1)[i];
    const double dd = (double)((const float*)in2)[i];
    double* o = (double*)out0 + 3 * i;
    o[0] = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn((double)u, (double)vw.cu), dd), (double)vw.fu), (double)vw.tx);
    o[1] = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn((double)v, (double)vw.cv), dd), (double)vw.fv), (double)vw.ty);
    o[2] = dd;
  } else if (op == 4) {
    const double* a = (const double*)in0 + 3 * i;
    double* o = (double*)out0 + 3 * i;
    for (int j = 0; j < 3; j++) o[j] = dfu_row4_f64(a[0], a[1], a[2], 1.0, vw.inv44, j, 4);
  } else if (op == 7) {
    // corners3d_to_img_boxes (:146-164): float64 corner [x y z 1] . P2^T, then x / w, y / w (divides by the HOMOGENEOUS
    // coordinate, unlike rect_to_img)
    const double* a = (const double*)in0 + 3 * i;
    double* o = (double*)out0 + 2 * i;
    const double h0 = dfu_row4_f64(a[0], a[1], a[2], 1.0, vw.P2T, 0, 3), h1 = dfu_row4_f64(a[0], a[1], a[2], 1.0, vw.P2T, 1, 3),
                 h2 = dfu_row4_f64(a[0], a[1], a[2], 1.0, vw.P2T, 2, 3);
    o[0] = __ddiv_rn(h0, h2);
    o[1] = __ddiv_rn(h1, h2);
  }
}

extern "C" int dfu_calib_apply(dfu_ctx* ctx, void* stream, const dfu_view* view_host, int op, const void* in0,
                               const void* in1, const void* in2, int64_t n, void* out0, void* out1) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !view_host || op < 0 || op > 7 || n < 0 || (n > 0 && (!in0 || !out0)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_calib_apply: bad argument");
  if (n > 0 && (op == 2 || op == 5) && !out1) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_calib_apply: op needs out1");
  if (n > 0 && op == 6 && !in1) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_calib_apply: op 6 needs the image shape");
  if (n > 0 && op == 3 && (!in1 || !in2)) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_calib_apply: op 3 needs u, v, depth");
  if (n == 0) return DFU_OK;
  DFU_PROF(ctx, (cudaStream_t)stream, PT_PROJECT);
  k_calib_apply<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*view_host, op, in0, in1, in2, n, out0, out1);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

// ---------------------------------------------------------------- batched pipeline: begin + project
extern "C" int dfu_fuse_begin(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_host,
                              const dfu_fusion_params* prm, const void* lidar_xyzi, const float* depth,
                              const void* image, const uint32_t* labelmap) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || V <= 0 || !views_host || !prm || !lidar_xyzi || !depth)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_begin: bad argument");
  if (prm->H <= 0 || prm->W <= 0 || prm->Hm < prm->H || prm->Wm < prm->W)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_begin: need 0 < H <= Hm and 0 < W <= Wm");
  ctx->stream = (cudaStream_t)stream;
  ctx->V = V;
  ctx->prm = *prm;
  ctx->lidar = lidar_xyzi;
  ctx->depth = depth;
  ctx->image = image;
  ctx->labelmap = labelmap;
  ctx->views_host = (dfu_view*)realloc(ctx->views_host, sizeof(dfu_view) * V);
  memcpy(ctx->views_host, views_host, sizeof(dfu_view) * V);
  int mx = 0, ninst = 0;
  int64_t slots = 0;
  for (int v = 0; v < V; v++) {
    dfu_view& w = ctx->views_host[v];
    if (w.n_inst < 0 || w.n_inst > DFU_MAX_INSTANCES)
      return dfu_set_err(ctx, DFU_ERR_DATA, "view %d has %d instances (max %d)", v, w.n_inst, DFU_MAX_INSTANCES);
    if (w.n_inst > 0 && !labelmap) return dfu_set_err(ctx, DFU_ERR_ARG, "instances present but labelmap is NULL");
    w.slot_off = slots;
    w.inst_off = ninst;
    slots += w.lidar_cnt;
    ninst += w.n_inst;
    if (w.lidar_cnt > mx) mx = w.lidar_cnt;
  }
  ctx->max_lidar_cnt = mx;
  ctx->total_slots = slots;
  ctx->total_inst = ninst;
  DFU_SLOT(ctx, vd, dfu_view*, SL_VIEWS, sizeof(dfu_view) * V);
  // pageable -> device async copy stages through the driver; views_host is our own stable copy
  DFU_CUDA(ctx, cudaMemcpyAsync(vd, ctx->views_host, sizeof(dfu_view) * V, cudaMemcpyHostToDevice, ctx->stream));
  ctx->cnt_project = (int64_t*)realloc(ctx->cnt_project, sizeof(int64_t) * 3 * V);
  ctx->off_project = (int64_t*)realloc(ctx->off_project, sizeof(int64_t) * 3 * (V + 1));
  ctx->cnt_vox = (int64_t*)realloc(ctx->cnt_vox, sizeof(int64_t) * V);
  ctx->off_vox = (int64_t*)realloc(ctx->off_vox, sizeof(int64_t) * (V + 1));
  ctx->stage = 1;  // bit0 bound, bit1 projected, bit2 densified
  return DFU_OK;
}

extern "C" int dfu_fuse_project_sync(dfu_ctx* ctx, int64_t* counts_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !(ctx->stage & 1)) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_project_sync: no batch bound");
  const int V = ctx->V;
  cudaStream_t st = ctx->stream;
  const int bpv = (ctx->max_lidar_cnt + S1_THREADS - 1) / S1_THREADS > 0 ? (ctx->max_lidar_cnt + S1_THREADS - 1) / S1_THREADS : 1;
  ctx->bpv_pts = bpv;
  const int64_t VB = (int64_t)V * bpv;
  const int64_t ns = ctx->total_slots > 0 ? ctx->total_slots : 1;
  dfu_view* vd = (dfu_view*)ctx->slot[SL_VIEWS];
  DFU_SLOT(ctx, flags, uint8_t*, SL_FLAGS, ns);
  DFU_SLOT(ctx, pix, int2*, SL_PIX, ns * sizeof(int2));
  DFU_SLOT(ctx, label, uint32_t*, SL_LABEL, ns * sizeof(uint32_t));
  DFU_SLOT(ctx, blkcnt, int32_t*, SL_BLKCNT, 3 * VB * sizeof(int32_t));
  DFU_SLOT(ctx, blkoff, int64_t*, SL_BLKOFF, (3 * VB + 1) * sizeof(int64_t));
  DFU_SLOT(ctx, flist, int32_t*, SL_FLIST, ns * sizeof(int32_t));
  DFU_SLOT(ctx, alist, int32_t*, SL_ALIST, ns * sizeof(int32_t));
  DFU_SLOT(ctx, blabel, uint32_t*, SL_BLABEL, ns * sizeof(uint32_t));
  DFU_SLOT(ctx, voff, int64_t*, SL_STATS, 3 * (V + 1) * sizeof(int64_t) + 64);
  dim3 grid(bpv, V);
  {
    DFU_PROF(ctx, st, PT_PROJECT);
    k_project<<<grid, S1_THREADS, 0, st>>>(vd, make_s1(&ctx->prm), (const float4*)ctx->lidar, ctx->labelmap, nullptr,
                                           nullptr, pix, flags, label, blkcnt);
    DFU_LAUNCH_CHECK(ctx);
  }
  {
    DFU_PROF(ctx, st, PT_LISTS);
    int rc = dfu_exclusive_sum_i32(ctx, st, blkcnt, blkoff, 3 * VB);
    if (rc) return rc;
    k_lists_scatter<<<grid, S1_THREADS, 0, st>>>(vd, flags, label, blkoff, flist, alist, blabel);
    DFU_LAUNCH_CHECK(ctx);
    k_lists_view_offsets<<<(3 * (V + 1) + 127) / 128, 128, 0, st>>>(blkoff, V, bpv, voff);
    DFU_LAUNCH_CHECK(ctx);
  }
  int64_t* pin = (int64_t*)dfu_pinned(ctx, 3 * (V + 1) * sizeof(int64_t));
  if (!pin) return DFU_ERR_CUDA;
  DFU_CUDA(ctx, cudaMemcpyAsync(pin, voff, 3 * (V + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  memcpy(ctx->off_project, pin, 3 * (V + 1) * sizeof(int64_t));
  for (int v = 0; v < V; v++)
    for (int j = 0; j < 3; j++) {
      ctx->cnt_project[v * 3 + j] = ctx->off_project[j * (V + 1) + v + 1] - ctx->off_project[j * (V + 1) + v];
      if (counts_host) counts_host[v * 3 + j] = ctx->cnt_project[v * 3 + j];
    }
  ctx->stage |= 2;
  return DFU_OK;
}

extern "C" int dfu_fuse_emit_project(dfu_ctx* ctx, int32_t* out_pix, uint8_t* out_flags, uint32_t* out_label) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !(ctx->stage & 2)) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_emit_project: project stage not run");
  const int64_t ns = ctx->total_slots;
  cudaStream_t st = ctx->stream;
  if (out_pix) DFU_CUDA(ctx, cudaMemcpyAsync(out_pix, ctx->slot[SL_PIX], ns * sizeof(int2), cudaMemcpyDeviceToDevice, st));
  if (out_flags) DFU_CUDA(ctx, cudaMemcpyAsync(out_flags, ctx->slot[SL_FLAGS], ns, cudaMemcpyDeviceToDevice, st));
  if (out_label) DFU_CUDA(ctx, cudaMemcpyAsync(out_label, ctx->slot[SL_LABEL], ns * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
  return DFU_OK;
}
