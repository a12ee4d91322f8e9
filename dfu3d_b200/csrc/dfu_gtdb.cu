// GT-database tooling: points-in-rotated-boxes and the per-box crop that
// pcdet/datasets/kitti/kitti_dataset.py:284-338 (create_groundtruth_database) performs with
// roiaware_pool3d_utils.points_in_boxes_cpu, plus the concatenation + offsets of
// tools/process_tools/create_integrated_database.py:9-55 — here one ordered compaction on the GPU.
// Arithmetic of the test: pcdet/ops/roiaware_pool3d/src/roiaware_pool3d.cpp:121-140 (float32 rotation by -heading,
// comparisons promoted to double, MARGIN = 1e-2f on x and y only).
#include "dfu_common.cuh"

#define GT_THREADS 256
// own scratch slots: a crop between two stages of a bound fusion batch must not touch the pipeline's scratch
enum { GT_BLKCNT = 210, GT_BLKOFF, GT_BOXOFF };
static_assert(GT_BOXOFF < DFU_NSLOTS, "slot ids");

// Per-box constants.  The reference compares float values against double thresholds (`fabs(z - cz) > dz / 2.0`,
// `fabs(local_x) < dx / 2.0 + MARGIN`): for a float a and a double T,  a > T <=> a > round_down_to_float(T)  and
// a < T <=> a < round_up_to_float(T), so the thresholds are rounded once per box and the per-pair test is all float32
// with identical outcomes.
struct GtBox { float cx, cy, cz, hx, hy, hz, cosa, sina; };

__device__ __forceinline__ GtBox gt_load_box(const float* __restrict__ b) {
  GtBox g;
  g.cx = b[0]; g.cy = b[1]; g.cz = b[2];
  const double margin = (double)1e-2f;                               // roiaware_pool3d.cpp:126 `float MARGIN = 1e-2`
  g.hx = __double2float_ru((double)b[3] / 2.0 + margin);             // :138
  g.hy = __double2float_ru((double)b[4] / 2.0 + margin);
  g.hz = __double2float_rd((double)b[5] / 2.0);                      // :136
  // :122 `cos(-rot_angle)` of the host libm is correctly rounded in practice (double cos, or glibc cosf at < 0.6 ulp);
  // CUDA's cosf is up to 2 ulp off, so the once-per-box value goes through the double routine and is rounded once
  g.cosa = (float)cos((double)(-b[6]));
  g.sina = (float)sin((double)(-b[6]));
  return g;
}

__device__ __forceinline__ bool gt_inside(const GtBox& g, float x, float y, float z) {
  if (fabsf(__fsub_rn(z, g.cz)) > g.hz) return false;                                              // :136
  const float sx = __fsub_rn(x, g.cx), sy = __fsub_rn(y, g.cy);
  const float lx = __fadd_rn(__fmul_rn(sx, g.cosa), __fmul_rn(sy, -g.sina));                       // :123
  const float ly = __fadd_rn(__fmul_rn(sx, g.sina), __fmul_rn(sy, g.cosa));                        // :124
  return (fabsf(lx) < g.hx) && (fabsf(ly) < g.hy);                                                 // :138
}

// One thread per point, the point read ONCE; a block walks a chunk of GT_CHUNK boxes held in shared memory (cos/sin
// evaluated once per box and block).  grid (blocks over N, chunks over M).  mask [M][N] (nullable) is written one
// coalesced row segment per box; blkcnt [M * gridDim.x] (nullable) gets the per-(box, block) hit counts.
#define GT_CHUNK 64
#define GT_WARPS (GT_THREADS / 32)

__global__ void __launch_bounds__(GT_THREADS) k_points_in_boxes(const float* __restrict__ pts, int C, int64_t n,
                                                                const float* __restrict__ boxes, int M,
                                                                int32_t* __restrict__ mask, int32_t* __restrict__ blkcnt) {
  __shared__ GtBox sb[GT_CHUNK];
  __shared__ int wcnt[GT_CHUNK][GT_WARPS];
  const int b0 = blockIdx.y * GT_CHUNK, nb = min(GT_CHUNK, M - b0);
  if (threadIdx.x < nb) sb[threadIdx.x] = gt_load_box(boxes + 7 * (int64_t)(b0 + threadIdx.x));
  __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * GT_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const bool valid = i < n;
  float x = 0.f, y = 0.f, z = 0.f;
  if (valid) { const float* p = pts + i * C; x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2); }
  for (int k = 0; k < nb; k++) {
    const bool in = valid && gt_inside(sb[k], x, y, z);
    if (mask && valid) __stcs(mask + (int64_t)(b0 + k) * n + i, in ? 1 : 0);
    if (blkcnt) {
      const unsigned bal = __ballot_sync(0xffffffffu, in);
      if (lane == 0) wcnt[k][w] = __popc(bal);
    }
  }
  if (blkcnt) {
    __syncthreads();
    if (threadIdx.x < nb) {
      int c = 0;
#pragma unroll
      for (int q = 0; q < GT_WARPS; q++) c += wcnt[threadIdx.x][q];
      blkcnt[(int64_t)(b0 + threadIdx.x) * gridDim.x + blockIdx.x] = c;
    }
  }
}

// ordered scatter: rows of box b, in point order, at blkoff[b * nblk + blk] + rank; xyz shifted by the box centre
// (kitti_dataset.py:318 `gt_points[:, :3] -= gt_boxes[i, :3]`)
__global__ void __launch_bounds__(GT_THREADS) k_crop_scatter(const float* __restrict__ pts, int C, int64_t n,
                                                             const float* __restrict__ boxes, int M, const double* __restrict__ c64,
                                                             const int64_t* __restrict__ blkoff, float* __restrict__ out) {
  __shared__ GtBox sb[GT_CHUNK];
  __shared__ int wcnt[GT_CHUNK][GT_WARPS];
  const int b0 = blockIdx.y * GT_CHUNK, nb = min(GT_CHUNK, M - b0);
  if (threadIdx.x < nb) sb[threadIdx.x] = gt_load_box(boxes + 7 * (int64_t)(b0 + threadIdx.x));
  __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * GT_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const bool valid = i < n;
  const float* p = pts + (valid ? i : 0) * C;
  float x = 0.f, y = 0.f, z = 0.f;
  if (valid) { x = __ldg(p); y = __ldg(p + 1); z = __ldg(p + 2); }
  unsigned long long bits = 0ull;
  for (int k = 0; k < nb; k++) {
    const bool in = valid && gt_inside(sb[k], x, y, z);
    bits |= (unsigned long long)in << k;
    const unsigned bal = __ballot_sync(0xffffffffu, in);
    if (lane == 0) wcnt[k][w] = __popc(bal);
  }
  __syncthreads();
  for (int k = 0; k < nb; k++) {
    const bool in = (bits >> k) & 1ull;
    const unsigned bal = __ballot_sync(0xffffffffu, in);
    if (!in) continue;
    int base = 0;
    for (int q = 0; q < w; q++) base += wcnt[k][q];
    const int b = b0 + k;
    const int64_t pos = blkoff[(int64_t)b * gridDim.x + blockIdx.x] + base + dfu_warp_excl(bal, lane);
    float* o = out + pos * C;
    if (c64) {   // float64 annotation boxes: numpy's in-place `float32 -= float64` computes in double, then rounds once
      o[0] = (float)((double)x - c64[3 * (int64_t)b]); o[1] = (float)((double)y - c64[3 * (int64_t)b + 1]);
      o[2] = (float)((double)z - c64[3 * (int64_t)b + 2]);
    } else {
      o[0] = __fsub_rn(x, sb[k].cx); o[1] = __fsub_rn(y, sb[k].cy); o[2] = __fsub_rn(z, sb[k].cz);
    }
    for (int c = 3; c < C; c++) o[c] = __ldg(p + c);
  }
}

__global__ void k_crop_box_offsets(const int64_t* __restrict__ blkoff, int M, int nblk, int64_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b <= M) out[b] = blkoff[(int64_t)b * nblk];
}

extern "C" int dfu_points_in_boxes(dfu_ctx* ctx, void* stream, const float* points, int C, int64_t n, const float* boxes,
                                   int M, int32_t* out_mask) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || C < 3 || n < 0 || M < 0 || (n > 0 && M > 0 && (!points || !boxes || !out_mask)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_points_in_boxes: bad argument");
  if (n == 0 || M == 0) return DFU_OK;
  dim3 grid((unsigned)((n + GT_THREADS - 1) / GT_THREADS), (M + GT_CHUNK - 1) / GT_CHUNK);
  DFU_PROF(ctx, (cudaStream_t)stream, PT_MISC);
  k_points_in_boxes<<<grid, GT_THREADS, 0, (cudaStream_t)stream>>>(points, C, n, boxes, M, out_mask, nullptr);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

extern "C" int dfu_crop_boxes_sync(dfu_ctx* ctx, void* stream, const float* points, int C, int64_t n, const float* boxes,
                                   int M, const double* centres64, float* out_rows, int64_t capacity,
                                   int64_t* offsets_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || C < 3 || n < 0 || M < 0 || !offsets_host || (n > 0 && M > 0 && (!points || !boxes)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_crop_boxes_sync: bad argument");
  for (int b = 0; b <= M; b++) offsets_host[b] = 0;
  if (n == 0 || M == 0) return DFU_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int nblk = (int)((n + GT_THREADS - 1) / GT_THREADS);
  const int64_t nb = (int64_t)M * nblk;
  DFU_SLOT(ctx, blkcnt, int32_t*, GT_BLKCNT, nb * sizeof(int32_t));
  DFU_SLOT(ctx, blkoff, int64_t*, GT_BLKOFF, (nb + 1) * sizeof(int64_t));
  DFU_SLOT(ctx, boff, int64_t*, GT_BOXOFF, (size_t)(M + 1) * sizeof(int64_t));
  dim3 grid(nblk, (M + GT_CHUNK - 1) / GT_CHUNK);
  k_points_in_boxes<<<grid, GT_THREADS, 0, st>>>(points, C, n, boxes, M, nullptr, blkcnt);
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_exclusive_sum_i32(ctx, st, blkcnt, blkoff, nb);
  if (rc) return rc;
  k_crop_box_offsets<<<(M + 1 + 127) / 128, 128, 0, st>>>(blkoff, M, nblk, boff);
  DFU_LAUNCH_CHECK(ctx);
  int64_t* pin = (int64_t*)dfu_pinned(ctx, (size_t)(M + 1) * sizeof(int64_t));
  if (!pin) return DFU_ERR_CUDA;
  DFU_CUDA(ctx, cudaMemcpyAsync(pin, boff, (size_t)(M + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  memcpy(offsets_host, pin, (size_t)(M + 1) * sizeof(int64_t));
  const int64_t total = pin[M];
  if (!out_rows) return DFU_OK;    // size query
  if (capacity < total)
    return dfu_set_err(ctx, DFU_ERR_CAPACITY, "crop needs %lld rows, capacity %lld", (long long)total, (long long)capacity);
  if (total == 0) return DFU_OK;
  k_crop_scatter<<<grid, GT_THREADS, 0, st>>>(points, C, n, boxes, M, centres64, blkoff, out_rows);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}
