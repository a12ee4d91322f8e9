// Consumer side of the 8-column cloud: modality split + range mask and the hard voxelisation that
// feed the detector, so that the dense cloud can stay on the device from the generator to the network input.
//   split:     pcdet/datasets/dataset_virtual.py:270-296 (`points[:, -1] == 1` / `== 2`),
//              pcdet/datasets/processor/data_processor_virtual.py:69-77 + pcdet/utils/common_utils.py:78-81
//              (mask_points_by_range: x and y only, inclusive on both sides)
//   voxelise:  data_processor_virtual.py:133-185 (transform_points_to_voxels -> VoxelGeneratorWrapper.generate, :16-58);
//              spconv v1 points_to_voxel semantics as restated in oracle/fakes.py:voxel_assign (spconv itself is not
//              in the reference tree: parity unpinned) — float32 points are binned with float32 arithmetic
//              c_j = floor((p_j - lo_j) / vsize_j); a point outside [0, grid_j) on any axis is dropped; voxels are
//              numbered in first-appearance order; a voxel keeps its first max_points points in input order; points
//              that would open voxel number max_voxels or later are dropped; coordinates are (z, y, x).
// The CPU routine is a sequential loop with a dense coor_to_voxelidx array.  Here: one stable radix sort of the linear
// cell ids (ties stay in input order, so a segment's head is the voxel's opening point), a scan over the opening flags
// in POINT order (= first-appearance numbering), a max-scan for the segment starts, one scatter.
#include <cub/cub.cuh>
#include "dfu_common.cuh"

#define CS_THREADS 256
#define VOX_NONE 0xFFFFFFFFu
// own scratch slots (the pipeline's scratch of a bound fusion batch stays intact)
enum { CS_BLKCNT = 214, CS_BLKOFF, CS_RANGE };

// ---------------------------------------------------------------- split
// pass 0: per-block counts of the two modalities; pass 1: ordered scatter
__device__ __forceinline__ int cs_class(const float* __restrict__ p, int C, const float* __restrict__ range) {
  const float tag = p[C - 1];
  int cls = (tag == 2.0f) ? 0 : (tag == 1.0f) ? 1 : -1;          // 0 = LiDAR rows, 1 = pseudo (camera) rows
  if (cls >= 0 && range) {
    const float x = p[0], y = p[1];
    if (!(x >= range[0] && x <= range[3] && y >= range[1] && y <= range[4])) cls = -1;
  }
  return cls;
}

__global__ void __launch_bounds__(CS_THREADS) k_split_count(const float* __restrict__ cloud, int C, int64_t n,
                                                            const float* __restrict__ range, int32_t* __restrict__ blkcnt) {
  const int64_t i = (int64_t)blockIdx.x * CS_THREADS + threadIdx.x;
  const int cls = (i < n) ? cs_class(cloud + i * C, C, range) : -1;
  const int n0 = __syncthreads_count(cls == 0), n1 = __syncthreads_count(cls == 1);
  if (threadIdx.x == 0) { blkcnt[blockIdx.x] = n0; blkcnt[gridDim.x + blockIdx.x] = n1; }
}

__global__ void __launch_bounds__(CS_THREADS) k_split_scatter(const float* __restrict__ cloud, int C, int64_t n,
                                                              const float* __restrict__ range,
                                                              const int64_t* __restrict__ blkoff, float* __restrict__ out0,
                                                              float* __restrict__ out1) {
  __shared__ int wsum[2][CS_THREADS / 32];
  const int64_t i = (int64_t)blockIdx.x * CS_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int cls = (i < n) ? cs_class(cloud + i * C, C, range) : -1;
  const unsigned b0 = __ballot_sync(0xffffffffu, cls == 0), b1 = __ballot_sync(0xffffffffu, cls == 1);
  if (lane == 0) { wsum[0][w] = __popc(b0); wsum[1][w] = __popc(b1); }
  __syncthreads();
  if (cls < 0) return;
  int base = 0;
  for (int k = 0; k < w; k++) base += wsum[cls][k];
  const int64_t first1 = blkoff[gridDim.x];          // rows of class 0 in total = offset of class 1 in the joint scan
  const int64_t pos = cls == 0 ? blkoff[blockIdx.x] + base + dfu_warp_excl(b0, lane)
                               : blkoff[gridDim.x + blockIdx.x] - first1 + base + dfu_warp_excl(b1, lane);
  float* o = (cls == 0 ? out0 : out1) + pos * C;
  const float* p = cloud + i * C;
  for (int c = 0; c < C; c++) o[c] = p[c];
}

extern "C" int dfu_split_cloud_sync(dfu_ctx* ctx, void* stream, const float* cloud, int C, int64_t n, const float* range_host,
                                    float* out_lidar, float* out_pseudo, int64_t* counts_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || C < 3 || n < 0 || !counts_host || (n > 0 && (!cloud || !out_lidar || !out_pseudo)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_split_cloud_sync: bad argument");
  counts_host[0] = counts_host[1] = 0;
  if (n == 0) return DFU_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t nblk = (n + CS_THREADS - 1) / CS_THREADS;
  if (nblk > 0x3fffffff) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_split_cloud_sync: too many rows");
  DFU_SLOT(ctx, blkcnt, int32_t*, CS_BLKCNT, 2 * nblk * sizeof(int32_t));
  DFU_SLOT(ctx, blkoff, int64_t*, CS_BLKOFF, (2 * nblk + 1) * sizeof(int64_t));
  float* range = nullptr;
  if (range_host) {
    DFU_SLOT(ctx, rd, float*, CS_RANGE, 6 * sizeof(float));
    DFU_CUDA(ctx, cudaMemcpyAsync(rd, range_host, 6 * sizeof(float), cudaMemcpyHostToDevice, st));
    range = rd;
  }
  DFU_PROF(ctx, st, PT_MISC);
  k_split_count<<<(unsigned)nblk, CS_THREADS, 0, st>>>(cloud, C, n, range, blkcnt);
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_exclusive_sum_i32(ctx, st, blkcnt, blkoff, 2 * nblk);
  if (rc) return rc;
  k_split_scatter<<<(unsigned)nblk, CS_THREADS, 0, st>>>(cloud, C, n, range, blkoff, out_lidar, out_pseudo);
  DFU_LAUNCH_CHECK(ctx);
  int64_t* pin = (int64_t*)dfu_pinned(ctx, 2 * sizeof(int64_t));
  if (!pin) return DFU_ERR_CUDA;
  DFU_CUDA(ctx, cudaMemcpyAsync(pin, blkoff + nblk, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(pin + 1, blkoff + 2 * nblk, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));   // also covers the pageable range_host copy
  counts_host[0] = pin[0];
  counts_host[1] = pin[1] - pin[0];
  return DFU_OK;
}

// ---------------------------------------------------------------- hard voxelisation
enum { CV_KEY = 200, CV_IDX, CV_SKEY, CV_SIDX, CV_OPENS, CV_RANK, CV_HEAD, CV_TMP, CV_PRM };
static_assert(CV_PRM < DFU_NSLOTS, "slot ids");

struct VoxPrm { float lo[3], vs[3]; int grid[3]; };

__global__ void __launch_bounds__(CS_THREADS) k_vox_keys(const float* __restrict__ pts, int C, int n, VoxPrm q,
                                                         unsigned* __restrict__ key, int32_t* __restrict__ idx,
                                                         int32_t* __restrict__ opens) {
  const int i = blockIdx.x * CS_THREADS + threadIdx.x;
  if (i >= n) return;
  const float* p = pts + (int64_t)i * C;
  unsigned lin = 0;
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const float c = floorf(__fdiv_rn(__fsub_rn(p[j], q.lo[j]), q.vs[j]));    // float32, as numpy / spconv evaluate it
    ok = ok && (c >= 0.0f) && (c < (float)q.grid[j]);
    lin = lin * (unsigned)q.grid[j] + (ok ? (unsigned)(int)c : 0u);
  }
  key[i] = ok ? lin : VOX_NONE;
  idx[i] = i;
  opens[i] = 0;
}

// sorted position j: segment heads (the head of a stable sort is the voxel's first point in input order)
__global__ void __launch_bounds__(CS_THREADS) k_vox_heads(const unsigned* __restrict__ skey, const int32_t* __restrict__ sidx,
                                                          int n, int32_t* __restrict__ headpos, int32_t* __restrict__ opens) {
  const int j = blockIdx.x * CS_THREADS + threadIdx.x;
  if (j >= n) return;
  const unsigned k = skey[j];
  const bool head = (k != VOX_NONE) && (j == 0 || skey[j - 1] != k);
  headpos[j] = head ? j : 0;
  if (head) opens[sidx[j]] = 1;
}

struct MaxOp { __device__ __forceinline__ int32_t operator()(int32_t a, int32_t b) const { return a > b ? a : b; } };

__global__ void __launch_bounds__(CS_THREADS) k_vox_scatter(const float* __restrict__ pts, int C, int n, VoxPrm q,
                                                            const unsigned* __restrict__ skey, const int32_t* __restrict__ sidx,
                                                            const int32_t* __restrict__ headpos, const int64_t* __restrict__ openrank,
                                                            int max_points, int max_voxels, float* __restrict__ voxels,
                                                            int32_t* __restrict__ coords, int32_t* __restrict__ num_points) {
  const int j = blockIdx.x * CS_THREADS + threadIdx.x;
  if (j >= n) return;
  const unsigned k = skey[j];
  if (k == VOX_NONE) return;
  const int h = headpos[j];
  const int64_t vid = openrank[sidx[h]];
  if (vid >= max_voxels) return;                       // the voxel would have been number max_voxels or later: dropped
  const int rank = j - h;
  if (rank < max_points) {
    const float* p = pts + (int64_t)sidx[j] * C;
    float* o = voxels + (vid * max_points + rank) * C;
    for (int c = 0; c < C; c++) o[c] = p[c];
  }
  if (rank == 0) {
    const unsigned cz = k % (unsigned)q.grid[2], r = k / (unsigned)q.grid[2];
    coords[vid * 3 + 0] = (int)cz;
    coords[vid * 3 + 1] = (int)(r % (unsigned)q.grid[1]);
    coords[vid * 3 + 2] = (int)(r / (unsigned)q.grid[1]);
  }
  if (j + 1 == n || skey[j + 1] != k) num_points[vid] = rank + 1 < max_points ? rank + 1 : max_points;
}

extern "C" int dfu_voxelize_sync(dfu_ctx* ctx, void* stream, const float* points, int C, int64_t n,
                                 const float* voxel_size_host, const float* range_host, int max_points, int max_voxels,
                                 float* voxels, int32_t* coords, int32_t* num_points, int64_t* nvox_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || C < 3 || n < 0 || !voxel_size_host || !range_host || max_points <= 0 || max_voxels <= 0 || !voxels || !coords ||
      !num_points || !nvox_host || (n > 0 && !points))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_voxelize_sync: bad argument");
  if (n >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_voxelize_sync: too many points");
  cudaStream_t st = (cudaStream_t)stream;
  VoxPrm q;
  double cells = 1.0;
  for (int j = 0; j < 3; j++) {
    q.lo[j] = range_host[j];
    q.vs[j] = voxel_size_host[j];
    // grid_size = round((hi - lo) / vsize) on the float32 constants (numpy rounds half to even)
    q.grid[j] = (int)nearbyintf((range_host[3 + j] - range_host[j]) / voxel_size_host[j]);
    if (q.grid[j] <= 0) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_voxelize_sync: empty grid on axis %d", j);
    cells *= (double)q.grid[j];
  }
  if (cells >= 4294967295.0) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_voxelize_sync: more than 2^32 - 1 cells");
  *nvox_host = 0;
  DFU_PROF(ctx, st, PT_MISC);
  DFU_CUDA(ctx, cudaMemsetAsync(voxels, 0, (size_t)max_voxels * max_points * C * sizeof(float), st));
  DFU_CUDA(ctx, cudaMemsetAsync(num_points, 0, (size_t)max_voxels * sizeof(int32_t), st));
  DFU_CUDA(ctx, cudaMemsetAsync(coords, 0, (size_t)max_voxels * 3 * sizeof(int32_t), st));
  if (n == 0) return DFU_OK;
  const int N = (int)n;
  DFU_SLOT(ctx, key, unsigned*, CV_KEY, n * sizeof(unsigned));
  DFU_SLOT(ctx, idx, int32_t*, CV_IDX, n * sizeof(int32_t));
  DFU_SLOT(ctx, skey, unsigned*, CV_SKEY, n * sizeof(unsigned));
  DFU_SLOT(ctx, sidx, int32_t*, CV_SIDX, n * sizeof(int32_t));
  DFU_SLOT(ctx, opens, int32_t*, CV_OPENS, n * sizeof(int32_t));
  DFU_SLOT(ctx, openrank, int64_t*, CV_RANK, (n + 1) * sizeof(int64_t));
  DFU_SLOT(ctx, headpos, int32_t*, CV_HEAD, n * sizeof(int32_t));
  const unsigned nb = (unsigned)((n + CS_THREADS - 1) / CS_THREADS);
  k_vox_keys<<<nb, CS_THREADS, 0, st>>>(points, C, N, q, key, idx, opens);
  DFU_LAUNCH_CHECK(ctx);
  size_t tb = 0, tb2 = 0;
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, skey, idx, sidx, N, 0, 32, st));
  DFU_CUDA(ctx, cub::DeviceScan::InclusiveScan(nullptr, tb2, headpos, headpos, MaxOp(), N, st));
  if (tb2 > tb) tb = tb2;
  DFU_SLOT(ctx, tmp, void*, CV_TMP, tb);
  DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, skey, idx, sidx, N, 0, 32, st));   // stable
  k_vox_heads<<<nb, CS_THREADS, 0, st>>>(skey, sidx, N, headpos, opens);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cub::DeviceScan::InclusiveScan(tmp, tb, headpos, headpos, MaxOp(), N, st));
  ctx->launches += 10;
  int rc = dfu_exclusive_sum_i32(ctx, st, opens, openrank, n);
  if (rc) return rc;
  k_vox_scatter<<<nb, CS_THREADS, 0, st>>>(points, C, N, q, skey, sidx, headpos, openrank, max_points, max_voxels, voxels,
                                           coords, num_points);
  DFU_LAUNCH_CHECK(ctx);
  int64_t* pin = (int64_t*)dfu_pinned(ctx, sizeof(int64_t));
  if (!pin) return DFU_ERR_CUDA;
  DFU_CUDA(ctx, cudaMemcpyAsync(pin, openrank + n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  *nvox_host = pin[0] < max_voxels ? pin[0] : max_voxels;
  return DFU_OK;
}
