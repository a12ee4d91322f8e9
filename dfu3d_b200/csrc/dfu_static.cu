// Allocation-free variant of the modality split (dfu_split_cloud_sync, dfu_consumer.cu) for callers that cannot call
// cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid allocation.  It has NO counterpart
// in the reference.
//
// The scratch of the split (per-block counts of the two modalities and their scanned offsets) lives in ONE file-scope
// `static __device__` array of fixed size (256 KB for DFU_SPLIT_STATIC_MAX_ROWS = 4 194 304 rows); every GPU that loads the
// library holds it until the process exits.  In host code the array's NAME is a host-side stand-in, not a device
// address, so the host resolves the address once per device with cudaGetSymbolAddress (that device current) and hands
// the pointer to the kernels.  The entry point calls no cudaMalloc / cudaFree / cub / torch routine, does not touch a
// dfu_ctx scratch slot or pinned buffer, does not synchronise and copies nothing to the host: the two row counts go to a
// caller-owned DEVICE buffer.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
// zero: k_ss_count overwrites every word that k_ss_scan / k_ss_scatter of the same call read (entries [0, 2*nblk)), and
// nothing beyond them is read.  Two calls that OVERLAP on one GPU (different streams, or replays of two captured graphs
// at once) race on the array: serialise them (one stream, or an event between them).
#include <mutex>
#include "dfu_common.cuh"

#define SS_THREADS 256
#define SS_SCAN_THREADS 1024
#define SS_MAX_ROWS DFU_SPLIT_STATIC_MAX_ROWS
#define SS_MAX_BLK (SS_MAX_ROWS / SS_THREADS)          // 16384 blocks of 256 rows
#define SS_MAX_DEVICES 64

struct SsScratch {
  int32_t cnt[2 * SS_MAX_BLK];      // [0, nblk): LiDAR rows of each block, [nblk, 2 nblk): pseudo rows
  int32_t off[2 * SS_MAX_BLK];      // exclusive scan of cnt over the joint [0, 2 nblk) sequence
};
static_assert(SS_MAX_ROWS % SS_THREADS == 0, "whole blocks");
static_assert(2 * SS_MAX_BLK <= SS_SCAN_THREADS * 32, "k_ss_scan: at most 32 items per thread");

static __device__ __align__(256) unsigned char g_split_scratch[sizeof(SsScratch)];

struct SsRange { float lo_x, lo_y, hi_x, hi_y; int on; };

// same classification as cs_class of dfu_consumer.cu (dataset_virtual.py:270-296 tags, mask_points_by_range on x / y)
__device__ __forceinline__ int ss_class(const float* __restrict__ p, int C, const SsRange& r) {
  const float tag = p[C - 1];
  int cls = (tag == 2.0f) ? 0 : (tag == 1.0f) ? 1 : -1;
  if (cls >= 0 && r.on) {
    const float x = p[0], y = p[1];
    if (!(x >= r.lo_x && x <= r.hi_x && y >= r.lo_y && y <= r.hi_y)) cls = -1;
  }
  return cls;
}

__global__ void __launch_bounds__(SS_THREADS) k_ss_count(const float* __restrict__ cloud, int C, int64_t n, SsRange r,
                                                         SsScratch* __restrict__ s) {
  const int64_t i = (int64_t)blockIdx.x * SS_THREADS + threadIdx.x;
  const int cls = (i < n) ? ss_class(cloud + i * C, C, r) : -1;
  const int n0 = __syncthreads_count(cls == 0), n1 = __syncthreads_count(cls == 1);
  if (threadIdx.x == 0) { s->cnt[blockIdx.x] = n0; s->cnt[gridDim.x + blockIdx.x] = n1; }
}

// one CTA: exclusive scan of the 2*nblk block counts (<= 32 768 words), totals to the caller's device buffer
__global__ void __launch_bounds__(SS_SCAN_THREADS) k_ss_scan(SsScratch* __restrict__ s, int nblk, int64_t* __restrict__ counts_out) {
  __shared__ int wsum[SS_SCAN_THREADS / 32];
  const int m = 2 * nblk, t = threadIdx.x, lane = t & 31, w = t >> 5;
  const int chunk = (m + SS_SCAN_THREADS - 1) / SS_SCAN_THREADS;
  const int b = min(t * chunk, m), e = min(b + chunk, m);
  int sum = 0;
  for (int i = b; i < e; i++) sum += s->cnt[i];
  int inc = sum;
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += v;
  }
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    int v = wsum[lane], iv = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, iv, d);
      if (lane >= d) iv += u;
    }
    wsum[lane] = iv - v;                 // exclusive offset of each warp
  }
  __syncthreads();
  int run = wsum[w] + inc - sum;
  for (int i = b; i < e; i++) {
    const int c = s->cnt[i];
    s->off[i] = run;
    run += c;
  }
  __syncthreads();
  if (t == SS_SCAN_THREADS - 1) {        // the last thread's running sum is the grand total
    const int64_t first1 = s->off[nblk];
    counts_out[0] = first1;
    counts_out[1] = (int64_t)run - first1;
  }
}

__global__ void __launch_bounds__(SS_THREADS) k_ss_scatter(const float* __restrict__ cloud, int C, int64_t n, SsRange r,
                                                           const SsScratch* __restrict__ s, float* __restrict__ out0,
                                                           float* __restrict__ out1) {
  __shared__ int wsum[2][SS_THREADS / 32];
  const int64_t i = (int64_t)blockIdx.x * SS_THREADS + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int cls = (i < n) ? ss_class(cloud + i * C, C, r) : -1;
  const unsigned b0 = __ballot_sync(0xffffffffu, cls == 0), b1 = __ballot_sync(0xffffffffu, cls == 1);
  if (lane == 0) { wsum[0][w] = __popc(b0); wsum[1][w] = __popc(b1); }
  __syncthreads();
  if (cls < 0) return;
  int base = 0;
  for (int k = 0; k < w; k++) base += wsum[cls][k];
  const int first1 = s->off[gridDim.x];
  const int64_t pos = cls == 0 ? (int64_t)s->off[blockIdx.x] + base + dfu_warp_excl(b0, lane)
                               : (int64_t)s->off[gridDim.x + blockIdx.x] - first1 + base + dfu_warp_excl(b1, lane);
  float* o = (cls == 0 ? out0 : out1) + pos * C;
  const float* p = cloud + i * C;
  for (int c = 0; c < C; c++) o[c] = p[c];
}

// Device address of the array, resolved once per device with that device current.
static std::mutex g_ss_mu;
static SsScratch* g_ss_ptr[SS_MAX_DEVICES];

static cudaError_t ss_scratch(int dev, SsScratch** out) {
  std::lock_guard<std::mutex> lk(g_ss_mu);
  if (!g_ss_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_split_scratch);
    if (e != cudaSuccess) return e;
    g_ss_ptr[dev] = (SsScratch*)p;
  }
  *out = g_ss_ptr[dev];
  return cudaSuccess;
}

extern "C" int dfu_split_cloud_static(dfu_ctx* ctx, void* stream, const float* cloud, int C, int64_t n, const float* range_host,
                                      float* out_lidar, float* out_pseudo, int64_t* counts_dev) {
  DFU_ON_DEVICE(ctx);                   // ctx may be NULL: the calling thread's current device is used
  if (C < 3 || n < 0 || !counts_dev || (n > 0 && (!cloud || !out_lidar || !out_pseudo)))
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_split_cloud_static: bad argument");
  if (n > (int64_t)SS_MAX_ROWS)
    return dfu_set_err(ctx, DFU_ERR_CAPACITY,
                       "dfu_split_cloud_static: %lld rows exceed the fixed scratch of this entry point "
                       "(DFU_SPLIT_STATIC_MAX_ROWS = %d rows); use dfu_split_cloud_sync or split the cloud",
                       (long long)n, (int)SS_MAX_ROWS);
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    DFU_CUDA(ctx, cudaMemsetAsync(counts_dev, 0, 2 * sizeof(int64_t), st));
    return DFU_OK;
  }
  int dev = -1;
  DFU_CUDA(ctx, cudaGetDevice(&dev));
  if (dev < 0 || dev >= SS_MAX_DEVICES) return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "dfu_split_cloud_static: device ordinal %d", dev);
  SsScratch* s = nullptr;
  DFU_CUDA(ctx, ss_scratch(dev, &s));
  SsRange r = {0.f, 0.f, 0.f, 0.f, 0};
  if (range_host) { r.lo_x = range_host[0]; r.lo_y = range_host[1]; r.hi_x = range_host[3]; r.hi_y = range_host[4]; r.on = 1; }
  const int nblk = (int)((n + SS_THREADS - 1) / SS_THREADS);
  k_ss_count<<<nblk, SS_THREADS, 0, st>>>(cloud, C, n, r, s);
  k_ss_scan<<<1, SS_SCAN_THREADS, 0, st>>>(s, nblk, counts_dev);
  k_ss_scatter<<<nblk, SS_THREADS, 0, st>>>(cloud, C, n, r, s, out_lidar, out_pseudo);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return dfu_set_err(ctx, DFU_ERR_CUDA, "dfu_split_cloud_static: launch: %s", cudaGetErrorString(e));
  if (ctx) ctx->launches += 3;
  return DFU_OK;
}
