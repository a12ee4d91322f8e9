// Context, scratch slots, error reporting, scan helper.
#include <stdarg.h>
#include <stdlib.h>
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>
#include "dfu_common.cuh"

int dfu_set_err(dfu_ctx* ctx, int code, const char* fmt, ...) {
  if (ctx) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

void* dfu_slot(dfu_ctx* ctx, int id, size_t bytes) {
  if (bytes == 0) bytes = 16;
  if (ctx->slot_bytes[id] >= bytes) return ctx->slot[id];
  // growing: everything queued on the device may still read the old block
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess && ctx->slot[id]) e = cudaFree(ctx->slot[id]);
  ctx->slot[id] = nullptr;
  ctx->slot_bytes[id] = 0;
  size_t want = bytes + bytes / 4 + 256;
  if (e == cudaSuccess) e = cudaMalloc(&ctx->slot[id], want);
  if (e != cudaSuccess) {
    dfu_set_err(ctx, DFU_ERR_CUDA, "scratch slot %d: cudaMalloc(%zu) failed: %s", id, want, cudaGetErrorString(e));
    return nullptr;
  }
  ctx->slot_bytes[id] = want;
  return ctx->slot[id];
}

void* dfu_pinned(dfu_ctx* ctx, size_t bytes) {
  if (ctx->pinned_bytes >= bytes) return ctx->pinned;
  cudaDeviceSynchronize();
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  ctx->pinned = nullptr;
  ctx->pinned_bytes = 0;
  size_t want = bytes * 2 + 4096;
  if (cudaMallocHost(&ctx->pinned, want) != cudaSuccess) {
    dfu_set_err(ctx, DFU_ERR_CUDA, "cudaMallocHost(%zu) failed", want);
    return nullptr;
  }
  ctx->pinned_bytes = want;
  return ctx->pinned;
}

__global__ void __launch_bounds__(256) k_upload_words(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, size_t n) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) dst[i] = src[i];
}

int dfu_upload_pinned(dfu_ctx* ctx, cudaStream_t stream, void* dst_dev, const void* src_pinned, size_t bytes) {
  if (bytes == 0) return DFU_OK;
  if (bytes & 3) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_upload_pinned: %zu bytes", bytes);
  const void* dsrc = src_pinned;          // unified addressing: cudaMallocHost memory has one address on host and device
  const size_t n = bytes >> 2;
  unsigned nb = (unsigned)((n + 255) / 256);
  if (nb > 296) nb = 296;
  k_upload_words<<<nb, 256, 0, stream>>>((uint32_t*)dst_dev, (const uint32_t*)dsrc, n);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

cudaError_t dfu_stream_sync(dfu_ctx* ctx, cudaStream_t stream) {
  if (!ctx->blocking_sync) return cudaStreamSynchronize(stream);
  if (!ctx->sync_ev) {
    cudaError_t e = cudaEventCreateWithFlags(&ctx->sync_ev, cudaEventBlockingSync | cudaEventDisableTiming);
    if (e != cudaSuccess) return e;
  }
  cudaError_t e = cudaEventRecord(ctx->sync_ev, stream);
  return e != cudaSuccess ? e : cudaEventSynchronize(ctx->sync_ev);
}

struct I32ToI64 {
  __host__ __device__ int64_t operator()(const int32_t& v) const { return (int64_t)v; }
};

__global__ void k_scan_total(const int32_t* in, int64_t* out, int64_t n) {
  // out[n] = out[n-1] + in[n-1]
  if (threadIdx.x == 0 && blockIdx.x == 0) out[n] = (n > 0) ? out[n - 1] + (int64_t)in[n - 1] : 0;
}

int dfu_exclusive_sum_i32(dfu_ctx* ctx, cudaStream_t stream, const int32_t* in, int64_t* out, int64_t n) {
  if (n > 0) {
    cub::TransformInputIterator<int64_t, I32ToI64, const int32_t*> it(in, I32ToI64());
    size_t tmp_bytes = 0;
    DFU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, it, out, n, stream));
    void* tmp = dfu_slot(ctx, SL_CUBTMP, tmp_bytes);
    if (!tmp) return DFU_ERR_CUDA;
    DFU_CUDA(ctx, cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, it, out, n, stream));
    ctx->launches += 2;
  }
  k_scan_total<<<1, 32, 0, stream>>>(in, out, n);
  DFU_LAUNCH_CHECK(ctx);
  return DFU_OK;
}

extern "C" int dfu_create(dfu_ctx** out, int device) {
  if (!out) return DFU_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return DFU_ERR_CUDA;
  if (cudaSetDevice(device) != cudaSuccess) return DFU_ERR_CUDA;
  dfu_ctx* c = (dfu_ctx*)calloc(1, sizeof(dfu_ctx));
  if (!c) return DFU_ERR_CUDA;
  c->device = device;
  const char* bs = getenv("DFU3D_BLOCKING_SYNC");
  c->blocking_sync = bs && bs[0] == '1';
  *out = c;
  return DFU_OK;
}

extern "C" void dfu_destroy(dfu_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  for (int i = 0; i < DFU_NSLOTS; i++)
    if (ctx->slot[i]) cudaFree(ctx->slot[i]);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  if (ctx->sync_ev) cudaEventDestroy(ctx->sync_ev);
  if (ctx->prof_ev) {
    for (int i = 0; i < 2 * DFU_PROF_POOL; i++) cudaEventDestroy(ctx->prof_ev[i]);
    free(ctx->prof_ev);
    free(ctx->prof_tag);
  }
  free(ctx->views_host);
  free(ctx->s2_consts);
  free(ctx->cnt_project);
  free(ctx->off_project);
  free(ctx->cnt_vox);
  free(ctx->off_vox);
  free(ctx);
}

extern "C" const char* dfu_last_error(const dfu_ctx* ctx) { return ctx ? ctx->err : "null context"; }
extern "C" int64_t dfu_launch_count(const dfu_ctx* ctx) { return ctx ? ctx->launches : 0; }

// ---------------------------------------------------------------- per-kernel device timing
extern "C" int dfu_profile_enable(dfu_ctx* ctx, int on) {
  DFU_ON_DEVICE(ctx);
  if (!ctx) return DFU_ERR_ARG;
  if (on && !ctx->prof_ev) {
    ctx->prof_ev = (cudaEvent_t*)calloc(2 * DFU_PROF_POOL, sizeof(cudaEvent_t));
    ctx->prof_tag = (int*)calloc(DFU_PROF_POOL, sizeof(int));
    for (int i = 0; i < 2 * DFU_PROF_POOL; i++) DFU_CUDA(ctx, cudaEventCreate(&ctx->prof_ev[i]));
  }
  ctx->prof_on = on;
  return DFU_OK;
}

extern "C" int dfu_profile_read(dfu_ctx* ctx, double* ms_out, int64_t* calls_out, int n, int reset) {
  DFU_ON_DEVICE(ctx);
  if (!ctx) return DFU_ERR_ARG;
  if (ctx->prof_used > 0) {
    DFU_CUDA(ctx, cudaDeviceSynchronize());
    for (int i = 0; i < ctx->prof_used; i++) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ctx->prof_ev[2 * i], ctx->prof_ev[2 * i + 1]) == cudaSuccess) {
        ctx->prof_ms[ctx->prof_tag[i]] += ms;
        ctx->prof_calls[ctx->prof_tag[i]] += 1;
      }
    }
    ctx->prof_used = 0;
  }
  for (int i = 0; i < n && i < DFU_PROF_NTAGS; i++) {
    if (ms_out) ms_out[i] = ctx->prof_ms[i];
    if (calls_out) calls_out[i] = ctx->prof_calls[i];
  }
  if (reset) {
    memset(ctx->prof_ms, 0, sizeof(ctx->prof_ms));
    memset(ctx->prof_calls, 0, sizeof(ctx->prof_calls));
  }
  return DFU_OK;
}
