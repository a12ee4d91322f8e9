// Shared declarations of libdfu3d.so (sm_100a).  See include/dfu3d.h for the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/dfu3d.h"

#define DFU_NSLOTS 256
#define DFU_PROF_POOL 8192
#ifndef DFU_PROF_NTAGS
#define DFU_PROF_NTAGS 24
#endif

struct dfu_ctx {
  int device;
  char err[512];
  int64_t launches;
  // growable device scratch slots (never shrink); slot ids are private to each .cu file range
  void* slot[DFU_NSLOTS];
  size_t slot_bytes[DFU_NSLOTS];
  // pinned host staging for small read-backs
  void* pinned;
  size_t pinned_bytes;
  // stream syncs of the fused pipeline: spin (default) or sleep on a blocking event (DFU3D_BLOCKING_SYNC=1 at dfu_create:
  // for hosts with fewer free cores than waiting threads, e.g. 8 ranks x 3 threads on 32 cores)
  int blocking_sync;
  cudaEvent_t sync_ev;
  // ---- bound batch (dfu_fuse_begin) ----
  cudaStream_t stream;
  int V;
  dfu_view* views_host;       // malloc'd copy
  void* s2_consts;            // malloc'd float64 view constants of the densify passes (dfu_densify.cu)
  dfu_fusion_params prm;
  const void* lidar;
  const float* depth;
  const void* image;
  const uint32_t* labelmap;
  int max_lidar_cnt;
  int64_t total_slots;
  int bpv_pts;                // S1 blocks per view
  int bpv_pix;                // S2 blocks per view
  int64_t* cnt_project;       // host [V][3]
  int64_t* off_project;       // host [3][V+1]  list offsets (F, A, B)
  int64_t* cnt_vox;           // host [V]
  int64_t* off_vox;           // host [V+1]
  int stage;                  // 0 none, 1 begun, 2 projected, 3 densified
  int total_inst;
  void* table_clean_ptr;      // angular voxel table known to be all-ones (pass C resets what pass A touched)
  size_t table_clean_bytes;
  int64_t sor_deferred, sor_points, sor_cause[3];   // last dfu_sor_filter call with out_stats_host: queries handled one by one
  // ---- per-kernel device timing (dfu_profile_*) ----
  int prof_on;
  int prof_used;
  cudaEvent_t* prof_ev;       // [2 * DFU_PROF_POOL]
  int* prof_tag;              // [DFU_PROF_POOL]
  double prof_ms[DFU_PROF_NTAGS];
  int64_t prof_calls[DFU_PROF_NTAGS];
};

// kernel tags of dfu_profile_read (order is part of the ABI, see include/dfu3d.h)
enum {
  PT_PACK_MASKS = 0, PT_PROJECT, PT_LISTS, PT_DENSIFY_A, PT_DENSIFY_B, PT_DENSIFY_C, PT_VOX_GATHER, PT_EMIT_CLOUD,
  PT_INST_PAIRS, PT_INST_SORT, PT_GRID_BUILD, PT_RADIUS, PT_GATE, PT_CLUSTER, PT_FIT, PT_SOR_KNN, PT_SOR_MASK,
  PT_SPARSE_DEPTH, PT_MISC
};

struct dfu_prof_scope {
  dfu_ctx* c; cudaStream_t st; int slot;
  dfu_prof_scope(dfu_ctx* ctx, cudaStream_t s, int tag) : c(ctx), st(s), slot(-1) {
    if (ctx->prof_on && ctx->prof_used < DFU_PROF_POOL) {
      slot = ctx->prof_used++;
      ctx->prof_tag[slot] = tag;
      cudaEventRecord(ctx->prof_ev[2 * slot], s);
    }
  }
  ~dfu_prof_scope() { if (slot >= 0) cudaEventRecord(c->prof_ev[2 * slot + 1], st); }
};
#define DFU_PROF(ctx, stream, tag) dfu_prof_scope prof_scope__##tag(ctx, stream, tag)

// scratch slot ids
enum {
  SL_VIEWS = 0, SL_FLAGS, SL_PIX, SL_LABEL, SL_BLKCNT, SL_BLKOFF, SL_CUBTMP, SL_FLIST, SL_ALIST, SL_BLABEL,
  SL_KEY, SL_TABLE, SL_VOXCNT, SL_VOXOFF, SL_REPZ, SL_REPY, SL_REPY_XYZ, SL_REPY_LABEL, SL_REPZ_XYZ, SL_STATS,
  SL_SEGCNT, SL_SEGOFF, SL_IPARAMS, SL_ANGLES, SL_ISCRATCH_D, SL_ISCRATCH_I, SL_BOXES, SL_BOXCNT, SL_SEGOUT,
  SL_OP0, SL_OP1, SL_OP2, SL_OP3, SL_OP4, SL_OP5, SL_OP6, SL_OP7, SL_OP8, SL_OP9, SL_OP10, SL_OP11, SL_VOXCELL, SL_WARPSTAT, SL_AMBQ
};
static_assert(SL_AMBQ < 44, "SL_* ids must stay below the operator slot ranges");
// Scratch slot ranges of the files (one table, so that no two users share an id):
//   0..43   SL_* (this header)                      44..93  OPS_A / OPS_B / OPS_C (dfu_ops.cuh: grid, clustering, fit)
//   100..131 IS_* (dfu_instances.cu)                140..152 SA_* (dfu_standalone.cu)     160..166 GA_* (dfu_instances.cu)
//   170..189 SO_* (dfu_sor.cu)                      200..208 CV_*, 214..216 CS_* (dfu_consumer.cu)   210..212 GT_* (dfu_gtdb.cu)
//   224..233 EV_* (dfu_evolve.cu)
#define DFU_SLOTS_OPS_FIRST 44
#define DFU_SLOTS_OPS_END 100

int dfu_set_err(dfu_ctx* ctx, int code, const char* fmt, ...);

// Every entry point runs on the context's device whatever the calling thread's current device is (one context per
// device in one process: torch's current device may differ), and restores the caller's device on return.
struct dfu_dev_guard {
  int prev;
  bool sw;
  explicit dfu_dev_guard(const dfu_ctx* c) : prev(-1), sw(false) {
    if (c && cudaGetDevice(&prev) == cudaSuccess && prev != c->device) sw = cudaSetDevice(c->device) == cudaSuccess;
  }
  ~dfu_dev_guard() { if (sw) cudaSetDevice(prev); }
};
#define DFU_ON_DEVICE(ctx) dfu_dev_guard dev_guard__(ctx)
void* dfu_slot(dfu_ctx* ctx, int id, size_t bytes);      // NULL on failure (error recorded)
void* dfu_pinned(dfu_ctx* ctx, size_t bytes);

#define DFU_CUDA(ctx, call)                                                                   \
  do {                                                                                        \
    cudaError_t e__ = (call);                                                                 \
    if (e__ != cudaSuccess)                                                                   \
      return dfu_set_err(ctx, DFU_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
  } while (0)

#define DFU_LAUNCH_CHECK(ctx)                                                                 \
  do {                                                                                        \
    (ctx)->launches++;                                                                        \
    cudaError_t e__ = cudaGetLastError();                                                     \
    if (e__ != cudaSuccess)                                                                   \
      return dfu_set_err(ctx, DFU_ERR_CUDA, "%s:%d launch: %s", __FILE__, __LINE__, cudaGetErrorString(e__)); \
  } while (0)

#define DFU_SLOT(ctx, var, type, id, bytes)                 \
  type var = (type)dfu_slot(ctx, id, bytes);                \
  if (!var) return DFU_ERR_CUDA;

// exclusive sum of n int32 -> int64 offsets (n+1 outputs, last = total), on `stream`
int dfu_exclusive_sum_i32(dfu_ctx* ctx, cudaStream_t stream, const int32_t* in, int64_t* out, int64_t n);
// Small host -> device upload that does NOT use a copy engine: a kernel reads the words from the context's pinned buffer
// over PCIe.  A copy engine keeps serving the stream that feeds it: behind another stream's bulk upload (the next frame
// group of FusionPipeline.run_host) a 90 KB cudaMemcpyAsync from pinned memory waited until that upload had finished
// (measured: 42 ms instead of 4 ms for the instance stage).  src must lie inside dfu_pinned() memory, bytes % 4 == 0.
int dfu_upload_pinned(dfu_ctx* ctx, cudaStream_t stream, void* dst_dev, const void* src_pinned, size_t bytes);
cudaError_t dfu_stream_sync(dfu_ctx* ctx, cudaStream_t stream);

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ unsigned long long dfu_order_f64(double v) {
  unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}

__device__ __forceinline__ double dfu_unorder_f64(unsigned long long b) {
  b = (b & 0x8000000000000000ull) ? (b & 0x7fffffffffffffffull) : ~b;
  return __longlong_as_double((long long)b);
}

// The reference's float32 [x y z 1]·M (K = 4) as OpenBLAS evaluates it: one accumulator, k ascending,
// FMA, first term a rounded product (measured; see oracle/restate.py:_sgemm_fma).
__device__ __forceinline__ float dfu_row4_f32(float a0, float a1, float a2, float a3, const float* M, int j, int ld) {
  float acc = __fmul_rn(a0, M[0 * ld + j]);
  acc = __fmaf_rn(a1, M[1 * ld + j], acc);
  acc = __fmaf_rn(a2, M[2 * ld + j], acc);
  acc = __fmaf_rn(a3, M[3 * ld + j], acc);
  return acc;
}
__device__ __forceinline__ double dfu_row4_f64(double a0, double a1, double a2, double a3, const float* M, int j, int ld) {
  double acc = __dmul_rn(a0, (double)M[0 * ld + j]);
  acc = __fma_rn(a1, (double)M[1 * ld + j], acc);
  acc = __fma_rn(a2, (double)M[2 * ld + j], acc);
  acc = __fma_rn(a3, (double)M[3 * ld + j], acc);
  return acc;
}

// calibration_kitti.py:134-144 then :89-102 — pixel (col,row,depth) -> LiDAR frame, float64.
__device__ __forceinline__ void dfu_backproject(const dfu_view& vw, int col, int row, float d, double& X, double& Y,
                                                double& Z) {
  double dd = (double)d;
  double xr = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn((double)col, (double)vw.cu), dd), (double)vw.fu), (double)vw.tx);
  double yr = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn((double)row, (double)vw.cv), dd), (double)vw.fv), (double)vw.ty);
  X = dfu_row4_f64(xr, yr, dd, 1.0, vw.inv44, 0, 4);
  Y = dfu_row4_f64(xr, yr, dd, 1.0, vw.inv44, 1, 4);
  Z = dfu_row4_f64(xr, yr, dd, 1.0, vw.inv44, 2, 4);
}

// a / d, correctly rounded, for a divisor whose correctly rounded reciprocal y = RN(1/d) is known (one per view):
// two Markstein refinements (q faithful after the first, RN(a/d) after the second; exact remainders through FMA) —
// 1 multiply + 4 FMA instead of the ~25-instruction generic DDIV sequence with its slow-path call.
// Checked against __ddiv_rn on 2^32 operand pairs by tools/divcheck.cu.
__device__ __forceinline__ double dfu_div_by(double a, double d, double y) {
  double q = __dmul_rn(a, y);
  double r = __fma_rn(-q, d, a);
  q = __fma_rn(r, y, q);
  r = __fma_rn(-q, d, a);
  return __fma_rn(r, y, q);
}

// same result as dfu_backproject, with the per-view reciprocals rfu = RN(1/fu), rfv = RN(1/fv) supplied
__device__ __forceinline__ void dfu_backproject_r(const dfu_view& vw, double rfu, double rfv, int col, int row, float d,
                                                  double& X, double& Y, double& Z) {
  double dd = (double)d;
  double xr = __dadd_rn(dfu_div_by(__dmul_rn(__dsub_rn((double)col, (double)vw.cu), dd), (double)vw.fu, rfu), (double)vw.tx);
  double yr = __dadd_rn(dfu_div_by(__dmul_rn(__dsub_rn((double)row, (double)vw.cv), dd), (double)vw.fv, rfv), (double)vw.ty);
  X = dfu_row4_f64(xr, yr, dd, 1.0, vw.inv44, 0, 4);
  Y = dfu_row4_f64(xr, yr, dd, 1.0, vw.inv44, 1, 4);
  Z = dfu_row4_f64(xr, yr, dd, 1.0, vw.inv44, 2, 4);
}

__device__ __forceinline__ int dfu_warp_excl(unsigned ballot, int lane) { return __popc(ballot & ((1u << lane) - 1u)); }

// ---------------------------------------------------------------- TMA 1-D bulk copies (global -> shared) + mbarrier
// cp.async.bulk needs 16-byte aligned source, destination and size; completion is signalled on an mbarrier through
// its transaction count (SASS: UBLKCP.S.G + SYNCS.ARRIVE.TRANS64).  Used to stage contiguous ranges of the sorted
// point arrays (grid cells / BVH leaves) in shared memory.
__device__ __forceinline__ uint32_t dfu_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void dfu_mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(dfu_smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void dfu_mbar_arrive_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dfu_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dfu_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dfu_smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(dfu_smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void dfu_mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(dfu_smem_u32(bar)),
      "r"(parity)
      : "memory");
}
