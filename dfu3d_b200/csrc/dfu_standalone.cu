// Stand-alone operators on caller-provided point sets (include/dfu3d.h, last section): radius filter, NN gate,
// L-shape fitting.  They share the segmented kernels of dfu_segops.cu / dfu_cluster.cu with the batched pipeline.
#include <vector>
#include "dfu_ops.cuh"

#define OPS_A 44
#define OPS_B (OPS_A + GRID_SLOTS + 1)
#define OPS_C (OPS_B + CLUSTER_SLOTS)
enum { SA_X = 140, SA_Y, SA_Z, SA_SEG, SA_SEGOFF, SA_PRM, SA_QX, SA_QY, SA_QZ, SA_QSEG, SA_MISC, SA_RECT, SA_ANG };

// rows [n][D] -> SoA, seg id by binary search in seg_off [nseg+1]
__global__ void __launch_bounds__(256) k_rows_to_soa(const double* __restrict__ rows, int D, int64_t n,
                                                     const int64_t* __restrict__ seg_off, int nseg, double* __restrict__ x,
                                                     double* __restrict__ y, double* __restrict__ z, int32_t* __restrict__ seg) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  x[i] = rows[i * D];
  y[i] = rows[i * D + 1];
  if (z) z[i] = D > 2 ? rows[i * D + 2] : 0.0;
  int lo = 0, hi = nseg;
  if (seg_off)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (seg_off[mid] <= i) lo = mid; else hi = mid;
    }
  seg[i] = lo;
}

__global__ void __launch_bounds__(256) k_count_keep(const int32_t* __restrict__ seg, const uint8_t* __restrict__ keep, int64_t n,
                                                    int32_t* __restrict__ cnt) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  // one atomic per distinct segment in the warp (segments are contiguous: per-lane atomics would hit one address)
  const int s = (i < n && keep[i]) ? seg[i] : -1;
  const unsigned m = __match_any_sync(0xffffffffu, s);
  if (s >= 0 && (threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(cnt + s, __popc(m));
}

static int check_grid_err(dfu_ctx* ctx, cudaStream_t st, int sb, const char* what) {
  int32_t e = 0;
  DFU_CUDA(ctx, cudaMemcpyAsync(&e, dfu_grid_err_ptr(ctx, sb), 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  if (e) return dfu_set_err(ctx, DFU_ERR_DATA, "%s: a coordinate is NaN or the cloud spans more than 65533 cells", what);
  return DFU_OK;
}

extern "C" int dfu_radius_filter(dfu_ctx* ctx, void* stream, const double* xyz, int nseg, const int64_t* seg_off_host,
                                 const double* radius_host, const int32_t* nb_host, const int32_t* every_k_host,
                                 uint8_t* out_keep, int32_t* out_count) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || nseg <= 0 || !seg_off_host || !radius_host || !nb_host || !out_keep)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_radius_filter: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = seg_off_host[nseg];
  if (out_count) DFU_CUDA(ctx, cudaMemsetAsync(out_count, 0, (size_t)nseg * 4, st));
  if (n == 0) return DFU_OK;
  if (!xyz) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_radius_filter: xyz is NULL");
  DFU_SLOT(ctx, x, double*, SA_X, n * 8);
  DFU_SLOT(ctx, y, double*, SA_Y, n * 8);
  DFU_SLOT(ctx, z, double*, SA_Z, n * 8);
  DFU_SLOT(ctx, seg, int32_t*, SA_SEG, n * 4);
  DFU_SLOT(ctx, seg_off, int64_t*, SA_SEGOFF, (size_t)(nseg + 1) * 8);
  DFU_SLOT(ctx, prm, SegPrm*, SA_PRM, (size_t)nseg * sizeof(SegPrm));
  std::vector<SegPrm> ph(nseg);
  for (int s = 0; s < nseg; s++) ph[s] = dfu_seg_prm(radius_host[s], nb_host[s], every_k_host ? every_k_host[s] : 1, 1);
  DFU_CUDA(ctx, cudaMemcpyAsync(seg_off, seg_off_host, (size_t)(nseg + 1) * 8, cudaMemcpyHostToDevice, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(prm, ph.data(), (size_t)nseg * sizeof(SegPrm), cudaMemcpyHostToDevice, st));
  k_rows_to_soa<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(xyz, 3, n, seg_off, nseg, x, y, z, seg);
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_grid_err_reset(ctx, st, OPS_A);
  if (rc) return rc;
  rc = dfu_radius_keep(ctx, st, OPS_A, x, y, z, seg, seg_off, n, nseg, prm, out_keep);
  if (rc) return rc;
  if (out_count) {
    k_count_keep<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(seg, out_keep, n, out_count);
    DFU_LAUNCH_CHECK(ctx);
  }
  return check_grid_err(ctx, st, OPS_A, "dfu_radius_filter");
}

extern "C" int dfu_nn_gate(dfu_ctx* ctx, void* stream, const double* query_xyz, int64_t nq, const double* ref_xyz, int64_t nr,
                           double max_dist, uint8_t* out_keep) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || nq < 0 || nr < 0 || !out_keep) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_nn_gate: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  if (nq == 0) return DFU_OK;
  if (nr == 0) {
    DFU_CUDA(ctx, cudaMemsetAsync(out_keep, 0, nq, st));
    return DFU_OK;
  }
  DFU_SLOT(ctx, x, double*, SA_X, nr * 8);
  DFU_SLOT(ctx, y, double*, SA_Y, nr * 8);
  DFU_SLOT(ctx, z, double*, SA_Z, nr * 8);
  DFU_SLOT(ctx, seg, int32_t*, SA_SEG, nr * 4);
  DFU_SLOT(ctx, qx, double*, SA_QX, nq * 8);
  DFU_SLOT(ctx, qy, double*, SA_QY, nq * 8);
  DFU_SLOT(ctx, qz, double*, SA_QZ, nq * 8);
  DFU_SLOT(ctx, qseg, int32_t*, SA_QSEG, nq * 4);
  DFU_SLOT(ctx, md, double*, SA_MISC, 64);
  DFU_CUDA(ctx, cudaMemcpyAsync(md, &max_dist, 8, cudaMemcpyHostToDevice, st));
  k_rows_to_soa<<<(unsigned)((nr + 255) / 256), 256, 0, st>>>(ref_xyz, 3, nr, nullptr, 1, x, y, z, seg);
  DFU_LAUNCH_CHECK(ctx);
  k_rows_to_soa<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(query_xyz, 3, nq, nullptr, 1, qx, qy, qz, qseg);
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_grid_err_reset(ctx, st, OPS_A);
  if (rc) return rc;
  rc = dfu_gate_keep(ctx, st, OPS_A, qx, qy, qz, qseg, nq, x, y, z, seg, nullptr, nr, 1, md, out_keep);
  if (rc) return rc;
  return check_grid_err(ctx, st, OPS_A, "dfu_nn_gate");
}

extern "C" int dfu_lshape_fit_sync(dfu_ctx* ctx, void* stream, const double* xy, int64_t n, double R0, double Rd,
                                   const double* angles_cs_host, int na, int32_t* out_cluster, double* rect_out_host,
                                   int64_t capacity, int64_t* n_clusters_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || n < 0 || !angles_cs_host || na <= 0 || na > DFU_N_ANGLES_MAX || !n_clusters_host)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_lshape_fit_sync: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  *n_clusters_host = 0;
  if (n == 0) return DFU_OK;
  DFU_SLOT(ctx, x, double*, SA_X, n * 8);
  DFU_SLOT(ctx, y, double*, SA_Y, n * 8);
  DFU_SLOT(ctx, seg, int32_t*, SA_SEG, n * 4);
  DFU_SLOT(ctx, misc, double*, SA_MISC, 64);
  DFU_SLOT(ctx, ang, double*, SA_ANG, (size_t)na * 16);
  const double rr[2] = {R0, Rd};
  DFU_CUDA(ctx, cudaMemcpyAsync(misc, rr, 16, cudaMemcpyHostToDevice, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(ang, angles_cs_host, (size_t)na * 16, cudaMemcpyHostToDevice, st));
  k_rows_to_soa<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(xy, 2, n, nullptr, 1, x, y, nullptr, seg);
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_grid_err_reset(ctx, st, OPS_B);
  if (rc) return rc;
  ClusterOut co;
  rc = dfu_cluster_points(ctx, st, OPS_B, x, y, seg, nullptr, n, 1, misc, misc + 1, &co);
  if (rc) return rc;
  int32_t hk[2] = {0, 0};
  DFU_CUDA(ctx, cudaMemcpyAsync(&hk[0], co.ncluster, 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(&hk[1], dfu_grid_err_ptr(ctx, OPS_B), 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  if (hk[1] & 2) return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "dfu_lshape_fit_sync: R0 <= 0, Rd < 0 or R0 + Rd*|p| > sqrt(2)*R0 for some point (the cell neighbourhood of the clustering would miss pairs)");
  if (hk[1]) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_lshape_fit_sync: a coordinate is NaN or the set spans more than 65533 cells");
  const int K = hk[0];
  *n_clusters_host = K;
  if (out_cluster) DFU_CUDA(ctx, cudaMemcpyAsync(out_cluster, co.cluster_of_point, n * 4, cudaMemcpyDeviceToDevice, st));
  if (!rect_out_host) return DFU_OK;
  if (K > capacity) return dfu_set_err(ctx, DFU_ERR_CAPACITY, "rect output needs %d records, capacity %lld", K, (long long)capacity);
  DFU_SLOT(ctx, rect, double*, SA_RECT, (size_t)K * 16 * 8);
  rc = dfu_fit_clusters(ctx, st, OPS_C, x, y, co.cluster_of_point, n, K, ang, na, rect);
  if (rc) return rc;
  DFU_CUDA(ctx, cudaMemcpyAsync(rect_out_host, rect, (size_t)K * 16 * 8, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaStreamSynchronize(st));
  return DFU_OK;
}
