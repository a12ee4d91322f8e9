// Per-instance stage of the batched pipeline: gather of inherited LiDAR points and masked pseudo points, radius
// filters, BallQuery gate, fuse, adaptive-range clustering, rectangle search, GenerateAnns arithmetic.
// Reference: tools/PENet/dataloaders/my_loader.py:522-530 (inherited LiDAR points per instance, quirk Q1),
// :547-557 (pseudo points of an instance = la_sampling20 representatives whose mask pixel is set),
// :579-599 (radius filters), :602-605 (gate + fuse), :633-684 (GenerateAnns).
//
// All instances of all views are processed together as SEGMENTS of one point array: segment 2g is instance g's
// LiDAR set, segment 2g+1 its pseudo set.  The (element, instance) membership pairs are generated in element order
// and stably radix-sorted by segment, which reproduces the reference's per-instance point order.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "dfu_ops.cuh"

#define I_THREADS 256
#define OPS_A 44                         // grid operators: GRID_SLOTS + 1 slots
#define OPS_B (OPS_A + GRID_SLOTS + 1)   // clustering: CLUSTER_SLOTS slots
#define OPS_C (OPS_B + CLUSTER_SLOTS)    // fit: FIT_SLOTS slots
static_assert(OPS_A == DFU_SLOTS_OPS_FIRST && OPS_C + FIT_SLOTS <= DFU_SLOTS_OPS_END, "operator slot ranges must stay inside 44..99");

struct InstTables {
  const dfu_view* views;
  const int64_t* offA;       // [V+1] above-plane list offsets
  const int64_t* offB;       // [V+1] in-bounds list offsets
  const int64_t* voxoff;     // [V+1]
  const int32_t* a_list;     // above-plane rows (per view, relative to the view's sweep)
  const uint32_t* b_label;   // label word of the k-th in-bounds point
  const uint32_t* repy_label;
  const double* repy_xyz;    // SoA x[nvox] y[nvox] z[nvox]
  const float4* lidar;
  int V;
  int64_t EL, EP;
};

__device__ __forceinline__ int view_of(const int64_t* __restrict__ off, int V, int64_t e) {
  int lo = 0, hi = V;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (off[mid] <= e) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ uint32_t elem_word(const InstTables& T, int64_t e, int& v) {
  uint32_t w;
  if (e < T.EL) {
    v = view_of(T.offB, T.V, e);
    w = T.b_label[e];
  } else {
    v = view_of(T.voxoff, T.V, e - T.EL);
    w = T.repy_label[e - T.EL];
  }
  const int n = T.views[v].n_inst;
  return n >= 32 ? w : (w & ((1u << n) - 1u));
}

__global__ void __launch_bounds__(I_THREADS) k_inst_count(InstTables T, int32_t* __restrict__ cnt) {
  const int64_t e = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  if (e >= T.EL + T.EP) return;
  int v;
  cnt[e] = __popc(elem_word(T, e, v));
}

__global__ void __launch_bounds__(I_THREADS) k_inst_emit(InstTables T, const int64_t* __restrict__ poff,
                                                         uint32_t* __restrict__ key, int32_t* __restrict__ val) {
  const int64_t e = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  if (e >= T.EL + T.EP) return;
  int v;
  uint32_t w = elem_word(T, e, v);
  const uint32_t side = e < T.EL ? 0u : 1u;
  const uint32_t base = (uint32_t)T.views[v].inst_off;
  int64_t pos = poff[e];
  while (w) {
    const int j = __ffs(w) - 1;
    w &= w - 1;
    key[pos] = 2u * (base + (uint32_t)j) + side;
    val[pos] = (int32_t)e;
    pos++;
  }
}

// seg_off [nseg+1] from the sorted keys
__global__ void __launch_bounds__(I_THREADS) k_seg_bounds(const uint32_t* __restrict__ skey, int64_t np, int nseg,
                                                          int64_t* __restrict__ seg_off) {
  const int64_t i = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  if (i >= np) return;
  const int k = (int)skey[i];
  const int prev = i == 0 ? -1 : (int)skey[i - 1];
  for (int s = prev + 1; s <= k; s++) seg_off[s] = i;
  if (i == np - 1)
    for (int s = k + 1; s <= nseg; s++) seg_off[s] = np;
}

__global__ void __launch_bounds__(I_THREADS) k_inst_gather(InstTables T, const uint32_t* __restrict__ skey,
                                                           const int32_t* __restrict__ sval, int64_t np,
                                                           double* __restrict__ px, double* __restrict__ py,
                                                           double* __restrict__ pz, int32_t* __restrict__ seg,
                                                           int32_t* __restrict__ inst) {
  const int64_t i = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  if (i >= np) return;
  const int64_t e = sval[i];
  seg[i] = (int32_t)skey[i];
  inst[i] = (int32_t)(skey[i] >> 1);
  if (e < T.EL) {
    const int v = view_of(T.offB, T.V, e);
    // quirk Q1: the k-th in-bounds label is attached to the k-th above-plane row (my_loader.py:526-530)
    const int row = T.a_list[T.offA[v] + (e - T.offB[v])];
    const float4 p = __ldg(T.lidar + T.views[v].lidar_off + row);
    px[i] = (double)p.x; py[i] = (double)p.y; pz[i] = (double)p.z;
  } else {
    const int64_t t = e - T.EL;
    px[i] = T.repy_xyz[t]; py[i] = T.repy_xyz[T.EP + t]; pz[i] = T.repy_xyz[2 * T.EP + t];
  }
}

// cnt[s] += 1 for every lane with s >= 0, one atomic per distinct segment in the warp (points of a segment are
// contiguous, so a warp almost always holds one or two segments: per-lane atomics would serialise on one address)
__device__ __forceinline__ void seg_warp_add(int32_t* __restrict__ cnt, int s) {
  const unsigned m = __match_any_sync(0xffffffffu, s);
  if (s >= 0 && (threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(cnt + s, __popc(m));
}

// per-segment count of kept points
__global__ void __launch_bounds__(I_THREADS) k_seg_count(const int32_t* __restrict__ seg, const uint8_t* __restrict__ keep,
                                                         int64_t np, int32_t* __restrict__ cnt) {
  const int64_t i = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  seg_warp_add(cnt, (i < np && keep[i]) ? seg[i] : -1);
}

// ractive: kept LiDAR points (the gate's reference set); qactive: kept pseudo points (the only queries whose answer
// k_fuse_flags reads)
__global__ void __launch_bounds__(I_THREADS) k_ref_active(const int32_t* __restrict__ seg, const uint8_t* __restrict__ keep,
                                                          int64_t np, uint8_t* __restrict__ ractive, uint8_t* __restrict__ qactive) {
  const int64_t i = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  if (i < np) {
    ractive[i] = (keep[i] && !(seg[i] & 1)) ? 1 : 0;
    qactive[i] = (keep[i] && (seg[i] & 1)) ? 1 : 0;
  }
}

// final membership: LiDAR side = radius keep; pseudo side = radius keep and (gate or no kept LiDAR point)  (:602-605)
__global__ void __launch_bounds__(I_THREADS) k_fuse_flags(const int32_t* __restrict__ seg, const uint8_t* __restrict__ keep,
                                                          const uint8_t* __restrict__ gate, const int32_t* __restrict__ cnt1,
                                                          int64_t np, int32_t* __restrict__ flag, int32_t* __restrict__ cnt2) {
  const int64_t i = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  int s = -1, f = 0;
  if (i < np) {
    s = seg[i];
    f = keep[i];
    if (f && (s & 1) && cnt1[s - 1] > 0) f = gate[i];
    flag[i] = f;
  }
  seg_warp_add(cnt2, f ? s : -1);
}

__global__ void __launch_bounds__(I_THREADS) k_fuse_scatter(const int64_t* __restrict__ flagscan /*exclusive, [np+1]*/,
                                                            const uint8_t* __restrict__ keep, const uint8_t* __restrict__ gate,
                                                            const int32_t* __restrict__ seg, const int32_t* __restrict__ cnt1,
                                                            const double* __restrict__ px, const double* __restrict__ py,
                                                            const double* __restrict__ pz, int64_t np, double* __restrict__ fx,
                                                            double* __restrict__ fy, double* __restrict__ fz,
                                                            int32_t* __restrict__ finst, unsigned long long* __restrict__ zmax) {
  const int64_t i = (int64_t)blockIdx.x * I_THREADS + threadIdx.x;
  int g = -1;
  unsigned long long zb = 0ull;
  if (i < np) {
    const int s = seg[i];
    int f = keep[i];
    if (f && (s & 1) && cnt1[s - 1] > 0) f = gate[i];
    if (f) {
      const int64_t o = flagscan[i];
      const double z = pz[i];
      fx[o] = px[i]; fy[o] = py[i]; fz[o] = z;
      finst[o] = s >> 1;
      g = s >> 1;
      zb = dfu_order_f64(z);
    }
  }
  // max z per instance: the warp almost always holds ONE instance -> shuffle reduction + one atomic
  const unsigned have = __ballot_sync(0xffffffffu, g >= 0);
  if (have == 0u) return;
  const int g0 = __shfl_sync(0xffffffffu, g, __ffs(have) - 1);
  if (__all_sync(0xffffffffu, g < 0 || g == g0)) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long ob = __shfl_xor_sync(0xffffffffu, zb, o);
      zb = ob > zb ? ob : zb;
    }
    if ((threadIdx.x & 31) == 0) atomicMax(zmax + g0, zb);
  } else if (g >= 0) {
    atomicMax(zmax + g, zb);
  }
}

// inst_off [G+1] in the fused array = flagscan at the instance's first pair (segment 2g)
__global__ void k_inst_offsets(const int64_t* __restrict__ seg_off, const int64_t* __restrict__ flagscan, int64_t np, int G,
                               int64_t* __restrict__ inst_off) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g > G) return;
  const int64_t p = g == G ? np : seg_off[2 * g];
  inst_off[g] = flagscan[p > np ? np : p];
}

// ---------------------------------------------------------------- GenerateAnns arithmetic (my_loader.py:633-684)
struct InstDev {
  int32_t view, cls, car_rule, anns_mode;
};

__device__ __forceinline__ void cross_point(double a0, double a1, double b0, double b1, double c0, double c1, double& x,
                                            double& y) {
  // calc_cross_point (my_loader.py:699-702)
  const double den = __dsub_rn(__dmul_rn(a0, b1), __dmul_rn(a1, b0));
  x = __ddiv_rn(__dsub_rn(__dmul_rn(b0, -c1), __dmul_rn(b1, -c0)), den);
  y = __ddiv_rn(__dsub_rn(__dmul_rn(a1, -c0), __dmul_rn(a0, -c1)), den);
}

// rect [K][16] -> box [K][DFU_BOX_DOUBLES] (layout: include/dfu3d.h)
// mode 0: my_loader.py:633-684; mode 1: my_loader0.py:811-930 (center_type 'prior_linecorner')
__device__ void anns_record(const double* __restrict__ r, double zmax, const float* M43 /*nullable*/, int car_rule, int mode,
                            double* __restrict__ o) {
  const double *a = r, *b = r + 4, *c = r + 8;
  double X[4], Y[4];
  cross_point(a[0], a[1], b[0], b[1], c[0], c[1], X[0], Y[0]);
  cross_point(a[1], a[2], b[1], b[2], c[1], c[2], X[1], Y[1]);
  cross_point(a[2], a[3], b[2], b[3], c[2], c[3], X[2], Y[2]);
  cross_point(a[3], a[0], b[3], b[0], c[3], c[0], X[3], Y[3]);
  double cx = (X[0] + X[2]) / 2, cy = (Y[0] + Y[2]) / 2;
  double cz = zmax / 2 - 1.5, height = zmax;
  if (mode == 1 && height < 1.0) height = 1.56;                            // my_loader0.py:832
  const double d3x = X[0] - X[3], d3y = Y[0] - Y[3], d1x = X[0] - X[1], d1y = Y[0] - Y[1];
  const double l1 = __dsqrt_rn(__dadd_rn(__dmul_rn(d3x, d3x), __dmul_rn(d3y, d3y)));
  const double l2 = __dsqrt_rn(__dadd_rn(__dmul_rn(d1x, d1x), __dmul_rn(d1y, d1y)));
  bool skip = (car_rule || mode == 1) && (l1 / l2 > 5 || l2 / l1 > 5);     // my_loader.py:651 (Car only) / my_loader0.py:835 (all)
  double length, width, rot;
  const double rot3 = atan((Y[3] - Y[0]) / (X[3] - X[0] + 1e-8)), rot1 = atan((Y[1] - Y[0]) / (X[1] - X[0] + 1e-8));
  if (l1 >= l2) {
    length = l1; width = l2;
    rot = rot3;
  } else {
    length = l2; width = l1;
    rot = rot1;
  }
  if (mode == 1) {
    // 'prior_linecorner' (my_loader0.py:848-857, :889-898): a short rectangle is taken for one end of a 3.9 m car seen
    // end-on: heading along the OTHER side, centre pushed along it; then every box becomes 3.9 x 1.6 (:876, :916) and
    // boxes whose centre is closer than 10 m are dropped (:918)
    if (width < 1.5 && length < 2.0) {
      rot = l1 >= l2 ? rot1 : rot3;
      cx = (3.9 - length) / 2 * cos(rot) + cx;
      cy = (3.9 - length) / 2 * sin(rot) + cy;
    }
    length = 3.9; width = 1.6;
    if (sqrt(cx * cx + cy * cy) < 10.0) skip = true;
  }
  rot = -rot - 3.141592653589793 / 2;
  const double theta = atan(-cx / (cy + 1e-8));
  const double alpha = rot - theta;
  double rx = cx, ry = cy, rz = cz;
  if (M43) {   // calib.lidar_to_rect of the float64 centre (float64 row times the float32 matrix)
    rx = dfu_row4_f64(cx, cy, cz, 1.0, M43, 0, 3);
    ry = dfu_row4_f64(cx, cy, cz, 1.0, M43, 1, 3);
    rz = dfu_row4_f64(cx, cy, cz, 1.0, M43, 2, 3);
  }
  o[4] = alpha; o[5] = height; o[6] = width; o[7] = length; o[8] = rx; o[9] = ry; o[10] = rz; o[11] = rot;
  o[12] = r[15];
  const double best = r[13], run = r[14];
  o[13] = (run > -INFINITY && (best - run) <= 1e-9 * fmax(1.0, fabs(best))) ? 1.0 : 0.0;
  o[14] = r[12]; o[15] = cx; o[16] = cy; o[17] = cz; o[18] = best; o[19] = run; o[20] = skip ? 1.0 : 0.0;
  o[21] = 0; o[22] = 0; o[23] = 0;
}

__global__ void __launch_bounds__(128) k_anns(const double* __restrict__ rect, int K, const int32_t* __restrict__ first_point,
                                              const int32_t* __restrict__ finst, const int32_t* __restrict__ cop,
                                              const int64_t* __restrict__ inst_off, const InstDev* __restrict__ idev,
                                              const dfu_view* __restrict__ views, const unsigned long long* __restrict__ zmax,
                                              double* __restrict__ boxes) {
  const int k = blockIdx.x * 128 + threadIdx.x;
  if (k >= K) return;
  const int g = finst[first_point[k]];
  const InstDev d = idev[g];
  double* o = boxes + (int64_t)k * DFU_BOX_DOUBLES;
  o[0] = d.view; o[1] = g; o[2] = k - cop[inst_off[g]]; o[3] = d.cls;
  anns_record(rect + (int64_t)k * 16, dfu_unorder_f64(zmax[g]), views[d.view].M43, d.car_rule, d.anns_mode, o);
}

// ---------------------------------------------------------------- stand-alone GenerateAnns (my_loader.py:633-684)
__global__ void __launch_bounds__(256) k_anns_prepare(const double* __restrict__ pts, int64_t n, double* __restrict__ x,
                                                      double* __restrict__ y, int32_t* __restrict__ seg,
                                                      unsigned long long* __restrict__ zmax) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  unsigned long long b = 0;
  if (i < n) {
    x[i] = pts[3 * i]; y[i] = pts[3 * i + 1]; seg[i] = 0;
    b = dfu_order_f64(pts[3 * i + 2]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long ob = __shfl_xor_sync(0xffffffffu, b, o);
    b = ob > b ? ob : b;
  }
  if ((threadIdx.x & 31) == 0 && b) atomicMax(zmax, b);
}

__global__ void __launch_bounds__(128) k_anns_single(const double* __restrict__ rect, int K, const dfu_view vw, int car_rule,
                                                     const unsigned long long* __restrict__ zmax, double* __restrict__ boxes) {
  const int k = blockIdx.x * 128 + threadIdx.x;
  if (k >= K) return;
  double* o = boxes + (int64_t)k * DFU_BOX_DOUBLES;
  o[0] = 0; o[1] = 0; o[2] = k; o[3] = 0;
  anns_record(rect + (int64_t)k * 16, dfu_unorder_f64(*zmax), vw.M43, car_rule & 1, car_rule >> 1, o);
}

#define GA_CLUSTER OPS_B
#define GA_FIT OPS_C
enum { GA_X = 160, GA_Y, GA_SEG, GA_MISC, GA_ANG, GA_RECT, GA_BOX };

extern "C" int dfu_generate_anns_sync(dfu_ctx* ctx, void* stream, const dfu_view* view_host, const double* points, int64_t n,
                                      int car_aspect_rule, double R0, double Rd, const double* angles_cs_host, int na,
                                      double* boxes_out_host, int64_t capacity, int64_t* n_boxes_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || !view_host || n < 0 || !angles_cs_host || na <= 0 || na > DFU_N_ANGLES_MAX || !n_boxes_host)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_generate_anns_sync: bad argument");
  *n_boxes_host = 0;
  if (n == 0) return DFU_OK;        // my_loader.py:635: empty objects write nothing
  if (!points) return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_generate_anns_sync: points is NULL");
  cudaStream_t st = (cudaStream_t)stream;
  DFU_SLOT(ctx, x, double*, GA_X, n * 8);
  DFU_SLOT(ctx, y, double*, GA_Y, n * 8);
  DFU_SLOT(ctx, seg, int32_t*, GA_SEG, n * 4);
  DFU_SLOT(ctx, misc, double*, GA_MISC, 64);
  DFU_SLOT(ctx, ang, double*, GA_ANG, (size_t)na * 16);
  const double rr[3] = {R0, Rd, 0.0};
  DFU_CUDA(ctx, cudaMemcpyAsync(misc, rr, 24, cudaMemcpyHostToDevice, st));   // misc[2] = zmax bits (0 = none yet)
  DFU_CUDA(ctx, cudaMemcpyAsync(ang, angles_cs_host, (size_t)na * 16, cudaMemcpyHostToDevice, st));
  k_anns_prepare<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(points, n, x, y, seg, (unsigned long long*)(misc + 2));
  DFU_LAUNCH_CHECK(ctx);
  int rc = dfu_grid_err_reset(ctx, st, GA_CLUSTER);
  if (rc) return rc;
  ClusterOut co;
  rc = dfu_cluster_points(ctx, st, GA_CLUSTER, x, y, seg, nullptr, n, 1, misc, misc + 1, &co);
  if (rc) return rc;
  int32_t hk[2] = {0, 0};
  DFU_CUDA(ctx, cudaMemcpyAsync(&hk[0], co.ncluster, 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(&hk[1], dfu_grid_err_ptr(ctx, GA_CLUSTER), 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  if (hk[1] & 2) return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "dfu_generate_anns_sync: R0 <= 0, Rd < 0 or R0 + Rd*|p| > sqrt(2)*R0 for some point (the cell neighbourhood of the clustering would miss pairs)");
  if (hk[1]) return dfu_set_err(ctx, DFU_ERR_DATA, "dfu_generate_anns_sync: a coordinate is NaN or the set spans more than 65533 cells");
  const int K = hk[0];
  *n_boxes_host = K;
  if (K > capacity || !boxes_out_host)
    return dfu_set_err(ctx, DFU_ERR_CAPACITY, "box output needs %d records, capacity %lld", K, (long long)capacity);
  DFU_SLOT(ctx, rect, double*, GA_RECT, (size_t)K * 16 * 8);
  DFU_SLOT(ctx, boxes, double*, GA_BOX, (size_t)K * DFU_BOX_DOUBLES * 8);
  rc = dfu_fit_clusters(ctx, st, GA_FIT, x, y, co.cluster_of_point, n, K, ang, na, rect);
  if (rc) return rc;
  k_anns_single<<<(K + 127) / 128, 128, 0, st>>>(rect, K, *view_host, car_aspect_rule, (const unsigned long long*)(misc + 2), boxes);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cudaMemcpyAsync(boxes_out_host, boxes, (size_t)K * DFU_BOX_DOUBLES * 8, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  return DFU_OK;
}

// ---------------------------------------------------------------- host orchestration
enum {
  IS_CNT = 100, IS_POFF, IS_KEY, IS_VAL, IS_SKEY, IS_SVAL, IS_TMP, IS_SEGOFF, IS_PX, IS_PY, IS_PZ, IS_SEG, IS_INST,
  IS_PRM, IS_KEEP, IS_RACT, IS_GATE, IS_CNT12, IS_FLAG, IS_FSCAN, IS_FX, IS_FY, IS_FZ, IS_FINST, IS_ZMAX, IS_INSTOFF,
  IS_GPRM, IS_IDEV, IS_ANGLES, IS_RECT, IS_BOXES, IS_XY
};

#define NBLK(n) (unsigned)(((n) + I_THREADS - 1) / I_THREADS)

extern "C" int dfu_fuse_instances_sync(dfu_ctx* ctx, const dfu_inst_params* ip, int G, const double* angles_cs_host,
                                       int na, double* boxes_out_host, int64_t capacity_boxes, int64_t* n_boxes_host,
                                       int64_t* seg_counts_host) {
  DFU_ON_DEVICE(ctx);
  if (!ctx || (ctx->stage & 6) != 6)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_instances_sync: project and densify stages must have run");
  if (G != ctx->total_inst || (G > 0 && !ip) || !angles_cs_host || na <= 0 || na > DFU_N_ANGLES_MAX || !n_boxes_host)
    return dfu_set_err(ctx, DFU_ERR_ARG, "dfu_fuse_instances_sync: bad argument (instances %d, bound batch has %d)", G,
                       ctx->total_inst);
  *n_boxes_host = 0;
  if (seg_counts_host) memset(seg_counts_host, 0, sizeof(int64_t) * 6 * (size_t)G);
  if (G == 0) return DFU_OK;
  const int V = ctx->V;
  cudaStream_t st = ctx->stream;
  const int nseg = 2 * G;
  InstTables T;
  T.views = (const dfu_view*)ctx->slot[SL_VIEWS];
  T.offA = (const int64_t*)ctx->slot[SL_STATS] + (V + 1);
  T.offB = (const int64_t*)ctx->slot[SL_STATS] + 2 * (V + 1);
  T.voxoff = (const int64_t*)ctx->slot[SL_VOXOFF];
  T.a_list = (const int32_t*)ctx->slot[SL_ALIST];
  T.b_label = (const uint32_t*)ctx->slot[SL_BLABEL];
  T.repy_label = (const uint32_t*)ctx->slot[SL_REPY_LABEL];
  T.repy_xyz = (const double*)ctx->slot[SL_REPY_XYZ];
  T.lidar = (const float4*)ctx->lidar;
  T.V = V;
  T.EL = ctx->off_project[2 * (V + 1) + V];
  T.EP = ctx->off_vox[V];
  const int64_t E = T.EL + T.EP;
  if (E >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_DATA, "too many elements");
  // pinned staging: [0,4K) read-backs, then the parameter uploads
  const size_t up_bytes = (size_t)nseg * sizeof(SegPrm) + (size_t)G * (3 * 8 + sizeof(InstDev)) + (size_t)na * 16 + 256;
  const size_t rb_bytes = 4096 + (size_t)(nseg + 1) * 8 + (size_t)nseg * 8;
  char* pin = (char*)dfu_pinned(ctx, rb_bytes + up_bytes);
  if (!pin) return DFU_ERR_CUDA;
  int64_t* rb = (int64_t*)pin;
  char* up = pin + rb_bytes;
  // ---- membership pairs
  int64_t NP = 0;
  if (E > 0) {
    DFU_PROF(ctx, st, PT_INST_PAIRS);
    DFU_SLOT(ctx, cnt, int32_t*, IS_CNT, E * 4);
    DFU_SLOT(ctx, poff, int64_t*, IS_POFF, (E + 1) * 8);
    k_inst_count<<<NBLK(E), I_THREADS, 0, st>>>(T, cnt);
    DFU_LAUNCH_CHECK(ctx);
    int rc = dfu_exclusive_sum_i32(ctx, st, cnt, poff, E);
    if (rc) return rc;
    DFU_CUDA(ctx, cudaMemcpyAsync(rb, poff + E, 8, cudaMemcpyDeviceToHost, st));
    DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
    NP = rb[0];
  }
  if (NP >= 0x7fffffffLL) return dfu_set_err(ctx, DFU_ERR_DATA, "too many (point, instance) pairs");
  if (NP == 0) return DFU_OK;
  DFU_SLOT(ctx, key, uint32_t*, IS_KEY, NP * 4);
  DFU_SLOT(ctx, val, int32_t*, IS_VAL, NP * 4);
  DFU_SLOT(ctx, skey, uint32_t*, IS_SKEY, NP * 4);
  DFU_SLOT(ctx, sval, int32_t*, IS_SVAL, NP * 4);
  DFU_SLOT(ctx, seg_off, int64_t*, IS_SEGOFF, (size_t)(nseg + 1) * 8);
  DFU_SLOT(ctx, px, double*, IS_PX, NP * 8);
  DFU_SLOT(ctx, py, double*, IS_PY, NP * 8);
  DFU_SLOT(ctx, pz, double*, IS_PZ, NP * 8);
  DFU_SLOT(ctx, seg, int32_t*, IS_SEG, NP * 4);
  DFU_SLOT(ctx, inst, int32_t*, IS_INST, NP * 4);
  {
    DFU_PROF(ctx, st, PT_INST_SORT);
    k_inst_emit<<<NBLK(E), I_THREADS, 0, st>>>(T, (const int64_t*)ctx->slot[IS_POFF], key, val);
    DFU_LAUNCH_CHECK(ctx);
    int bits = 1;
    while ((1ll << bits) < nseg) bits++;
    size_t tb = 0;
    DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tb, key, skey, val, sval, (int)NP, 0, bits, st));
    DFU_SLOT(ctx, tmp, void*, IS_TMP, tb);
    DFU_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tb, key, skey, val, sval, (int)NP, 0, bits, st));
    ctx->launches += 4;
    k_seg_bounds<<<NBLK(NP), I_THREADS, 0, st>>>(skey, NP, nseg, seg_off);
    DFU_LAUNCH_CHECK(ctx);
    k_inst_gather<<<NBLK(NP), I_THREADS, 0, st>>>(T, skey, sval, NP, px, py, pz, seg, inst);
    DFU_LAUNCH_CHECK(ctx);
  }
  // ---- parameter upload
  SegPrm* sp_h = (SegPrm*)up;
  double* g_h = (double*)(sp_h + nseg);            // gate[G] R0[G] Rd[G]
  InstDev* id_h = (InstDev*)(g_h + 3 * G);
  double* ang_h = (double*)(id_h + G);
  {
    int g = 0;
    for (int v = 0; v < V; v++)
      for (int j = 0; j < ctx->views_host[v].n_inst; j++, g++) {
        const dfu_inst_params& q = ip[g];
        sp_h[2 * g] = dfu_seg_prm(q.lidar_r, q.lidar_nb, q.lidar_every_k, q.lidar_enabled);
        sp_h[2 * g].n0 = q.lidar_n0; sp_h[2 * g].r_small = q.lidar_r_small;
        sp_h[2 * g + 1] = dfu_seg_prm(q.pseudo_r, q.pseudo_nb, q.pseudo_every_k, 1);
        sp_h[2 * g + 1].n_hi = q.pseudo_n_hi; sp_h[2 * g + 1].k_hi = q.pseudo_k_hi;
        sp_h[2 * g + 1].n_mid = q.pseudo_n_mid; sp_h[2 * g + 1].k_mid = q.pseudo_k_mid;
        g_h[g] = q.gate_dist; g_h[G + g] = q.cluster_R0; g_h[2 * G + g] = q.cluster_Rd;
        id_h[g] = InstDev{v, q.cls, q.car_aspect_rule, q.anns_mode};
      }
  }
  memcpy(ang_h, angles_cs_host, (size_t)na * 16);
  DFU_SLOT(ctx, prm, SegPrm*, IS_PRM, (size_t)nseg * sizeof(SegPrm));
  DFU_SLOT(ctx, gprm, double*, IS_GPRM, (size_t)G * 24);
  DFU_SLOT(ctx, idev, InstDev*, IS_IDEV, (size_t)G * sizeof(InstDev));
  DFU_SLOT(ctx, angles, double*, IS_ANGLES, (size_t)na * 16);
  // by kernel, not by copy engine: the engine may be busy with another stream's bulk upload for tens of milliseconds
  int urc;
  if ((urc = dfu_upload_pinned(ctx, st, prm, sp_h, (size_t)nseg * sizeof(SegPrm)))) return urc;
  if ((urc = dfu_upload_pinned(ctx, st, gprm, g_h, (size_t)G * 24))) return urc;
  if ((urc = dfu_upload_pinned(ctx, st, idev, id_h, (size_t)G * sizeof(InstDev)))) return urc;
  if ((urc = dfu_upload_pinned(ctx, st, angles, ang_h, (size_t)na * 16))) return urc;
  // ---- radius filters, gate, fuse
  DFU_SLOT(ctx, keep, uint8_t*, IS_KEEP, NP);
  DFU_SLOT(ctx, ract, uint8_t*, IS_RACT, NP);
  DFU_SLOT(ctx, gate, uint8_t*, IS_GATE, NP);
  DFU_SLOT(ctx, cnt12, int32_t*, IS_CNT12, (size_t)nseg * 8);
  DFU_SLOT(ctx, flag, int32_t*, IS_FLAG, NP * 4);
  DFU_SLOT(ctx, fscan, int64_t*, IS_FSCAN, (NP + 1) * 8);
  DFU_SLOT(ctx, zmax, unsigned long long*, IS_ZMAX, (size_t)G * 8);
  DFU_SLOT(ctx, inst_off, int64_t*, IS_INSTOFF, (size_t)(G + 1) * 8);
  int32_t *cnt1 = cnt12, *cnt2 = cnt12 + nseg;
  int rc = dfu_grid_err_reset(ctx, st, OPS_A);
  if (rc) return rc;
  rc = dfu_grid_err_reset(ctx, st, OPS_B);
  if (rc) return rc;
  rc = dfu_radius_keep(ctx, st, OPS_A, px, py, pz, seg, seg_off, NP, nseg, prm, keep);
  if (rc) return rc;
  DFU_CUDA(ctx, cudaMemsetAsync(cnt12, 0, (size_t)nseg * 8, st));
  DFU_CUDA(ctx, cudaMemsetAsync(zmax, 0, (size_t)G * 8, st));
  k_seg_count<<<NBLK(NP), I_THREADS, 0, st>>>(seg, keep, NP, cnt1);
  DFU_LAUNCH_CHECK(ctx);
  k_ref_active<<<NBLK(NP), I_THREADS, 0, st>>>(seg, keep, NP, ract, (uint8_t*)flag);      // flag (int32 [NP]) is written later: its bytes hold qactive until then
  DFU_LAUNCH_CHECK(ctx);
  rc = dfu_gate_keep(ctx, st, OPS_A, px, py, pz, inst, NP, px, py, pz, inst, ract, NP, G, gprm, gate, (const uint8_t*)flag);
  if (rc) return rc;
  k_fuse_flags<<<NBLK(NP), I_THREADS, 0, st>>>(seg, keep, gate, cnt1, NP, flag, cnt2);
  DFU_LAUNCH_CHECK(ctx);
  rc = dfu_exclusive_sum_i32(ctx, st, flag, fscan, NP);
  if (rc) return rc;
  k_inst_offsets<<<(G + 1 + 127) / 128, 128, 0, st>>>(seg_off, fscan, NP, G, inst_off);
  DFU_LAUNCH_CHECK(ctx);
  // read-back: NF, seg_off, cnt1/cnt2, grid error flag
  int64_t* rb_seg = rb + 8;
  int32_t* rb_cnt = (int32_t*)(rb_seg + nseg + 1);
  DFU_CUDA(ctx, cudaMemcpyAsync(rb, fscan + NP, 8, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(rb + 1, dfu_grid_err_ptr(ctx, OPS_A), 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(rb_seg, seg_off, (size_t)(nseg + 1) * 8, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(rb_cnt, cnt12, (size_t)nseg * 8, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  const int64_t NF = rb[0];
  if ((int32_t)rb[1] != 0) return dfu_set_err(ctx, DFU_ERR_DATA, "radius/gate grid: a coordinate is NaN or spans more than 65533 cells");
  if (seg_counts_host)
    for (int g = 0; g < G; g++) {
      seg_counts_host[6 * g + 0] = rb_seg[2 * g + 1] - rb_seg[2 * g];
      seg_counts_host[6 * g + 1] = rb_seg[2 * g + 2] - rb_seg[2 * g + 1];
      seg_counts_host[6 * g + 2] = rb_cnt[2 * g];
      seg_counts_host[6 * g + 3] = rb_cnt[2 * g + 1];
      seg_counts_host[6 * g + 4] = rb_cnt[nseg + 2 * g + 1];
    }
  if (NF == 0) return DFU_OK;
  DFU_SLOT(ctx, fx, double*, IS_FX, NF * 8);
  DFU_SLOT(ctx, fy, double*, IS_FY, NF * 8);
  DFU_SLOT(ctx, fz, double*, IS_FZ, NF * 8);
  DFU_SLOT(ctx, finst, int32_t*, IS_FINST, NF * 4);
  k_fuse_scatter<<<NBLK(NP), I_THREADS, 0, st>>>(fscan, keep, gate, seg, cnt1, px, py, pz, NP, fx, fy, fz, finst, zmax);
  DFU_LAUNCH_CHECK(ctx);
  // ---- clustering + fit + annotations
  ClusterOut co;
  rc = dfu_cluster_points(ctx, st, OPS_B, fx, fy, finst, inst_off, NF, G, gprm + G, gprm + 2 * G, &co);
  if (rc) return rc;
  DFU_CUDA(ctx, cudaMemcpyAsync(rb, co.ncluster, 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, cudaMemcpyAsync(rb + 1, dfu_grid_err_ptr(ctx, OPS_B), 4, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  const int K = *(int32_t*)rb;
  if (*(int32_t*)(rb + 1) & 2) return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "cluster: R0 <= 0, Rd < 0 or R0 + Rd*|p| > sqrt(2)*R0 for some point (the cell neighbourhood of the clustering would miss pairs)");
  if (*(int32_t*)(rb + 1) != 0) return dfu_set_err(ctx, DFU_ERR_DATA, "cluster grid: a coordinate is NaN or spans more than 65533 cells");
  *n_boxes_host = K;
  if (K == 0) return DFU_OK;
  if (K > capacity_boxes || !boxes_out_host)
    return dfu_set_err(ctx, DFU_ERR_CAPACITY, "box output needs %d records, capacity %lld", K, (long long)capacity_boxes);
  DFU_SLOT(ctx, rect, double*, IS_RECT, (size_t)K * 16 * 8);
  DFU_SLOT(ctx, boxes, double*, IS_BOXES, (size_t)K * DFU_BOX_DOUBLES * 8);
  rc = dfu_fit_clusters(ctx, st, OPS_C, fx, fy, co.cluster_of_point, NF, K, angles, na, rect);
  if (rc) return rc;
  k_anns<<<(K + 127) / 128, 128, 0, st>>>(rect, K, co.first_point, finst, co.cluster_of_point, inst_off, idev, T.views, zmax, boxes);
  DFU_LAUNCH_CHECK(ctx);
  DFU_CUDA(ctx, cudaMemcpyAsync(boxes_out_host, boxes, (size_t)K * DFU_BOX_DOUBLES * 8, cudaMemcpyDeviceToHost, st));
  DFU_CUDA(ctx, dfu_stream_sync(ctx, st));
  if (seg_counts_host)
    for (int k = 0; k < K; k++) seg_counts_host[6 * (int64_t)boxes_out_host[(size_t)k * DFU_BOX_DOUBLES + 1] + 5] += 1;
  return DFU_OK;
}
