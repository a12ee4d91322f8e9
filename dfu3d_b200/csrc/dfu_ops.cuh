// Internal (non-ABI) interfaces shared by the segmented operators: uniform hash grid, radius count, NN gate,
// clustering + rectangle fit.  "Segment" = one independent point set (an instance's LiDAR or pseudo points, or one
// caller-provided cloud); points of a segment are contiguous and segments ascend.
#pragma once
#include "dfu_common.cuh"

struct SegPrm {
  double r;          // radius (radius filter) / max distance (gate)
  int32_t nb;        // keep iff count > nb
  int32_t every_k;   // uniform_down_sample stride applied before the filter (1 = none)
  int32_t enabled;   // 0: the segment is emptied (quirk Q5)
  // count-dependent schedules of the KITTI loaders (my_loader0.py:739-766): resolved on the device from the
  // segment's point count n by k_seg_resolve, BEFORE anything else reads r / every_k
  int32_t n0;        // >= 0: r = (n > n0) ? r : r_small
  double r_small;
  int32_t n_hi, k_hi, n_mid, k_mid;   // n_hi >= 0: every_k = n > n_hi ? k_hi : (n > n_mid ? k_mid : every_k)
};
static inline SegPrm dfu_seg_prm(double r, int nb, int every_k, int enabled) {
  SegPrm p;
  p.r = r; p.nb = nb; p.every_k = every_k; p.enabled = enabled; p.n0 = -1; p.r_small = r; p.n_hi = -1; p.k_hi = 1; p.n_mid = -1;
  p.k_mid = 1;
  return p;
}

// Uniform grid over the ACTIVE points of all segments.  Cell coordinates are relative to the segment's bounding-box
// minimum, 16 bits per axis, key = seg:16 | cx:16 | cy:16 | cz:16 (cx,cy,cz offset by +1 so that -1 is representable).
struct DfuGrid {
  int64_t n;
  int nseg;
  const double* cell;            // [nseg] cell edge
  unsigned long long* segmin;    // [nseg*3] orderable bits of the per-segment minimum
  unsigned long long* skey;      // [n] sorted keys (inactive points: ~0, sorted last)
  int32_t* sidx;                 // [n] original index of the sorted point
  double *sx, *sy, *sz;          // [n] sorted coordinates
  int32_t* pcell;                // [n] cell ordinal of the sorted point
  unsigned long long* ukey;      // [ncell] unique keys
  int32_t* ustart;               // [ncell+1] first sorted position of each cell
  int32_t* segcell;              // [nseg+1] first cell of each segment (cells of a segment are contiguous)
  unsigned long long* hkey;      // [hmask+1] open-addressing hash of the cell keys (empty: ~0) -> hval = cell ordinal
  int32_t* hval;
  unsigned hmask;
  int dims;                      // 2 (z == NULL: every cz is 1) or 3
  int32_t* ncell;                // device scalar
  int32_t* errflag;              // device scalar: != 0 if a coordinate overflowed 16 bits
};

#define GRID_SLOTS 16   /* + 1 (cell edges) used by the operators: reserve GRID_SLOTS + 1 */
// slot_base .. slot_base+GRID_SLOTS-1 are used.  z may be NULL (2-D grid).  active may be NULL (all active).
int dfu_grid_build(dfu_ctx* ctx, cudaStream_t st, int slot_base, const double* x, const double* y, const double* z,
                   const int32_t* seg, const uint8_t* active, int64_t n, int nseg, const double* cell_dev, DfuGrid* g);


__device__ __forceinline__ unsigned long long dfu_grid_key(int seg, int cx, int cy, int cz) {
  return ((unsigned long long)(unsigned)seg << 48) | ((unsigned long long)(unsigned)cx << 32) |
         ((unsigned long long)(unsigned)cy << 16) | (unsigned long long)(unsigned)cz;
}

// cell ordinal of a key, or -1: one or two probes instead of a 20-step binary search of the cell table
__device__ __forceinline__ unsigned dfu_hash_u64(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return (unsigned)k;
}
__device__ __forceinline__ int dfu_grid_find(const DfuGrid& g, unsigned long long key) {
  unsigned h = dfu_hash_u64(key) & g.hmask;
  while (true) {
    const unsigned long long k = __ldg(g.hkey + h);
    if (k == key) return __ldg(g.hval + h);
    if (k == ~0ull) return -1;
    h = (h + 1) & g.hmask;
  }
}

// first index in a[0..n) with a[i] >= v
__device__ __forceinline__ int dfu_lower_bound_u64(const unsigned long long* __restrict__ a, int n, unsigned long long v) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (__ldg(a + mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// cell coordinate of a value relative to the segment minimum (+1 offset); -1000 if far outside
__device__ __forceinline__ int dfu_cell_coord(double v, double mn, double cell) {
  double c = floor((v - mn) / cell);
  if (!(c >= -1.0 && c < 65533.0)) return -1000;
  return (int)c + 1;
}

// radius filter (Open3D remove_radius_outlier semantics) over all segments; keep [n] u8 (0 for inactive points)
int dfu_radius_keep(dfu_ctx* ctx, cudaStream_t st, int slot_base, const double* x, const double* y, const double* z,
                    const int32_t* seg, const int64_t* seg_off_dev, int64_t n, int nseg, SegPrm* prm_dev,
                    uint8_t* keep);

// keep[i] = exists reference point j of the same segment with sqrt(|q_i - r_j|^2) < maxdist[seg]
int dfu_gate_keep(dfu_ctx* ctx, cudaStream_t st, int slot_base, const double* qx, const double* qy, const double* qz,
                  const int32_t* qseg, int64_t nq, const double* rx, const double* ry, const double* rz,
                  const int32_t* rseg, const uint8_t* ractive, int64_t nr, int nseg, const double* maxdist_dev,
                  uint8_t* keep, const uint8_t* qactive = nullptr);   // qactive (nullable): queries whose answer is needed

// Adaptive-range clustering + rectangle search on point sets ("instances"): x,y [n], inst [n] ascending,
// inst_off_dev [ninst+1].  Results stay in ctx slots; see dfu_cluster.cu.
struct ClusterOut {
  int64_t n;
  int32_t* cluster_of_point;   // [n] global cluster number (numbered by (instance, min point index))
  int32_t* ncluster;           // device scalar
  int32_t* first_point;        // [ncluster] min point index of the cluster
};
int dfu_cluster_points(dfu_ctx* ctx, cudaStream_t st, int slot_base, const double* x, const double* y, const int32_t* inst,
                       const int64_t* inst_off_dev, int64_t n, int ninst, const double* R0_dev, const double* Rd_dev,
                       ClusterOut* out);

// Rectangle search per cluster; rect_dev [K][16] = a[4] b[4] c[4] theta best runner-up n.  K = host copy of *ncluster.
#define FIT_SLOTS 8
int dfu_fit_clusters(dfu_ctx* ctx, cudaStream_t st, int slot_base, const double* x, const double* y, const int32_t* cop,
                     int64_t n, int K, const double* angles_dev, int na, double* rect_dev);
#define CLUSTER_SLOTS (GRID_SLOTS + 1 + 10)

// grid coordinate-overflow flag of the operators sharing slot base sb (sticky until reset)
int dfu_grid_err_reset(dfu_ctx* ctx, cudaStream_t st, int sb);
const int32_t* dfu_grid_err_ptr(dfu_ctx* ctx, int sb);
