/*
 * dfu3d.h — C ABI of libdfu3d.so: the sm_100a CUDA implementation of DFU3D's data-level
 * fusion / noise-suppression / pseudo-box path.
 *
 * The reference has no FFI: its boundary is a set of Python functions under
 * /root/reference/tools/PENet.  Each entry point below names the reference
 * function(s) whose arithmetic it replaces (file:line, relative to tools/PENet).  All pointers
 * that are not marked "host" are DEVICE pointers owned by the caller (torch allocates); the
 * library never frees them.  Scratch lives in the context.  All launches go to the caller's
 * cudaStream_t (passed as void*).  Return value: 0 = ok, negative = dfu_status; the message of
 * the last failure is returned by dfu_last_error().  Capacity overflow is an error, never a
 * silent truncation.  Functions whose names end in _sync return host-visible counts and
 * therefore synchronise the stream once before returning.
 */
#ifndef DFU3D_H
#define DFU3D_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dfu_ctx dfu_ctx;

enum dfu_status {
  DFU_OK = 0,
  DFU_ERR_ARG = -1,        /* bad argument */
  DFU_ERR_CUDA = -2,       /* CUDA runtime error (see dfu_last_error) */
  DFU_ERR_CAPACITY = -3,   /* a caller-provided buffer is too small */
  DFU_ERR_UNSUPPORTED = -4,
  DFU_ERR_DATA = -5        /* input violates a documented precondition (e.g. > 32 instances) */
};

#define DFU_MAX_INSTANCES 32   /* instance masks per view packed into one uint32 label word */
#define DFU_N_ANGLES_MAX 96    /* rectangle search angles (reference: 89) */
#define DFU_BOX_DOUBLES 24     /* doubles per box record, see dfu_instances_run */

/* Per-view constants.  Everything here is derived on the host exactly as the reference derives it
 * (float32 products / float32 inverse: dataloaders/calibration_kitti.py:62-79,95-101,110). */
typedef struct dfu_view {
  float M43[12];       /* V2C^T · R0^T, row-major 4x3      (lidar_to_rect, calibration_kitti.py:110) */
  float P2T[12];       /* P2^T, row-major 4x3              (rect_to_img,   calibration_kitti.py:120) */
  float inv44[16];     /* inv((R0ext · V2Cext)^T), 4x4     (rect_to_lidar, calibration_kitti.py:101) */
  float cu, cv, fu, fv, tx, ty;                         /* calibration_kitti.py:74-79 */
  int32_t lidar_cnt;   /* LiDAR rows of this view (before the FOV cull) */
  int32_t n_inst;      /* instance masks of this view (<= DFU_MAX_INSTANCES) */
  double plane[4];     /* ground plane a,b,c,d (dataloaders/my_loader.py:448-469 output) */
  int64_t lidar_off;   /* first row of this view's sweep in lidar_xyzi (views may share a sweep) */
  int64_t slot_off;    /* first per-point output slot of this view (exclusive sum of lidar_cnt) */
  int32_t inst_off;    /* first instance of this view in the flat per-instance arrays */
  int32_t pad0;
} dfu_view;

typedef struct dfu_fusion_params {
  int32_t H, W;              /* depth map rows / cols (reference crop: vis_utils.py:147,161) */
  int32_t Hm, Wm;            /* label-map (mask / image) rows / cols, Hm >= H, Wm >= W */
  int32_t bound_x, bound_y;  /* pixel bounds of the label-inheritance loop (my_loader.py:526: 1600, 900) */
  int32_t round_mode;        /* 0 = np.round half-to-even (my_loader.py:520); 1 = truncate (my_loader.py:113) */
  int32_t image_is_i32;      /* image dtype: 0 = uint8 [.,3], 1 = int32 [.,3] */
  float   min_depth;         /* my_loader.py:507 (0.001) */
  int32_t max_voxels;        /* my_loader.py:74 (1000000): angular voxels kept per view, in first-appearance order; 0 = no limit */
  double  max_z;             /* height gate of my_loader.py:540 (1.0) */
  double  plane_offset;      /* my_loader.py:471 (0.05) */
  double  plane_range[4];    /* x lo, x hi, y lo, y hi (my_loader.py:471: -54, 54, -54, 54) */
  double  min_theta;         /* to_sphere_coords gate, my_loader.py:176 (1.5) */
  int32_t max_per_voxel;     /* my_loader.py:73 (100) */
  int32_t no_fov_cull;       /* 1: the sweep is already FOV-culled by the caller (vis_utils.py:152-154); every row is kept */
} dfu_fusion_params;

/* Per-instance filter schedule (my_loader.py:581-599; variants my_loader_KittiPandasetWaymo.py:780-823). */
typedef struct dfu_inst_params {
  int32_t cls;               /* index into thing_classes, copied into the box record */
  int32_t lidar_enabled;     /* 0 reproduces quirk Q5 (LiDAR filter on an empty cloud, my_loader.py:592-594) */
  int32_t lidar_nb, lidar_every_k;
  int32_t pseudo_nb, pseudo_every_k;
  int32_t car_aspect_rule;   /* 1: skip boxes with aspect > 5 (my_loader.py:651) */
  int32_t anns_mode;         /* 0: GenerateAnns of my_loader.py:633-684; 1: of my_loader0.py:811-930 (kitti_legacy: height
                                floor 1.56, 'prior_linecorner' centre shift, every box forced to 3.9 x 1.6, boxes closer
                                than 10 m skipped) */
  double  lidar_r, pseudo_r;
  double  gate_dist;         /* BallQuery C (my_loader.py:603: 0.1) */
  double  cluster_R0, cluster_Rd;   /* rectangle_fitting.py:36-37 */
  /* count-dependent schedules (my_loader0.py:739-766), evaluated on the device from the instance's point counts:
   * LiDAR radius = lidar_r if n > lidar_n0 else lidar_r_small (lidar_n0 < 0: always lidar_r);
   * pseudo stride = pseudo_k_hi if n > pseudo_n_hi, pseudo_k_mid if n > pseudo_n_mid, else pseudo_every_k
   * (pseudo_n_hi < 0: always pseudo_every_k) */
  double  lidar_r_small;
  int32_t lidar_n0;
  int32_t pseudo_n_hi, pseudo_k_hi, pseudo_n_mid, pseudo_k_mid;
  int32_t pad;
} dfu_inst_params;

int  dfu_create(dfu_ctx** out, int device);
void dfu_destroy(dfu_ctx* ctx);
const char* dfu_last_error(const dfu_ctx* ctx);
/* number of kernels this context has launched since creation (bench.py: gpu_launches) */
int64_t dfu_launch_count(const dfu_ctx* ctx);

/* Per-kernel device timing with CUDA events on the launching stream (bench.py: roofline.achieved).
 * While enabled every instrumented launch is bracketed by two events (pool of 8192 launches between
 * reads).  dfu_profile_read synchronises the device, accumulates, and returns the totals per tag:
 *  0 pack_masks  1 project  2 lists  3 densify_a  4 densify_b  5 densify_c  6 vox_gather  7 emit_cloud
 *  8 inst_pairs  9 inst_sort  10 grid_build  11 radius  12 gate  13 cluster  14 fit  15 sor_knn
 *  16 sor_mask  17 sparse_depth  18 misc */
#define DFU_PROF_NTAGS 24
int dfu_profile_enable(dfu_ctx* ctx, int on);
int dfu_profile_read(dfu_ctx* ctx, double* ms_out, int64_t* calls_out, int n, int reset);

/* ---- S1: LiDAR -> image projection with label inheritance --------------------------------------
 * Replaces Calibration.lidar_to_rect / rect_to_img / lidar_to_img (calibration_kitti.py:104-132),
 * get_fov_flag (vis_utils.py:108-123), above_plane (my_loader.py:471-484) and the per-instance
 * getpixel loop (my_loader.py:518-530).  One thread per (view, LiDAR row).
 *   lidar_xyzi  float4 rows [x y z i]
 *   labelmap    uint32 [V][Hm][Wm], bit j = instance j of the view (dfu_pack_masks); may be NULL
 * Per-slot outputs (slot = view.slot_off + row), any may be NULL:
 *   out_uv   float2  un-rounded u,v      out_depth float   rect depth
 *   out_pix  int2    rounded/truncated x,y
 *   out_flags uint8  bit0 FOV (un-rounded u,v inside H x W and depth >= 0), bit1 above-plane,
 *                    bit2 rounded pixel inside bound_x x bound_y
 *   out_label uint32 label word at the rounded pixel (0 if outside the label map or bounds)
 */
int dfu_project_label(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_dev, int max_lidar_cnt,
                      const dfu_fusion_params* prm_host, const void* lidar_xyzi, const uint32_t* labelmap,
                      float* out_uv, float* out_depth, int32_t* out_pix, uint8_t* out_flags, uint32_t* out_label);

/* The Calibration methods (calibration_kitti.py:89-144) on caller arrays, one thread per row, with the reference's
 * dtypes and operation order.  view_host: only the calibration fields are read.
 *   op 0 lidar_to_rect  in0 float  [n][3]                          -> out0 float  [n][3]
 *   op 1 lidar_to_rect  in0 double [n][3] (my_loader.py:679)       -> out0 double [n][3]
 *   op 2 rect_to_img    in0 float  [n][3]                          -> out0 float  [n][2], out1 float [n] (depth)
 *   op 3 img_to_rect    in0 int64 u[n], in1 int64 v[n], in2 float d[n] -> out0 double [n][3]
 *   op 4 rect_to_lidar  in0 double [n][3]                          -> out0 double [n][3]
 *   op 5 lidar_to_img   in0 float  [n][3]                          -> out0 float  [n][2], out1 float [n]
 *   op 6 get_fov_flag   in0 float rect [n][3], in1 int32 {H, W} (device) -> out0 uint8 [n]   (vis_utils.py:108-123)
 *   op 7 corners3d_to_img_boxes  in0 double corners [n][3] (rect frame) -> out0 double [n][2] image x, y (:146-164) */
int dfu_calib_apply(dfu_ctx* ctx, void* stream, const dfu_view* view_host, int op, const void* in0, const void* in1,
                    const void* in2, int64_t n, void* out0, void* out1);

/* masks [sum n_inst][Hm][Wm] (uint8 when is_f32 == 0, float when 1; != 0 means set) -> labelmap.
 * Replaces Image.fromarray(np.uint8(mask)) (my_loader.py:525). */
int dfu_pack_masks(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_dev, const void* masks, int is_f32,
                   int Hm, int Wm, uint32_t* labelmap);

/* Compact alternative to dfu_pack_masks for callers that already hold BIT-PACKED instance masks (bit j of a pixel's word =
 * instance j of the view): src uint8 (<= 8 instances per view) or uint16 (<= 16) [V][Hm][Wm] -> labelmap uint32.  The
 * reference layout is one full plane per instance (vis_utils.py:218, my_loader.py:522-530); packed words are what travels
 * over PCIe in the compact end-to-end entry (pipeline.HostBatch.from_device(compact=True)). */
int dfu_widen_labelmap(dfu_ctx* ctx, void* stream, const void* src, int bytes_per_pixel, int64_t n, uint32_t* labelmap);

/* load_depth_input (my_loader.py:102-122): FOV cull, truncating pixel index, last-writer-wins
 * scatter of the rect depth into out_depth [V][H][W] (float64, zero elsewhere). */
int dfu_sparse_depth(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_dev, int max_lidar_cnt,
                     const dfu_fusion_params* prm_host, const void* lidar_xyzi, double* out_depth);

/* ---- the batched pipeline --------------------------------------------------------------------
 * dfu_fuse_begin binds a batch of V views; the stage calls below then run on the bound batch and
 * keep their intermediates in the context.  Stage order: project -> densify -> instances.
 */
int dfu_fuse_begin(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_host, const dfu_fusion_params* prm_host,
                   const void* lidar_xyzi, const float* depth /*[V][H][W]*/, const void* image /*[V][Hm][Wm][3]*/,
                   const uint32_t* labelmap /*[V][Hm][Wm]*/);

/* S1 + ordered compaction (vis_utils.py:152-154, my_loader.py:511-530).  counts_host [V][3]:
 * rows after the FOV cull, rows above the plane, of those rows with an in-bounds pixel. */
int dfu_fuse_project_sync(dfu_ctx* ctx, int64_t* counts_host);

/* S2: pixel -> 3-D back-projection + height gate + angular-voxel dedup evaluated once
 * (my_loader.py:507-509,532-542,554-557,166-180,247-275; calibration_kitti.py:89-102,134-144).
 * counts_host [V][4]: valid pixels M, points passing the height gate, angular voxels, voxels with
 * more than max_per_voxel points (must be 0: DFU_ERR_DATA otherwise unless allow_overflow). */
int dfu_fuse_densify_sync(dfu_ctx* ctx, int64_t* counts_host, int allow_overflow);

/* Copy-out of the fused cloud (my_loader.py:611-617 + vis_utils.py:164): rows [x y z i*10 0 0 0 2]
 * for the FOV-culled LiDAR then [x y z 0 R/3 G/3 B/3 1] per angular voxel (first-appearance
 * order).  out is [sum over views (nF + nVox)][8]; dtype 0 = float64, 1 = float32, 2 = float16.
 * row_off_host [V+1] receives the first row of each view.   `out` must be 16-byte aligned (rows are written with 128-bit stores). */
int dfu_fuse_emit_cloud(dfu_ctx* ctx, void* out, int dtype, int64_t capacity_rows, int64_t* row_off_host);

/* Per-voxel debug / parity outputs: linear pixel index (row*W+col) of the min-z and min-y
 * representative of each angular voxel, and its xyz (float64).  Any pointer may be NULL. */
int dfu_fuse_emit_voxels(dfu_ctx* ctx, int32_t* repz_pix, int32_t* repy_pix, double* repz_xyz, double* repy_xyz,
                         int64_t capacity);

/* Per-slot S1 results of the bound batch (same layout as dfu_project_label).  Debug / parity output: the fused pipeline
 * evaluates the plane flag, the bounds flag, the pixel index and the label of FOV points only (the three lists nest);
 * a slot outside the FOV reads flags = 0, pix = (0, 0), label = 0 unless the batch was bound with no_fov_cull. */
int dfu_fuse_emit_project(dfu_ctx* ctx, int32_t* out_pix, uint8_t* out_flags, uint32_t* out_label);

/* Instances: gather of inherited LiDAR points and masked pseudo points, radius filters
 * (Open3D remove_radius_outlier / uniform_down_sample call sites my_loader.py:569-599), BallQuery
 * gate (:489-494,602-604), adaptive-range clustering (rectangle_fitting.py:161-191), rectangle
 * search (:113-159,83-111) and the GenerateAnns arithmetic (my_loader.py:633-684).
 *   inst_params_host [total instances]; boxes2d is copied through by the host wrapper.
 *   angles_cs_host   [n_angles][2] cos, sin of the search angles (host-evaluated np.cos/np.sin)
 * Box record (DFU_BOX_DOUBLES doubles): 0 view, 1 instance (global), 2 cluster ordinal, 3 cls,
 *   4 alpha, 5 h, 6 w, 7 l, 8 x, 9 y, 10 z (rect frame), 11 ry, 12 n points, 13 near-tie flag,
 *   14 theta_fit, 15 cx, 16 cy, 17 cz (lidar frame), 18 best cost, 19 runner-up cost,
 *   20 skipped-by-aspect flag, 21..23 reserved.
 * Records arrive sorted by (view, instance, cluster).  n_boxes_host receives the count.
 * seg_counts_host [total instances][6] (may be NULL): nL, nP, nL', nP', nP'', clusters. */
int dfu_fuse_instances_sync(dfu_ctx* ctx, const dfu_inst_params* inst_params_host, int n_inst_total,
                            const double* angles_cs_host, int n_angles, double* boxes_out_host, int64_t capacity_boxes,
                            int64_t* n_boxes_host, int64_t* seg_counts_host);

/* ---- stand-alone operators on caller-provided point sets (segments) -----------------------------
 * xyz: float64 [n][3] rows.  seg_off_host [nseg+1].  Used by the drop-in Python functions and by
 * the filter benchmarks. */

/* Back-projection of depth maps into the LiDAR frame (my_loader.py:507-509 depth threshold + row-major nonzero,
 * calibration_kitti.py:134-144 img_to_rect and :89-102 rect_to_lidar in float64; height_gate != 0 also applies
 * new_p[:, 2] < max_z of my_loader.py:540).  depth [V][H][W]; rows come out per view in row-major pixel order.
 * counts_host [V] receives the rows per view.  out_xyz == NULL: size query only.  out_xyz float64 [capacity][3],
 * out_pix int32 [capacity] linear pixel index (nullable).  Only H, W, min_depth, max_z of prm_host are read. */
int dfu_backproject_sync(dfu_ctx* ctx, void* stream, int V, const dfu_view* views_host, const dfu_fusion_params* prm_host,
                         const float* depth, int height_gate, double* out_xyz, int32_t* out_pix, int64_t capacity,
                         int64_t* counts_host);

/* Open3D remove_radius_outlier semantics (oracle/fakes.py): keep iff #{q: |p-q|^2 < r^2} > nb.
 * every_k_host (nullable, [nseg]): uniform_down_sample(every_k) applied to the segment first (call sites
 * my_loader.py:571, my_loader_KittiPandasetWaymo.py:792-821): only rows 0, k, 2k, ... of the segment take part
 * (as queries and as neighbours); all other rows get keep = 0. */
int dfu_radius_filter(dfu_ctx* ctx, void* stream, const double* xyz, int nseg, const int64_t* seg_off_host,
                      const double* radius_host, const int32_t* nb_host, const int32_t* every_k_host,
                      uint8_t* out_keep, int32_t* out_count /*nullable*/);

/* Open3D remove_statistical_outlier semantics: mean distance to the k nearest (self included),
 * mu / Bessel sigma over the segment, keep iff 0 < mean < mu + ratio*sigma.
 * out_mean float64 [n] (nullable), out_stats_host [nseg][2] = mu, sigma (nullable; forces a sync). */
int dfu_sor_filter(dfu_ctx* ctx, void* stream, const double* xyz, int nseg, const int64_t* seg_off_host,
                   int k, double std_ratio, uint8_t* out_keep, double* out_mean, double* out_stats_host);

/* Diagnostics of the last dfu_sor_filter call that returned statistics (out_stats_host != NULL):
 * which = 0: queries that were handled one by one (k_sor_single) instead of by the tile kernel; 1: points of that call;
 * 2..4: of those queries, deferred for a far bound / a search box meeting too many nodes / ties at the k-th distance. */
int64_t dfu_sor_counter(const dfu_ctx* ctx, int which);

/* BallQuery gate: keep[i] = min_j |query_i - ref_j| < max_dist. */
int dfu_nn_gate(dfu_ctx* ctx, void* stream, const double* query_xyz, int64_t nq, const double* ref_xyz, int64_t nr,
                double max_dist, uint8_t* out_keep);

/* LShapeFitting.fitting (rectangle_fitting.py:39-53) on one point set: cluster id per point
 * (numbered by minimum point index) and one rectangle a[4] b[4] c[4] + theta + 2 costs per cluster
 * (rect_out_host [cap][16]). */
int dfu_lshape_fit_sync(dfu_ctx* ctx, void* stream, const double* xy /*[n][2]*/, int64_t n, double R0, double Rd,
                        const double* angles_cs_host, int n_angles, int32_t* out_cluster /*dev [n]*/,
                        double* rect_out_host, int64_t capacity, int64_t* n_clusters_host);

/* GenerateAnns (my_loader.py:633-684) on one object's fused points (float64 [n][3]): adaptive-range clustering,
 * rectangle search and the box arithmetic, one record per cluster in cluster order (fields 4..20 as in
 * dfu_fuse_instances_sync; field 20 = skipped by the Car aspect rule of :651).  view_host supplies the calibration
 * for the lidar_to_rect of :679.  car_aspect_rule: bit 0 = the rule of :651, bit 1 = anns_mode 1 (my_loader0.py:811-930). */
int dfu_generate_anns_sync(dfu_ctx* ctx, void* stream, const dfu_view* view_host, const double* points, int64_t n,
                           int car_aspect_rule, double R0, double Rd, const double* angles_cs_host, int n_angles,
                           double* boxes_out_host, int64_t capacity, int64_t* n_boxes_host);

/* The 2-D box tokens of the label lines: "xmin ymin xmax ymax" of every instance as str(np.float32(.)) prints them
 * (my_loader.py:680; box2d holds float32), NUL-terminated, off_host[i] = start of instance i -- the b2txt / b2_off table of
 * dfu_format_label_lines.  Host routine.  Returns the bytes written, or -needed when cap is too small. */
int64_t dfu_box2d_tokens(const float* b2_host, int64_t n_inst, char* out_host, int64_t cap, int32_t* off_host);

/* ---- label text (host only; my_loader.py:680) --------------------------------------------------------------
 * dfu_repr_f64: Python's str() / repr() of float64 values (shortest round-trip digits, fixed notation for
 * 1e-4 <= |v| < 1e16), one per line.  dfu_format_label_lines: every KITTI label line of a batch of box records
 * ("cls 0 0 alpha xmin ymin xmax ymax h w l x y z ry"), names / b2txt are NUL-separated string tables addressed by
 * name_off [n_names] / b2_off [n_inst].  Both return the bytes written, or -(bytes needed) if cap is too small. */
int64_t dfu_repr_f64(const double* vals_host, int64_t n, char* out_host, int64_t cap);
int64_t dfu_format_label_lines(const double* boxes_host, int64_t n, int V, const char* names, const int32_t* name_off,
                               int n_names, const char* b2txt, const int32_t* b2_off, int64_t n_inst, char* out_host,
                               int64_t cap, int64_t* lines_per_view_host);

/* ---- GT-database tooling (the consumer of the pseudo-boxes) ---------------------------------------------
 * points-in-rotated-boxes with the arithmetic of the reference's CPU routine
 * (pcdet/ops/roiaware_pool3d/src/roiaware_pool3d.cpp:121-168, called by
 * pcdet/datasets/kitti/kitti_dataset.py:310-312): points float [n][C] (xyz first, C >= 3), boxes float [M][7]
 * = x y z dx dy dz heading (centre).  out_mask int32 [M][n], 1 = inside. */
int dfu_points_in_boxes(dfu_ctx* ctx, void* stream, const float* points, int C, int64_t n, const float* boxes, int M,
                        int32_t* out_mask);

/* The per-box crop of create_groundtruth_database (kitti_dataset.py:314-319: rows inside box i in point order, xyz
 * shifted by the box centre) for all boxes at once, concatenated box by box = the integrated database layout of
 * tools/process_tools/create_integrated_database.py:26-45 (offsets_host [M+1] are its global_data_offset pairs).
 * out_rows float [capacity][C]; out_rows == NULL: size query (offsets only).  centres64 (nullable, device double
 * [M][3]): the annotation's float64 box centres — numpy's in-place `float32 -= float64` of :318 subtracts in double
 * and rounds once; without it the float32 centres of `boxes` are subtracted in float32. */
int dfu_crop_boxes_sync(dfu_ctx* ctx, void* stream, const float* points, int C, int64_t n, const float* boxes, int M,
                        const double* centres64, float* out_rows, int64_t capacity, int64_t* offsets_host);

/* ---- consumer side of the 8-column cloud (detector input) -----------------------------------------------
 * Modality split of pcdet/datasets/dataset_virtual.py:270-296 (last column == 2: LiDAR rows, == 1: pseudo rows), each
 * in input order, optionally with mask_points_by_range (pcdet/utils/common_utils.py:78-81 as applied by
 * processor/data_processor_virtual.py:69-77: x, y inside [lo, hi], both ends inclusive).  cloud float [n][C];
 * range_host float[6] = x y z lo, x y z hi, or NULL; out_* float [n][C] (capacity n rows each); counts_host[2]. */
int dfu_split_cloud_sync(dfu_ctx* ctx, void* stream, const float* cloud, int C, int64_t n, const float* range_host,
                         float* out_lidar, float* out_pseudo, int64_t* counts_host);

/* Allocation-free variant of dfu_split_cloud_sync.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * allocation).  Same rows, same order and same counts as dfu_split_cloud_sync, but: its scratch is a fixed 256 KB
 * `static __device__` array inside the library (held by every GPU that loads the library until exit), so
 * n <= DFU_SPLIT_STATIC_MAX_ROWS (more: DFU_ERR_CAPACITY, nothing launched); it calls no cudaMalloc / cudaFree / allocating
 * library routine, uses no context scratch, does not synchronise and copies nothing to the host: counts_dev is a
 * caller-owned DEVICE int64[2] (LiDAR rows, pseudo rows) valid once `stream` reaches this point.  ctx may be NULL (error
 * text is then not recorded and the calling thread's current device is used).  All calls on a device share the array:
 * calls that overlap on the GPU (different streams without an event between them) race.  The first call on a device
 * resolves the array's device address (cudaGetSymbolAddress): make it before starting a graph capture. */
#define DFU_SPLIT_STATIC_MAX_ROWS 4194304
int dfu_split_cloud_static(dfu_ctx* ctx, void* stream, const float* cloud, int C, int64_t n, const float* range_host,
                           float* out_lidar, float* out_pseudo, int64_t* counts_dev);

/* Hard voxelisation of transform_points_to_voxels (processor/data_processor_virtual.py:133-185; VoxelGeneratorWrapper
 * :16-58 -> spconv points_to_voxel): float32 bins floor((p - lo) / vsize), voxels numbered by first appearance, the
 * first max_points points of a voxel in input order, voxels beyond max_voxels dropped.  voxels float
 * [max_voxels][max_points][C] (zero padded), coords int32 [max_voxels][3] = z y x, num_points int32 [max_voxels]. */
int dfu_voxelize_sync(dfu_ctx* ctx, void* stream, const float* points, int C, int64_t n, const float* voxel_size_host,
                      const float* range_host, int max_points, int max_voxels, float* voxels, int32_t* coords,
                      int32_t* num_points, int64_t* nvox_host);

/* ---- self-evolution glue (detector predictions -> refreshed pseudo-boxes) -------------------------------
 * Rotated bird's-eye-view box overlap of pcdet/ops/iou3d_nms (iou3d_nms_utils.py:31-42 boxes_iou_bev, :45-77
 * boxes_iou3d_gpu; src/iou3d_nms_kernel.cu:128-234,279-293; float32 arithmetic in the order of the reference's CPU build
 * src/iou3d_cpu.cpp:59-230).  boxes_* float [n][stride], stride >= 7, a row = x y z dx dy dz heading (+ ignored columns);
 * out float [na][nb].  mode 0: BEV IoU, 1: BEV overlap area, 2: 3-D IoU; mode | 4: paired (na == nb, row i against row i
 * only, out float [na]: paired_boxes_iou3d_gpu / boxes_aligned_iou3d_gpu, iou3d_nms_utils.py:80-117,155-188). */
int dfu_boxes_iou_sync(dfu_ctx* ctx, void* stream, const float* boxes_a, int stride_a, int na, const float* boxes_b,
                       int stride_b, int nb, int mode, float* out);

/* class_agnostic_nms (pcdet/models/model_utils/model_nms_utils.py:6-27) with NMS_TYPE nms_gpu (iou3d_nms_utils.py:120-136,
 * src/iou3d_nms_kernel.cu:295-339, src/iou3d_nms.cpp:137-183): rows with score >= score_thresh (all rows when
 * use_score_thresh == 0), the pre_maxsize best by score (descending, ties by row index; < 0: no cut), greedy rotated-BEV
 * suppression at IoU > iou_thresh, the first post_maxsize survivors (< 0: all).  keep_out int64 [>= min(n, post_maxsize)]
 * (device) = indices into the ORIGINAL rows in score order; *n_keep_host = their number.  At most 65536 candidates. */
int dfu_nms_sync(dfu_ctx* ctx, void* stream, const float* boxes, int stride, const float* scores, int n,
                 int use_score_thresh, float score_thresh, float iou_thresh, int pre_maxsize, int post_maxsize,
                 int64_t* keep_out, int64_t* n_keep_host);

/* The same with NMS_TYPE nms_normal_gpu (iou3d_nms_utils.py:139-152, src/iou3d_nms_kernel.cu:341-400): suppression by the
 * axis-aligned overlap of the x / y extents, heading ignored. */
int dfu_nms_normal_sync(dfu_ctx* ctx, void* stream, const float* boxes, int stride, const float* scores, int n,
                        int use_score_thresh, float score_thresh, float iou_thresh, int pre_maxsize, int post_maxsize,
                        int64_t* keep_out, int64_t* n_keep_host);

#ifdef __cplusplus
}
#endif
#endif /* DFU3D_H */
