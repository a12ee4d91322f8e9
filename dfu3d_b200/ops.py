"""Stand-alone operators of libdfu3d.so on caller-provided point sets (include/dfu3d.h, last section).

Thin argument marshalling only: numpy or torch in, the same kind out.  Segments (``seg_off``) are independent
point sets processed in one launch sequence.  No CPU fallback: every function needs the CUDA library.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


def _dev_f64(ctx, a, cols):
    """(device float64 [n, cols] contiguous, was_numpy)"""
    if isinstance(a, torch.Tensor):
        t = a.to(device=ctx.device, dtype=torch.float64).reshape(-1, cols).contiguous()
        return t, False
    t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float64).reshape(-1, cols))).to(ctx.device)
    return t, True


def _seg(n, seg_off):
    if seg_off is None:
        return np.array([0, n], np.int64)
    so = np.ascontiguousarray(np.asarray(seg_off, np.int64))
    assert so[0] == 0 and so[-1] == n and np.all(np.diff(so) >= 0), 'seg_off must ascend from 0 to n'
    return so


def _out(t, as_numpy, dtype=None):
    if as_numpy:
        torch.cuda.current_stream(t.device).synchronize()
        a = t.cpu().numpy()
        return a.astype(dtype) if dtype is not None else a
    return t.to(dtype=torch.bool) if dtype is bool else t


def radius_filter(points, nb_points, radius, seg_off=None, ctx=None, return_counts=False, every_k=None):
    """Open3D ``remove_radius_outlier(nb_points, radius)`` keep mask (call sites
    tools/PENet/dataloaders/my_loader.py:569-599): keep iff #{q : |p - q|^2 < r^2} > nb_points, self included.
    ``nb_points`` / ``radius`` / ``every_k`` are scalars or one value per segment.  ``every_k``:
    ``uniform_down_sample(every_k)`` first (my_loader.py:571): rows 0, k, 2k, ... of the segment take part, the others
    get keep = False."""
    ctx = ctx or _lib.default_context(0)
    xyz, as_np = _dev_f64(ctx, points, 3)
    n = xyz.shape[0]
    so = _seg(n, seg_off)
    nseg = so.shape[0] - 1
    rad = np.ascontiguousarray(np.broadcast_to(np.asarray(radius, np.float64), (nseg,)))
    nbs = np.ascontiguousarray(np.broadcast_to(np.asarray(nb_points, np.int32), (nseg,)))
    evk = None if every_k is None else np.ascontiguousarray(np.broadcast_to(np.asarray(every_k, np.int32), (nseg,)))
    keep = torch.zeros(n, dtype=torch.uint8, device=ctx.device)
    cnt = torch.zeros(nseg, dtype=torch.int32, device=ctx.device)
    ctx.check(ctx.lib.dfu_radius_filter(ctx.h, ctx.stream, _lib.dptr(xyz) if n else None, nseg, _lib.hptr(so),
                                        _lib.hptr(rad), _lib.hptr(nbs), _lib.hptr(evk),
                                        _lib.dptr(keep) if n else _lib.dptr(cnt), _lib.dptr(cnt)))
    k = _out(keep, as_np, bool)
    return (k, _out(cnt, as_np)) if return_counts else k


def sor_filter(points, nb_neighbors, std_ratio, seg_off=None, ctx=None, return_stats=False):
    """Open3D ``remove_statistical_outlier(nb_neighbors, std_ratio)`` keep mask (commented-out call
    tools/PENet/dataloaders/my_loader0.py:735; BASELINE.json config 3).  With ``return_stats``: (keep, mean k-NN
    distance per point, [nseg, 2] mu / sigma)."""
    ctx = ctx or _lib.default_context(0)
    xyz, as_np = _dev_f64(ctx, points, 3)
    n = xyz.shape[0]
    so = _seg(n, seg_off)
    nseg = so.shape[0] - 1
    keep = torch.zeros(max(n, 1), dtype=torch.uint8, device=ctx.device)
    mean = torch.zeros(max(n, 1), dtype=torch.float64, device=ctx.device) if return_stats else None
    stats = np.zeros((nseg, 2), np.float64) if return_stats else None
    ctx.check(ctx.lib.dfu_sor_filter(ctx.h, ctx.stream, _lib.dptr(xyz) if n else None, nseg, _lib.hptr(so),
                                     int(nb_neighbors), C.c_double(float(std_ratio)), _lib.dptr(keep), _lib.dptr(mean),
                                     _lib.hptr(stats)))
    k = _out(keep[:n], as_np, bool)
    if return_stats:
        return k, _out(mean[:n], as_np), stats
    return k


def nn_gate(query, ref, max_dist, ctx=None):
    """BallQuery (tools/PENet/dataloaders/my_loader.py:489-494 with quirk Q4): keep[i] = min_j |q_i - r_j| < max_dist."""
    ctx = ctx or _lib.default_context(0)
    q, as_np = _dev_f64(ctx, query, 3)
    r, _ = _dev_f64(ctx, ref, 3)
    keep = torch.zeros(max(q.shape[0], 1), dtype=torch.uint8, device=ctx.device)
    ctx.check(ctx.lib.dfu_nn_gate(ctx.h, ctx.stream, _lib.dptr(q) if q.shape[0] else None, q.shape[0],
                                  _lib.dptr(r) if r.shape[0] else None, r.shape[0], C.c_double(float(max_dist)), _lib.dptr(keep)))
    return _out(keep[:q.shape[0]], as_np, bool)


_ANGLES = None


def search_angles_cs(dtheta_deg=1.0):
    """cos / sin of the search angles of rectangle_fitting.py:117-122 (89 for the reference's 1 degree step; host numpy,
    as the reference evaluates them)."""
    global _ANGLES
    if dtheta_deg != 1.0:
        dtheta = np.deg2rad(dtheta_deg)
        th = np.arange(0.0, np.pi / 2.0 - dtheta, dtheta)
        return np.ascontiguousarray(np.stack([np.cos(th), np.sin(th)], 1)), th
    if _ANGLES is None:
        dtheta = np.deg2rad(1.0)
        th = np.arange(0.0, np.pi / 2.0 - dtheta, dtheta)
        _ANGLES = (np.ascontiguousarray(np.stack([np.cos(th), np.sin(th)], 1)), th)
    return _ANGLES


def lshape_fit(ox, oy, R0=3.0, Rd=0.001, ctx=None, dtheta_deg=1.0):
    """LShapeFitting.fitting arithmetic (rectangle_fitting.py:39-191): (cluster id per point numbered by minimum point
    index [n] int32 numpy, rect [K, 16] float64 numpy: a[4] b[4] c[4] theta best-cost runner-up n)."""
    ctx = ctx or _lib.default_context(0)
    xy = np.ascontiguousarray(np.stack([np.asarray(ox, np.float64).reshape(-1), np.asarray(oy, np.float64).reshape(-1)], 1))
    n = xy.shape[0]
    if n == 0:
        return np.zeros(0, np.int32), np.zeros((0, 16))
    ang, _ = search_angles_cs(dtheta_deg)
    if ang.shape[0] < 1 or ang.shape[0] > 96:
        raise ValueError('dtheta_deg_for_serarch = %r gives %d search angles (supported: 1..96)' % (dtheta_deg, ang.shape[0]))
    xyd = torch.from_numpy(xy).to(ctx.device)
    cl = torch.zeros(n, dtype=torch.int32, device=ctx.device)
    rect = np.zeros((n, 16))
    nk = np.zeros(1, np.int64)
    ctx.check(ctx.lib.dfu_lshape_fit_sync(ctx.h, ctx.stream, _lib.dptr(xyd), n, C.c_double(R0), C.c_double(Rd),
                                          _lib.hptr(ang), ang.shape[0], _lib.dptr(cl), _lib.hptr(rect), n, _lib.hptr(nk)))
    torch.cuda.current_stream(ctx.device).synchronize()
    return cl.cpu().numpy(), rect[:int(nk[0])].copy()


def backproject(depth, calibs, min_depth=0.001, max_z=None, ctx=None):
    """Depth maps -> LiDAR-frame points (my_loader.py:507-509,535-536 [+ the height gate of :540 when ``max_z`` is
    given]; calibration_kitti.py:134-144,89-102, float64).  ``depth`` [V,H,W] or [H,W] float32 (NOT mutated — quirk
    Q10 is the drop-in wrapper's business); ``calibs``: one (P2, R0, V2C) per view.
    Returns (xyz [M,3] float64, pix [M] int32 linear pixel index, counts [V] int64); rows are ordered by (view, pixel)."""
    from . import pipeline
    ctx = ctx or _lib.default_context(0)
    as_np = not isinstance(depth, torch.Tensor)
    d = torch.as_tensor(np.asarray(depth, np.float32) if as_np else depth)
    d = d.to(device=ctx.device, dtype=torch.float32)
    if d.dim() == 2:
        d = d[None]
    d = d.contiguous()
    V, H, W = d.shape
    if len(calibs) == 3 and not isinstance(calibs[0], (tuple, list)):
        calibs = [calibs]      # a single (P2, R0, V2C)
    assert len(calibs) == V, 'one calibration per view'
    views = np.zeros(V, _lib.VIEW_DTYPE)
    for v in range(V):
        pipeline.fill_view(views[v], pipeline.calib_constants(*calibs[v]))
    prm = np.zeros(1, _lib.FUSION_DTYPE)
    prm['H'], prm['W'], prm['Hm'], prm['Wm'] = H, W, H, W
    prm['min_depth'] = min_depth
    prm['max_z'] = 0.0 if max_z is None else max_z
    counts = np.zeros(V, np.int64)
    gate = int(max_z is not None)
    ctx.check(ctx.lib.dfu_backproject_sync(ctx.h, ctx.stream, V, _lib.hptr(views), _lib.hptr(prm), _lib.dptr(d), gate,
                                           None, None, 0, _lib.hptr(counts)))
    M = int(counts.sum())
    xyz = torch.empty((max(M, 1), 3), dtype=torch.float64, device=ctx.device)
    pix = torch.empty(max(M, 1), dtype=torch.int32, device=ctx.device)
    ctx.check(ctx.lib.dfu_backproject_sync(ctx.h, ctx.stream, V, _lib.hptr(views), _lib.hptr(prm), _lib.dptr(d), gate,
                                           _lib.dptr(xyz), _lib.dptr(pix), xyz.shape[0], _lib.hptr(counts)))
    return _out(xyz[:M], as_np), _out(pix[:M], as_np), counts


def _dev_f32(ctx, a, cols=None):
    if isinstance(a, torch.Tensor):
        t, was_np = a.to(device=ctx.device, dtype=torch.float32), False
    else:
        t, was_np = torch.from_numpy(np.ascontiguousarray(np.asarray(a, np.float32))).to(ctx.device), True
    if cols is not None:
        t = t.reshape(-1, cols)
    return t.contiguous(), was_np


def points_in_boxes(points, boxes, ctx=None):
    """roiaware_pool3d_utils.points_in_boxes_cpu (pcdet/ops/roiaware_pool3d/roiaware_pool3d_utils.py:9-25): points
    [N, C >= 3] float32, boxes [M, 7] = x y z dx dy dz heading -> int32 [M, N]."""
    ctx = ctx or _lib.default_context(0)
    b, _ = _dev_f32(ctx, boxes, 7)
    p, as_np = _dev_f32(ctx, points)
    assert p.dim() == 2 and p.shape[1] >= 3
    out = torch.empty((b.shape[0], p.shape[0]), dtype=torch.int32, device=ctx.device)   # every element is written
    ctx.check(ctx.lib.dfu_points_in_boxes(ctx.h, ctx.stream, _lib.dptr(p) if p.numel() else None, p.shape[1], p.shape[0],
                                          _lib.dptr(b) if b.numel() else None, b.shape[0], _lib.dptr(out) if out.numel() else None))
    return _out(out, as_np)


def crop_boxes(points, boxes, ctx=None):
    """create_groundtruth_database's per-box crop (pcdet/datasets/kitti/kitti_dataset.py:310-319) for all boxes at once:
    returns (rows [total, C] float32: the points inside box 0, then box 1, ... in point order, xyz minus the box centre;
    offsets [M+1] int64) — the `global_data_offset` layout of tools/process_tools/create_integrated_database.py.
    float64 boxes (what the annotation pickles hold) keep their float64 centres for the subtraction, as numpy does."""
    ctx = ctx or _lib.default_context(0)
    b, _ = _dev_f32(ctx, boxes, 7)
    p, as_np = _dev_f32(ctx, points)
    M, (n, Cc) = b.shape[0], p.shape
    off = np.zeros(M + 1, np.int64)
    c64 = None
    if (boxes.dtype == torch.float64) if isinstance(boxes, torch.Tensor) else (np.asarray(boxes).dtype == np.float64):
        c64 = (boxes if isinstance(boxes, torch.Tensor) else torch.from_numpy(np.asarray(boxes))).reshape(-1, 7)[:, :3]
        c64 = c64.to(ctx.device).contiguous()
    args = (ctx.h, ctx.stream, _lib.dptr(p) if p.numel() else None, Cc, n, _lib.dptr(b) if b.numel() else None, M,
            _lib.dptr(c64) if c64 is not None and c64.numel() else None)
    # one call: ground-truth boxes seldom overlap, so n rows nearly always suffice; the offsets are filled in even when the
    # capacity does not, and only then is the call repeated with the exact size (no separate size query)
    rows = torch.empty((max(n, 1), Cc), dtype=torch.float32, device=ctx.device)
    rc = ctx.lib.dfu_crop_boxes_sync(*args, _lib.dptr(rows), rows.shape[0], _lib.hptr(off))
    if rc == -3 and int(off[-1]) > rows.shape[0]:           # DFU_ERR_CAPACITY
        rows = torch.empty((int(off[-1]), Cc), dtype=torch.float32, device=ctx.device)
        rc = ctx.lib.dfu_crop_boxes_sync(*args, _lib.dptr(rows), rows.shape[0], _lib.hptr(off))
    ctx.check(rc)
    return _out(rows[:int(off[-1])], as_np), off


def boxes_iou(boxes_a, boxes_b, mode='bev', paired=False, ctx=None):
    """pcdet/ops/iou3d_nms/iou3d_nms_utils.py:31-42 (`boxes_iou_bev`, mode 'bev'), the `boxes_overlap_bev_gpu` area
    (mode 'overlap') and :45-77 (`boxes_iou3d_gpu`, mode '3d'): boxes [N, >= 7] / [M, >= 7] = x y z dx dy dz heading
    -> float32 [N, M]; paired=True (:80-117, :155-188): N == M, row i against row i only -> float32 [N]."""
    ctx = ctx or _lib.default_context(0)
    a, as_np = _dev_f32(ctx, boxes_a)
    b, _ = _dev_f32(ctx, boxes_b)
    assert a.dim() == 2 and b.dim() == 2 and a.shape[1] >= 7 and b.shape[1] >= 7
    assert not paired or a.shape[0] == b.shape[0]
    out = torch.empty((a.shape[0],) if paired else (a.shape[0], b.shape[0]), dtype=torch.float32, device=ctx.device)
    ctx.check(ctx.lib.dfu_boxes_iou_sync(ctx.h, ctx.stream, _lib.dptr(a) if a.numel() else None, a.shape[1], a.shape[0],
                                         _lib.dptr(b) if b.numel() else None, b.shape[1], b.shape[0],
                                         {'bev': 0, 'overlap': 1, '3d': 2}[mode] | (4 if paired else 0), _lib.dptr(out) if out.numel() else None))
    return _out(out, as_np)


def nms(boxes, scores, iou_thresh, pre_maxsize=None, post_maxsize=None, score_thresh=None, normal=False, ctx=None):
    """class_agnostic_nms with NMS_TYPE nms_gpu (pcdet/models/model_utils/model_nms_utils.py:6-27,
    pcdet/ops/iou3d_nms/iou3d_nms_utils.py:120-136): int64 indices into `boxes` of the survivors, best score first
    (equal scores: lower index first).  normal=True: NMS_TYPE nms_normal_gpu (:139-152), axis-aligned overlap."""
    ctx = ctx or _lib.default_context(0)
    b, as_np = _dev_f32(ctx, boxes)
    s, _ = _dev_f32(ctx, scores)
    s = s.reshape(-1)
    assert b.dim() == 2 and b.shape[1] >= 7 and s.shape[0] == b.shape[0]
    n = b.shape[0]
    keep = torch.empty((max(n, 1),), dtype=torch.int64, device=ctx.device)
    nk = np.zeros(1, np.int64)
    ctx.check((ctx.lib.dfu_nms_normal_sync if normal else ctx.lib.dfu_nms_sync)(ctx.h, ctx.stream, _lib.dptr(b) if n else None, b.shape[1], _lib.dptr(s) if n else None, n,
                                   int(score_thresh is not None), float(score_thresh or 0.0), float(iou_thresh),
                                   -1 if pre_maxsize is None else int(pre_maxsize),
                                   -1 if post_maxsize is None else int(post_maxsize), _lib.dptr(keep), _lib.hptr(nk)))
    return _out(keep[:int(nk[0])], as_np)


def split_cloud(cloud, point_cloud_range=None, ctx=None):
    """The modality split the detector's dataset does on the 8-column cloud (pcdet/datasets/dataset_virtual.py:270-296)
    with the x/y range mask of mask_points_and_boxes_outside_range: returns (lidar_rows, pseudo_rows), input order."""
    ctx = ctx or _lib.default_context(0)
    c, as_np = _dev_f32(ctx, cloud)
    n, Cc = c.shape
    o0 = torch.empty((max(n, 1), Cc), dtype=torch.float32, device=ctx.device)
    o1 = torch.empty((max(n, 1), Cc), dtype=torch.float32, device=ctx.device)
    cnt = np.zeros(2, np.int64)
    rg = None if point_cloud_range is None else np.ascontiguousarray(np.asarray(point_cloud_range, np.float32).reshape(6))
    ctx.check(ctx.lib.dfu_split_cloud_sync(ctx.h, ctx.stream, _lib.dptr(c) if n else None, Cc, n,
                                           _lib.hptr(rg) if rg is not None else None, _lib.dptr(o0), _lib.dptr(o1), _lib.hptr(cnt)))
    return _out(o0[:int(cnt[0])], as_np), _out(o1[:int(cnt[1])], as_np)


def voxelize(points, voxel_size, point_cloud_range, max_points_per_voxel, max_voxels, ctx=None):
    """VoxelGeneratorWrapper.generate (pcdet/datasets/processor/data_processor_virtual.py:16-58): returns
    (voxels [V, max_points, C] float32, coordinates [V, 3] int32 (z, y, x), num_points [V] int32)."""
    ctx = ctx or _lib.default_context(0)
    p, as_np = _dev_f32(ctx, points)
    n, Cc = p.shape
    vs = np.ascontiguousarray(np.asarray(voxel_size, np.float32).reshape(3))
    rg = np.ascontiguousarray(np.asarray(point_cloud_range, np.float32).reshape(6))
    vox = torch.empty((max_voxels, max_points_per_voxel, Cc), dtype=torch.float32, device=ctx.device)
    coords = torch.empty((max_voxels, 3), dtype=torch.int32, device=ctx.device)
    num = torch.empty((max_voxels,), dtype=torch.int32, device=ctx.device)
    nv = np.zeros(1, np.int64)
    ctx.check(ctx.lib.dfu_voxelize_sync(ctx.h, ctx.stream, _lib.dptr(p) if n else None, Cc, n, _lib.hptr(vs), _lib.hptr(rg),
                                        int(max_points_per_voxel), int(max_voxels), _lib.dptr(vox), _lib.dptr(coords),
                                        _lib.dptr(num), _lib.hptr(nv)))
    v = int(nv[0])
    return _out(vox[:v], as_np), _out(coords[:v], as_np), _out(num[:v], as_np)
