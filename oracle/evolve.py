"""TEST INFRASTRUCTURE ONLY — oracle of the self-evolution glue: rotated BEV IoU and greedy NMS.

`reference_iou_bev(a, b)` calls the reference's own compiled C++ (oracle/_ref/iou3d_nms_cuda*, built by
oracle/build_ref.py:build_iou from /root/reference/pcdet/ops/iou3d_nms/src/iou3d_cpu.cpp).  `restated_iou_bev` is the
float32 restatement of that routine (iou3d_cpu.cpp:59-230) in scalar numpy — small cases only; it is pinned against the
compiled reference by tests/test_oracle_evolve.py and the committed fixture tests/golden/evolve_iou_nms.npz.
`restated_nms` restates class_agnostic_nms + nms_gpu (pcdet/models/model_utils/model_nms_utils.py:6-27,
pcdet/ops/iou3d_nms/iou3d_nms_utils.py:120-136, src/iou3d_nms.cpp:166-181) over an IoU matrix: the reference's own NMS
entry point is CUDA-only and cannot run where this oracle is built.  Never imported by dfu3d_b200."""
import importlib.util
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None
F = np.float32


def _load():
    global _mod
    if _mod is not None:
        return _mod
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    d = os.path.join(HERE, '_ref')
    cands = [f for f in (os.listdir(d) if os.path.isdir(d) else []) if f.startswith('iou3d_nms_cuda') and f.endswith('.so')]
    if not cands:
        return None
    spec = importlib.util.spec_from_file_location('iou3d_nms_cuda', os.path.join(d, cands[0]))
    flags = sys.getdlopenflags()
    sys.setdlopenflags(os.RTLD_LAZY | os.RTLD_LOCAL)      # the CUDA entry points of the pybind table stay unresolved
    try:
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
    finally:
        sys.setdlopenflags(flags)
    _mod = m
    return m


def available():
    try:
        return _load() is not None
    except Exception:
        return False


def reference_iou_bev(boxes_a, boxes_b):
    """iou3d_nms_utils.py:12-28 (boxes_bev_iou_cpu) around the reference's compiled boxes_iou_bev_cpu."""
    import torch
    m = _load()
    a = torch.from_numpy(np.ascontiguousarray(boxes_a, np.float32))
    b = torch.from_numpy(np.ascontiguousarray(boxes_b, np.float32))
    out = a.new_zeros((a.shape[0], b.shape[0]))
    m.boxes_iou_bev_cpu(a.contiguous(), b.contiguous(), out)
    return out.numpy()


# ---------------------------------------------------------------- float32 restatement (iou3d_cpu.cpp)
EPS = F(1e-8)          # :38
MARGIN = F(1e-2)       # :77


def _cross3(p1, p2, p0):                                   # :63-65
    return (p1[0] - p0[0]) * (p2[1] - p0[1]) - (p2[0] - p0[0]) * (p1[1] - p0[1])


def _corners(box):                                         # :131-166
    cx, cy, dx, dy, ang = box[0], box[1], box[3], box[4], box[6]
    hx, hy = dx / F(2), dy / F(2)
    x1, y1, x2, y2 = cx - hx, cy - hy, cx + hx, cy + hy
    c, s = F(math.cos(float(ang))), F(math.sin(float(ang)))     # cosf / sinf of the host libm: correctly rounded in practice
    out = []
    for px, py in ((x1, y1), (x2, y1), (x2, y2), (x1, y2)):     # rotate_around_center :118-122
        nx = (px - cx) * c + (py - cy) * (-s) + cx
        ny = (px - cx) * s + (py - cy) * c + cy
        out.append((nx, ny))
    return out, c, s


def _inside(box, c, s, p):                                 # check_in_box2d :75-85 (cos(-a) = c, sin(-a) = -s)
    ac, as_ = c, -s
    rx = (p[0] - box[0]) * ac + (p[1] - box[1]) * (-as_)
    ry = (p[0] - box[0]) * as_ + (p[1] - box[1]) * ac
    return abs(rx) < box[3] / F(2) + MARGIN and abs(ry) < box[4] / F(2) + MARGIN


def _intersection(p1, p0, q1, q0):                         # :87-116
    if not (min(p0[0], p1[0]) <= max(q0[0], q1[0]) and min(q0[0], q1[0]) <= max(p0[0], p1[0]) and
            min(p0[1], p1[1]) <= max(q0[1], q1[1]) and min(q0[1], q1[1]) <= max(p0[1], p1[1])):
        return None
    s1, s2, s3, s4 = _cross3(q0, p1, p0), _cross3(p1, q1, p0), _cross3(p0, q1, q0), _cross3(q1, p1, q0)
    if not (s1 * s2 > 0 and s3 * s4 > 0):
        return None
    s5 = _cross3(q1, p1, p0)
    if abs(s5 - s1) > EPS:
        return ((s5 * q0[0] - s1 * q1[0]) / (s5 - s1), (s5 * q0[1] - s1 * q1[1]) / (s5 - s1))
    a0, b0, c0 = p0[1] - p1[1], p1[0] - p0[0], p0[0] * p1[1] - p1[0] * p0[1]
    a1, b1, c1 = q0[1] - q1[1], q1[0] - q0[0], q0[0] * q1[1] - q1[0] * q0[1]
    D = a0 * b1 - a1 * b0
    return ((b0 * c1 - b1 * c0) / D, (a1 * c0 - a0 * c1) / D)


def _overlap(box_a, box_b):                                # box_overlap :128-220
    ca, cosa, sina = _corners(box_a)
    cb, cosb, sinb = _corners(box_b)
    pts = []
    sx, sy = F(0), F(0)
    for i in range(4):
        for j in range(4):
            r = _intersection(ca[(i + 1) % 4], ca[i], cb[(j + 1) % 4], cb[j])
            if r is not None:
                sx, sy = sx + r[0], sy + r[1]
                pts.append(r)
    for k in range(4):
        if _inside(box_a, cosa, sina, cb[k]):
            sx, sy = sx + cb[k][0], sy + cb[k][1]
            pts.append(cb[k])
        if _inside(box_b, cosb, sinb, ca[k]):
            sx, sy = sx + ca[k][0], sy + ca[k][1]
            pts.append(ca[k])
    n = len(pts)
    if n < 3:
        return F(0)
    mx, my = sx / F(n), sy / F(n)
    ang = [F(math.atan2(float(p[1] - my), float(p[0] - mx))) for p in pts]      # atan2f
    order = sorted(range(n), key=lambda k: ang[k])          # stable, as the reference's bubble sort with a strict test
    pts = [pts[k] for k in order]
    area = F(0)
    for k in range(n - 1):
        ax, ay = pts[k][0] - pts[0][0], pts[k][1] - pts[0][1]
        bx, by = pts[k + 1][0] - pts[0][0], pts[k + 1][1] - pts[0][1]
        area = area + (ax * by - ay * bx)
    return F(abs(area) / F(2))


def restated_overlap_bev(boxes_a, boxes_b):
    a, b = np.asarray(boxes_a, F), np.asarray(boxes_b, F)
    out = np.zeros((a.shape[0], b.shape[0]), F)
    with np.errstate(all='ignore'):
        for i in range(a.shape[0]):
            for j in range(b.shape[0]):
                out[i, j] = _overlap(a[i], b[j])
    return out


def restated_iou_bev(boxes_a, boxes_b):                    # iou_bev :222-230
    a, b = np.asarray(boxes_a, F), np.asarray(boxes_b, F)
    ov = restated_overlap_bev(a, b)
    sa, sb = (a[:, 3] * a[:, 4])[:, None], (b[:, 3] * b[:, 4])[None, :]
    return (ov / np.maximum((sa + sb) - ov, EPS)).astype(F)


def restated_iou3d(boxes_a, boxes_b, overlap_bev=None):    # iou3d_nms_utils.py:45-77 (float32 torch elementwise ops)
    a, b = np.asarray(boxes_a, F), np.asarray(boxes_b, F)
    ov = restated_overlap_bev(a, b) if overlap_bev is None else np.asarray(overlap_bev, F)
    amax, amin = (a[:, 2] + a[:, 5] / F(2))[:, None], (a[:, 2] - a[:, 5] / F(2))[:, None]
    bmax, bmin = (b[:, 2] + b[:, 5] / F(2))[None, :], (b[:, 2] - b[:, 5] / F(2))[None, :]
    oh = np.maximum(np.minimum(amax, bmax) - np.maximum(amin, bmin), F(0))
    o3 = ov * oh
    va, vb = (a[:, 3] * a[:, 4] * a[:, 5])[:, None], (b[:, 3] * b[:, 4] * b[:, 5])[None, :]
    return (o3 / np.maximum(va + vb - o3, F(1e-6))).astype(F)


def restated_iou_normal(boxes_a, boxes_b):
    """iou_normal of iou3d_nms_kernel.cu:341-353 (axis-aligned x / y extents, float32).  The reference has no CPU build of
    it (CUDA only), so this restatement is unpinned; the value does not depend on FMA contraction (halving is exact)."""
    a, b = np.asarray(boxes_a, F)[:, None, :], np.asarray(boxes_b, F)[None, :, :]
    left = np.maximum(a[..., 0] - a[..., 3] / F(2), b[..., 0] - b[..., 3] / F(2))
    right = np.minimum(a[..., 0] + a[..., 3] / F(2), b[..., 0] + b[..., 3] / F(2))
    top = np.maximum(a[..., 1] - a[..., 4] / F(2), b[..., 1] - b[..., 4] / F(2))
    bottom = np.minimum(a[..., 1] + a[..., 4] / F(2), b[..., 1] + b[..., 4] / F(2))
    inter = np.maximum(right - left, F(0)) * np.maximum(bottom - top, F(0))
    return (inter / np.maximum(a[..., 3] * a[..., 4] + b[..., 3] * b[..., 4] - inter, EPS)).astype(F)


def restated_nms(iou_fn, boxes, scores, iou_thresh, pre_maxsize=None, post_maxsize=None, score_thresh=None,
                 return_margin=False):
    """Indices into `boxes` of the survivors.  iou_fn(a, b) -> IoU matrix (reference_iou_bev or restated_iou_bev).
    Ties between equal scores are broken by the lower index (torch's topk / sort leave them unspecified).
    With return_margin also min |IoU - thresh| over the pairs the greedy pass consulted."""
    boxes, scores = np.asarray(boxes, F), np.asarray(scores, F).reshape(-1)
    idx = np.arange(boxes.shape[0])
    if score_thresh is not None:
        idx = idx[scores >= F(score_thresh)]
    else:
        idx = idx[~np.isnan(scores)]
    idx = idx[np.argsort(-scores[idx].astype(np.float64), kind='stable')]
    if pre_maxsize is not None and pre_maxsize >= 0:
        idx = idx[:pre_maxsize]
    iou = iou_fn(boxes[idx, :7], boxes[idx, :7]) if len(idx) else np.zeros((0, 0), F)
    removed = np.zeros(len(idx), bool)
    keep, margin = [], np.inf
    for i in range(len(idx)):
        if removed[i]:
            continue
        keep.append(idx[i])
        if i + 1 < len(idx):
            margin = min(margin, float(np.min(np.abs(iou[i, i + 1:].astype(np.float64) - float(F(iou_thresh))))))
        removed[i + 1:] |= iou[i, i + 1:] > F(iou_thresh)
    keep = np.asarray(keep, np.int64)
    if post_maxsize is not None and post_maxsize >= 0:
        keep = keep[:post_maxsize]
    return (keep, margin) if return_margin else keep


def synth_boxes(n, seed=0, extent=20.0, clustered=True):
    """Detector-like predictions: cars / pedestrians / cyclists around a few centres so that many pairs overlap; includes
    exact duplicates, axis-aligned and 90-degree-rotated copies (collinear edges: the degenerate branches)."""
    rng = np.random.default_rng(seed)
    sizes = np.array([[3.9, 1.6, 1.56], [0.8, 0.6, 1.73], [1.76, 0.6, 1.73]], F)
    k = rng.integers(0, 3, n)
    if clustered:
        centres = rng.uniform(-extent, extent, (max(n // 6, 1), 2))
        xy = centres[rng.integers(0, len(centres), n)] + rng.normal(0, 0.6, (n, 2))
    else:
        xy = rng.uniform(-extent, extent, (n, 2))
    b = np.zeros((n, 7), F)
    b[:, :2] = xy
    b[:, 2] = rng.normal(-1.0, 0.3, n)
    b[:, 3:6] = sizes[k] * rng.uniform(0.85, 1.15, (n, 3))
    b[:, 6] = rng.uniform(-np.pi, np.pi, n)
    if n >= 8:
        b[1] = b[0]                                   # exact duplicate
        b[3] = b[2]; b[3, 6] = b[2, 6] + F(np.pi / 2)  # same centre, rotated a quarter turn
        b[4, 6] = 0.0; b[5] = b[4]; b[5, 0] += F(0.5)   # axis-aligned, shifted: collinear edges
        b[6, 3:5] = 0.0                               # degenerate (zero area)
    scores = rng.uniform(0.05, 0.99, n).astype(F)
    if n >= 8:
        scores[7] = scores[0]                         # tie
    return b, scores
