"""TEST INFRASTRUCTURE ONLY — points-in-boxes oracle for the GT-database row.

`reference(points, boxes)` calls the reference's own compiled C++ (oracle/_ref, built by oracle/build_ref.py from
/root/reference/pcdet/ops/roiaware_pool3d/src/roiaware_pool3d.cpp); `restated(points, boxes)` is the numpy restatement
that works everywhere (roiaware_pool3d.cpp:121-140: float32 rotation by -heading, |z - cz| <= dz/2, |local| < d/2 + 1e-2
with the comparisons in double as the C++ promotes them).  Never imported by dfu3d_b200."""
import importlib.util
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_mod = None


def _load():
    global _mod
    if _mod is not None:
        return _mod
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    d = os.path.join(HERE, '_ref')
    cands = [f for f in (os.listdir(d) if os.path.isdir(d) else []) if f.startswith('roiaware_pool3d_cuda') and f.endswith('.so')]
    if not cands:
        return None
    spec = importlib.util.spec_from_file_location('roiaware_pool3d_cuda', os.path.join(d, cands[0]))
    import sys
    flags = sys.getdlopenflags()
    sys.setdlopenflags(os.RTLD_LAZY | os.RTLD_LOCAL)      # the CUDA launchers the file declares stay unresolved (never called)
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


def reference(points, boxes):
    """pcdet/ops/roiaware_pool3d/roiaware_pool3d_utils.py:9-25 around the reference's compiled points_in_boxes_cpu."""
    import torch
    m = _load()
    pts = torch.from_numpy(np.ascontiguousarray(points[:, :3], np.float32))
    bx = torch.from_numpy(np.ascontiguousarray(boxes, np.float32))
    out = pts.new_zeros((bx.shape[0], pts.shape[0]), dtype=torch.int)
    m.points_in_boxes_cpu(bx.contiguous(), pts.contiguous(), out)
    return out.numpy()


def restated(points, boxes, return_margin=False):
    """int32 [M, N]; with return_margin also the distance (metres) of every (box, point) pair to the nearest decision
    boundary, so that tests can set pairs within 1e-5 m aside (cosf / sinf differ in the last ulp between libm and CUDA)."""
    p = np.asarray(points, np.float32)[:, :3]
    b = np.asarray(boxes, np.float32)
    cosa = np.cos(-b[:, 6]).astype(np.float32)[:, None]
    sina = np.sin(-b[:, 6]).astype(np.float32)[:, None]
    sx = p[None, :, 0] - b[:, None, 0]
    sy = p[None, :, 1] - b[:, None, 1]
    lx = (sx * cosa).astype(np.float32) + (sy * (-sina)).astype(np.float32)
    ly = (sx * sina).astype(np.float32) + (sy * cosa).astype(np.float32)
    dzv = np.abs(p[None, :, 2] - b[:, None, 2]).astype(np.float64) - b[:, None, 5].astype(np.float64) / 2.0
    mx = np.abs(lx).astype(np.float64) - (b[:, None, 3].astype(np.float64) / 2.0 + np.float64(np.float32(1e-2)))
    my = np.abs(ly).astype(np.float64) - (b[:, None, 4].astype(np.float64) / 2.0 + np.float64(np.float32(1e-2)))
    inside = (~(dzv > 0)) & (mx < 0) & (my < 0)
    if return_margin:
        return inside.astype(np.int32), np.minimum(np.minimum(np.abs(dzv), np.abs(mx)), np.abs(my))
    return inside.astype(np.int32)
