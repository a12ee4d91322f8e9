"""GPU parity of the GT-database row: points-in-rotated-boxes and the per-box crop, against the
numpy restatement of the reference's C++ (oracle/ref_points_in_boxes.py; itself checked against the reference's own
compiled source in tests/test_oracle_points_in_boxes.py)."""
import numpy as np
import pytest

from dfu3d_b200 import ops
from oracle import ref_points_in_boxes as rpb

pytestmark = pytest.mark.gpu


def scene(seed, n, m, C=4):
    rng = np.random.default_rng(seed)
    boxes = np.stack([rng.uniform(-30, 30, m), rng.uniform(-30, 30, m), rng.uniform(-1.5, 0.5, m), rng.uniform(0.5, 12, m),
                      rng.uniform(0.4, 3, m), rng.uniform(0.5, 3.5, m), rng.uniform(-7, 7, m)], 1).astype(np.float32)
    pts = rng.uniform(-35, 35, size=(n, C)).astype(np.float32)
    pts[:, 2] = rng.uniform(-3, 3, size=n)
    k = n // 2                                 # half of the points sit on / near the boxes, some exactly on faces
    bi = rng.integers(0, m, k)
    loc = rng.uniform(-0.55, 0.55, size=(k, 3)).astype(np.float32) * boxes[bi, 3:6]
    c, s = np.cos(boxes[bi, 6]), np.sin(boxes[bi, 6])
    pts[:k, 0] = boxes[bi, 0] + loc[:, 0] * c - loc[:, 1] * s
    pts[:k, 1] = boxes[bi, 1] + loc[:, 0] * s + loc[:, 1] * c
    pts[:k, 2] = boxes[bi, 2] + loc[:, 2]
    return pts, boxes


@pytest.mark.parametrize('seed,n,m', [(0, 50000, 40), (1, 1000, 1), (2, 257, 300), (3, 1, 5)])
def test_points_in_boxes_matches_oracle(seed, n, m):
    pts, boxes = scene(seed, n, m)
    want, margin = rpb.restated(pts, boxes, return_margin=True)
    got = ops.points_in_boxes(pts, boxes)
    assert got.dtype == np.int32 and got.shape == (m, n)
    near = margin < 1e-5                       # cosf / sinf of CUDA and libm differ in the last ulp: reported separately
    assert np.array_equal(got[~near], want[~near])
    print('pairs', m * n, 'inside', int(want.sum()), 'within 1e-5 m of a face (excluded)', int(near.sum()),
          'of which differ', int((got != want)[near].sum()))
    assert want.sum() > 0 or n == 1


def test_crop_is_the_integrated_database_layout():
    pts, boxes = scene(5, 80000, 25, C=8)
    mask = ops.points_in_boxes(pts, boxes)
    rows, off = ops.crop_boxes(pts, boxes)
    assert off[0] == 0 and np.array_equal(np.diff(off), mask.sum(1)) and rows.shape == (off[-1], 8)
    for i in range(boxes.shape[0]):            # kitti_dataset.py:314-319
        gt = pts[mask[i] > 0].copy()
        gt[:, :3] -= boxes[i, :3]
        assert np.array_equal(rows[off[i]:off[i + 1]], gt)
    b64 = boxes.astype(np.float64) + 1e-9       # float64 annotation boxes: centre subtraction in double, one rounding
    rows64, off64 = ops.crop_boxes(pts, b64)
    mask64 = ops.points_in_boxes(pts, b64)
    assert np.array_equal(np.diff(off64), mask64.sum(1))
    for i in (0, 7, 24):
        gt = pts[mask64[i] > 0].copy()
        gt[:, :3] -= b64[i, :3]
        assert np.array_equal(rows64[off64[i]:off64[i + 1]], gt)
    # more rows than points (every box taken three times: the first call's capacity does not suffice and it is repeated)
    inside = pts[mask.any(0)]
    rows3, off3 = ops.crop_boxes(inside, np.concatenate([boxes] * 3, 0))
    assert off3[-1] > inside.shape[0] and rows3.shape[0] == off3[-1]
    m3 = ops.points_in_boxes(inside, boxes)
    assert np.array_equal(np.diff(off3), np.tile(m3.sum(1), 3))
    k = boxes.shape[0] + 3
    gt = inside[m3[3] > 0].copy()
    gt[:, :3] -= boxes[3, :3]
    assert np.array_equal(rows3[off3[k]:off3[k + 1]], gt)
    # empty inputs
    r0, o0 = ops.crop_boxes(pts[:0], boxes)
    assert r0.shape == (0, 8) and not o0.any()
    assert ops.points_in_boxes(pts, boxes[:0]).shape == (0, pts.shape[0])


def test_points_in_boxes_matches_reference_golden():
    """tests/golden/gtdb_points_in_boxes.npz was produced by the reference's own compiled C++ (tests/golden/make_gtdb.py)."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'gtdb_points_in_boxes.npz'))
    pts, boxes = g['points'], g['boxes']
    mask = np.unpackbits(g['mask'], axis=1)[:, :int(g['n'])].astype(np.int32)
    _, margin = rpb.restated(pts, boxes, return_margin=True)
    got = ops.points_in_boxes(pts, boxes)
    near = margin < 1e-5
    assert np.array_equal(got[~near], mask[~near]) and int((got != mask).sum()) <= int(near.sum())
