"""CPU: the numpy restatement of points_in_boxes_cpu against the reference's OWN C++ compiled in place
(oracle/build_ref.py -> oracle/_ref/; skipped where neither /root/reference nor the built module exists)."""
import numpy as np
import pytest

from oracle import build_ref, ref_points_in_boxes as rpb


def test_restatement_matches_compiled_reference():
    build_ref.build()
    if not rpb.available():
        pytest.skip('reference source not present and oracle/_ref not built')
    rng = np.random.default_rng(7)
    m, n = 60, 40000
    boxes = np.stack([rng.uniform(-30, 30, m), rng.uniform(-30, 30, m), rng.uniform(-1.5, 0.5, m), rng.uniform(0.5, 12, m),
                      rng.uniform(0.4, 3, m), rng.uniform(0.5, 3.5, m), rng.uniform(-7, 7, m)], 1).astype(np.float32)
    pts = rng.uniform(-35, 35, size=(n, 3)).astype(np.float32)
    pts[:, 2] = rng.uniform(-3, 3, size=n)
    bi = rng.integers(0, m, n // 2)
    pts[:n // 2] = boxes[bi, :3] + rng.uniform(-0.55, 0.55, size=(n // 2, 3)).astype(np.float32) * boxes[bi, 3:6]
    ref = rpb.reference(pts, boxes)
    got, margin = rpb.restated(pts, boxes, return_margin=True)
    near = margin < 1e-6
    assert ref.sum() > 1000 and np.array_equal(ref[~near], got[~near])
    assert int((ref != got).sum()) <= int(near.sum())


def golden():
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'gtdb_points_in_boxes.npz'))
    mask = np.unpackbits(g['mask'], axis=1)[:, :int(g['n'])].astype(np.int32)
    return g['points'], g['boxes'], mask


def test_restatement_matches_golden_from_reference():
    pts, boxes, mask = golden()
    got, margin = rpb.restated(pts, boxes, return_margin=True)
    near = margin < 1e-6
    assert mask.sum() > 1000 and np.array_equal(mask[~near], got[~near])
