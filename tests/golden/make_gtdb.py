"""Writes tests/golden/gtdb_points_in_boxes.npz with the reference's OWN compiled points_in_boxes_cpu
(oracle/build_ref.py; needs /root/reference) — small enough to commit, so the check travels to the GPU box.
    python tests/golden/make_gtdb.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import build_ref, ref_points_in_boxes as rpb   # noqa: E402

build_ref.build()
rng = np.random.default_rng(20240)
m, n = 24, 6000
boxes = np.stack([rng.uniform(-30, 30, m), rng.uniform(-30, 30, m), rng.uniform(-1.5, 0.5, m), rng.uniform(0.5, 12, m),
                  rng.uniform(0.4, 3, m), rng.uniform(0.5, 3.5, m), rng.uniform(-7, 7, m)], 1).astype(np.float32)
pts = rng.uniform(-35, 35, size=(n, 4)).astype(np.float32)
pts[:, 2] = rng.uniform(-3, 3, size=n)
k = 2 * n // 3
bi = rng.integers(0, m, k)
loc = rng.uniform(-0.55, 0.55, size=(k, 3)).astype(np.float32) * boxes[bi, 3:6]
c, s = np.cos(boxes[bi, 6]), np.sin(boxes[bi, 6])
pts[:k, 0] = boxes[bi, 0] + loc[:, 0] * c - loc[:, 1] * s
pts[:k, 1] = boxes[bi, 1] + loc[:, 0] * s + loc[:, 1] * c
pts[:k, 2] = boxes[bi, 2] + loc[:, 2]
mask = rpb.reference(pts, boxes)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'gtdb_points_in_boxes.npz')
np.savez_compressed(out, points=pts, boxes=boxes, mask=np.packbits(mask.astype(np.uint8), axis=1), n=n)
print(out, mask.shape, int(mask.sum()))
