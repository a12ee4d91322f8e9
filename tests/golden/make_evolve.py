"""Writes tests/golden/evolve_iou_nms.npz with the reference's OWN compiled boxes_iou_bev_cpu
(oracle/build_ref.py:build_iou; needs /root/reference): 160 detector-like boxes, their 160 x 160 rotated BEV IoU matrix,
and the survivors of the greedy suppression (model_nms_utils.py:6-27 semantics restated over that matrix — the
reference's own nms_gpu is CUDA-only) for the DFU3D configs' settings (centerpoint_nuscenes2kitti.yaml:83-93).
    python tests/golden/make_evolve.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import build_ref, evolve as ev   # noqa: E402

build_ref.build_iou()
boxes, scores = ev.synth_boxes(160, seed=5)
iou = ev.reference_iou_bev(boxes, boxes)
cases, keeps, margins = [], [], []
for thr, pre, post, sthr in ((0.6, 1000, 83, 0.1), (0.1, 1000, 83, 0.1), (0.01, 100, 20, 0.3), (0.2, -1, -1, -1.0)):
    k, mg = ev.restated_nms(ev.reference_iou_bev, boxes, scores, thr, pre, post, None if sthr < 0 else sthr, return_margin=True)
    cases.append((thr, pre, post, sthr)); keeps.append(k); margins.append(mg)
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'evolve_iou_nms.npz')
np.savez_compressed(out, boxes=boxes, scores=scores, iou=iou, cases=np.asarray(cases, np.float64),
                    margins=np.asarray(margins), **{'keep%d' % i: k for i, k in enumerate(keeps)})
print(out, iou.shape, float((iou > 0).mean()), [len(k) for k in keeps], margins)
