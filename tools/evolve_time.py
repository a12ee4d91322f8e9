"""Timing of the self-evolution glue (row f4) on one GPU: rotated BEV IoU matrix and class-agnostic rotated NMS, next to
the reference's compiled CPU IoU (oracle/_ref, one core) on a bounded sample.   python tools/evolve_time.py > out.json"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import ops          # noqa: E402
from oracle import evolve as ev     # noqa: E402  (bench-side CPU baseline only)


def gpu_ms(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


out = {}
b, s = ev.synth_boxes(20000, seed=21, extent=60.0)
tb, ts = torch.from_numpy(b).cuda(), torch.from_numpy(s).cuda()
for pre in (1000, 4096):
    keep = ops.nms(tb, ts, 0.1, pre_maxsize=pre, post_maxsize=500, score_thresh=0.1)
    out['nms_pre%d' % pre] = {'predictions': 20000, 'candidates': pre, 'kept': int(keep.numel()),
                              'ms': gpu_ms(lambda: ops.nms(tb, ts, 0.1, pre_maxsize=pre, post_maxsize=500, score_thresh=0.1))}
per_frame, _ = ev.synth_boxes(500, seed=3)
tp = torch.from_numpy(per_frame).cuda()
sp = torch.rand(500, device='cuda')
out['nms_frame500'] = {'predictions': 500, 'ms': gpu_ms(lambda: ops.nms(tp, sp, 0.6, 1000, 83, 0.1))}
q = tb[:4096].contiguous()
ms = gpu_ms(lambda: ops.boxes_iou(q, q), reps=5)
out['iou_4096x4096'] = {'ms': ms, 'pairs_per_s': 4096 * 4096 / ms * 1e3, 'out_gbs': 4096 * 4096 * 4 / ms / 1e6}
if ev.available():
    t = time.time()
    ev.reference_iou_bev(b[:1024], b[:1024])
    dt = time.time() - t
    out['cpu_reference_iou_1024x1024'] = {'ms': dt * 1e3, 'pairs_per_s': 1024 * 1024 / dt, 'cores': 1,
                                          'kind': 'reference (iou3d_cpu.cpp compiled in place)'}
print(json.dumps(out))
