"""Times the GT-database row on the GPU (resident, CUDA events) and the reference's compiled C++ beside it (one core,
as the reference runs it) when oracle/_ref is present.   python tools/gtdb_time.py > gpurun_out/gtdb.json"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import _lib, ops   # noqa: E402


def scene(n, m, C=4, seed=0):
    rng = np.random.default_rng(seed)
    boxes = np.stack([rng.uniform(-50, 50, m), rng.uniform(-50, 50, m), rng.uniform(-1.5, 0.5, m), rng.uniform(0.5, 12, m),
                      rng.uniform(0.4, 3, m), rng.uniform(0.5, 3.5, m), rng.uniform(-7, 7, m)], 1).astype(np.float32)
    pts = rng.uniform(-55, 55, size=(n, C)).astype(np.float32)
    pts[:, 2] = rng.uniform(-3, 3, size=n)
    return pts, boxes


def main():
    ctx = _lib.default_context(0)
    out = []
    for n, m in ((120000, 12), (250000, 40), (4000000, 64)):
        pts, boxes = scene(n, m)
        p = torch.from_numpy(pts).cuda()
        b = torch.from_numpy(boxes).cuda()
        for fn, name in ((lambda: ops.points_in_boxes(p, b), 'points_in_boxes'), (lambda: ops.crop_boxes(p, b), 'crop_boxes')):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(torch.cuda.current_stream())
            reps = 10
            for _ in range(reps):
                fn()
            e1.record(torch.cuda.current_stream())
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            out.append({'op': name, 'points': n, 'boxes': m, 'ms': round(ms, 4), 'pairs_per_s': round(n * m / ms * 1e3, 1),
                        'mask_GBps': round(4.0 * n * m / ms / 1e6, 1) if name == 'points_in_boxes' else None})
        try:
            from oracle import ref_points_in_boxes as rpb
            if rpb.available() and n <= 250000:
                t = time.perf_counter()
                rpb.reference(pts[:, :3].copy(), boxes)
                out.append({'op': 'reference points_in_boxes_cpu (compiled, 1 core)', 'points': n, 'boxes': m,
                            'ms': round((time.perf_counter() - t) * 1e3, 3)})
        except Exception as e:
            out.append({'reference': 'unavailable: %s' % e})
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
