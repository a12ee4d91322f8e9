"""Times the consumer-side row (split + voxelise) on the GPU (resident, CUDA events) with the numpy oracle beside it.
    python tools/consumer_time.py > gpurun_out/consumer.json"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import ops   # noqa: E402
from oracle import consumer   # noqa: E402

RANGE = [0, -40, -3, 70.4, 40, 1]


def ev_time(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream())
    for _ in range(reps):
        fn()
    e1.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    out = []
    for n in (300000, 1000000, 4000000):
        rng = np.random.default_rng(n)
        c = np.empty((n, 8), np.float32)
        c[:, 0] = rng.uniform(-2, 75, n); c[:, 1] = rng.uniform(-42, 42, n); c[:, 2] = rng.uniform(-3.2, 1.2, n)
        c[:, 3:7] = rng.uniform(0, 85, (n, 4)); c[:, 7] = rng.choice([1.0, 2.0], size=n, p=[0.75, 0.25])
        d = torch.from_numpy(c).cuda()
        ms = ev_time(lambda: ops.split_cloud(d, RANGE))
        out.append({'op': 'split_cloud + range mask', 'rows': n, 'ms': round(ms, 4), 'GBps_read+write': round(n * 32 * 1.9 / ms / 1e6, 1)})
        lid, pse = ops.split_cloud(d, RANGE)
        for name, p, mv in (('voxelize pseudo rows', pse, 150000), ('voxelize lidar rows', lid, 40000)):
            ms = ev_time(lambda: ops.voxelize(p, [0.05, 0.05, 0.1], RANGE, 5, mv))
            v = ops.voxelize(p, [0.05, 0.05, 0.1], RANGE, 5, mv)[0].shape[0]
            out.append({'op': name, 'points': int(p.shape[0]), 'voxels': v, 'max_voxels': mv, 'ms': round(ms, 4),
                        'Mpoints_per_s': round(p.shape[0] / ms / 1e3, 1)})
        if n <= 1000000:
            t = time.perf_counter(); wl, wp = consumer.split_cloud(c, RANGE); t1 = time.perf_counter() - t
            t = time.perf_counter(); consumer.voxelize(wp, [0.05, 0.05, 0.1], RANGE, 5, 150000); t2 = time.perf_counter() - t
            out.append({'op': 'numpy oracle (1 core)', 'rows': n, 'split_ms': round(t1 * 1e3, 2), 'voxelize_pseudo_ms': round(t2 * 1e3, 2)})
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
