"""BASELINE.json configs[1]: KITTI-shaped frame (64 beams x 1875 azimuths, 375x1242 image), S1 only (mask packing +
projection + label inheritance through dfu_pack_masks / dfu_project_label): single-frame latency and batch-64
throughput, CUDA events, inputs resident.  python tools/s1_kitti_time.py"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import synth, pipeline, _lib   # noqa: E402


def run(nframes, iters=20):
    pipe = pipeline.FusionPipeline(0, config=pipeline.FusionConfig.preset('kitti'))
    ctx = pipe.ctx
    distinct = [synth.make_frame(2000 + i, 'kitti', n_cams=1, sweeps=1, device='cuda') for i in range(min(nframes, 4))]
    frames = [distinct[i % len(distinct)] for i in range(nframes)]
    vb = pipeline.batch_from_frames(frames, pipe.device)
    views = pipe.build_views(vb)
    V = len(vb.view_sweep)
    vd = torch.from_numpy(views.view(np.uint8).reshape(V, -1)).to(pipe.device)
    H, W = vb.depth.shape[1:3]
    Hm, Wm = vb.image.shape[1:3]
    prm = pipeline.fusion_params(pipe.cfg, H, W, Hm, Wm)
    slots = int(views['lidar_cnt'].sum())
    pix = torch.empty((slots, 2), dtype=torch.int32, device='cuda')
    flags = torch.empty(slots, dtype=torch.uint8, device='cuda')
    label = torch.empty(slots, dtype=torch.int32, device='cuda')
    uv = torch.empty((slots, 2), dtype=torch.float32, device='cuda')
    dep = torch.empty(slots, dtype=torch.float32, device='cuda')

    def step():
        lm = pipe.pack_masks(vb, views)
        ctx.check(ctx.lib.dfu_project_label(ctx.h, ctx.stream, V, _lib.dptr(vd), int(views['lidar_cnt'].max()), _lib.hptr(prm),
                                            _lib.dptr(vb.lidar), _lib.dptr(lm), _lib.dptr(uv), _lib.dptr(dep), _lib.dptr(pix),
                                            _lib.dptr(flags), _lib.dptr(label)))
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    nb = 45 * slots + int(vb.masks.numel()) + 4 * V * Hm * Wm
    return {'frames': nframes, 'lidar_rows': slots, 'image': [int(Hm), int(Wm)], 'instances': int(sum(vb.n_inst)), 'ms': ms,
            'frames_per_s': nframes / ms * 1e3, 'algorithmic_GBps': nb / 1e9 / (ms / 1e3)}


if __name__ == '__main__':
    print(json.dumps({'workload': 'configs[1]: KITTI-shaped, S1 only', 'single_frame': run(1), 'batch_64': run(64)}))
