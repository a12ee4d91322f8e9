"""Wall-clock of every C-ABI call of one resident pipeline step against the GPU time of its kernels: shows where the
GPU idles between stages (host syncs, Python, launch latency).   python tools/host_gaps.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import pipeline, synth   # noqa: E402


def main():
    pipe = pipeline.FusionPipeline(0)
    dev = pipe.device
    frames = [synth.make_frame(3000 + i, 'nuscenes', n_cams=6, sweeps=10, device=dev, max_inst=8) for i in range(8)]
    vb = pipeline.batch_from_frames([frames[i % 8] for i in range(64)], dev)
    del frames
    lib = pipe.lib
    log = []
    names = ['dfu_pack_masks', 'dfu_fuse_begin', 'dfu_fuse_project_sync', 'dfu_fuse_densify_sync', 'dfu_fuse_emit_cloud',
             'dfu_fuse_instances_sync']

    class Wrap:
        def __init__(self, lib):
            self._lib = lib

        def __getattr__(self, n):
            f = getattr(self._lib, n)
            if n not in names:
                return f

            def g(*args):
                t = time.perf_counter()
                r = f(*args)
                log.append((n, time.perf_counter() - t))
                return r
            return g
    pipe.lib = Wrap(lib)
    for it in range(4):
        vb.labelmap = None
        log.clear()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pipe.run(vb, cloud_dtype=torch.float16, want_lines=False)
        torch.cuda.synchronize()
        tot = time.perf_counter() - t0
    print('step wall %.3f ms' % (tot * 1e3))
    acc = 0.0
    for n, t in log:
        print('  %-26s %.3f ms' % (n, t * 1e3))
        acc += t
    print('  python between calls       %.3f ms' % ((tot - acc) * 1e3))


if __name__ == '__main__':
    main()
