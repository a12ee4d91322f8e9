import sys, time, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import synth, ops, _lib
ctx = _lib.default_context(0)
fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device='cuda')
depth = torch.stack([v['depth'] for v in fr['views']])
calibs = [(v['P2'], v['R0'], v['V2C']) for v in fr['views']]
xyz, pix, counts = ops.backproject(depth, calibs)
print('points', xyz.shape, counts)
def timeit(fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): r = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, r
for nb in (1, 16):
    ms, keep = timeit(lambda: ops.radius_filter(xyz, nb, 0.2))
    print('radius r=0.2 nb=%d' % nb, ms, 'ms', 'kept', float(keep.float().mean()))
ctx.profile(True)
ms, keep = timeit(lambda: ops.sor_filter(xyz, 20, 2.0), n=2)
print('sor k=20', ms, 'ms', 'kept', float(keep.float().mean()), ctx.profile_read())
