"""Timing of the config-3 filters (radius r = 0.2, SOR k = 20) on the 4.0 M / 8.64 M point clouds; per-kernel CUDA events."""
import sys, time, json, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import synth, ops, _lib
ctx = _lib.default_context(0)
out = {}
for name, backdrop in (('4.0M', None), ('8.64M', 90.0)):
    fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device='cuda', backdrop=backdrop)
    depth = torch.stack([v['depth'] for v in fr['views']])
    calibs = [(v['P2'], v['R0'], v['V2C']) for v in fr['views']]
    xyz, pix, counts = ops.backproject(depth, calibs)
    n = int(xyz.shape[0])

    def timeit(fn, it=3):
        fn(); torch.cuda.synchronize()
        ctx.profile(True); ctx.profile_read(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(it): r = fn()
        e1.record(); torch.cuda.synchronize()
        prof = ctx.profile_read(reset=True); ctx.profile(False)
        return e0.elapsed_time(e1) / it, r, {k: v[0] / it for k, v in prof.items()}
    res = {'points': n}
    for nb in (1, 16):
        ms, keep, prof = timeit(lambda: ops.radius_filter(xyz, nb, 0.2))
        res['radius_nb%d' % nb] = {'ms': ms, 'kept': float(keep.float().mean()), 'kernels_ms': prof}
    ms, keep, prof = timeit(lambda: ops.sor_filter(xyz, 20, 2.0))
    ops.sor_filter(xyz, 20, 2.0, return_stats=True)
    res['sor_k20'] = {'ms': ms, 'kept': float(keep.float().mean()), 'kernels_ms': prof,
                      'deferred': int(ctx.lib.dfu_sor_counter(ctx.h, 0)), 'points': int(ctx.lib.dfu_sor_counter(ctx.h, 1)),
                      'deferred_far_overflow_ties': [int(ctx.lib.dfu_sor_counter(ctx.h, c)) for c in (2, 3, 4)]}
    out[name] = res
    print(name, json.dumps(res))
    del xyz, depth, fr
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'gpurun_out', 'filt_time.json'), 'w'), indent=1)
