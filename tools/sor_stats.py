import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import synth, ops, _lib
ctx = _lib.default_context(0)
fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device='cuda')
depth = torch.stack([v['depth'] for v in fr['views']])
calibs = [(v['P2'], v['R0'], v['V2C']) for v in fr['views']]
xyz, pix, counts = ops.backproject(depth, calibs)
ops.sor_filter(xyz, 20, 2.0, return_stats=True)
torch.cuda.synchronize()
names = ['tiles', 'groups_swept', 'groups_with_a_hit', 'tightenings_in_sweep', 'final_entries', 'final_lanes', 'lanes_hit_per_group_sum', 'appends']
n = xyz.shape[0]; tiles = n / 32
for i, nm in enumerate(names):
    v = int(ctx.lib.dfu_sor_counter(ctx.h, 10 + i))
    print(nm, v, 'per tile %.2f' % (v / tiles))
print('deferred', [int(ctx.lib.dfu_sor_counter(ctx.h, c)) for c in (0, 2, 3, 4)])
