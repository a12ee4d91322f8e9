#!/bin/bash
# SOR (k = 20) on the config-3 cloud once per library variant in gpurun_variants/ (and the in-tree build as "base")
for lib in base gpurun_variants/libdfu3d_*.so; do
  if [ "$lib" = base ]; then unset DFU3D_LIB; n=base; else export DFU3D_LIB=$PWD/$lib; n=$(basename $lib .so); fi
  python - <<P
import sys, os, torch, json
sys.path.insert(0, '.')
from dfu3d_b200 import synth, ops, _lib
ctx = _lib.default_context(0)
fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device='cuda')
depth = torch.stack([v['depth'] for v in fr['views']])
xyz, _, _ = ops.backproject(depth, [(v['P2'], v['R0'], v['V2C']) for v in fr['views']])
ops.sor_filter(xyz, 20, 2.0); torch.cuda.synchronize()
ctx.profile(True); ctx.profile_read(reset=True)
for _ in range(3): k = ops.sor_filter(xyz, 20, 2.0)
torch.cuda.synchronize()
pr = ctx.profile_read()
ops.sor_filter(xyz, 20, 2.0, return_stats=True)
print('$n', 'sor_knn ms', round(pr['sor_knn'][0] / 3, 3), 'kept', float(k.float().mean()), 'deferred', [int(ctx.lib.dfu_sor_counter(ctx.h, c)) for c in (0, 2, 3, 4)])
P
done
