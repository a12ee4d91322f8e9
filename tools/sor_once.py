"""One SOR call (k = 20, 2 sigma) on the config-3 cloud: the command the ncu captures of k_sor_* are taken from."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import synth, ops, _lib
backdrop = float(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1] != 'none' else None
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device='cuda', backdrop=backdrop)
depth = torch.stack([v['depth'] for v in fr['views']])
calibs = [(v['P2'], v['R0'], v['V2C']) for v in fr['views']]
xyz, pix, counts = ops.backproject(depth, calibs)
for _ in range(reps):
    keep = ops.sor_filter(xyz, 20, 2.0)
    k1 = ops.radius_filter(xyz, 1, 0.2)
torch.cuda.synchronize()
print('points', xyz.shape[0], 'kept', float(keep.float().mean()), float(k1.float().mean()))
