import sys, os, torch, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from dfu3d_b200 import synth, pipeline, _lib
dev = torch.device('cuda', 0)
pipe = pipeline.FusionPipeline(0)
frames = [synth.make_frame(100 + i % 4, 'nuscenes', n_cams=6, sweeps=10, device=dev, max_inst=8) for i in range(16)]
vb = pipeline.batch_from_frames(frames, dev)
for _ in range(2): pipe.run(vb); vb.labelmap = None
torch.cuda.synchronize()
# a big H2D on a side stream while the pipeline runs on the main stream, with host time stamps after every stage
x = torch.empty(1 << 30, dtype=torch.uint8).pin_memory(); d = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
side = torch.cuda.Stream()
lib, ctx = pipe.lib, pipe.ctx
orig = {}
stamps = []
def wrap(name):
    f = getattr(lib, name)
    def g(*a):
        r = f(*a)
        stamps.append((name, time.time()))
        return r
    return g
class L:  # proxy
    def __getattr__(self, n):
        if n.startswith('dfu_'):
            return wrap(n)
        return getattr(lib, n)
pipe.lib = L()
for mode in ('alone', 'with big H2D in flight'):
    torch.cuda.synchronize()
    stamps.clear()
    if mode != 'alone':
        with torch.cuda.stream(side):
            for i in range(2): d.copy_(x, non_blocking=True)      # ~39 ms
    t0 = time.time()
    vb.labelmap = None
    pipe.run(vb)
    t1 = time.time()
    torch.cuda.synchronize()
    print(mode, 'run() host time %.1f ms' % ((t1 - t0) * 1e3))
    prev = t0
    for n, t in stamps:
        if (t - prev) * 1e3 > 0.3: print('    %-28s +%.2f ms' % (n, (t - prev) * 1e3))
        prev = t
