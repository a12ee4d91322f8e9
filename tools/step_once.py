"""One resident step of the bench workload (64 frames x 6 cams, 10 sweeps) between cudaProfilerStart/Stop: the command the
per-kernel ncu table of the pipeline is taken from.
    ncu --set full --clock-control none --profile-from-start off -k regex:"^k_" -o out python tools/step_once.py [frames]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import pipeline, synth      # noqa: E402

dev = torch.device('cuda', 0)
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 64
pipe = pipeline.FusionPipeline(0)
nd = int(sys.argv[2]) if len(sys.argv) > 2 else 8          # distinct frames, repeated (as bench.py --distinct)
fr = [synth.make_frame(1000 * 3 + i, 'nuscenes', n_cams=6, sweeps=10, device=dev, max_inst=8) for i in range(min(nd, nf))]
frames = [fr[i % len(fr)] for i in range(nf)]
vb = pipeline.batch_from_frames(frames, dev)
for _ in range(2):
    vb.labelmap = None
    pipe.run(vb, cloud_dtype=torch.float16)
torch.cuda.synchronize()
torch.cuda.profiler.start()
vb.labelmap = None
res = pipe.run(vb, cloud_dtype=torch.float16)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print('frames', nf, 'cloud rows', int(res.cloud.shape[0]), 'boxes', int(res.boxes.shape[0]))
