import sys, os, torch, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from dfu3d_b200 import synth, pipeline
dev = torch.device('cuda', 0)
pipe = pipeline.FusionPipeline(0)
frames = [synth.make_frame(100 + i % 4, 'nuscenes', n_cams=6, sweeps=10, device=dev, max_inst=8) for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 64)]
vb = pipeline.batch_from_frames(frames, dev)
hb = pipeline.HostBatch.from_device(vb)
for _ in range(2): pipe.run_host(hb, prefetch=None)
torch.cuda.synchronize()
st = pipe.__dict__['_stage']
st['trace'] = []
ref = torch.cuda.Event(enable_timing=True); ref.record(); 
t0 = time.time()
N = 4
for i in range(N): pipe.run_host(hb, prefetch=hb if i + 1 < N else None)
torch.cuda.synchronize()
print('wall per call ms', (time.time() - t0) / N * 1e3)
for rec in st['trace']:
    if rec[0] == 'compute':
        print('   compute', rec[1], 'start %.1f end %.1f  host returned at %.1f ms' % (ref.elapsed_time(rec[2]), ref.elapsed_time(rec[3]), (rec[4] - t0) * 1e3))
        continue
    g, a, b, nb = rec
    print('group', g, 'start %.1f end %.1f dur %.1f ms  %.1f GB/s' % (ref.elapsed_time(a), ref.elapsed_time(b), a.elapsed_time(b), nb / a.elapsed_time(b) / 1e6))
