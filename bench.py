"""bench.py — pseudo-box frames/s of the DFU3D fusion / noise-suppression / pseudo-box path on B200.

One "step" = one pass of the full path (mask packing, S1 projection + label inheritance, S2
densification + angular dedup, per-instance radius filters + BallQuery gate + clustering + box
fitting, fp16 cloud emit) over one batch of synthetic nuScenes-shaped frames
(BASELINE.json configs[3]: 10-sweep frames, 6 cameras of 1600x900, batch 64, one B200).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--frames F] [--impl reference]

N > 1: launched by torchrun, one rank per GPU, frames sharded frame-wise (weak scaling: every rank
processes its own F frames), no data-path collective; the timed region is bracketed by a barrier and
torch.cuda.synchronize() and the max over ranks is reported.

The line carries: value (inputs resident in HBM), e2e (host pinned buffers in, cloud + boxes back to
the host, through dfu3d_b200.pipeline.FusionPipeline — the public API), roofline (dominant kernel,
CUDA-event time inside the timed region, algorithmic bytes of SURVEY.md §8(d); roofline.filters = the
worst filter kernel of the configs[2] leg), cpu_baseline (the oracle port on the host cores, bounded
sample; rank 0, N=1 only), clocks, gpu_launches.

What is inside every timed step: the host-side marshalling of the inputs (pipeline.forget_marshalling
drops every memoised record first: the synthetic batch is the same object each step, a sweep never
is), all kernels, the label text of every box and, at N > 1, the gather of the box lists to rank 0's
host (on a helper thread that overlaps the next step and is joined before the region ends).  The e2e
legs are a sweep: each call names the next batch (run_host(prefetch=...)), whose first frame group is
uploaded during the call before; exactly `steps` batches cross PCIe inside the region, and
e2e.single_call is the same call without a successor.

--impl reference: the reference's CPU implementation of the path.  The reference is pure Python over
numpy / Open3D / spconv / sklearn (no native code) and /root/reference does not exist on the GPU box,
so this arm times the oracle port (oracle/restate.py: the reference's first-party arithmetic restated
in numpy + the restated third-party filters) on all host cores, one whole view per worker process.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = 'pseudo-box frames/s (nuScenes-shaped)'
UNIT = 'frames/s'


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--frames', type=int, default=int(os.environ.get('DFU_BENCH_FRAMES', 64)), help='frames per step per GPU')
    ap.add_argument('--cams', type=int, default=6)
    ap.add_argument('--sweeps', type=int, default=10)
    ap.add_argument('--distinct', type=int, default=int(os.environ.get('DFU_BENCH_DISTINCT', 8)),
                    help='distinct synthetic frames generated per rank (tiled to --frames)')
    ap.add_argument('--max-inst', type=int, default=8)
    ap.add_argument('--parts', type=int, default=2, help='frame groups of a step that run concurrently (1 = plain run)')
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--cpu-sample-views', type=int, default=0, help='views of the CPU sample (0 = 24 for the baseline leg, 1 per worker for --impl reference)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-filters', action='store_true')
    return ap.parse_args()


def workload_config(a):
    return {'workload': 'configs[3]: full pipeline (S1 projection+label inheritance, S2 densification+angular dedup, '
                        'radius filters, BallQuery gate, clustering, box fitting, fp16 cloud emit)',
            'frames_per_step_per_gpu': a.frames, 'cams_per_frame': a.cams, 'sweeps': a.sweeps,
            'lidar_rays_per_frame': 34720 * a.sweeps, 'lidar_note': 'rays without a return (sky) and 2 % dropouts are not in the sweep: ~246 k rows per frame', 'image': '1600x900', 'max_instances_per_view': a.max_inst,
            'distinct_frames': min(a.distinct, a.frames), 'l2': 'inputs per step >> 126 MB L2 (no flush needed)',
            'parallelism': 'frame-wise sharding, 1 process per GPU, no collective'}


# ----------------------------------------------------------------------------- clocks
class ClockSampler(object):
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            f = [x.strip() for x in r.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


# ----------------------------------------------------------------------------- CPU side (oracle port)
def _cpu_view_worker(job):
    seed, cams, sweeps, cam, max_inst = job
    import torch
    torch.set_num_threads(1)
    from dfu3d_b200 import synth
    from oracle import restate
    fr = synth.make_frame(seed, 'nuscenes', n_cams=cams, sweeps=sweeps, max_inst=max_inst, only_cam=cam)
    inp = synth.view_as_reference_inputs(fr, 0)
    t0 = time.perf_counter()
    out = restate.run_view(inp)
    return time.perf_counter() - t0, len(out[2])


def cpu_baseline(a, max_views, budget_s=15.0):
    """Times the oracle port on whole views on ONE host core until `budget_s` seconds of CPU work are spent (at least
    one view, at most `max_views`): a bounded sample, so that the default run ends within minutes.  A full-resolution
    view costs 0.5 - 30 s depending on how many points its instances hold (the reference's clustering is quadratic).
    Returns (frames/s, seconds, sample text)."""
    wall, n = 0.0, 0
    while n < max_views and (n == 0 or wall < budget_s):
        job = (910000 + n // a.cams, a.cams, a.sweeps, n % a.cams, a.max_inst)
        wall += _cpu_view_worker(job)[0]
        n += 1
    fps = (n / float(a.cams)) / wall
    return fps, wall, ('%d view(s) of %d-cam %d-sweep nuScenes-shaped frames on 1 core (sample bounded at ~%.0f s of CPU '
                       'work), compute only: inputs generated outside the timed part' % (n, a.cams, a.sweeps, budget_s))


def run_reference(a):
    """--impl reference: the reference path's CPU implementation (oracle port, see the module docstring) on all host
    cores, one whole view per worker process; a step = one bounded sample of `procs * views_per_proc` views."""
    rank = int(os.environ.get('RANK', 0))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 64))
    per_step = procs * max(1, a.cpu_sample_views)
    times = []
    with mp.get_context('spawn').Pool(procs) as pool:
        def step(k):
            jobs = [(910000 + (k * per_step + i) // a.cams, a.cams, a.sweeps, (k * per_step + i) % a.cams, a.max_inst)
                    for i in range(per_step)]
            res = pool.map(_cpu_view_worker, jobs, chunksize=1)
            # step time = MEAN worker: summed compute seconds / workers (a lower bound of the wall time of the parallel
            # step, i.e. it favours the CPU arm)
            return sum(r[0] for r in res) / float(procs)
        for w in range(1):                  # ONE warm-up step whatever --warmup says: a CPU step costs ~20 s
            step(1000 + w)                  # imports + first-touch in every worker
        for k in range(a.steps):
            times.append(step(k))
    tot = sum(times)
    value = (per_step * a.steps / float(a.cams)) / tot
    sample = ('%d views per step (%d worker processes x %d), %d-cam %d-sweep nuScenes-shaped frames, one whole view per '
              'worker, inputs generated outside the timed part; step time = summed compute seconds / workers (the mean worker, not the busiest)' % (per_step, procs, max(1, a.cpu_sample_views), a.cams, a.sweeps))
    line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup,
            'ms_per_step': 1000.0 * tot / a.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic', 'impl': 'reference', 'config': workload_config(a),
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': procs, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------- algorithmic bytes
def algorithmic_bytes(tag, st):
    """Compulsory bytes of one launch of kernel `tag` on the benchmark batch (DESIGN.md 'Kernels')."""
    if tag == 'pack_masks':
        return st['mask_bytes'] + 4 * st['pixels_m']
    if tag == 'project':
        return 45 * st['lidar_slots']
    if tag == 'densify_a':                       # S2: 4 B depth per pixel + key write 4 B; per valid point table atomics not counted
        return 8 * st['pixels'] + 0 * st['gated']
    if tag == 'densify_b':
        return 8 * st['pixels']
    if tag == 'densify_c':
        return 4 * st['pixels'] + 8 * st['voxels']
    if tag == 'emit_cloud':
        return st['cloud_rows'] * (16 + 16)
    if tag == 'vox_gather':
        return st['voxels'] * (8 + 4 + 24 + 4 + 4)
    if tag == 'radius':
        return 13 * st.get('inst_points', 0)
    if tag == 'grid_build':
        return 28 * st.get('inst_points', 0)
    if tag == 'cluster':
        return 20 * st.get('fused_points', 0)
    if tag == 'fit':
        return 20 * st.get('fused_points', 0) + 120 * st.get('boxes', 0)
    if tag == 'gate':
        return 13 * st.get('inst_points', 0)
    return None


def filters_leg(pipe, dev, peak, iters=3):
    """BASELINE.json configs[2]: one nuScenes-shaped frame, 6 x 1600x900 full-resolution densification (height gate
    off) through the radius filter (r = 0.2 m, nb_points = 1) and SOR (k = 20, 2 sigma); per-kernel CUDA-event times
    (dfu_profile_*) against the algorithmic bytes of SURVEY.md 8(d): grid build 28 B/pt, radius 13 B/pt, SOR 29 B/pt."""
    import torch
    from dfu3d_b200 import synth, ops
    # sky-free views (rays that hit nothing land on a 90 m backdrop): every one of the 6 x 1600 x 900 pixels back-projects,
    # which is the ~8.6 M points BASELINE.json quotes; the same cloud is compared with the oracle in
    # tests/test_gpu_fullsize_oracle.py::test_config3_radius_and_sor_full_size_vs_oracle
    fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device=dev, backdrop=90.0)
    depth = torch.stack([v['depth'] for v in fr['views']])
    calibs = [(v['P2'], v['R0'], v['V2C']) for v in fr['views']]
    xyz, _, counts = ops.backproject(depth, calibs, ctx=pipe.ctx)
    n = int(xyz.shape[0])
    out = {'workload': 'configs[2]: 6 x 1600x900 depth maps -> %d back-projected points; radius (r=0.2, nb=1); SOR (k=20, 2.0)' % n,
           'points': n}
    for name, fn, tags, bpp in (('radius', lambda: ops.radius_filter(xyz, 1, 0.2, ctx=pipe.ctx), ('grid_build', 'radius'), (28, 13)),
                                ('sor', lambda: ops.sor_filter(xyz, 20, 2.0, ctx=pipe.ctx), ('grid_build', 'sor_knn', 'sor_mask'), (28, 21, 8))):
        keep = fn()
        torch.cuda.synchronize()
        pipe.ctx.profile(True)
        pipe.ctx.profile_read(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            keep = fn()
        e1.record()
        torch.cuda.synchronize()
        prof = pipe.ctx.profile_read(reset=True)
        pipe.ctx.profile(False)
        leg = {'ms': e0.elapsed_time(e1) / iters, 'kept_fraction': float(keep.float().mean()), 'kernels': {}}
        for tag, b in zip(tags, bpp):
            if tag in prof:
                ms = prof[tag][0] / iters
                gbs = n * b / 1e9 / (ms / 1e3)
                leg['kernels'][tag] = {'ms': ms, 'algorithmic_bytes': n * b, 'achieved_gbs': gbs, 'frac_of_peak': gbs / peak}
        out[name] = leg
    return out


def bind_to_gpu_numa(index):
    """Pin this rank's host threads (and hence its first-touch pinned allocations) to the NUMA node of its GPU, so that
    the per-step host->device traffic does not cross the socket interconnect when 8 ranks copy at once.  Best effort."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(index)
        bdf = '%04x:%02x:%02x.0' % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open('/sys/bus/pci/devices/%s/numa_node' % bdf).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open('/sys/devices/system/node/node%d/cpulist' % node).read().strip().split(','):
            lo, _, hi = part.partition('-')
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        return None
    return None


def main():
    a = parse()
    if a.impl == 'reference':
        return run_reference(a)
    import torch
    import torch.distributed as dist
    from dfu3d_b200 import synth, pipeline
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    json_fd = None
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else None
    # Host cores per rank (reported): with 8 ranks on a 32-core host the host side of a step stretches (two group threads,
    # the gather thread and NCCL's threads per rank on 4 cores).  Measured there: 23.5 ms per step as is, 25.0 ms with the
    # library sleeping in its stream syncs (DFU3D_BLOCKING_SYNC=1), 28.2 ms with one frame group in flight (--parts 1).
    try:
        cores_per_rank = len(os.sched_getaffinity(0)) / float(world)
    except Exception:
        cores_per_rank = (os.cpu_count() or 1) / float(world)

    dev = torch.device('cuda', local)
    if world > 1:
        # an externally set NCCL_DEBUG is honoured (the driver counts ranks from the INFO lines); NCCL writes to fd 1,
        # which is pointed at stderr while the communicator is created so that stdout stays the ONE JSON line
        if 'DFU_NCCL_DEBUG' in os.environ:
            os.environ['NCCL_DEBUG'] = os.environ['DFU_NCCL_DEBUG']
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        # NCCL logs to fd 1 (banner at communicator creation, INFO lines at every collective and at destroy when
        # NCCL_DEBUG is set): fd 1 is pointed at stderr for the whole run and the ONE JSON line goes to the saved descriptor
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group('nccl', device_id=dev)
        dist.barrier()
        torch.cuda.synchronize()
    pipe = pipeline.FusionPipeline(local)

    # ---- synthetic batch, generated on the device (identical generator for the CPU arm)
    nd = max(1, min(a.distinct, a.frames))
    frames = [synth.make_frame(1000 * 3 + rank * 100 + i, 'nuscenes', n_cams=a.cams, sweeps=a.sweeps, device=dev,
                               max_inst=a.max_inst) for i in range(nd)]
    frames = [frames[i % nd] for i in range(a.frames)]
    vb = pipeline.batch_from_frames(frames, dev)
    del frames
    host = pipeline.HostBatch.from_device(vb)          # pinned host copies of every input tensor

    from dfu3d_b200 import shard
    # global frame id / camera ordinal of every local view: rank r owns frames {i : i mod world == r}
    frame_of_view = [(rank + world * (v // a.cams), v % a.cams) for v in range(a.frames * a.cams)]
    gathered = {'boxes': 0}

    def finish(res):
        # the step delivers what depth2pointsrgbpm delivers: the label text of every box (formatted on the host) and, at
        # N > 1, the host-side gather of the pseudo-box lists (north star; dfu3d_b200/shard.py) -- both inside the timed region
        # The gather of step n runs on a helper thread while step n + 1 computes (it mostly WAITS for the slowest rank); it is
        # joined before the next gather starts and before the timed region ends (drain), so every step's gather is inside it.
        if world > 1 and not os.environ.get('DFU_DIAG_NO_GATHER'):       # diagnosis only: a line without the gather is not a bench line
            drain()
            boxes = shard.globalise(res.boxes, frame_of_view)

            def job():
                torch.cuda.set_device(local)
                t0 = time.perf_counter()
                allb = shard.gather_boxes(boxes, device=dev, dst=0)     # the lists of all ranks, ordered, on rank 0's host
                gathered['boxes'] = int(allb.shape[0])
                gathered['seconds'] = gathered.get('seconds', 0.0) + time.perf_counter() - t0
                gathered['calls'] = gathered.get('calls', 0) + 1
            gathered['fut'] = gather_pool.submit(job)
        return res

    import concurrent.futures
    gather_pool = concurrent.futures.ThreadPoolExecutor(max_workers=1)

    def drain():
        f = gathered.pop('fut', None)
        if f is not None:
            f.result()

    parts = a.parts if a.parts > 0 else 2

    def step_resident():
        pipeline.forget_marshalling(vb)        # nothing derived from the inputs on the host survives from step to step
        # two frame groups in flight (own stream / context / host thread each): the host-side sizing between the stages of
        # one group overlaps the kernels of the other
        vb.labelmap = None
        return finish(pipe.run_split(vb, cloud_dtype=torch.float16, instances=True, want_lines=True, parts=parts))

    def step_profiled():
        pipeline.forget_marshalling(vb)
        vb.labelmap = None
        return finish(pipe.run(vb, cloud_dtype=torch.float16, instances=True, want_lines=True))

    # The end-to-end legs are a SWEEP, the way the path is used (one call per batch of a dataset): every call names the
    # batch of the next call, whose first frame group then crosses PCIe while this call's last group is computed, read
    # back and formatted.  The last call of a timed run names nothing, so exactly `steps` batches are uploaded inside
    # the timed region (and none before it: the warm-up's last call names nothing either).
    sweep = {'left': 0}

    def step_e2e():
        pipeline.forget_marshalling(host)
        sweep['left'] -= 1
        return finish(pipe.run_host(host, cloud_dtype=torch.float16, want_lines=True, prefetch=host if sweep['left'] > 0 else None))

    def step_e2e_single():
        pipeline.forget_marshalling(host)
        return finish(pipe.run_host(host, cloud_dtype=torch.float16, want_lines=True))

    host_c = {}

    def step_e2e_compact():
        pipeline.forget_marshalling(host_c['hb'])
        sweep['left'] -= 1
        return finish(pipe.run_host(host_c['hb'], cloud_dtype=torch.float16, want_lines=True,
                                    prefetch=host_c['hb'] if sweep['left'] > 0 else None))

    def barrier():
        drain()                    # never two threads in a collective at once
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, profile=False):
        barrier()
        l0 = pipe.total_launches()
        if profile:
            pipe.ctx.profile(True)
            pipe.ctx.profile_read(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for _ in range(steps):
            last = fn()
        drain()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        prof = pipe.ctx.profile_read(reset=True) if profile else None
        if profile:
            pipe.ctx.profile(False)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), pipe.total_launches() - l0, prof, last

    for _ in range(max(a.warmup, 0)):
        step_resident()
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
    ms, launches, _, res = timed(step_resident, a.steps)
    clocks = clk.stop() if rank == 0 else None
    # same steps again with per-kernel events (the events perturb the step a little, so they are kept out of `value`)
    for _ in range(2):                 # the plain run uses its own context: let it size its scratch before it is timed
        step_profiled()
    ms_prof, _, prof, res = timed(step_profiled, a.steps, profile=True)
    value = world * a.frames * a.steps / (ms / 1000.0)

    e2e = None
    if not a.no_e2e:
        for _ in range(max(min(a.warmup, 2), 1)):
            step_e2e_single()
        ms_1, _, _, _ = timed(step_e2e_single, a.steps)
        sweep['left'] = a.steps
        ms_e, _, _, res_e = timed(step_e2e, a.steps)
        e2e = {'value': world * a.frames * a.steps / (ms_e / 1000.0), 'unit': UNIT,
               'h2d_bytes_per_step': int(host.h2d_bytes), 'd2h_bytes_per_step': int(res_e.d2h_bytes),
               'ms_per_step': ms_e / a.steps,
               'entry': 'reference layout: float32 depth, uint8 image, one uint8 mask plane per instance; sweep of %d calls, '
                        'each naming the next batch (run_host(prefetch=...)): its first frame group is uploaded during the call before' % a.steps,
               'single_call': {'value': world * a.frames * a.steps / (ms_1 / 1000.0), 'ms_per_step': ms_1 / a.steps,
                               'entry': 'the same call without naming the next batch (nothing overlaps between calls)'}}
        # second host entry: the instance masks arrive bit-packed (one byte per pixel); everything else as above.  Reported
        # next to the headline, never instead of it.
        host_c['hb'] = pipeline.HostBatch.from_device(vb, compact=True, pipe=pipe)
        for _ in range(max(min(a.warmup, 2), 1)):
            sweep['left'] = 0
            step_e2e_compact()
        sweep['left'] = a.steps
        ms_c, _, _, res_c = timed(step_e2e_compact, a.steps)
        e2e['compact_masks'] = {'value': world * a.frames * a.steps / (ms_c / 1000.0), 'unit': UNIT,
                                'h2d_bytes_per_step': int(host_c['hb'].h2d_bytes), 'd2h_bytes_per_step': int(res_c.d2h_bytes),
                                'ms_per_step': ms_c / a.steps, 'entry': 'bit-packed instance masks (uint8 word per pixel) instead of one plane per instance'}
        host_c.clear()

    if rank == 0:
        st = pipe.batch_stats(vb, res)
        roof = None
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        peak = float(peaks.get('hbm_gbs', 6650.0))
        peak_src = 'measured (MEASURED_PEAKS.json hbm_gbs)' if 'hbm_gbs' in peaks else 'fallback 6650 GB/s'
        kernels = {}
        for tag, (tms, calls) in (prof or {}).items():
            ab = algorithmic_bytes(tag, st)
            kernels[tag] = {'ms_per_step': tms / a.steps, 'launches_per_step': calls / a.steps,
                            'algorithmic_bytes_per_step': ab,
                            'achieved_gbs': (ab / 1e9) / (tms / a.steps / 1000.0) if ab and tms > 0 else None}
        cand = [(v['ms_per_step'], k) for k, v in kernels.items() if v['achieved_gbs'] is not None]
        if cand:
            _, top = max(cand)
            k = kernels[top]
            roof = {'bound': 'hbm', 'kernel': top, 'achieved': k['achieved_gbs'], 'peak': peak, 'unit': 'GB/s',
                    'frac': k['achieved_gbs'] / peak, 'traffic': None, 'peak_source': peak_src,
                    'share_of_step': k['ms_per_step'] / (ms_prof / a.steps)}
        if roof is not None:
            try:      # dram bytes of one launch from the committed `ncu --set full` capture of the same command
                tr = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json'))).get(roof['kernel'])
                if tr and tr.get('frames') == a.frames and tr.get('cams') == a.cams:
                    roof['traffic'] = tr['dram_bytes_per_launch']
                    roof['traffic_source'] = tr.get('source')
            except Exception:
                pass
        filt = None
        if world == 1 and not a.no_filters:
            filt = filters_leg(pipe, dev, peak)
            # BASELINE.json's second metric is the FILTER kernels' GB/s: the worst of them rides in `roofline` as well
            worst = None
            for leg in ('radius', 'sor'):
                for tag, kv in filt[leg]['kernels'].items():
                    if tag in ('radius', 'sor_knn') and (worst is None or kv['frac_of_peak'] < worst[2]['frac_of_peak']):
                        worst = (leg, tag, kv)
            if worst is not None and roof is not None:
                leg, tag, kv = worst
                roof['filters'] = {'kernel': tag, 'leg': leg, 'points': filt['points'], 'ms': kv['ms'],
                                   'achieved': kv['achieved_gbs'], 'frac': kv['frac_of_peak'], 'unit': 'GB/s', 'traffic': None}
                try:
                    tr = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json'))).get('filters.' + tag)
                    if tr:
                        roof['filters']['traffic'] = tr['dram_bytes_per_launch']
                        roof['filters']['traffic_source'] = tr.get('source')
                        roof['filters']['traffic_points'] = tr.get('points')    # the capture's cloud, not this leg's
                except Exception:
                    pass
        cpu = None
        if world == 1 and not a.no_cpu_baseline:
            fps, wall, sample = cpu_baseline(a, a.cpu_sample_views or 24)
            cpu = {'value': fps, 'unit': UNIT, 'cores': 1, 'kind': 'port', 'sample': sample, 'seconds': wall}
        line = {'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': a.steps, 'warmup': a.warmup,
                'ms_per_step': ms / a.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
                'dtype': 'f64', 'data': 'synthetic', 'config': workload_config(a), 'e2e': e2e,
                'gpu_launches': int(launches), 'roofline': roof, 'cpu_baseline': cpu, 'clocks': clocks,
                'kernels': kernels, 'batch': st, 'filters': filt, 'numa_node_rank0': numa,
                'frame_groups_in_flight': parts, 'host_cores_per_rank': cores_per_rank,
                'blocking_stream_syncs': os.environ.get('DFU3D_BLOCKING_SYNC') == '1',
                'box_gather': ({'records': gathered['boxes'], 'helper_thread_wall_ms_per_call': 1e3 * gathered.get('seconds', 0.0) / max(gathered.get('calls', 1), 1),
                                'how': 'helper thread, overlapped with the next step, joined inside the timed region'}
                               if world > 1 else None)}
        if json_fd is None:
            print(json.dumps(line))
        else:
            os.write(json_fd, (json.dumps(line) + '\n').encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
