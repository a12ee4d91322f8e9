"""BASELINE.json configs[4]-shaped run: a sweep of distinct frame ids sharded frame-wise over the GPUs of one box
(`i mod world`), batches of 64 frames per rank, box lists gathered on the host every batch, order and coverage checked.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/config5_run.py --frames 1024 [--scale 0.25] [--cams 6]

The full train_28130 sweep (tools/PENet/main.py:238,449-457; README.md:65,77) is 28 130 frames; the default here is 1024
distinct ids at reduced image scale so that the run fits a few GPU-minutes (the generator, not the pipeline, is the
slow part: every frame id is ray-cast afresh).  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dfu3d_b200 import pipeline, shard, synth      # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--frames', type=int, default=1024)
    ap.add_argument('--batch', type=int, default=64)
    ap.add_argument('--cams', type=int, default=6)
    ap.add_argument('--sweeps', type=int, default=1)
    ap.add_argument('--scale', type=float, default=0.25)
    ap.add_argument('--check', type=int, default=64, help='frames re-run on rank 0 alone and compared with the gathered records')
    a = ap.parse_args()
    rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)              # NCCL prints to fd 1: keep stdout for the JSON line
        try:
            dist.init_process_group('nccl', device_id=dev)
            dist.barrier()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    pipe = pipeline.FusionPipeline(local)
    cfg = synth.SHAPES['nuscenes']
    Hm, Wm = int(round(cfg['img'][0] * a.scale)), int(round(cfg['img'][1] * a.scale))

    def make(i):
        return synth.make_frame(500000 + i, 'nuscenes', n_cams=a.cams, sweeps=a.sweeps, scale=a.scale, device=dev,
                                max_inst=8, crop=(Hm - 2, Wm - 2) if a.scale != 1.0 else None)

    clock = {'pipe': 0.0}

    def run_ids(ids):
        vb = pipeline.batch_from_frames([make(i) for i in ids], dev)
        torch.cuda.synchronize()
        t1 = time.time()
        res = pipe.run(vb, cloud_dtype=torch.float16)
        torch.cuda.synchronize()
        clock['pipe'] += time.time() - t1
        fov = [(f, c) for f in ids for c in range(a.cams)]
        return shard.globalise(res.boxes, fov), int(res.cloud.shape[0])

    torch.cuda.synchronize()
    t0 = time.time()
    gathered, rows, t_pipe = [], 0, 0.0
    my_batches = shard.batches_of_rank(a.frames, rank, world, a.batch)
    nb = max(len(shard.batches_of_rank(a.frames, r, world, a.batch)) for r in range(world))
    for b in range(nb):
        ids = my_batches[b] if b < len(my_batches) else []
        boxes, r = run_ids(ids) if ids else (np.zeros((0, shard.BOX_DOUBLES)), 0)
        rows += r
        t1 = time.time()
        gathered.append(shard.gather_boxes(boxes, device=dev))        # every rank takes part in every gather
        t_pipe += time.time() - t1
    wall = time.time() - t0
    t_pipe += clock['pipe']
    allb = shard.sort_boxes(np.concatenate(gathered, 0))
    ok_order = bool(np.array_equal(allb, allb[np.lexsort((allb[:, 2], allb[:, 1], allb[:, 21], allb[:, 0]))]))
    frames_seen = np.unique(allb[:, 0].astype(np.int64))
    ok_check = None
    if rank == 0 and a.check > 0:
        ids = list(range(min(a.check, a.frames)))
        cols = [c for c in range(shard.BOX_DOUBLES) if c != 1]
        one = shard.sort_boxes(np.concatenate([run_ids(ids[i:i + a.batch])[0] for i in range(0, len(ids), a.batch)], 0))
        sub = allb[allb[:, 0] < len(ids)]
        ok_check = bool(sub.shape == one.shape and np.array_equal(sub[:, cols], one[:, cols]))
    if world > 1:
        t = torch.tensor([wall, t_pipe], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall, t_pipe = float(t[0]), float(t[1])
    if rank == 0:
        print(json.dumps({'workload': 'configs[4]-shaped sweep: %d distinct frame ids, %d cams of %dx%d, frame i on rank i mod %d, batches of %d'
                                      % (a.frames, a.cams, Wm, Hm, world, a.batch),
                          'n_gpus': world, 'frames': a.frames, 'boxes_gathered': int(allb.shape[0]),
                          'frames_with_boxes': int(frames_seen.size), 'order_ok': ok_order,
                          'gathered_equals_one_rank_on_first_%d_frames' % min(a.check, a.frames): ok_check,
                          'wall_s_incl_generation': wall, 'pipeline_plus_gather_s': t_pipe,
                          'frames_per_s_pipeline': a.frames / t_pipe if t_pipe > 0 else None}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
