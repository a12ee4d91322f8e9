"""GPU: frame-wise sharding end to end -- "N rank shards + ordered gather == the 1-process result" with the real kernels.
Two rank processes (gloo rendezvous on 127.0.0.1; each computes on cuda:(rank mod device_count), so the test also runs on
a 1-GPU box), frames i mod world, batches of 4 frames, shard.globalise + shard.gather_boxes, compared record for record
with one process that ran every frame (tools/PENet/main.py:238,449-457 iterates the frames one by one)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

N_FRAMES, BATCH, CAMS = 22, 4, 2


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close()
    return p


def _frames(ids):
    from tests import helpers
    return [helpers.small_frame(100 + i, 'nuscenes', scale=0.2, n_cams=CAMS, n_obj=6) for i in ids]


def _run_batches(pipe, batches):
    from dfu3d_b200 import pipeline, shard
    out = []
    for ids in batches:
        vb = pipeline.batch_from_frames(_frames(ids), pipe.device)
        res = pipe.run(vb, cloud_dtype=torch.float16)
        fov = [(f, c) for f in ids for c in range(CAMS)]
        out.append(shard.globalise(res.boxes, fov))
    return np.concatenate(out + [np.zeros((0, shard.BOX_DOUBLES))], 0)


def _worker(rank, world, port, out_dir):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from dfu3d_b200 import pipeline, shard
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    pipe = pipeline.FusionPipeline(dev)
    local = _run_batches(pipe, shard.batches_of_rank(N_FRAMES, rank, world, BATCH))
    allb = shard.gather_boxes(local)
    if rank == 0:
        np.save(os.path.join(out_dir, 'gathered.npy'), allb)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_and_gather_equal_one_process(tmp_path):
    from dfu3d_b200 import pipeline, shard
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = np.load(os.path.join(str(tmp_path), 'gathered.npy'))
    pipe = pipeline.FusionPipeline(0)
    one = shard.sort_boxes(_run_batches(pipe, [list(range(i, min(i + BATCH, N_FRAMES))) for i in range(0, N_FRAMES, BATCH)]))
    assert got.shape == one.shape and got.shape[0] > N_FRAMES
    # field 1 (instance ordinal) is local to a batch: compare everything else, in (frame, camera, instance, cluster) order
    cols = [c for c in range(shard.BOX_DOUBLES) if c != 1]
    assert np.array_equal(got[:, cols], one[:, cols])
    key = got[:, [0, 21]]
    assert np.array_equal(key, key[np.lexsort((key[:, 1], key[:, 0]))])
    assert sorted(set(got[:, 0].astype(int))) == sorted(set(one[:, 0].astype(int)))
