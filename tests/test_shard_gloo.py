"""CPU, world_size 2 over gloo: the frame-wise partition and the host-side gather of the pseudo-box lists
(dfu3d_b200/shard.py).  No kernel runs here: the box records are synthetic."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dfu3d_b200 import shard


def _free_port():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); p = s.getsockname()[1]; s.close()
    return p


def _fake_boxes(frame, cams=2):
    """deterministic records of one frame: `frame % 3` boxes per camera"""
    rows = []
    for cam in range(cams):
        for j in range(frame % 3):
            r = np.zeros(shard.BOX_DOUBLES)
            r[0] = cam                      # LOCAL view id inside a one-frame batch
            r[1] = 10 * frame + j           # instance
            r[2] = j % 2                    # cluster
            r[5:8] = frame + 0.25 * j
            rows.append(r)
    return np.array(rows).reshape(-1, shard.BOX_DOUBLES)


def _worker(rank, world, port, n_frames, out_dir):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    local = []
    for batch in shard.batches_of_rank(n_frames, rank, world, batch=2):
        for f in batch:
            local.append(shard.globalise(_fake_boxes(f), [(f, 0), (f, 1)]))
    local = np.concatenate(local + [np.zeros((0, shard.BOX_DOUBLES))], 0)
    allb = shard.gather_boxes(local)
    np.save(os.path.join(out_dir, 'rank%d.npy' % rank), allb)
    np.save(os.path.join(out_dir, 'dst0_rank%d.npy' % rank), shard.gather_boxes(local, dst=0))     # delivery to rank 0 only
    dist.barrier()
    dist.destroy_process_group()


def test_partition_covers_every_frame_once():
    for world in (1, 2, 4, 8):
        seen = sorted(f for r in range(world) for f in shard.frames_of_rank(37, r, world))
        assert seen == list(range(37))
        sizes = [len(shard.frames_of_rank(37, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
    assert shard.batches_of_rank(10, 1, 2, 2) == [[1, 3], [5, 7], [9]]


def test_gather_boxes_world2(tmp_path):
    n_frames, world = 9, 2
    mp.spawn(_worker, args=(world, _free_port(), n_frames, str(tmp_path)), nprocs=world, join=True)
    a = np.load(os.path.join(str(tmp_path), 'rank0.npy'))
    b = np.load(os.path.join(str(tmp_path), 'rank1.npy'))
    assert np.array_equal(a, b)
    want = shard.sort_boxes(np.concatenate([shard.globalise(_fake_boxes(f), [(f, 0), (f, 1)]) for f in range(n_frames)], 0))
    assert np.array_equal(a, want)
    d0 = np.load(os.path.join(str(tmp_path), 'dst0_rank0.npy'))
    d1 = np.load(os.path.join(str(tmp_path), 'dst0_rank1.npy'))
    assert np.array_equal(d0, want) and d1.shape == (0, shard.BOX_DOUBLES)
    key = a[:, [0, 21, 1, 2]]
    assert np.array_equal(key, key[np.lexsort((key[:, 3], key[:, 2], key[:, 1], key[:, 0]))])


def test_gather_without_process_group_is_local_sort():
    b = np.concatenate([shard.globalise(_fake_boxes(f), [(f, 0), (f, 1)]) for f in (5, 2, 4)], 0)
    out = shard.gather_boxes(b)
    assert out.shape == b.shape and np.all(np.diff(out[:, 0]) >= 0)
