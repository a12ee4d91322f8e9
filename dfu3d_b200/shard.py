"""Frame-wise sharding of the path across the GPUs of one box.

Frames share nothing (tools/PENet/main.py:238,452-457 iterates them one by one), so rank r owns frames
{i : i mod world == r}; there is NO collective on the data path.  The only exchange is a host-side gather of the
variable-length pseudo-box lists (~200 B per box): all_gather of the counts, then of zero-padded [n, 24] float64
records, re-ordered by (frame, instance, cluster) = the order the reference writes label lines in.  Works with the
`nccl` backend (tensors on the rank's GPU) and with `gloo` (CPU tensors; used by the CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import _lib

BOX_DOUBLES = _lib.BOX_DOUBLES


def frames_of_rank(n_frames, rank, world):
    """Global frame ids owned by `rank` (interleaved, so that the imbalance is at most one frame)."""
    return list(range(rank, n_frames, world))


def batches_of_rank(n_frames, rank, world, batch):
    """The rank's frames grouped into batches of `batch` frames (last one ragged)."""
    mine = frames_of_rank(n_frames, rank, world)
    return [mine[i:i + batch] for i in range(0, len(mine), batch)]


def globalise(boxes, frame_of_view):
    """Box records of one local batch -> records whose field 0 is the GLOBAL frame id and field 21 the local view
    (camera) ordinal inside the frame.  frame_of_view[v] = (global frame id, camera ordinal) of local view v."""
    out = np.array(boxes, dtype=np.float64, copy=True).reshape(-1, BOX_DOUBLES)
    if out.shape[0]:
        fv = np.asarray(frame_of_view, np.int64).reshape(-1, 2)
        v = out[:, 0].astype(np.int64)
        out[:, 0] = fv[v, 0]
        out[:, 21] = fv[v, 1]
    return out


def gather_boxes(local_boxes, device=None, group=None, dst=None):
    """All ranks' box records, sorted by (frame, camera, instance, cluster): on every rank (dst=None) or on rank `dst`
    only (the other ranks get an empty list and skip the read-back and the sort).  No-op without an initialised
    process group."""
    local = np.ascontiguousarray(np.asarray(local_boxes, np.float64).reshape(-1, BOX_DOUBLES))
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return sort_boxes(local)
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device('cpu') if backend == 'gloo' else (device or torch.device('cuda', torch.cuda.current_device()))
    if backend != 'gloo':
        return sort_boxes(_gather_one_collective(local, world, dev, group, dst))
    cnt = torch.tensor([local.shape[0]], dtype=torch.int64, device=dev)
    cnts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(cnts, cnt, group=group)
    counts = [int(c.item()) for c in cnts]
    mx = max(max(counts), 1)
    pad = torch.zeros((mx, BOX_DOUBLES), dtype=torch.float64, device=dev)
    if local.shape[0]:
        pad[:local.shape[0]] = torch.from_numpy(local).to(dev)
    bufs = [torch.zeros_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    if dst is not None and dist.get_rank(group) != dst:
        return np.zeros((0, BOX_DOUBLES), np.float64)
    parts = [b[:c].cpu().numpy() for b, c in zip(bufs, counts)]
    return sort_boxes(np.concatenate(parts, 0))


_CAP = {}


def _gather_one_collective(local, world, dev, group, dst=None):
    """NCCL path: ONE all_gather_into_tensor of [count | records padded to a capacity every rank agrees on] and one
    device->host copy (the list form needs two collectives, world + 1 host syncs for the counts and a copy per rank).
    The capacity is remembered; a rank that holds more records than it makes every rank repeat with the larger size."""
    n = int(local.shape[0])
    cap = _CAP.get(id(group), 2048)
    while True:
        host = np.zeros((cap + 1, BOX_DOUBLES), np.float64)
        host[0, 0] = n
        m = min(n, cap)
        host[1:1 + m] = local[:m]
        buf = torch.from_numpy(host).to(dev)
        out = torch.empty((world * (cap + 1), BOX_DOUBLES), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(out, buf, group=group)
        # every rank needs the counts (they decide whether the round is repeated); the records only where they are wanted
        counts = out.view(world, cap + 1, BOX_DOUBLES)[:, 0, 0].cpu().numpy().astype(np.int64)
        if int(counts.max()) <= cap:
            if dst is not None and dist.get_rank(group) != dst:
                return np.zeros((0, BOX_DOUBLES), np.float64)
            h = out.cpu().numpy().reshape(world, cap + 1, BOX_DOUBLES)
            return np.concatenate([h[r, 1:1 + int(counts[r])] for r in range(world)], 0)
        while cap < int(counts.max()):
            cap *= 2
        _CAP[id(group)] = cap


def sort_boxes(boxes):
    """(frame, camera, instance, cluster) order: the reference writes one label file per frame, instance by instance,
    cluster by cluster (my_loader.py:547-607, 642)."""
    b = np.asarray(boxes, np.float64).reshape(-1, BOX_DOUBLES)
    if b.shape[0] == 0:
        return b
    order = np.lexsort((b[:, 2], b[:, 1], b[:, 21], b[:, 0]))
    return b[order]
