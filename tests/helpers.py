"""Shared helpers of the parity tests (oracle = checker only)."""
import numpy as np
import torch

from dfu3d_b200 import synth, pipeline, _lib
from oracle import restate


def small_frame(seed, shape='nuscenes', scale=0.25, n_obj=8, n_cams=1, sweeps=1, crop_margin=2):
    cfg = synth.SHAPES[shape]
    Hm, Wm = int(round(cfg['img'][0] * scale)), int(round(cfg['img'][1] * scale))
    return synth.make_frame(seed, shape, n_cams=n_cams, sweeps=sweeps, scale=scale, n_obj=n_obj,
                            crop=(Hm - crop_margin, Wm - crop_margin))


def oracle_view(frame, v=0, debug=True):
    inp = synth.view_as_reference_inputs(frame, v)
    return inp, restate.run_view(inp, return_debug=debug)


def emit_project(pipe, total_slots):
    dev = pipe.device
    pix = torch.empty((total_slots, 2), dtype=torch.int32, device=dev)
    flags = torch.empty(total_slots, dtype=torch.uint8, device=dev)
    label = torch.empty(total_slots, dtype=torch.int32, device=dev)
    pipe.ctx.check(pipe.lib.dfu_fuse_emit_project(pipe.ctx.h, _lib.dptr(pix), _lib.dptr(flags), _lib.dptr(label)))
    torch.cuda.synchronize()
    return pix.cpu().numpy(), flags.cpu().numpy(), label.cpu().numpy().view(np.uint32)


def emit_voxels(pipe, n):
    dev = pipe.device
    rz = torch.empty(n, dtype=torch.int32, device=dev)
    ry = torch.empty(n, dtype=torch.int32, device=dev)
    xz = torch.empty((n, 3), dtype=torch.float64, device=dev)
    xy = torch.empty((n, 3), dtype=torch.float64, device=dev)
    pipe.ctx.check(pipe.lib.dfu_fuse_emit_voxels(pipe.ctx.h, _lib.dptr(rz), _lib.dptr(ry), _lib.dptr(xz), _lib.dptr(xy), n))
    torch.cuda.synchronize()
    return rz.cpu().numpy(), ry.cpu().numpy(), xz.cpu().numpy(), xy.cpu().numpy()
