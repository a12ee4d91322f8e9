"""Generates tests/golden/*.npz by EXECUTING THE REFERENCE (oracle/ref_harness.py imports
/root/reference/tools/PENet unchanged, third-party pieces replaced by oracle/fakes.py).

    python tests/golden/make_golden.py

Each fixture holds the inputs of one small synthetic view (so that nothing has to be re-generated on the machine
that checks it) and the reference's outputs: the FOV flag of vis_utils.py:152-154, the fused cloud returned by
depth2pointsrgbpm (my_loader.py:617) and the label file it wrote (my_loader.py:680).  The ground plane is injected
(quirk Q8: the reference's RANSAC is unseeded) and the 15 s clustering abort is frozen (Q6).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from dfu3d_b200 import synth            # noqa: E402
from oracle import ref_harness          # noqa: E402

CASES = [  # name, seed, shape, scale, n_obj
    ('nusc_a', 21, 'nuscenes', 0.125, 6),
    ('nusc_b', 22, 'nuscenes', 0.125, 9),
    ('kitti_a', 23, 'kitti', 0.25, 6),
]


def make(name, seed, shape, scale, n_obj):
    cfg = synth.SHAPES[shape]
    Hm, Wm = int(round(cfg['img'][0] * scale)), int(round(cfg['img'][1] * scale))
    fr = synth.make_frame(seed, shape, n_cams=1, scale=scale, n_obj=n_obj, crop=(Hm - 2, Wm - 2))
    inp = synth.view_as_reference_inputs(fr, 0)
    # thin the sweep: keep every point that projects near the image plus one in four of the rest
    calib = ref_harness.make_calib(inp['P2'], inp['R0'], inp['V2C'])
    img, depth = calib.lidar_to_img(inp['lidar'][:, :3])
    near = (depth > -1) & (img[:, 0] > -50) & (img[:, 0] < Wm + 50) & (img[:, 1] > -50) & (img[:, 1] < Hm + 50)
    keep = near | (np.arange(inp['lidar'].shape[0]) % 4 == 0)
    inp['lidar'] = np.ascontiguousarray(inp['lidar'][keep])
    out = ref_harness.run_depth2pointsrgbpm(inp, plane=inp['plane'])
    H, W = inp['depth'].shape
    pts_rect = calib.lidar_to_rect(inp['lidar'][:, :3])
    fov = ref_harness.get_fov_flag(pts_rect, (H, W, 3), calib)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), name + '.npz')
    np.savez_compressed(
        path, lidar=inp['lidar'], P2=inp['P2'], R0=inp['R0'], V2C=inp['V2C'], depth=inp['depth'],
        image=inp['image'].astype(np.uint8), masks=inp['masks'].astype(np.uint8), classes=inp['classes'],
        scores=inp['scores'], boxes2d=inp['boxes2d'], plane=inp['plane'], thing_classes=np.array(inp['thing_classes']),
        ref_fov=fov, ref_all_points=out['all_points'], ref_label_text=np.array(out['label_text']))
    print(name, 'lidar', inp['lidar'].shape, 'fov', int(fov.sum()), 'cloud', out['all_points'].shape, 'labels',
          out['label_text'].count('\n'), '%.0f kB' % (os.path.getsize(path) / 1e3))


if __name__ == '__main__':
    for c in CASES:
        make(*c)
