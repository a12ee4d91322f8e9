"""TEST INFRASTRUCTURE ONLY — run the reference's own Python for the path, unchanged.

Imports ``dataloaders.calibration_kitti``, ``dataloaders.my_loader`` and
``rectangle_fitting.rectangle_fitting`` from ``/root/reference/tools/PENet`` after installing
``sys.modules`` stand-ins for the third-party packages that cannot be installed here
(open3d, spconv, skimage, matplotlib, cumm).  open3d and spconv get FUNCTIONAL fakes
(``oracle/fakes.py``); the others are only imported, never called, on the path.

Patches applied to the imported reference modules (each is a documented quirk):
  * Q6  the 15 s wall-clock abort of ``_adoptive_range_segmentation``
        (rectangle_fitting.py:175-188) is disabled by freezing ``time.time`` inside that module;
  * Q8  ``estimate_plane`` uses an unseeded sklearn RANSAC (my_loader.py:456); the harness passes a
        fixed plane by replacing ``estimate_plane`` for the duration of a call;
  * the hard-coded absolute label / calib directories of ``depth2pointsrgbpm``
        (my_loader.py:503-504) do not exist here, so ``os.path.join`` *inside the module* is
        redirected into a temp directory that holds the frame's calib file.

Only usable where ``/root/reference`` exists.  Never imported by ``dfu3d_b200``.
"""
import contextlib
import importlib
import os
import sys
import tempfile
import types

import numpy as np

from . import fakes

REF_ROOT = os.environ.get('DFU3D_REFERENCE', '/root/reference')
PENET = os.path.join(REF_ROOT, 'tools', 'PENet')

_loaded = {}


def available():
    return os.path.isfile(os.path.join(PENET, 'dataloaders', 'my_loader.py'))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _install_stubs():
    if 'open3d' not in sys.modules or not getattr(sys.modules['open3d'], '_dfu_fake', False):
        geometry = types.SimpleNamespace(PointCloud=fakes.FakePointCloud)
        utility = types.SimpleNamespace(Vector3dVector=lambda a: fakes._Vector3dVector(a))
        _stub('open3d', geometry=geometry, utility=utility, _dfu_fake=True)
    if 'spconv' not in sys.modules:
        sp = _stub('spconv')
        sp.__path__ = []
        utils = _stub('spconv.utils', VoxelGeneratorV2=fakes.FakeVoxelGeneratorV2)
        sp.utils = utils
        sp.pytorch = _stub('spconv.pytorch')
    for name in ('matplotlib', 'skimage'):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                m = _stub(name)
                m.__path__ = []
    if 'matplotlib.pyplot' not in sys.modules:
        try:
            importlib.import_module('matplotlib.pyplot')
        except Exception:
            sys.modules['matplotlib'].pyplot = _stub('matplotlib.pyplot')
    if 'skimage.io' not in sys.modules:
        try:
            importlib.import_module('skimage.io')
        except Exception:
            sys.modules['skimage'].io = _stub('skimage.io')


def load(variant='my_loader'):
    """Returns (calibration_kitti, loader_module, rectangle_fitting) of the reference."""
    if variant in _loaded:
        return _loaded[variant]
    if not available():
        raise RuntimeError('reference tree not present at %s' % PENET)
    _install_stubs()
    if PENET not in sys.path:
        sys.path.insert(0, PENET)
    ck = importlib.import_module('dataloaders.calibration_kitti')
    rf = importlib.import_module('rectangle_fitting.rectangle_fitting')
    ml = importlib.import_module('dataloaders.' + variant)
    # Q6: freeze the clock seen by the clustering loop only
    rf.time = types.SimpleNamespace(time=lambda: 0.0)
    _loaded[variant] = (ck, ml, rf)
    return _loaded[variant]


def write_calib_file(path, P2, R0, V2C):
    """KITTI calib text that ``get_calib_from_file`` (calibration_kitti.py:23-60) parses back
    to the same float32 values (repr round-trips float32 exactly)."""
    def row(a):
        return ' '.join(repr(float(np.float32(x))) for x in np.asarray(a).reshape(-1))
    with open(path, 'w') as f:
        f.write('P0: ' + row(P2) + '\n')
        f.write('P1: ' + row(P2) + '\n')
        f.write('P2: ' + row(P2) + '\n')
        f.write('P3: ' + row(P2) + '\n')
        f.write('R0_rect: ' + row(R0) + '\n')
        f.write('Tr_velo_to_cam: ' + row(V2C) + '\n')
        f.write('Tr_imu_to_velo: ' + row(V2C) + '\n')


def make_calib(P2, R0, V2C):
    ck, _, _ = load()
    return ck.Calibration({'P2': np.asarray(P2, np.float32).reshape(3, 4),
                           'R0': np.asarray(R0, np.float32).reshape(3, 3),
                           'Tr_velo2cam': np.asarray(V2C, np.float32).reshape(3, 4)})


def get_fov_flag(pts_rect, img_shape, calib):
    """vis_utils.py:108-123 cannot be imported (XDecoder / detectron2 at module level,
    vis_utils.py:125-133); my_loader.py:85-100 is the identical function and is used instead."""
    _, ml, _ = load()
    return ml.get_fov_flag(pts_rect, img_shape, calib)


@contextlib.contextmanager
def _patched(ml, plane, tmpdir):
    real_join = os.path.join
    real_est = ml.estimate_plane

    def join(a, *rest):
        if isinstance(a, str) and a.startswith('/opt/data/private'):
            sub = 'calib' if a.rstrip('/').endswith('calib') else 'label'
            return real_join(tmpdir, sub, *rest)
        return real_join(a, *rest)

    fake_os = types.SimpleNamespace(**{k: getattr(os, k) for k in dir(os) if not k.startswith('__')})
    fake_os.path = types.SimpleNamespace(**{k: getattr(os.path, k) for k in dir(os.path) if not k.startswith('__')})
    fake_os.path.join = join
    real_os = ml.os
    ml.os = fake_os
    if plane is not None:
        ml.estimate_plane = lambda *a, **k: np.array(plane, dtype=np.float64)
    try:
        yield
    finally:
        ml.os = real_os
        ml.estimate_plane = real_est


def run_depth2pointsrgbpm(frame, plane=None, quiet=True, variant='my_loader'):
    """Runs the reference ``depth2pointsrgbpm`` (my_loader.py:502-617) on a synthetic view.

    ``frame``: dict with lidar [N,4] f32 (NOT yet FOV culled), P2/R0/V2C, depth [H,W] f32,
    image [Hm,Wm,3] int32, masks [n,Hm,Wm] {0,1}, classes [n], scores [n], boxes2d [n,4],
    thing_classes.  Follows vis_utils.save_depth_as_points (vis_utils.py:142-163): FOV cull with
    the cropped image shape, then the call.  Returns dict(all_points, label_text, lidar_fov, depth).
    """
    import torch
    ck, ml, rf = load(variant)
    H, W = frame['depth'].shape[:2]
    calib = make_calib(frame['P2'], frame['R0'], frame['V2C'])
    lidar = np.asarray(frame['lidar'], np.float32)
    image1 = np.asarray(frame['image'], np.int32)
    image = image1[:H, :W]
    pts_rect = calib.lidar_to_rect(lidar[:, 0:3])
    fov_flag = ml.get_fov_flag(pts_rect, image.shape, calib)
    lidar = lidar[fov_flag]
    depth = np.array(frame['depth'], np.float32).reshape(H, W, 1)
    masks = torch.as_tensor(np.asarray(frame['masks'], np.float32))
    classes = torch.as_tensor(np.asarray(frame['classes'], np.int64))
    scores = torch.as_tensor(np.asarray(frame['scores'], np.float32))
    boxes = fakes.FakeBoxes(frame['boxes2d'])
    ld = None if variant == 'my_loader' else np.zeros((H, W), np.float32)      # my_loader0.py:632 prints its shape
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, 'calib'))
        os.makedirs(os.path.join(tmp, 'label'))
        write_calib_file(os.path.join(tmp, 'calib', '000000.txt'), frame['P2'], frame['R0'], frame['V2C'])
        with _patched(ml, plane, tmp):
            if quiet:
                with open(os.devnull, 'w') as dn, contextlib.redirect_stdout(dn):
                    out = ml.depth2pointsrgbpm(depth, image, image1, calib, lidar, list(frame['thing_classes']),
                                               masks, classes, scores, boxes, ld, '000000')
            else:
                out = ml.depth2pointsrgbpm(depth, image, image1, calib, lidar, list(frame['thing_classes']),
                                           masks, classes, scores, boxes, ld, '000000')
        with open(os.path.join(tmp, 'label', '000000.txt')) as f:
            text = f.read()
    return {'all_points': out, 'label_text': text, 'lidar_fov': lidar, 'depth': depth}
