"""ctypes binding of libdfu3d.so (include/dfu3d.h).  There is no fallback: a missing library or
CUDA device raises."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# DFU3D_LIB: another build of the same library (kernel tuning variants, tools/variants.py)
LIB_PATH = os.environ.get('DFU3D_LIB') or os.path.join(_HERE, 'csrc', 'libdfu3d.so')

MAX_INSTANCES = 32
BOX_DOUBLES = 24

VIEW_DTYPE = np.dtype([
    ('M43', np.float32, (12,)), ('P2T', np.float32, (12,)), ('inv44', np.float32, (16,)),
    ('cu', np.float32), ('cv', np.float32), ('fu', np.float32), ('fv', np.float32), ('tx', np.float32), ('ty', np.float32),
    ('lidar_cnt', np.int32), ('n_inst', np.int32),
    ('plane', np.float64, (4,)),
    ('lidar_off', np.int64), ('slot_off', np.int64),
    ('inst_off', np.int32), ('pad0', np.int32),
], align=True)
assert VIEW_DTYPE.itemsize == 248, VIEW_DTYPE.itemsize

FUSION_DTYPE = np.dtype([
    ('H', np.int32), ('W', np.int32), ('Hm', np.int32), ('Wm', np.int32), ('bound_x', np.int32), ('bound_y', np.int32),
    ('round_mode', np.int32), ('image_is_i32', np.int32), ('min_depth', np.float32), ('max_voxels', np.int32),
    ('max_z', np.float64), ('plane_offset', np.float64), ('plane_range', np.float64, (4,)), ('min_theta', np.float64),
    ('max_per_voxel', np.int32), ('no_fov_cull', np.int32),
], align=True)
assert FUSION_DTYPE.itemsize == 104, FUSION_DTYPE.itemsize

INST_DTYPE = np.dtype([
    ('cls', np.int32), ('lidar_enabled', np.int32), ('lidar_nb', np.int32), ('lidar_every_k', np.int32),
    ('pseudo_nb', np.int32), ('pseudo_every_k', np.int32), ('car_aspect_rule', np.int32), ('anns_mode', np.int32),
    ('lidar_r', np.float64), ('pseudo_r', np.float64), ('gate_dist', np.float64),
    ('cluster_R0', np.float64), ('cluster_Rd', np.float64),
    ('lidar_r_small', np.float64), ('lidar_n0', np.int32),
    ('pseudo_n_hi', np.int32), ('pseudo_k_hi', np.int32), ('pseudo_n_mid', np.int32), ('pseudo_k_mid', np.int32),
    ('pad', np.int32),
], align=True)
assert INST_DTYPE.itemsize == 104, INST_DTYPE.itemsize

PROF_NTAGS = 24
PROF_TAGS = ['pack_masks', 'project', 'lists', 'densify_a', 'densify_b', 'densify_c', 'vox_gather', 'emit_cloud',
             'inst_pairs', 'inst_sort', 'grid_build', 'radius', 'gate', 'cluster', 'fit', 'sor_knn', 'sor_mask',
             'sparse_depth', 'misc']

_lib = None

_P = C.c_void_p
_SIGS = {
    'dfu_create': (C.c_int, [C.POINTER(_P), C.c_int]),
    'dfu_destroy': (None, [_P]),
    'dfu_last_error': (C.c_char_p, [_P]),
    'dfu_launch_count': (C.c_int64, [_P]),
    'dfu_profile_enable': (C.c_int, [_P, C.c_int]),
    'dfu_profile_read': (C.c_int, [_P, _P, _P, C.c_int, C.c_int]),
    'dfu_project_label': (C.c_int, [_P, _P, C.c_int, _P, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P]),
    'dfu_calib_apply': (C.c_int, [_P, _P, _P, C.c_int, _P, _P, _P, C.c_int64, _P, _P]),
    'dfu_pack_masks': (C.c_int, [_P, _P, C.c_int, _P, _P, C.c_int, C.c_int, C.c_int, _P]),
    'dfu_widen_labelmap': (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P]),
    'dfu_sparse_depth': (C.c_int, [_P, _P, C.c_int, _P, C.c_int, _P, _P, _P]),
    'dfu_fuse_begin': (C.c_int, [_P, _P, C.c_int, _P, _P, _P, _P, _P, _P]),
    'dfu_fuse_project_sync': (C.c_int, [_P, _P]),
    'dfu_fuse_densify_sync': (C.c_int, [_P, _P, C.c_int]),
    'dfu_fuse_emit_cloud': (C.c_int, [_P, _P, C.c_int, C.c_int64, _P]),
    'dfu_fuse_emit_voxels': (C.c_int, [_P, _P, _P, _P, _P, C.c_int64]),
    'dfu_fuse_emit_project': (C.c_int, [_P, _P, _P, _P]),
    'dfu_fuse_instances_sync': (C.c_int, [_P, _P, C.c_int, _P, C.c_int, _P, C.c_int64, _P, _P]),
    'dfu_backproject_sync': (C.c_int, [_P, _P, C.c_int, _P, _P, _P, C.c_int, _P, _P, C.c_int64, _P]),
    'dfu_split_cloud_sync': (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P, _P, _P, _P]),
    'dfu_split_cloud_static': (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P, _P, _P, _P]),
    'dfu_voxelize_sync': (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P, _P, C.c_int, C.c_int, _P, _P, _P, _P]),
    'dfu_points_in_boxes': (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P, C.c_int, _P]),
    'dfu_crop_boxes_sync': (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P, C.c_int, _P, _P, C.c_int64, _P]),
    'dfu_boxes_iou_sync': (C.c_int, [_P, _P, _P, C.c_int, C.c_int, _P, C.c_int, C.c_int, C.c_int, _P]),
    'dfu_nms_sync': (C.c_int, [_P, _P, _P, C.c_int, _P, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, _P, _P]),
    'dfu_nms_normal_sync': (C.c_int, [_P, _P, _P, C.c_int, _P, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, _P, _P]),
    'dfu_radius_filter': (C.c_int, [_P, _P, _P, C.c_int, _P, _P, _P, _P, _P, _P]),
    'dfu_sor_filter': (C.c_int, [_P, _P, _P, C.c_int, _P, C.c_int, C.c_double, _P, _P, _P]),
    'dfu_repr_f64': (C.c_int64, [_P, C.c_int64, _P, C.c_int64]),
    'dfu_box2d_tokens': (C.c_int64, [_P, C.c_int64, _P, C.c_int64, _P]),
    'dfu_format_label_lines': (C.c_int64, [_P, C.c_int64, C.c_int, _P, _P, C.c_int, _P, _P, C.c_int64, _P, C.c_int64, _P]),
    'dfu_sor_counter': (C.c_int64, [_P, C.c_int]),
    'dfu_nn_gate': (C.c_int, [_P, _P, _P, C.c_int64, _P, C.c_int64, C.c_double, _P]),
    'dfu_generate_anns_sync': (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_int, C.c_double, C.c_double, _P, C.c_int, _P, C.c_int64, _P]),
    'dfu_lshape_fit_sync': (C.c_int, [_P, _P, _P, C.c_int64, C.c_double, C.c_double, _P, C.c_int, _P, _P, C.c_int64, _P]),
}
EXPORTED = sorted(_SIGS)


def load():
    """dlopen the library and bind every entry point of include/dfu3d.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError('libdfu3d.so is not built (%s); run `python -c "import __graft_entry__ as g; g.build()"` — '
                           'there is no CPU fallback' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    missing = [name for name in _SIGS if not hasattr(lib, name)]
    if missing:
        raise RuntimeError('libdfu3d.so does not export %s (stale build?)' % ', '.join(missing))
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def hptr(a):
    """host pointer of a numpy array (None -> NULL)"""
    return None if a is None else a.ctypes.data_as(_P)


def dptr(t):
    """device pointer of a CUDA torch tensor (None -> NULL)"""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), 'expected a contiguous CUDA tensor'
    return _P(t.data_ptr())


class Context(object):
    """One dfu_ctx per (device); stream is torch's current stream at call time."""

    def __init__(self, device=0):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError('dfu3d_b200 needs a CUDA device (no CPU fallback)')
        self.lib = load()
        self.device = torch.device('cuda', device if isinstance(device, int) else torch.device(device).index or 0)
        h = _P()
        rc = self.lib.dfu_create(C.byref(h), self.device.index)
        if rc != 0:
            raise RuntimeError('dfu_create failed (%d)' % rc)
        self.h = h

    def check(self, rc):
        if rc != 0:
            raise RuntimeError('libdfu3d error %d: %s' % (rc, self.lib.dfu_last_error(self.h).decode()))

    @property
    def stream(self):
        import torch
        return _P(torch.cuda.current_stream(self.device).cuda_stream)

    @property
    def launches(self):
        return int(self.lib.dfu_launch_count(self.h))

    def profile(self, on=True):
        self.check(self.lib.dfu_profile_enable(self.h, int(on)))

    def profile_read(self, reset=True):
        """{tag: (total ms, launches)} of the instrumented kernels since the last reset."""
        ms = np.zeros(PROF_NTAGS, np.float64)
        calls = np.zeros(PROF_NTAGS, np.int64)
        self.check(self.lib.dfu_profile_read(self.h, hptr(ms), hptr(calls), PROF_NTAGS, int(reset)))
        return {PROF_TAGS[i]: (float(ms[i]), int(calls[i])) for i in range(len(PROF_TAGS)) if calls[i] > 0}

    def close(self):
        if getattr(self, 'h', None):
            self.lib.dfu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}


def default_context(device=0):
    import torch
    idx = device if isinstance(device, int) else (torch.device(device).index or 0)
    if idx not in _default:
        _default[idx] = Context(idx)
    return _default[idx]
