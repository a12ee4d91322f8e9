"""Batched fusion / noise-suppression / pseudo-box pipeline over libdfu3d.so.

Host-side orchestration of the stages of the reference's ``depth2pointsrgbpm``
(tools/PENet/dataloaders/my_loader.py:502-617) and ``save_depth_as_points``
(tools/PENet/vis_utils.py:136-166) for a BATCH of camera views.  All arithmetic on points and
pixels happens in the CUDA library; this module only derives per-view constants the way the
reference derives them (float32 numpy on 3x4 / 4x4 matrices), packs argument structs and formats
label text.
"""
import math
from dataclasses import dataclass, field

import os
import time

import numpy as np
import torch

from . import _lib

NUSC_NAMES = ['Car', 'Truck', 'Construction_vehicle', 'Bus', 'Trailer', 'Barrier', 'Motorcycle', 'Bicycle',
              'Pedestrian', 'Traffic_cone']                                  # my_loader.py:634
NUSC_VEHICLES = ('Car', 'Truck', 'Construction_vehicle', 'Bus', 'Trailer')   # my_loader.py:581
NUSC_SMALL = ('Pedestrian', 'Bicycle', 'Motorcycle', 'Barrier', 'Traffic_cone')  # my_loader.py:591


@dataclass
class FusionConfig:
    """Literals of the reference path, gathered in one place (SURVEY.md §5 'Config / flags').
    Defaults = the ``nuscenes`` preset, i.e. tools/PENet/dataloaders/my_loader.py as shipped."""
    bounds: tuple = (1600, 900)            # my_loader.py:526
    min_depth: float = 0.001               # :507
    max_z: float = 1.0                     # :540
    plane_offset: float = 0.05             # :471
    plane_range: tuple = (-54.0, 54.0, -54.0, 54.0)   # :471
    min_theta: float = 1.5                 # :176
    max_per_voxel: int = 100               # :73
    max_voxels: int = 1000000              # :74 (angular voxels kept per view, first-appearance order)
    gate_dist: float = 0.1                 # :603
    cluster_R0: float = 3.0                # rectangle_fitting.py:36
    cluster_Rd: float = 0.001              # rectangle_fitting.py:37
    label_names: list = field(default_factory=lambda: list(NUSC_NAMES))
    allow_voxel_overflow: bool = False
    instances: bool = True                 # False: the per-instance / GenerateAnns stage is not run
    legacy: bool = False                   # True: my_loader0.py schedule (:748-766) and GenerateAnns (:811-930)

    @staticmethod
    def preset(name):
        """Literal sets of the reference's loader variants (SURVEY.md 8(a) "Variants").
        'nuscenes' = my_loader.py (what vis_utils.py:14 imports); 'kitti' = my_loader_KittiPandasetWaymo.py: pixel
        bounds 1216x352 (:709), plane gate range x, y in (-30, 80) (:470); its GenerateAnns call is commented out
        (:831), so the preset stops after the fused cloud."""
        if name == 'nuscenes':
            return FusionConfig()
        if name == 'kitti':
            return FusionConfig(bounds=(1216, 352), plane_range=(-30.0, 80.0, -30.0, 80.0), instances=False)
        if name == 'kitti_legacy':
            # my_loader0.py: KITTI plane range (:470); no bounds test in the label loop (:680) -- PIL raises beyond the
            # 1216 x 352 mask, so the crop is the bound; class-agnostic count-dependent filters; the class name is
            # written as given (label_names = None: the batch's thing_classes)
            return FusionConfig(bounds=(1216, 352), plane_range=(-30.0, 80.0, -30.0, 80.0), legacy=True, label_names=None)
        raise KeyError(name)

    def instance_schedule(self, cls_name):
        """Per-class filter schedule of my_loader.py:581-599.
        Returns dict(lidar_enabled, lidar_nb, lidar_r, pseudo_nb, pseudo_r, car_aspect_rule)."""
        if self.legacy:
            # my_loader0.py:748-766, every class alike: LiDAR (20, r = 2 if n > 50 else 3); pseudo points
            # uniform_down_sample(8 if n > 500, 4 if n > 100, none) then (20, 3).  The counts are read on the device.
            return dict(lidar_enabled=1, lidar_nb=20, lidar_r=2.0, lidar_n0=50, lidar_r_small=3.0, pseudo_nb=20, pseudo_r=3.0,
                        pseudo_n_hi=500, pseudo_k_hi=8, pseudo_n_mid=100, pseudo_k_mid=4, car_aspect_rule=1, anns_mode=1)
        if cls_name in NUSC_VEHICLES:
            return dict(lidar_enabled=1, lidar_nb=1, lidar_r=3.0, pseudo_nb=1, pseudo_r=3.0,
                        car_aspect_rule=int(cls_name == 'Car'))
        if cls_name in NUSC_SMALL:
            # quirk Q5 (my_loader.py:592-594): the LiDAR filter runs on an EMPTY cloud
            return dict(lidar_enabled=0, lidar_nb=1, lidar_r=0.6, pseudo_nb=1, pseudo_r=0.6, car_aspect_rule=0)
        # quirk Q9: the reference raises AttributeError for any other class
        raise AttributeError("class %r is outside both class groups of my_loader.py:581,591 (the reference fails in "
                             "BallQuery with numpy inputs)" % (cls_name,))


def calib_constants(P2, R0, V2C):
    """The derived matrices of calibration_kitti.py:74-79,95-101,110,120, computed with the SAME numpy
    expressions the reference uses (float32 BLAS product, float32 inverse)."""
    P2 = np.asarray(P2, np.float32).reshape(3, 4)
    R0 = np.asarray(R0, np.float32).reshape(3, 3)
    V2C = np.asarray(V2C, np.float32).reshape(3, 4)
    M43 = np.dot(V2C.T, R0.T)
    R0e = np.hstack((R0, np.zeros((3, 1), dtype=np.float32)))
    R0e = np.vstack((R0e, np.zeros((1, 4), dtype=np.float32)))
    R0e[3, 3] = 1
    V2Ce = np.vstack((V2C, np.zeros((1, 4), dtype=np.float32)))
    V2Ce[3, 3] = 1
    inv44 = np.linalg.inv(np.dot(R0e, V2Ce).T)
    fu, fv = P2[0, 0], P2[1, 1]
    return dict(M43=np.ascontiguousarray(M43, np.float32), P2T=np.ascontiguousarray(P2.T, np.float32),
                inv44=np.ascontiguousarray(inv44, np.float32), cu=P2[0, 2], cv=P2[1, 2], fu=fu, fv=fv,
                tx=P2[0, 3] / (-fu), ty=P2[1, 3] / (-fv))


def fill_view(rec, consts, plane=None, lidar_off=0, lidar_cnt=0, n_inst=0, slot_off=0, inst_off=0):
    rec['M43'] = consts['M43'].reshape(-1)
    rec['P2T'] = consts['P2T'].reshape(-1)
    rec['inv44'] = consts['inv44'].reshape(-1)
    for k in ('cu', 'cv', 'fu', 'fv', 'tx', 'ty'):
        rec[k] = consts[k]
    rec['plane'] = np.asarray(plane if plane is not None else (0, 0, 1, 1e9), np.float64)
    rec['lidar_off'] = lidar_off
    rec['lidar_cnt'] = lidar_cnt
    rec['n_inst'] = n_inst
    rec['slot_off'] = slot_off
    rec['inst_off'] = inst_off


def fusion_params(cfg, H, W, Hm, Wm, image_is_i32=False, round_mode=0, no_fov_cull=False):
    p = np.zeros(1, _lib.FUSION_DTYPE)
    p['H'], p['W'], p['Hm'], p['Wm'] = H, W, Hm, Wm
    p['bound_x'], p['bound_y'] = cfg.bounds
    p['round_mode'] = round_mode
    p['image_is_i32'] = int(image_is_i32)
    p['min_depth'] = cfg.min_depth
    p['max_z'] = cfg.max_z
    p['plane_offset'] = cfg.plane_offset
    p['plane_range'] = np.asarray(cfg.plane_range, np.float64)
    p['min_theta'] = cfg.min_theta
    p['max_per_voxel'] = cfg.max_per_voxel
    p['max_voxels'] = cfg.max_voxels
    p['no_fov_cull'] = int(no_fov_cull)
    return p


def search_angles_cs():
    """cos / sin of the 89 search angles of rectangle_fitting.py:117-122, evaluated on the host with the
    same numpy calls (a 1.4 kB constant table)."""
    dtheta = np.deg2rad(1.0)
    th = np.arange(0.0, np.pi / 2.0 - dtheta, dtheta)
    return np.ascontiguousarray(np.stack([np.cos(th), np.sin(th)], 1)), th


@dataclass
class ViewBatch:
    """Device-resident inputs of V camera views."""
    lidar: torch.Tensor            # [sum N, 4] float32 cuda
    sweeps: list                   # [(offset, count)] rows of `lidar`
    view_sweep: list               # sweep index per view
    calibs: list                   # per view (P2, R0, V2C)
    planes: list                   # per view plane[4] float64
    depth: torch.Tensor            # [V, H, W] float32 cuda
    image: torch.Tensor            # [V, Hm, Wm, 3] uint8 (or int32) cuda
    masks: torch.Tensor            # [sum n_inst, Hm, Wm] uint8 (or float32) cuda
    n_inst: list                   # per view
    classes: list                  # per view int array [n_inst]
    boxes2d: list                  # per view float32 [n_inst, 4]
    thing_classes: list = field(default_factory=lambda: list(NUSC_NAMES))
    labelmap: torch.Tensor = None  # [V, Hm, Wm] uint32 as int32 storage (filled by pack)


@dataclass
class FusionResult:
    cloud: torch.Tensor            # [rows, 8]
    row_off: np.ndarray            # [V+1]
    counts_project: np.ndarray     # [V,3]  nFOV, nAbove, nInBounds
    counts_densify: np.ndarray     # [V,4]  valid pixels, gated, voxels, overflowing voxels
    boxes: np.ndarray = None       # [nbox, 24] float64 records (include/dfu3d.h)
    seg_counts: np.ndarray = None  # [n_inst_total, 6]
    box2d: np.ndarray = None       # [nbox, 4] float32
    label_lines: list = None       # per view list of KITTI label lines


class FusionPipeline(object):
    def __init__(self, device=0, config=None, ctx=None):
        self.cfg = config or FusionConfig()
        self.ctx = ctx or _lib.default_context(device)
        self.device = self.ctx.device
        self.lib = self.ctx.lib
        self._angles, _ = search_angles_cs()
        self._calib_cache = {}
        self._calib_ids = {}

    # ------------------------------------------------------------------ helpers
    def build_views(self, vb):
        """dfu_view records of the batch (cached on the ViewBatch: calibs / sweeps / planes are treated as immutable)."""
        cached = getattr(vb, '_views', None)
        if cached is not None:
            return cached
        V = len(vb.view_sweep)
        views = np.zeros(V, _lib.VIEW_DTYPE)
        # A rig has a handful of cameras: the derived constants are kept per calibration.  First by the identity of the three
        # arrays (the same Calibration objects come back frame after frame; the entry holds them, so an id cannot be reused
        # by other arrays), else by their bytes.  One template record per calibration, then one fancy-indexed copy.
        rows, index = [], np.empty(V, np.int64)
        seen = {}
        for v in range(V):
            cal = vb.calibs[v]
            ik = (id(cal[0]), id(cal[1]), id(cal[2]))
            j = seen.get(ik)
            if j is None:
                ent = self._calib_ids.get(ik)
                if ent is None or ent[0] is not cal[0] or ent[1] is not cal[1] or ent[2] is not cal[2]:
                    P2, R0, V2C = cal
                    ck = (np.asarray(P2, np.float32).tobytes(), np.asarray(R0, np.float32).tobytes(), np.asarray(V2C, np.float32).tobytes())
                    c = self._calib_cache.get(ck)
                    if c is None:
                        if len(self._calib_cache) > 4096:
                            self._calib_cache.clear()
                        c = self._calib_cache[ck] = calib_constants(P2, R0, V2C)
                    rec = np.zeros(1, _lib.VIEW_DTYPE)
                    rec['M43'], rec['P2T'], rec['inv44'] = c['M43'].reshape(-1), c['P2T'].reshape(-1), c['inv44'].reshape(-1)
                    for k in ('cu', 'cv', 'fu', 'fv', 'tx', 'ty'):
                        rec[k] = c[k]
                    if len(self._calib_ids) > 4096:
                        self._calib_ids.clear()
                    ent = self._calib_ids[ik] = (cal[0], cal[1], cal[2], rec)
                j = seen[ik] = len(rows)
                rows.append(ent[3])
            index[v] = j
        tab = np.zeros(len(rows), _lib.VIEW_DTYPE)      # (np.concatenate of structured arrays promotes field by field: 2 ms)
        for i, r in enumerate(rows):
            tab[i] = r[0]
        views[:] = tab[index]
        if all(p is not None for p in vb.planes):
            views['plane'] = np.asarray(vb.planes, np.float64).reshape(V, 4)
        else:
            views['plane'] = np.stack([np.asarray(p if p is not None else (0, 0, 1, 1e9), np.float64) for p in vb.planes])
        sw = np.asarray([vb.sweeps[s] for s in vb.view_sweep], np.int64).reshape(V, 2)
        views['lidar_off'] = sw[:, 0]
        views['lidar_cnt'] = sw[:, 1]
        ni = np.asarray(vb.n_inst, np.int64)
        views['n_inst'] = ni
        views['slot_off'] = np.cumsum(sw[:, 1]) - sw[:, 1]
        views['inst_off'] = np.cumsum(ni) - ni
        vb._views = views
        return views

    def pack_masks(self, vb, views=None):
        V = len(vb.view_sweep)
        Hm, Wm = vb.image.shape[1:3] if vb.image is not None else vb.masks.shape[1:3]
        views = self.build_views(vb) if views is None else views
        vd = getattr(vb, '_views_dev', None)            # device copy of the view constants, cached with the host one
        if vd is None or vd.device != self.device or vd.shape[0] != V:
            vd = torch.from_numpy(views.view(np.uint8).reshape(V, -1)).to(self.device)
            vb._views_dev = vd
        lm = torch.empty((V, Hm, Wm), dtype=torch.int32, device=self.device)
        masks = vb.masks
        is_f32 = int(masks.dtype == torch.float32)
        if not is_f32 and masks.dtype not in (torch.uint8, torch.bool):
            raise TypeError('masks must be uint8/bool or float32')
        if masks.dtype == torch.bool:
            masks = masks.view(torch.uint8)
        mptr = _lib.dptr(masks.contiguous()) if masks.numel() else None
        self.ctx.check(self.lib.dfu_pack_masks(self.ctx.h, self.ctx.stream, V, _lib.dptr(vd), mptr, is_f32, Hm, Wm,
                                               _lib.dptr(lm)))
        vb.labelmap = lm
        return lm

    # ------------------------------------------------------------------ the pipeline
    def run(self, vb, cloud_dtype=torch.float16, instances=True, want_lines=True, no_fov_cull=False):
        if instances and self.cfg.instances and len(vb.n_inst) and max(vb.n_inst) > _lib.MAX_INSTANCES:
            return self._run_wide(vb, cloud_dtype, want_lines, no_fov_cull)
        cfg = self.cfg
        V = len(vb.view_sweep)
        H, W = vb.depth.shape[1:3]
        Hm, Wm = vb.image.shape[1:3]
        views = self.build_views(vb)
        if vb.labelmap is None:
            self.pack_masks(vb, views)
        prm = fusion_params(cfg, H, W, Hm, Wm, image_is_i32=(vb.image.dtype == torch.int32), no_fov_cull=no_fov_cull)
        self.ctx.check(self.lib.dfu_fuse_begin(self.ctx.h, self.ctx.stream, V, _lib.hptr(views), _lib.hptr(prm),
                                               _lib.dptr(vb.lidar), _lib.dptr(vb.depth), _lib.dptr(vb.image),
                                               _lib.dptr(vb.labelmap)))
        cp = np.zeros((V, 3), np.int64)
        self.ctx.check(self.lib.dfu_fuse_project_sync(self.ctx.h, _lib.hptr(cp)))
        cd = np.zeros((V, 4), np.int64)
        self.ctx.check(self.lib.dfu_fuse_densify_sync(self.ctx.h, _lib.hptr(cd), int(cfg.allow_voxel_overflow)))
        row_off = np.zeros(V + 1, np.int64)
        self.ctx.check(self.lib.dfu_fuse_emit_cloud(self.ctx.h, None, 0, 0, _lib.hptr(row_off)))
        dt = {torch.float64: 0, torch.float32: 1, torch.float16: 2}[cloud_dtype]
        cloud = torch.empty((int(row_off[-1]), 8), dtype=cloud_dtype, device=self.device)
        self.ctx.check(self.lib.dfu_fuse_emit_cloud(self.ctx.h, _lib.dptr(cloud), dt, cloud.shape[0], _lib.hptr(row_off)))
        res = FusionResult(cloud=cloud, row_off=row_off, counts_project=cp, counts_densify=cd)
        if instances and cfg.instances:
            self.run_instances(vb, res, want_lines=want_lines)
        return res

    def _run_wide(self, vb, cloud_dtype, want_lines, no_fov_cull):
        """Views with more than 32 instances (the label word of a pixel holds 32; the reference keeps every SEEM instance,
        vis_utils.py:218).  The reference treats the instances of a view one by one (my_loader.py:547-607), nothing couples
        them, so such a view is run as ceil(n / 32) views that share its inputs and take 32 masks each; the cloud and the
        counts are taken from the first of them, the box records get the original view number back."""
        lim = _lib.MAX_INSTANCES
        idx, n_inst, classes, boxes2d, first = [], [], [], [], []
        for v, n in enumerate(vb.n_inst):
            first.append(len(idx))
            cls = np.asarray(vb.classes[v]).reshape(-1)
            b2 = np.asarray(vb.boxes2d[v], np.float32).reshape(-1, 4)
            for lo in range(0, max(int(n), 1), lim):
                hi = min(int(n), lo + lim)
                idx.append(v)
                n_inst.append(hi - lo)
                classes.append(cls[lo:hi])
                boxes2d.append(b2[lo:hi])
        sel = torch.as_tensor(idx, dtype=torch.int64, device=vb.depth.device)
        vx = ViewBatch(lidar=vb.lidar, sweeps=vb.sweeps, view_sweep=[vb.view_sweep[v] for v in idx],
                       calibs=[vb.calibs[v] for v in idx], planes=[vb.planes[v] for v in idx],
                       depth=vb.depth.index_select(0, sel), image=vb.image.index_select(0, sel), masks=vb.masks,
                       n_inst=n_inst, classes=classes, boxes2d=boxes2d, thing_classes=vb.thing_classes)
        r = self.run(vx, cloud_dtype=cloud_dtype, instances=True, want_lines=want_lines, no_fov_cull=no_fov_cull)
        V = len(vb.n_inst)
        rows = [r.cloud[int(r.row_off[i]):int(r.row_off[i + 1])] for i in first]
        row_off = np.zeros(V + 1, np.int64)
        row_off[1:] = np.cumsum([int(x.shape[0]) for x in rows])
        out = FusionResult(cloud=torch.cat(rows, 0) if rows else r.cloud[:0], row_off=row_off,
                           counts_project=r.counts_project[first], counts_densify=r.counts_densify[first])
        orig = np.asarray(idx, np.int64)
        out.boxes = r.boxes.copy()
        if out.boxes.shape[0]:
            out.boxes[:, 0] = orig[out.boxes[:, 0].astype(np.int64)]
        out.seg_counts, out.box2d = r.seg_counts, r.box2d
        if r.label_lines is not None:
            out.label_lines = [[] for _ in range(V)]
            for i, lines in enumerate(r.label_lines):
                out.label_lines[idx[i]].extend(lines)
        return out

    def inst_params(self, vb):
        """dfu_inst_params of every instance of the batch (cached on the ViewBatch)."""
        cached = getattr(vb, '_inst_params', None)
        if cached is not None:
            return cached
        cfg = self.cfg
        cls = np.concatenate(list(vb.classes) + [np.zeros(0, np.int64)]).astype(np.int64, copy=False).reshape(-1)
        assert cls.shape[0] == int(sum(vb.n_inst)), 'classes / n_inst mismatch'
        table = np.zeros(len(vb.thing_classes), _lib.INST_DTYPE)
        for c in np.unique(cls):
            sch = cfg.instance_schedule(vb.thing_classes[int(c)])
            t = table[int(c)]
            t['cls'] = int(c)
            t['lidar_enabled'] = sch['lidar_enabled']
            t['lidar_nb'] = sch['lidar_nb']
            t['lidar_every_k'] = sch.get('lidar_every_k', 1)
            t['pseudo_nb'] = sch['pseudo_nb']
            t['pseudo_every_k'] = sch.get('pseudo_every_k', 1)
            t['car_aspect_rule'] = sch['car_aspect_rule']
            t['anns_mode'] = sch.get('anns_mode', 0)
            t['lidar_n0'] = sch.get('lidar_n0', -1)
            t['lidar_r_small'] = sch.get('lidar_r_small', sch['lidar_r'])
            t['pseudo_n_hi'] = sch.get('pseudo_n_hi', -1)
            t['pseudo_k_hi'] = sch.get('pseudo_k_hi', 1)
            t['pseudo_n_mid'] = sch.get('pseudo_n_mid', -1)
            t['pseudo_k_mid'] = sch.get('pseudo_k_mid', 1)
            t['lidar_r'] = sch['lidar_r']
            t['pseudo_r'] = sch['pseudo_r']
            t['gate_dist'] = cfg.gate_dist
            t['cluster_R0'] = cfg.cluster_R0
            t['cluster_Rd'] = cfg.cluster_Rd
        ip = np.ascontiguousarray(table[cls])
        vb._inst_params = ip
        return ip

    def run_instances(self, vb, res, want_lines=True, capacity=None):
        V = len(vb.view_sweep)
        ip = self.inst_params(vb)
        tot = ip.shape[0]
        cap = capacity or max(64, 16 * tot)
        while True:
            boxes = np.empty((cap, _lib.BOX_DOUBLES), np.float64)     # rows [0, nb) are written by the call
            nb = np.zeros(1, np.int64)
            segc = np.zeros((max(tot, 1), 6), np.int64)
            rc = self.lib.dfu_fuse_instances_sync(self.ctx.h, _lib.hptr(ip), tot, _lib.hptr(self._angles),
                                                  self._angles.shape[0], _lib.hptr(boxes), cap, _lib.hptr(nb), _lib.hptr(segc))
            if rc == -3 and int(nb[0]) > cap:      # DFU_ERR_CAPACITY: retry with the reported size
                cap = int(nb[0])
                continue
            self.ctx.check(rc)
            break
        res.boxes = boxes[:int(nb[0])]
        res.seg_counts = segc[:tot]
        b2 = getattr(vb, '_boxes2d_flat', None)
        if b2 is None or b2.shape[0] != tot:
            b2 = np.concatenate([np.reshape(b, (-1, 4)) for b in vb.boxes2d] + [np.zeros((0, 4), np.float32)], 0).astype(np.float32, copy=False)
            vb._boxes2d_flat = b2
        res.box2d = b2[res.boxes[:, 1].astype(np.int64)] if res.boxes.shape[0] else np.zeros((0, 4), np.float32)
        if want_lines:
            res.label_lines = format_label_lines(self.cfg.label_names or vb.thing_classes, res.boxes, vb, b2, V)
        return res


def _string_table(strings):
    """NUL-separated byte table + int32 offsets (the layout dfu_format_label_lines reads)."""
    enc = [x.encode() for x in strings]
    off = np.zeros(max(len(enc), 1), np.int32)
    pos = 0
    for i, e in enumerate(enc):
        off[i] = pos
        pos += len(e) + 1
    return np.frombuffer(b'\0'.join(enc) + b'\0', np.uint8).copy(), off


def format_label_lines(names, boxes, vb, b2_flat, V):
    """All label lines of a batch (my_loader.py:680), per view: the text of format_label_line for every record that
    the Car aspect rule did not skip, produced by the library's host routines dfu_box2d_tokens + dfu_format_label_lines
    (Python spends ~1 us per repr(float); a 64-frame batch has ~36 k of them).  Nothing is cached between calls."""
    lines = [[] for _ in range(V)]
    n = int(boxes.shape[0])
    if n == 0:
        return lines
    lib = _lib.load()
    # class names and the instances' float32 2-D box tokens as NUL-separated tables; both are rebuilt on every call (the
    # tokens by the library: a batch has ~1000 instances, Python's str(np.float32) would take milliseconds)
    nt, no = _string_table(names)
    ni = int(b2_flat.shape[0])
    b2c = np.ascontiguousarray(b2_flat, np.float32).reshape(-1, 4)
    bt = np.empty(max(ni, 1) * 100, np.uint8)
    bo = np.zeros(max(ni, 1), np.int32)
    if lib.dfu_box2d_tokens(_lib.hptr(b2c), ni, _lib.hptr(bt), bt.shape[0], _lib.hptr(bo)) < 0:
        raise RuntimeError('dfu_box2d_tokens failed')
    boxes = np.ascontiguousarray(boxes, np.float64)
    out = np.empty(n * 512, np.uint8)
    per_view = np.zeros(V, np.int64)
    nb = lib.dfu_format_label_lines(_lib.hptr(boxes), n, V, _lib.hptr(nt), _lib.hptr(no), len(names), _lib.hptr(bt),
                                    _lib.hptr(bo), int(b2_flat.shape[0]), _lib.hptr(out), out.shape[0], _lib.hptr(per_view))
    if nb < 0:
        raise RuntimeError('dfu_format_label_lines failed (%d)' % nb)
    flat = out[:nb].tobytes().decode().splitlines(True)
    # records are sorted by view, so each view's lines are one slice
    pos = 0
    for v in range(V):
        c = int(per_view[v])
        lines[v] = flat[pos:pos + c]
        pos += c
    return lines


def format_label_line(name, rec, box2d):
    """The KITTI label line of my_loader.py:680 with Python's str() of float64 / float32 values."""
    f = lambda x: str(np.float64(x))
    g = lambda x: str(np.float32(x))
    return (name + ' ' + str(0) + ' ' + str(0) + ' ' + f(rec[4]) + ' ' + g(box2d[0]) + ' ' + g(box2d[1]) + ' ' +
            g(box2d[2]) + ' ' + g(box2d[3]) + ' ' + f(rec[5]) + ' ' + f(rec[6]) + ' ' + f(rec[7]) + ' ' + f(rec[8]) +
            ' ' + f(rec[9]) + ' ' + f(rec[10]) + ' ' + f(rec[11]) + '\n')


def batch_from_frames(frames, device, thing_classes=None):
    """ViewBatch from dfu3d_b200.synth frames (each frame: one sweep, n views)."""
    lidars, sweeps, view_sweep, calibs, planes, depth, image, masks, n_inst, classes, boxes2d = ([] for _ in range(11))
    off = 0
    for fi, fr in enumerate(frames):
        L = fr['lidar'].to(device=device, dtype=torch.float32)
        lidars.append(L)
        sweeps.append((off, L.shape[0]))
        off += L.shape[0]
        for vw in fr['views']:
            view_sweep.append(fi)
            calibs.append((vw['P2'], vw['R0'], vw['V2C']))
            planes.append(vw['plane'])
            depth.append(vw['depth'].to(device))
            image.append(vw['image'].to(device))
            masks.append(vw['masks'].to(device))
            n_inst.append(int(vw['masks'].shape[0]))
            classes.append(np.asarray(vw['classes'], np.int64))
            boxes2d.append(np.asarray(vw['boxes2d'], np.float32).reshape(-1, 4))
    tc = thing_classes or frames[0]['views'][0]['thing_classes']
    return ViewBatch(lidar=torch.cat(lidars, 0).contiguous(), sweeps=sweeps, view_sweep=view_sweep, calibs=calibs,
                     planes=planes, depth=torch.stack(depth, 0).contiguous(), image=torch.stack(image, 0).contiguous(),
                     masks=torch.cat(masks, 0).contiguous(), n_inst=n_inst, classes=classes, boxes2d=boxes2d,
                     thing_classes=list(tc))


@dataclass
class HostBatch:
    """Pinned host copies of a ViewBatch's tensors (what a caller of the reference holds: numpy arrays on
    the host).  FusionPipeline.run_host copies them to the device inside the call."""
    lidar: torch.Tensor
    depth: torch.Tensor
    image: torch.Tensor
    masks: torch.Tensor
    meta: ViewBatch            # everything that is not a tensor (tensor fields are None)
    h2d_bytes: int = 0
    labelmap: torch.Tensor = None   # compact entry: bit-packed masks [V, Hm, Wm] uint8 / int16 (bit j = instance j); masks is None

    @staticmethod
    def from_device(vb, compact=False, pipe=None):
        """Pinned host copies of the batch.  compact=False: the reference's layout (one uint8 plane per instance).
        compact=True: the instance masks as ONE bit-packed word per pixel (uint8 for <= 8 instances per view, int16 for
        <= 16) -- what a caller that already holds packed masks would pass; the per-instance planes are not uploaded."""
        def pin(t):
            h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            h.copy_(t)
            return h
        meta = ViewBatch(lidar=None, sweeps=vb.sweeps, view_sweep=vb.view_sweep, calibs=vb.calibs, planes=vb.planes,
                         depth=None, image=None, masks=None, n_inst=vb.n_inst, classes=vb.classes, boxes2d=vb.boxes2d,
                         thing_classes=vb.thing_classes)
        if not compact:
            hb = HostBatch(lidar=pin(vb.lidar), depth=pin(vb.depth), image=pin(vb.image), masks=pin(vb.masks), meta=meta)
            hb.h2d_bytes = sum(t.numel() * t.element_size() for t in (hb.lidar, hb.depth, hb.image, hb.masks))
            return hb
        mx = max(list(vb.n_inst) + [0])
        if mx > 16:
            raise ValueError('compact host batch: %d instances in a view (the packed word holds 16)' % mx)
        pipe = pipe or FusionPipeline(vb.depth.device.index or 0)
        lm = pipe.pack_masks(vb)                                      # [V, Hm, Wm] int32 on the device
        vb.labelmap = None
        packed = lm.to(torch.uint8) if mx <= 8 else lm.to(torch.int16)
        hb = HostBatch(lidar=pin(vb.lidar), depth=pin(vb.depth), image=pin(vb.image), masks=None, meta=meta, labelmap=pin(packed))
        hb.h2d_bytes = sum(t.numel() * t.element_size() for t in (hb.lidar, hb.depth, hb.image, hb.labelmap))
        return hb


def _chunk_metas(hb, nchunk):
    """Split a HostBatch frame-wise into `nchunk` contiguous groups of sweeps (frames); cached on the HostBatch.
    Each entry: (meta ViewBatch with chunk-local sweep offsets, lidar row range, view range, instance range)."""
    cache = hb.__dict__.setdefault('_chunks', {})
    if nchunk in cache:
        return cache[nchunk]
    m = hb.meta
    S = len(m.sweeps)
    nchunk = max(1, min(nchunk, S))
    bounds = [round(i * S / nchunk) for i in range(nchunk + 1)]
    vs = np.asarray(m.view_sweep, np.int64)
    assert np.all(np.diff(vs) >= 0), 'views must be grouped by frame'
    inst_off = np.concatenate([[0], np.cumsum(np.asarray(m.n_inst, np.int64))])
    out = []
    for c in range(nchunk):
        s0, s1 = bounds[c], bounds[c + 1]
        v0, v1 = int(np.searchsorted(vs, s0, 'left')), int(np.searchsorted(vs, s1, 'left'))
        l0 = m.sweeps[s0][0]
        l1 = m.sweeps[s1 - 1][0] + m.sweeps[s1 - 1][1]
        meta = ViewBatch(lidar=None, sweeps=[(o - l0, n) for (o, n) in m.sweeps[s0:s1]],
                         view_sweep=[int(x - s0) for x in vs[v0:v1]], calibs=m.calibs[v0:v1], planes=m.planes[v0:v1],
                         depth=None, image=None, masks=None, n_inst=m.n_inst[v0:v1], classes=m.classes[v0:v1],
                         boxes2d=m.boxes2d[v0:v1], thing_classes=m.thing_classes)
        out.append((meta, (l0, l1), (v0, v1), (int(inst_off[v0]), int(inst_off[v1]))))
    cache[nchunk] = out
    return out


def _run_host(self, hb, cloud_dtype=torch.float16, want_lines=True, chunks=4, prefetch=None):
    """The end-to-end call: pinned host buffers in, cloud + boxes back on the host.

    prefetch: the HostBatch of the NEXT call (a sweep over a dataset knows it).  Its first group is put on the copy
    stream while this call's last group is computed, read back, merged and formatted, so that the host->device link
    does not idle between calls; the next call picks the group up (identity of the HostBatch object is checked).

    The batch is cut frame-wise into `chunks` groups that flow through three streams: host->device copies of group
    i+1 (copy stream) overlap the kernels of group i (the caller's stream) and the device->host copy of the cloud of
    group i-1 (second copy stream); PCIe is full duplex, so the call costs about max(transfer, compute) instead of
    their sum.  Device staging buffers are double-buffered and kept between calls."""
    dev = self.device
    st = self.__dict__.setdefault('_stage', {})
    metas = _chunk_metas(hb, int(os.environ.get('DFU3D_HOST_CHUNKS', chunks)))
    main = torch.cuda.current_stream(dev)
    if 's_in' not in st:
        st['s_in'], st['s_out'] = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    s_in, s_out = st['s_in'], st['s_out']
    s_in.wait_stream(main)
    s_out.wait_stream(main)

    def dbuf(name, par, h):
        key = (name, par)
        d = st.get(key)
        if d is None or d.dtype != h.dtype or d.shape[1:] != h.shape[1:] or d.shape[0] < h.shape[0]:
            d = st[key] = torch.empty((int(h.shape[0] * 1.1) + 1,) + tuple(h.shape[1:]), dtype=h.dtype, device=dev)
        return d[:h.shape[0]]

    # groups are numbered through consecutive calls (base + i) so that the staging parity keeps alternating when a
    # group of the next call is uploaded early; done[g]: the kernels of group g have finished with their staging buffers
    pre = st.pop('prefetched', None)
    if pre is not None and pre['hb'] is not hb:
        pre['ev'].synchronize()
        pre = None
    base = pre['g'] if pre is not None else 0
    done = {base - 1: pre['prev_done']} if pre is not None and pre['prev_done'] is not None else {}

    def upload(h, mts, j, g):
        """Group j of batch h onto the copy stream; returns (device tensors, event).  The small parameter uploads of the
        group being computed do not queue behind these copies: the library sends them by kernel (dfu_upload_pinned)."""
        meta, (l0, l1), (v0, v1), (m0, m1) = mts[j]
        par = g & 1
        trace = st.get('trace')
        srcs = (h.lidar[l0:l1], h.depth[v0:v1], h.image[v0:v1], h.masks[m0:m1] if h.labelmap is None else h.labelmap[v0:v1])
        with torch.cuda.stream(s_in):
            if g - 2 in done:
                s_in.wait_event(done[g - 2])          # the staging buffers of this parity are free again
            dst = [dbuf(n, par, x) for n, x in zip(('lidar', 'depth', 'image', 'masks' if h.labelmap is None else 'packed'), srcs)]
            if trace is not None:
                t0 = torch.cuda.Event(enable_timing=True)
                t0.record(s_in)
            for d, x in zip(dst, srcs):
                d.copy_(x, non_blocking=True)
            ev = torch.cuda.Event(enable_timing=trace is not None)
            ev.record(s_in)
            if trace is not None:
                trace.append((g, t0, ev, sum(x.numel() * x.element_size() for x in srcs)))
        return dst, ev

    # two pinned output buffers used alternately: the result of a call stays valid until the call after the next one
    # (a saver thread may still be writing batch i to disk while batch i + 1 is computed)
    st['flip'] = st.get('flip', 0) ^ 1
    hkey = 'cloud_host%d' % st['flip']
    hc = st.get(hkey)
    results, pending = [], ((pre['dst'], pre['ev']) if pre is not None else upload(hb, metas, 0, base))
    row_base, rows_needed, d2h_ev = 0, 0, []
    n_groups = len(metas)
    for i in range(n_groups):
        dst, ev = pending
        if i + 1 < n_groups:
            pending = upload(hb, metas, i + 1, base + i + 1)
        else:
            pending = None
            if prefetch is not None:
                # first group of the next call: global group number base + n_groups; its parity was last used by group
                # base + n_groups - 2, which has been computed (or is the group before this call's first)
                nm = _chunk_metas(prefetch, int(os.environ.get('DFU3D_HOST_CHUNKS', chunks)))
                g = base + n_groups
                pdst, pev = upload(prefetch, nm, 0, g)
                st['prefetched'] = {'hb': prefetch, 'dst': pdst, 'ev': pev, 'g': g & 1, 'prev_done': None}
        meta = metas[i][0]
        main.wait_event(ev)
        meta.labelmap = None
        if hb.labelmap is None:
            meta.lidar, meta.depth, meta.image, meta.masks = dst
        else:
            # compact entry: widen the packed words to the uint32 label map on the device (no per-instance planes at all)
            meta.lidar, meta.depth, meta.image, packed = dst
            meta.masks = packed[:0]
            lm = torch.empty(packed.shape, dtype=torch.int32, device=dev)
            self.ctx.check(self.lib.dfu_widen_labelmap(self.ctx.h, self.ctx.stream, _lib.dptr(packed), packed.element_size(),
                                                       packed.numel(), _lib.dptr(lm)))
            meta.labelmap = lm
        trace = st.get('trace')
        if trace is not None:
            c0 = torch.cuda.Event(enable_timing=True)
            c0.record(main)
        res = self.run(meta, cloud_dtype=cloud_dtype, instances=True, want_lines=want_lines)
        e = torch.cuda.Event(enable_timing=trace is not None)
        e.record(main)
        done[base + i] = e
        if trace is not None:
            trace.append(('compute', base + i, c0, e, time.time()))
        n = int(res.cloud.shape[0])
        if hc is None or hc.dtype != res.cloud.dtype or hc.shape[0] < row_base + n:
            # first call (or a larger batch than ever seen): grow the pinned output, keeping what is already there
            for x in d2h_ev:
                x.synchronize()
            est = int((row_base + n) * max(1.0, len(metas) / float(i + 1)) * 1.25) + 1024
            new = torch.empty((est, 8), dtype=res.cloud.dtype, pin_memory=True)
            if hc is not None and row_base and hc.dtype == new.dtype:
                new[:row_base].copy_(hc[:row_base])
            hc = st[hkey] = new
        with torch.cuda.stream(s_out):
            s_out.wait_event(e)
            hc[row_base:row_base + n].copy_(res.cloud, non_blocking=True)
            res.cloud.record_stream(s_out)
            x = torch.cuda.Event()
            x.record(s_out)
            d2h_ev.append(x)
        results.append((res, row_base))
        row_base += n
    if 'prefetched' in st:
        st['prefetched']['prev_done'] = done.get(base + n_groups - 1)     # frees the other parity for the next call's group 1
    s_out.synchronize()
    main.synchronize()
    out = _merge_results(results, metas, want_lines)
    out.cloud_host = hc[:row_base]
    out.d2h_bytes = out.cloud_host.numel() * out.cloud_host.element_size() + out.boxes.nbytes
    return out


def _batch_stats(self, vb, res):
    """Counts the algorithmic-byte formulas of bench.py need (per step)."""
    V = len(vb.view_sweep)
    H, W = vb.depth.shape[1:3]
    Hm, Wm = vb.image.shape[1:3]
    st = {'views': V, 'pixels': int(V * H * W), 'pixels_m': int(V * Hm * Wm),
          'mask_bytes': int(vb.masks.numel() * vb.masks.element_size()),
          'lidar_slots': int(sum(vb.sweeps[s][1] for s in vb.view_sweep)),
          'valid_pixels': int(res.counts_densify[:, 0].sum()), 'gated': int(res.counts_densify[:, 1].sum()),
          'voxels': int(res.counts_densify[:, 2].sum()), 'cloud_rows': int(res.row_off[-1]),
          'instances': int(sum(vb.n_inst))}
    if res.seg_counts is not None and res.seg_counts.size:
        sc = res.seg_counts
        st['inst_points'] = int(sc[:, 0].sum() + sc[:, 1].sum())
        st['fused_points'] = int(sc[:, 2].sum() + sc[:, 4].sum())
        st['boxes'] = int(res.boxes.shape[0])
    return st


def _merge_results(results, metas, want_lines):
    """Per-group FusionResults -> one (views, instances and rows renumbered batch-wide).  results: [(res, row_base)]."""
    row_off = [np.zeros(1, np.int64)]
    boxes, segc, b2d, lines = [], [], [], []
    for (res, base), (meta, _, (v0, v1), (m0, m1)) in zip(results, metas):
        row_off.append(res.row_off[1:] + base)
        if res.boxes is not None and res.boxes.shape[0]:
            bx = res.boxes.copy()
            bx[:, 0] += v0
            bx[:, 1] += m0
            boxes.append(bx)
            b2d.append(res.box2d)
        if res.seg_counts is not None:
            segc.append(res.seg_counts)
        if res.label_lines is not None:
            lines += res.label_lines
    out = FusionResult(cloud=None, row_off=np.concatenate(row_off),
                       counts_project=np.concatenate([r.counts_project for r, _ in results], 0),
                       counts_densify=np.concatenate([r.counts_densify for r, _ in results], 0))
    out.boxes = np.concatenate(boxes, 0) if boxes else np.zeros((0, _lib.BOX_DOUBLES))
    out.box2d = np.concatenate(b2d, 0) if b2d else np.zeros((0, 4), np.float32)
    out.seg_counts = np.concatenate(segc, 0) if segc else None
    out.label_lines = lines if want_lines else None
    return out


class _Holder(object):
    def __init__(self, meta):
        self.meta = meta


def _run_split(self, vb, cloud_dtype=torch.float16, instances=True, want_lines=True, parts=2):
    """`run` on a device-resident batch cut frame-wise into `parts` groups that run CONCURRENTLY, each on its own
    stream, library context and host thread (ctypes releases the GIL inside the library): while one group's host code
    waits for a count or sizes the next stage, the other group's kernels keep the GPU busy.  Frames share nothing, so
    the result equals `run` on the whole batch (tests/test_gpu_instances.py::test_run_split_equals_run)."""
    import concurrent.futures as cf
    holder = vb.__dict__.setdefault('_split_holder', _Holder(vb))
    metas = _chunk_metas(holder, int(os.environ.get('DFU3D_SPLIT_PARTS', parts)))
    if len(metas) == 1:
        return self.run(vb, cloud_dtype=cloud_dtype, instances=instances, want_lines=want_lines)
    st = self.__dict__.setdefault('_split', {})
    while len(st.setdefault('workers', [])) < len(metas):
        st['workers'].append((FusionPipeline(self.device.index, config=self.cfg, ctx=_lib.Context(self.device.index)),
                              torch.cuda.Stream(self.device)))
    if 'pool' not in st:
        st['pool'] = cf.ThreadPoolExecutor(max_workers=8)
    main = torch.cuda.current_stream(self.device)
    dev = self.device

    def job(i):
        meta, (l0, l1), (v0, v1), (m0, m1) = metas[i]
        sub, stream = st['workers'][i]
        torch.cuda.set_device(dev)
        stream.wait_stream(main)
        with torch.cuda.stream(stream):
            meta.lidar, meta.depth, meta.image, meta.masks = vb.lidar[l0:l1], vb.depth[v0:v1], vb.image[v0:v1], vb.masks[m0:m1]
            meta.labelmap = None
            res = sub.run(meta, cloud_dtype=cloud_dtype, instances=instances, want_lines=want_lines)
            stream.synchronize()
        return res
    futs = [st['pool'].submit(job, i) for i in range(1, len(metas))]
    res0 = job(0)
    parts_res = [res0] + [f.result() for f in futs]
    results, base = [], 0
    for r in parts_res:
        results.append((r, base))
        base += int(r.cloud.shape[0])
    out = _merge_results(results, metas, want_lines and instances and self.cfg.instances)
    if not (instances and self.cfg.instances):
        out.boxes = None
        out.label_lines = None
    out.cloud = torch.cat([r.cloud for r in parts_res], 0)
    return out


def forget_marshalling(batch):
    """Drops everything the pipeline memoises on a ViewBatch / HostBatch between calls (view records and their device
    copy, instance parameter table, flat 2-D boxes, frame-group splits), so that the next call marshals its inputs from
    scratch as it does for a batch it has never seen.  bench.py calls it before every step: the synthetic batch is the
    same object step after step, a sweep over a dataset never is."""
    for k in ('_views', '_views_dev', '_inst_params', '_boxes2d_flat', '_split_holder', '_chunks'):
        batch.__dict__.pop(k, None)
    meta = getattr(batch, 'meta', None)
    if meta is not None:
        forget_marshalling(meta)


def _total_launches(self):
    n = self.ctx.launches
    for sub, _ in self.__dict__.get('_split', {}).get('workers', []):
        n += sub.ctx.launches
    return n


FusionPipeline.run_split = _run_split
FusionPipeline.total_launches = _total_launches
FusionPipeline.run_host = _run_host
FusionPipeline.batch_stats = _batch_stats
