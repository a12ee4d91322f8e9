"""Host side of the batched pipeline that needs no GPU: the view records / instance parameter table the library receives
(dfu3d_b200/pipeline.py: build_views, inst_params), the frame-wise split of a batch and forget_marshalling."""
import types

import numpy as np

from dfu3d_b200 import _lib, pipeline


def _pipe(cfg=None):
    p = object.__new__(pipeline.FusionPipeline)          # no context: only the host-side methods are used
    p.cfg = cfg or pipeline.FusionConfig()
    p._calib_cache = {}
    p._calib_ids = {}
    return p


def _batch(V=30, cams=3, seed=0, share=True):
    r = np.random.default_rng(seed)
    rig = [(r.normal(size=(3, 4)).astype(np.float32), np.linalg.qr(r.normal(size=(3, 3)))[0].astype(np.float32),
            np.hstack([np.linalg.qr(r.normal(size=(3, 3)))[0], r.normal(size=(3, 1))]).astype(np.float32)) for _ in range(cams)]
    n_inst = [int(x) for x in r.integers(0, 5, V)]
    calibs = [rig[v % cams] if share else tuple(a.copy() for a in rig[v % cams]) for v in range(V)]
    return types.SimpleNamespace(
        view_sweep=[v // cams for v in range(V)], calibs=calibs, planes=[r.normal(size=4) for _ in range(V)],
        sweeps=[(sum(1000 - j for j in range(i)), 1000 - i) for i in range(V // cams)],      # contiguous rows of `lidar`
        n_inst=n_inst, classes=[r.integers(0, 10, n) for n in n_inst],
        boxes2d=[r.uniform(0, 900, (n, 4)).astype(np.float32) for n in n_inst], thing_classes=list(pipeline.NUSC_NAMES))


def test_view_records_equal_the_per_view_constants():
    for share in (True, False):                           # calibration arrays shared between frames / fresh copies per view
        vb = _batch(share=share)
        views = _pipe().build_views(vb)
        assert views.dtype == _lib.VIEW_DTYPE and views.shape == (30,)
        for v in range(30):
            c = pipeline.calib_constants(*vb.calibs[v])
            rec = np.zeros(1, _lib.VIEW_DTYPE)[0]
            pipeline.fill_view(rec, c, plane=vb.planes[v], lidar_off=vb.sweeps[vb.view_sweep[v]][0],
                               lidar_cnt=vb.sweeps[vb.view_sweep[v]][1], n_inst=vb.n_inst[v],
                               slot_off=sum(vb.sweeps[vb.view_sweep[u]][1] for u in range(v)), inst_off=sum(vb.n_inst[:v]))
            assert views[v].tobytes() == rec.tobytes(), v


def test_calibration_memo_follows_identity_then_content():
    p = _pipe()
    vb = _batch()
    a = p.build_views(vb).copy()
    vb2 = _batch(share=False)                              # same numbers, different array objects
    assert np.array_equal(p.build_views(vb2).view(np.uint8), a.view(np.uint8))
    vb3 = _batch(seed=1)                                   # other numbers
    assert not np.array_equal(p.build_views(vb3).view(np.uint8), a.view(np.uint8))
    assert len(p._calib_cache) == 6                        # three cameras per rig, two rigs


def test_instance_parameter_table():
    cfg = pipeline.FusionConfig()
    vb = _batch()
    ip = _pipe(cfg).inst_params(vb)
    cls = np.concatenate(vb.classes)
    assert ip.dtype == _lib.INST_DTYPE and ip.shape == cls.shape and np.array_equal(ip['cls'], cls)
    for i in range(0, ip.shape[0], 7):
        sch = cfg.instance_schedule(vb.thing_classes[int(cls[i])])
        assert ip['lidar_r'][i] == sch['lidar_r'] and ip['pseudo_nb'][i] == sch['pseudo_nb'] and ip['gate_dist'][i] == cfg.gate_dist


def test_forget_marshalling_and_frame_groups():
    p = _pipe()
    vb = _batch()
    p.build_views(vb)
    p.inst_params(vb)
    assert hasattr(vb, '_views') and hasattr(vb, '_inst_params')
    holder = vb.__dict__.setdefault('_split_holder', pipeline._Holder(vb))
    metas = pipeline._chunk_metas(holder, 3)
    # contiguous frame groups that cover every sweep, view and instance once, offsets made local
    assert [m[2] for m in metas][0][0] == 0 and [m[2] for m in metas][-1][1] == 30
    assert sum(len(m[0].view_sweep) for m in metas) == 30 and sum(sum(m[0].n_inst) for m in metas) == sum(vb.n_inst)
    for meta, (l0, l1), (v0, v1), (i0, i1) in metas:
        assert meta.sweeps[0][0] == 0 and l1 - l0 == sum(n for _, n in meta.sweeps) and meta.view_sweep[0] == 0
        assert i1 - i0 == sum(meta.n_inst) and v1 - v0 == len(meta.view_sweep)
    pipeline.forget_marshalling(vb)
    assert not any(k in vb.__dict__ for k in ('_views', '_inst_params', '_split_holder'))
