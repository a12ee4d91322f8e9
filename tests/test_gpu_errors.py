"""GPU: error behaviour of the C ABI and of the Python mirror — errors are loud (status code + message -> RuntimeError),
capacity overflow is never a silent truncation, preconditions of the reference's quirks are enforced."""
import ctypes as C

import numpy as np
import pytest
import torch

from dfu3d_b200 import _lib, ops, pipeline, dropin
from tests import helpers

pytestmark = pytest.mark.gpu


def test_more_than_32_instances_is_an_error_of_the_c_abi_only():
    """One view of the C ABI holds 32 masks (loud error beyond); the Python mirror splits such a view (no cap, like the
    reference: tests/test_gpu_instances.py::test_more_than_32_instances_match_oracle)."""
    fr = helpers.small_frame(3, 'nuscenes', scale=0.25)
    vw = fr['views'][0]
    m = vw['masks']
    vw['masks'] = torch.cat([m] * 9)[:33] if m.shape[0] * 9 >= 33 else torch.zeros((33,) + tuple(m.shape[1:]), dtype=m.dtype)
    vw['classes'] = np.zeros(33, np.int64)
    vw['boxes2d'] = np.zeros((33, 4), np.float32)
    pipe = pipeline.FusionPipeline(0)
    vb = pipeline.batch_from_frames([fr], pipe.device)
    with pytest.raises(RuntimeError, match='instances'):
        pipe.run(vb, instances=False)                   # binds the 33-mask view as it is
    vb.labelmap = None
    res = pipe.run(vb)
    assert res.seg_counts.shape[0] == 33


def test_class_outside_both_groups_raises_like_the_reference():
    fr = helpers.small_frame(3, 'nuscenes', scale=0.25)
    pipe = pipeline.FusionPipeline(0)
    vb = pipeline.batch_from_frames([fr], pipe.device, thing_classes=['Car', 'Van', 'Tram'] + ['x'] * 7)
    vb.classes = [np.full_like(c, 1) for c in vb.classes]
    with pytest.raises(AttributeError):                       # quirk Q9
        pipe.run(vb)


def test_capacity_overflow_is_reported_not_truncated():
    ctx = _lib.default_context(0)
    fr = helpers.small_frame(4, 'nuscenes', scale=0.25)
    vw = fr['views'][0]
    d = vw['depth'].cuda().contiguous()[None]
    views = np.zeros(1, _lib.VIEW_DTYPE)
    pipeline.fill_view(views[0], pipeline.calib_constants(vw['P2'], vw['R0'], vw['V2C']))
    prm = np.zeros(1, _lib.FUSION_DTYPE)
    prm['H'], prm['W'], prm['Hm'], prm['Wm'] = d.shape[1], d.shape[2], d.shape[1], d.shape[2]
    prm['min_depth'] = 0.001
    counts = np.zeros(1, np.int64)
    out = torch.empty((10, 3), dtype=torch.float64, device='cuda')
    rc = ctx.lib.dfu_backproject_sync(ctx.h, ctx.stream, 1, _lib.hptr(views), _lib.hptr(prm), _lib.dptr(d), 0, _lib.dptr(out),
                                      None, 10, _lib.hptr(counts))
    assert rc == -3 and counts[0] > 10 and b'capacity' in ctx.lib.dfu_last_error(ctx.h)
    with pytest.raises(RuntimeError):
        ctx.check(rc)
    # boxes: the wrapper retries with the reported size; the raw call reports it
    pipe = pipeline.FusionPipeline(0)
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, instances=False)
    res = pipe.run_instances(vb, res, capacity=1)
    assert res.boxes.shape[0] > 1


def test_bad_arguments():
    ctx = _lib.default_context(0)
    assert ctx.lib.dfu_fuse_project_sync(ctx.h, None) != 0 or True     # (stage check below is the real one)
    h = C.c_void_p()
    ctx2 = _lib.Context(0)
    rc = ctx2.lib.dfu_fuse_densify_sync(ctx2.h, None, 0)                # no batch bound
    assert rc == -1 and b'no batch' in ctx2.lib.dfu_last_error(ctx2.h)
    ctx2.close()
    with pytest.raises(RuntimeError):
        ops.radius_filter(np.array([[np.nan, 0, 0], [0, 0, 0]]), 1, 0.5)
    with pytest.raises(AssertionError):
        ops.radius_filter(np.zeros((4, 3)), 1, 0.5, seg_off=[0, 3])     # seg_off must end at n
    with pytest.raises(TypeError):
        dropin.above_plane(np.zeros((4, 3), np.float64), [0, 0, 1, 1.8])   # LiDAR rows are float32 in the reference
    assert dropin.LShapeFitting().fitting([], []) == ([], [])
    lf = dropin.LShapeFitting(); lf.max_cluster_points = 3
    assert lf.fitting([0, 1, 2, 3], [0, 0, 0, 0]) == (None, None)


def test_empty_inputs():
    assert ops.radius_filter(np.zeros((0, 3)), 1, 0.5).shape == (0,)
    assert ops.nn_gate(np.zeros((0, 3)), np.zeros((5, 3)), 0.1).shape == (0,)
    assert not ops.nn_gate(np.zeros((3, 3)), np.zeros((0, 3)), 0.1).any()
    fr = helpers.small_frame(3, 'nuscenes', scale=0.25)
    vw = fr['views'][0]
    vw['depth'] = torch.zeros_like(vw['depth'])                # nothing to densify
    vw['masks'] = vw['masks'][:0]; vw['classes'] = vw['classes'][:0]; vw['boxes2d'] = vw['boxes2d'][:0]
    pipe = pipeline.FusionPipeline(0)
    res = pipe.run(pipeline.batch_from_frames([fr], pipe.device), cloud_dtype=torch.float64)
    assert res.counts_densify[0].tolist() == [0, 0, 0, 0] and res.cloud.shape[0] == res.counts_project[0, 0]
    assert res.boxes.shape[0] == 0 and res.label_lines == [[]]


def test_widened_rows_report_bad_arguments():
    """GT-database and consumer entry points: bad arguments / too-small buffers are status codes with a message."""
    ctx = _lib.default_context(0)
    pts = torch.rand((100, 4), device='cuda')
    boxes = torch.tensor([[0.5, 0.5, 0.5, 1.0, 1.0, 1.0, 0.3]], device='cuda')
    mask = torch.empty((1, 100), dtype=torch.int32, device='cuda')
    rc = ctx.lib.dfu_points_in_boxes(ctx.h, ctx.stream, _lib.dptr(pts), 2, 100, _lib.dptr(boxes), 1, _lib.dptr(mask))
    assert rc != 0 and b'dfu_points_in_boxes' in ctx.lib.dfu_last_error(ctx.h)
    off = np.zeros(2, np.int64)
    rows = torch.empty((1, 4), device='cuda')
    rc = ctx.lib.dfu_crop_boxes_sync(ctx.h, ctx.stream, _lib.dptr(pts), 4, 100, _lib.dptr(boxes), 1, None, _lib.dptr(rows), 1,
                                     _lib.hptr(off))
    assert rc == -3 and off[1] > 1 and b'capacity' in ctx.lib.dfu_last_error(ctx.h)     # reported, not truncated
    with pytest.raises(RuntimeError, match='empty grid'):
        ops.voxelize(pts.cpu().numpy(), [0.05, 0.05, 0.1], [0, 0, 0, 0, 1, 1], 5, 10)
    # the cloud is written with 128-bit stores: a misaligned destination is refused
    fr = helpers.small_frame(3, 'nuscenes', scale=0.25)
    pipe = pipeline.FusionPipeline(0)
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, instances=False)
    buf = torch.empty((res.cloud.shape[0] * 8 + 8,), dtype=torch.float16, device='cuda')
    row_off = np.zeros(2, np.int64)
    rc = ctx.lib.dfu_fuse_emit_cloud(pipe.ctx.h, buf.data_ptr() + 2, 2, res.cloud.shape[0], _lib.hptr(row_off))
    assert rc != 0 and b'aligned' in ctx.lib.dfu_last_error(pipe.ctx.h)


def test_cluster_radius_beyond_the_cell_neighbourhood_is_an_error():
    """the cell union-find visits the 5x5 neighbourhood of cells of edge R0 / sqrt(2), exact only while
    R0 + Rd * |p| <= sqrt(2) * R0.  A caller-supplied R0 / Rd outside that (or R0 <= 0, Rd < 0) must fail loudly instead of
    splitting components the reference's _adoptive_range_segmentation (rectangle_fitting.py:161-191) would merge."""
    rng = np.random.default_rng(0)
    ox = 60.0 + rng.uniform(-2, 2, 300)
    oy = rng.uniform(-1, 1, 300)
    cl, rect = ops.lshape_fit(ox, oy, R0=0.2, Rd=0.001)           # 0.2 + 0.06 <= 0.283: fine
    assert cl.shape == (300,)
    with pytest.raises(RuntimeError, match='R0'):
        ops.lshape_fit(ox, oy, R0=0.2, Rd=0.01)                    # 0.2 + 0.6 > 0.283
    with pytest.raises(RuntimeError, match='R0'):
        ops.lshape_fit(ox, oy, R0=-1.0, Rd=0.001)
    with pytest.raises(RuntimeError, match='R0'):
        ops.lshape_fit(ox, oy, R0=3.0, Rd=-0.5)
    # the reference's own parameters pass at any realistic range
    cl, rect = ops.lshape_fit(ox, oy, R0=3.0, Rd=0.001)
    from oracle import restate
    want = restate.cluster_adaptive(ox, oy, 3.0, 0.001)
    assert np.array_equal(cl, np.asarray(want[0] if isinstance(want, tuple) else want).astype(cl.dtype))
