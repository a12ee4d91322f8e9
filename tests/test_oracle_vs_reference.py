"""CPU, only where /root/reference exists (skipped on the GPU box): the travelling oracle against the reference's
OWN Python executed in place by oracle/ref_harness.py, on fresh seeded views and on the building blocks."""
import numpy as np
import pytest

from dfu3d_b200 import synth
from oracle import ref_harness, restate

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason='reference tree not present')


def _view(seed, shape, scale, n_obj=6, thin=8):
    cfg = synth.SHAPES[shape]
    Hm, Wm = int(round(cfg['img'][0] * scale)), int(round(cfg['img'][1] * scale))
    fr = synth.make_frame(seed, shape, n_cams=1, scale=scale, n_obj=n_obj, crop=(Hm - 2, Wm - 2))
    inp = synth.view_as_reference_inputs(fr, 0)
    inp['lidar'] = np.ascontiguousarray(inp['lidar'][::thin])
    return inp


# sizes keep every instance below ~600 fused points: the reference's set merging (rectangle_fitting.py:161-191) is
# O(n^3) Python and, with its 15 s abort frozen (Q6), takes minutes at n ~ 1000
@pytest.mark.parametrize('seed,shape,scale,thin', [(31, 'nuscenes', 0.05, 8), (32, 'kitti', 0.12, 6)])
def test_whole_view(seed, shape, scale, thin):
    inp = _view(seed, shape, scale, thin=thin)
    ref = ref_harness.run_depth2pointsrgbpm(inp, plane=inp['plane'])
    fov, allp, lines, recs = restate.run_view(inp)
    nF = int(fov.sum())
    assert np.array_equal(inp['lidar'][fov], ref['lidar_fov'])
    assert allp.shape == ref['all_points'].shape
    assert np.array_equal(allp[:nF], ref['all_points'][:nF])
    np.testing.assert_allclose(allp[nF:, :3], ref['all_points'][nF:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(allp[nF:, 3:], ref['all_points'][nF:, 3:])
    assert len(lines) == ref['label_text'].count('\n')


def test_calibration_methods_bit_exact():
    """calibration_kitti.py:89-144 — forward chain float32, backward chain float64, same bits."""
    inp = _view(33, 'nuscenes', 0.1)
    rc = ref_harness.make_calib(inp['P2'], inp['R0'], inp['V2C'])
    oc = restate.Calib(inp['P2'], inp['R0'], inp['V2C'])
    pts = inp['lidar'][:, :3]
    r_ref = rc.lidar_to_rect(pts)
    r_or = oc.lidar_to_rect(pts)
    assert r_ref.dtype == r_or.dtype == np.float32 and np.array_equal(r_ref, r_or)
    (i_ref, d_ref), (i_or, d_or) = rc.rect_to_img(r_ref), oc.rect_to_img(r_or)
    assert np.array_equal(i_ref, i_or) and np.array_equal(d_ref, d_or)
    u = np.arange(0, 150, dtype=np.int64); v = (u * 7) % 90; d = np.linspace(1, 60, 150).astype(np.float32)
    a, b = rc.img_to_rect(u, v, d), oc.img_to_rect(u, v, d)
    assert a.dtype == b.dtype == np.float64 and np.array_equal(a, b)
    assert np.array_equal(rc.rect_to_lidar(a), oc.rect_to_lidar(b))


@pytest.mark.parametrize('seed,n', [(0, 60), (1, 200), (2, 1), (3, 2)])
def test_lshape_fitting(seed, n):
    """rectangle_fitting.py:39-191 — clusters (ordered by minimum index) and rectangles."""
    _, _, rf = ref_harness.load()
    rng = np.random.default_rng(seed)
    xy = np.concatenate([rng.normal(size=(max(n // 3, 1), 2)) * 0.8 + c for c in ((0, 0), (5.5, 1), (30, -8))])[:n]
    rects, idsets = rf.LShapeFitting().fitting(list(xy[:, 0]), list(xy[:, 1]))
    lab, nc = restate.cluster_adaptive(xy[:, 0], xy[:, 1])
    assert nc == len(idsets)
    for k, s in enumerate(idsets):
        assert set(np.nonzero(lab == k)[0].tolist()) == set(s)
        a, b, c, th, costs = restate.rectangle_search(list(xy[lab == k, 0]), list(xy[lab == k, 1]), return_costs=True)
        srt = np.sort(costs)[::-1]
        if srt.size > 1 and (srt[0] - srt[1]) <= 1e-9 * max(1.0, abs(srt[0])):
            continue
        np.testing.assert_allclose(rects[k].a, a, atol=1e-12)
        np.testing.assert_allclose(rects[k].b, b, atol=1e-12)
        np.testing.assert_allclose(rects[k].c, c, atol=1e-9)


def test_plane_mask_and_sparse_depth():
    _, ml, _ = ref_harness.load()
    inp = _view(34, 'nuscenes', 0.1)
    pts = inp['lidar'][:, :3]
    assert np.array_equal(ml.above_plane(pts, inp['plane']), restate.above_plane(pts, inp['plane']))
    rc = ref_harness.make_calib(inp['P2'], inp['R0'], inp['V2C'])
    oc = restate.Calib(inp['P2'], inp['R0'], inp['V2C'])
    H, W = inp['depth'].shape
    img = inp['image'][:H, :W]
    rgb_r, d_r = ml.load_depth_input(rc, img, inp['lidar'])
    rgb_o, d_o = restate.load_depth_input(oc, img, inp['lidar'])
    assert np.array_equal(rgb_r, rgb_o) and np.array_equal(d_r, d_o)
