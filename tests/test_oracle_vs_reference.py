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


def test_earlier_variants_and_corner_boxes():
    """depth2points / depth2pointsrgb / depth2pointsrgbp (my_loader.py:142-164,415-442; imported by vis_utils.py:14) and
    Calibration.corners3d_to_img_boxes (calibration_kitti.py:146-164): the restatements equal the reference's own code."""
    inp = _view(35, 'nuscenes', 0.06)
    _, ml, _ = ref_harness.load('my_loader')
    rc = ref_harness.make_calib(inp['P2'], inp['R0'], inp['V2C'])
    oc = restate.Calib(inp['P2'], inp['R0'], inp['V2C'])
    H, W = inp['depth'].shape[:2]
    img = inp['image'][:H, :W]
    d = lambda: np.array(inp['depth'], np.float32).reshape(H, W, 1)
    assert np.array_equal(ml.depth2points(d(), rc), restate.depth2points(d(), oc))
    assert np.array_equal(ml.depth2pointsrgb(d(), img, rc), restate.depth2pointsrgb(d(), img, oc))
    a = ml.depth2pointsrgbp(d(), img, rc, inp['lidar'])
    b = restate.depth2pointsrgbp(d(), img, oc, inp['lidar'])
    assert a.shape == b.shape and a.shape[0] > inp['lidar'].shape[0]
    np.testing.assert_allclose(a, b, rtol=1e-12, atol=0)
    rng = np.random.default_rng(0)
    c3 = rng.uniform(-10, 10, size=(40, 8, 3)) + np.array([0, 0, 25.0])
    rb, rcn = rc.corners3d_to_img_boxes(c3)
    ob, ocn = restate.corners3d_to_img_boxes(inp['P2'], c3)
    assert np.array_equal(rb, ob) and np.array_equal(rcn, ocn)


@pytest.mark.parametrize('seed,scale,thin', [(43, 0.09, 8), (45, 0.08, 6)])
def test_kitti_legacy_whole_view(seed, scale, thin):
    """my_loader0.py (the older KITTI loader: count-dependent radius / stride schedule :748-766, GenerateAnns with
    'prior_linecorner', forced 3.9 x 1.6, < 10 m skip :811-930) executed in place vs the 'kitti_legacy' restatement."""
    inp = _view(seed, 'kitti', scale, thin=thin)
    inp['thing_classes'] = ['Car', 'Pedestrian', 'Cyclist']
    inp['classes'] = np.asarray(inp['classes']) % 3
    ref = ref_harness.run_depth2pointsrgbpm(inp, plane=inp['plane'], variant='my_loader0')
    fov, allp, lines, recs = restate.run_view(inp, preset='kitti_legacy')
    nF = int(fov.sum())
    assert allp.shape == ref['all_points'].shape
    assert np.array_equal(allp[:nF], ref['all_points'][:nF])
    np.testing.assert_allclose(allp[nF:, :3], ref['all_points'][nF:, :3], rtol=1e-5, atol=1e-9)
    ref_lines = ref['label_text'].splitlines(True)
    assert len(lines) == len(ref_lines), (len(lines), len(ref_lines))
    for a, b in zip(lines, ref_lines):
        fa, fb = a.split(), b.split()
        assert fa[:3] == fb[:3] and fa[4:8] == fb[4:8]
        np.testing.assert_allclose(np.array(fa[3:4] + fa[8:], float), np.array(fb[3:4] + fb[8:], float), rtol=1e-9, atol=1e-9)
    print('legacy boxes', len(lines))
