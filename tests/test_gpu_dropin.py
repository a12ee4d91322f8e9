"""GPU: the drop-in mirror of the reference's Python entry points (dfu3d_b200/dropin.py), called exactly as the
reference's callers call them, against (a) the committed golden outputs of the reference itself and (b) the oracle."""
import glob
import io
import os

import numpy as np
import pytest
import torch

from dfu3d_b200 import dropin
from oracle import restate, fakes
from tests.test_oracle_golden import GOLDEN, load_golden

pytestmark = pytest.mark.gpu


def _calib(inp):
    return dropin.Calibration({'P2': inp['P2'], 'R0': inp['R0'], 'Tr_velo2cam': inp['V2C']})


def _compare_lines(lines, ref_lines):
    assert len(lines) == len(ref_lines)
    for a, b in zip(lines, ref_lines):
        fa, fb = a.split(), b.split()
        assert fa[:3] == fb[:3] and fa[4:8] == fb[4:8]
        got = np.array([float(x) for x in fa[8:14]]); want = np.array([float(x) for x in fb[8:14]])
        assert np.all(np.abs(got - want) <= 1e-3), (a, b)
        assert abs(float(fa[14]) - float(fb[14])) <= 1e-3 and abs(float(fa[3]) - float(fb[3])) <= 1e-3


@pytest.mark.parametrize('path', GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_depth2pointsrgbpm_reproduces_reference_golden(path, tmp_path):
    """vis_utils.py:152-163 call sequence on the golden inputs; outputs vs what the reference produced."""
    inp, ref_fov, ref_all, ref_text = load_golden(path)
    calib = _calib(inp)
    H, W = inp['depth'].shape
    image1 = inp['image']
    image = image1[:H, :W]
    lidar = inp['lidar']
    pts_rect = calib.lidar_to_rect(lidar[:, 0:3])
    fov_flag = dropin.get_fov_flag(pts_rect, image.shape, calib)
    assert fov_flag.dtype == bool and np.array_equal(fov_flag, ref_fov)
    lidar = lidar[fov_flag]
    depth = np.array(inp['depth'], np.float32).reshape(H, W, 1)
    depth[0, 0, 0] = 0.0005                                    # below the threshold: must be zeroed in place (Q10)
    out = dropin.depth2pointsrgbpm(depth, image, image1, calib, lidar, inp['thing_classes'], torch.as_tensor(inp['masks']),
                                   torch.as_tensor(inp['classes']), torch.as_tensor(inp['scores']),
                                   fakes.FakeBoxes(inp['boxes2d']), None, '000007', plane=inp['plane'],
                                   label_root=str(tmp_path), quiet=True)
    assert depth[0, 0, 0] == 0.0
    assert out.dtype == np.float64 and out.shape == ref_all.shape
    nF = int(ref_fov.sum())
    assert np.array_equal(out[:nF], ref_all[:nF])
    np.testing.assert_allclose(out[nF:, :3], ref_all[nF:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(out[nF:, 3:], ref_all[nF:, 3:])
    text = open(os.path.join(str(tmp_path), '000007.txt')).read()
    _compare_lines(text.splitlines(True), ref_text.splitlines(True))
    assert len(dropin.last_forvis()) == len(ref_text.splitlines())


def test_calibration_methods_match_oracle():
    inp, _, _, _ = load_golden(GOLDEN[0])
    c, o = _calib(inp), restate.Calib(inp['P2'], inp['R0'], inp['V2C'])
    pts = inp['lidar'][:, :3]
    r = c.lidar_to_rect(pts)
    assert r.dtype == np.float32 and np.array_equal(r, o.lidar_to_rect(pts))
    (img, dep), (oi, od) = c.rect_to_img(r), o.rect_to_img(r)
    assert np.array_equal(img, oi) and np.array_equal(dep, od)
    (img2, dep2) = c.lidar_to_img(pts)
    assert np.array_equal(img2, oi) and np.array_equal(dep2, od)
    u = np.arange(0, 150, dtype=np.int64); v = (u * 7) % 90; d = np.linspace(1, 60, 150).astype(np.float32)
    a = c.img_to_rect(u, v, d)
    assert a.dtype == np.float64 and np.array_equal(a, o.img_to_rect(u, v, d))
    assert np.array_equal(c.rect_to_lidar(a), o.rect_to_lidar(a))
    p64 = np.array([[3.25, -1.5, 0.75], [10.0, 2.0, -1.0]])
    hom = np.hstack((p64, np.ones((2, 1), np.float32)))
    np.testing.assert_allclose(c.lidar_to_rect(p64), np.dot(hom, o.M43), rtol=1e-15, atol=1e-15)
    # CUDA tensors in -> CUDA tensors out
    t = c.lidar_to_rect(torch.from_numpy(pts).cuda())
    assert t.is_cuda and np.array_equal(t.cpu().numpy(), r)
    assert c.lidar_to_rect(np.zeros((0, 3), np.float32)).shape == (0, 3)


def test_plane_mask_and_sparse_depth_match_oracle():
    inp, _, _, _ = load_golden(GOLDEN[1])
    pts = inp['lidar'][:, :3]
    assert np.array_equal(dropin.above_plane(pts, inp['plane']), restate.above_plane(pts, inp['plane']))
    assert np.array_equal(dropin.above_plane(pts, inp['plane'], offset=0.2, only_range=None),
                          restate.above_plane(pts, inp['plane'], offset=0.2, only_range=None))
    H, W = inp['depth'].shape
    img = inp['image'][:H, :W]
    rgb, d = dropin.load_depth_input(_calib(inp), img, inp['lidar'])
    orgb, od = restate.load_depth_input(restate.Calib(inp['P2'], inp['R0'], inp['V2C']), img, inp['lidar'])
    assert rgb.dtype == np.uint8 and d.dtype == np.float64 and d.shape == (H, W, 1)
    assert np.array_equal(rgb, orgb) and np.array_equal(d, od)
    assert (d > 0).sum() > 50


def test_lshape_fitting_and_generate_anns_match_oracle(tmp_path):
    rng = np.random.default_rng(3)
    pts = np.concatenate([rng.normal(size=(150, 3)) * [0.9, 0.4, 0.3] + c for c in ((8, 2, 0.2), (8.5, 9.0, 0.1), (30, -8, 0.5))])
    rects, idsets = dropin.LShapeFitting().fitting(list(pts[:, 0]), list(pts[:, 1]))
    lab, nc = restate.cluster_adaptive(pts[:, 0], pts[:, 1])
    assert len(rects) == len(idsets) == nc
    for k in range(nc):
        assert idsets[k] == set(np.nonzero(lab == k)[0].tolist())
        a, b, c, th = restate.rectangle_search(list(pts[lab == k, 0]), list(pts[lab == k, 1]))
        np.testing.assert_allclose(rects[k].a, a, atol=1e-12); np.testing.assert_allclose(rects[k].c, c, atol=1e-9)
    inp, _, _, _ = load_golden(GOLDEN[0])
    calib = _calib(inp)
    buf = io.StringIO()
    forvis = dropin.GenerateAnns('Truck', pts, 1, 0.9, fakes.FakeBoxes(np.array([[1., 2., 30., 40.]], np.float32))[0],
                                 None, '000000', buf, [], calib=calib)
    recs, lines = restate.generate_anns('Truck', pts, 1, np.array([1., 2., 30., 40.], np.float32),
                                        restate.Calib(inp['P2'], inp['R0'], inp['V2C']))
    _compare_lines(buf.getvalue().splitlines(True), lines)
    assert len(forvis) == len(recs)
    for f, r in zip(forvis, recs):
        np.testing.assert_allclose(f, [r['x'], r['y'], r['z'], r['l'], r['w'], r['h'], r['ry']], atol=1e-3)
    assert dropin.GenerateAnns('Car', np.zeros((0, 3)), 0, 0.9, None, None, '0', buf, []) == []


def test_ball_query_matches_oracle():
    rng = np.random.default_rng(5)
    ref = rng.uniform(-5, 5, size=(2000, 3))
    q = np.concatenate([ref[:700] + rng.normal(size=(700, 3)) * 0.06, rng.uniform(-6, 6, size=(900, 3))])
    _, _, _, mask = dropin.BallQuery(torch.from_numpy(q), torch.from_numpy(ref), 0.1)
    assert np.array_equal(mask.cpu().numpy(), restate.ball_gate(q, ref, 0.1))


def test_save_depth_as_points_round_trip(tmp_path):
    """vis_utils.py:136-166 through real files: .bin / .png / calib .txt in, velodyne_depth/<idx>.npy + label out."""
    from PIL import Image
    from oracle import ref_harness
    inp, ref_fov, ref_all, ref_text = load_golden(GOLDEN[2])
    root = str(tmp_path)
    for d in ('image_2', 'velodyne', 'calib', 'labels'):
        os.makedirs(os.path.join(root, d))
    inp['lidar'].astype(np.float32).tofile(os.path.join(root, 'velodyne', '000003.bin'))
    Image.fromarray(inp['image'].astype(np.uint8)).save(os.path.join(root, 'image_2', '000003.png'))
    ref_harness.write_calib_file(os.path.join(root, 'calib', '000003.txt'), inp['P2'], inp['R0'], inp['V2C'])
    H, W = inp['depth'].shape
    seg = (inp['thing_classes'], torch.as_tensor(inp['masks']), torch.as_tensor(inp['classes']), torch.as_tensor(inp['scores']),
           fakes.FakeBoxes(inp['boxes2d']))
    dropin.save_depth_as_points(torch.from_numpy(inp['depth'].copy()).reshape(1, 1, H, W), 3, root, masks=seg, plane=inp['plane'],
                                label_root=os.path.join(root, 'labels'), crop=(H, W))
    out = np.load(os.path.join(root, 'velodyne_depth', '000003.npy'))
    assert out.dtype == np.float16 and out.shape == ref_all.shape
    assert np.array_equal(out, ref_all.astype(np.float16)) or \
        float((out == ref_all.astype(np.float16)).all(1).mean()) > 0.999   # 1e-5-relative coordinates may flip an fp16 ulp
    assert open(os.path.join(root, 'labels', '000003.txt')).read().count('\n') == ref_text.count('\n')
    loader = dropin.MyLoader(root)
    rgb, d = loader[0]
    assert len(loader) == 1 and rgb.shape[:2] == d.shape[:2]


def test_names_the_reference_callers_import_resolve():
    """`import dfu3d_b200.dropin as my_loader` must satisfy vis_utils.py:7-14 and kitti_loader.py:15."""
    for name in ('depth2pointsrgb', 'depth2pointsrgbp', 'depth2pointsrgbpm', 'depth2points', 'MyLoader', 'Calibration',
                 'get_calib_from_file', 'get_fov_flag', 'load_depth_input', 'GenerateAnns', 'LShapeFitting', 'BallQuery',
                 'estimate_plane', 'above_plane', 'calc_rect_contour', 'calc_cross_point', 'save_depth_as_points'):
        assert callable(getattr(dropin, name)), name
    xs, ys = dropin.calc_rect_contour([1.0, 0.0, 1.0, 0.0], [0.0, 1.0, 0.0, 1.0], [0.0, 0.0, 2.0, 3.0])
    assert len(xs) == 5 and len(ys) == 5 and xs[4] is None and ys[4] is None      # my_loader.py:687-688
    assert (xs[0], ys[0], xs[2], ys[2]) == (0.0, 0.0, 2.0, 3.0)


@pytest.mark.parametrize('path', GOLDEN[:2], ids=[os.path.basename(p)[:-4] for p in GOLDEN[:2]])
def test_earlier_variants_match_oracle(path):
    """depth2points / depth2pointsrgb / depth2pointsrgbp (my_loader.py:142-164,415-442) through the CUDA library vs the
    restatements that tests/test_oracle_vs_reference.py pins to the reference's own functions."""
    inp, ref_fov, ref_all, ref_text = load_golden(path)
    calib = _calib(inp)
    oc = restate.Calib(inp['P2'], inp['R0'], inp['V2C'])
    H, W = inp['depth'].shape
    img = inp['image'][:H, :W]
    d = lambda: np.array(inp['depth'], np.float32).reshape(H, W, 1)
    dd = d()
    got = dropin.depth2points(dd, calib)
    want = restate.depth2points(d(), oc)
    assert got.shape == want.shape and got.dtype == np.float64
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-9)
    assert float(dd[dd > 0].min()) >= 0.1                                       # thresholded in place like the reference
    got = dropin.depth2pointsrgb(d(), img, calib)
    want = restate.depth2pointsrgb(d(), img, oc)
    np.testing.assert_allclose(got[:, :3], want[:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(got[:, 3:], want[:, 3:])
    lidar = inp['lidar'][ref_fov]
    got = dropin.depth2pointsrgbp(d(), img, calib, lidar)
    want = restate.depth2pointsrgbp(d(), img, oc, lidar)
    assert got.shape == want.shape
    n = lidar.shape[0]
    assert np.array_equal(got[:n], want[:n])
    np.testing.assert_allclose(got[n:, :3], want[n:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(got[n:, 3:], want[n:, 3:])


def test_corners3d_to_img_boxes_matches_oracle():
    inp, _, _, _ = load_golden(GOLDEN[0])
    calib = _calib(inp)
    rng = np.random.default_rng(0)
    c3 = rng.uniform(-10, 10, size=(40, 8, 3)) + np.array([0, 0, 25.0])
    boxes, corners = calib.corners3d_to_img_boxes(c3)
    ob, oc = restate.corners3d_to_img_boxes(inp['P2'], c3)
    assert boxes.shape == (40, 4) and corners.shape == (40, 8, 2)
    np.testing.assert_allclose(corners, oc, rtol=1e-12, atol=1e-9)
    np.testing.assert_allclose(boxes, ob, rtol=1e-12, atol=1e-9)


def test_lshape_fitting_honours_its_attributes():
    rng = np.random.default_rng(2)
    ox = np.r_[rng.uniform(10, 14, 120), np.full(60, 14.0) + rng.normal(0, 0.02, 60)]
    oy = np.r_[np.full(120, 3.0) + rng.normal(0, 0.02, 120), rng.uniform(3, 5, 60)]
    f = dropin.LShapeFitting()
    r1, ids1 = f.fitting(ox, oy)
    f.dtheta_deg_for_serarch = 5.0                    # 17 angles instead of 89: the rectangle comes from the coarser table
    r5, ids5 = f.fitting(ox, oy)
    assert ids1 == ids5 and len(r1) == len(r5) == 1
    th5 = np.degrees(np.arctan2(r5[0].b[0], r5[0].a[0])) % 5.0
    assert min(th5, 5.0 - th5) < 1e-9
    f.criteria = f.Criteria.AREA
    with pytest.raises(NotImplementedError):
        f.fitting(ox, oy)
