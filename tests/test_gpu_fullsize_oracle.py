"""GPU vs ORACLE at BASELINE.json's stated sizes: every comparison below runs the CUDA path through
the C ABI and the oracle (oracle/restate.py, oracle/fakes.py) on the same full-size seeded inputs.

  config 2   KITTI-shaped frame, 64 x 1875 rays, 1242 x 375 image, crop 352 x 1216: S1 pixel indices, FOV / plane / bounds
             flags and per-instance labels bit-exact
  config 3   6 x 1600 x 900 depth maps back-projected without the height gate (4.0 M points with sky = 0, 8.64 M with the
             sky-free backdrop): radius (r = 0.2; nb = 1, 16) and SOR (k = 20, 2 sigma) keep masks, bit-exact outside the
             1e-6 m band the north star allows (band counts printed)
  config 4   one whole 6-camera x 10-sweep frame: cloud rows and boxes of every view vs restate.run_view
  a16        uniform_down_sample(every_k) + remove_radius_outlier vs oracle/fakes.FakePointCloud
"""
import numpy as np
import pytest
import torch

from dfu3d_b200 import synth, pipeline, ops
from oracle import restate, fakes
from tests import helpers
from tests.test_gpu_instances import compare_boxes

pytestmark = pytest.mark.gpu

BAND = 1e-6


# ----------------------------------------------------------------------------- config 2
@pytest.mark.parametrize('preset', ['kitti', 'nuscenes'])
def test_config2_kitti_full_frame_s1_bit_exact(preset):
    fr = synth.make_frame(2102, 'kitti', n_cams=1, sweeps=1, crop=(352, 1216), max_inst=16)
    N = fr['lidar'].shape[0]
    assert N > 100000 and fr['views'][0]['masks'].shape[1:] == (375, 1242)
    inp = synth.view_as_reference_inputs(fr, 0)
    fov, allp, lines, recs, dbg = restate.run_view(inp, return_debug=True, preset=preset)
    cfg = pipeline.FusionConfig.preset(preset)
    pipe = pipeline.FusionPipeline(0, config=cfg)
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, cloud_dtype=torch.float64, instances=False)
    pix, flags, label = helpers.emit_project(pipe, N)
    assert np.array_equal((flags & 1) != 0, fov)
    lid_fov = inp['lidar'][fov]
    rng = ((-30, 80), (-30, 80)) if preset == 'kitti' else ((-54, 54), (-54, 54))
    above = restate.above_plane(lid_fov[:, :3], inp['plane'], only_range=rng)
    assert np.array_equal(((flags[fov] & 2) != 0), above)
    sel = np.nonzero(fov)[0][above]
    assert np.array_equal(pix[sel, 0], dbg['px']) and np.array_equal(pix[sel, 1], dbg['py'])
    assert np.array_equal((flags[sel] & 4) != 0, dbg['inb'])
    n_inst = inp['masks'].shape[0]
    assert n_inst >= 1
    lab_gpu = label[sel][dbg['inb']]
    for j in range(n_inst):
        assert np.array_equal(((lab_gpu >> j) & 1).astype(bool), dbg['lab'][j]), 'instance %d labels' % j
    # pixels whose un-rounded coordinate sits within 1e-4 px of x.5 / an image edge may legitimately differ (north star);
    # none does: the comparison above is exact for every point.  Report how many were that close.
    calib = restate.Calib(inp['P2'], inp['R0'], inp['V2C'])
    uv, _ = calib.lidar_to_img(lid_fov[above][:, :3])
    frac = np.abs(uv - np.floor(uv) - 0.5)
    print('config 2 (%s): %d rays, %d in FOV, %d above plane, %d instances; %d coordinates within 1e-4 px of x.5 (all exact)'
          % (preset, N, int(fov.sum()), int(above.sum()), n_inst, int((frac < 1e-4).sum())))
    # the fused cloud of the full-size frame as well
    cloud = res.cloud.cpu().numpy()
    nF = int(fov.sum())
    assert cloud.shape == allp.shape
    assert np.array_equal(cloud[:nF], allp[:nF])
    np.testing.assert_allclose(cloud[nF:, :3], allp[nF:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(cloud[nF:, 3:], allp[nF:, 3:])


# ----------------------------------------------------------------------------- config 3
def _config3_cloud(backdrop):
    fr = synth.make_frame(3000, 'nuscenes', n_cams=6, sweeps=1, device='cuda', backdrop=backdrop)
    depth = torch.stack([v['depth'] for v in fr['views']])
    calibs = [(v['P2'], v['R0'], v['V2C']) for v in fr['views']]
    xyz, _, _ = ops.backproject(depth, calibs)
    return xyz


@pytest.mark.parametrize('backdrop,nmin', [(None, 3_900_000), (90.0, 8_600_000)])
def test_config3_radius_and_sor_full_size_vs_oracle(backdrop, nmin):
    xyz = _config3_cloud(backdrop)
    n = int(xyz.shape[0])
    assert n > nmin
    pts = xyz.cpu().numpy()
    for nb in ((1, 16) if backdrop is None else (1,)):
        got = ops.radius_filter(xyz, nb, 0.2).cpu().numpy()
        want = fakes.radius_outlier_mask(pts, nb, 0.2)
        diff = int((got != want).sum())
        print('config 3 (%d points) radius r=0.2 nb=%d: kept %.4f, mismatches %d' % (n, nb, want.mean(), diff))
        assert diff == 0
    keep, mean, st = ops.sor_filter(xyz, 20, 2.0, return_stats=True)
    wk, wm, mu, sd = fakes.statistical_outlier_mask(pts, 20, 2.0, return_stats=True)
    mean = mean.cpu().numpy()
    keep = keep.cpu().numpy()
    np.testing.assert_allclose(mean, wm, rtol=1e-13, atol=0)
    exact = float((mean == wm).mean())
    assert exact > 0.999, exact
    assert abs(st[0, 0] - mu) <= 1e-12 * mu and abs(st[0, 1] - sd) <= 1e-10 * sd
    near = np.abs(wm - (mu + 2.0 * sd)) < BAND
    print('config 3 (%d points) SOR k=20 2 sigma: kept %.4f, mean distances bit-exact %.6f, %d points within %.0e m of the '
          'threshold (excluded)' % (n, wk.mean(), exact, int(near.sum()), BAND))
    assert np.array_equal(keep[~near], wk[~near])


# ----------------------------------------------------------------------------- config 4
def test_config4_full_frame_all_views_vs_oracle():
    fr = synth.make_frame(4242, 'nuscenes', n_cams=6, sweeps=10, max_inst=8)
    pipe = pipeline.FusionPipeline(0)
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, cloud_dtype=torch.float64)
    cloud = res.cloud.cpu().numpy()
    assert fr['lidar'].shape[0] > 200000 and vb.depth.shape == (6, 900, 1600)
    nbox = ncmp = 0
    inst_base = 0
    for v in range(6):
        inp = synth.view_as_reference_inputs(fr, v)
        fov, allp, lines, recs, dbg = restate.run_view(inp, return_debug=True)
        rows = cloud[res.row_off[v]:res.row_off[v + 1]]
        nF = int(fov.sum())
        assert rows.shape == allp.shape, (v, rows.shape, allp.shape)
        assert np.array_equal(rows[:nF], allp[:nF])
        np.testing.assert_allclose(rows[nF:, :3], allp[nF:, :3], rtol=1e-5, atol=1e-9)
        assert np.array_equal(rows[nF:, 3:], allp[nF:, 3:])
        n_inst = inp['masks'].shape[0]
        for i, d in enumerate(dbg['inst']):
            got = res.seg_counts[inst_base + i]
            assert (got[0], got[1], got[2], got[3], got[4]) == (d['nL'], d['nP'], d['nL1'], d['nP1'], d['nP2']), (v, i, got, d)

        class R(object):
            pass
        sub = R()
        sub.boxes = res.boxes[res.boxes[:, 0] == v].copy()
        sub.boxes[:, 1] -= inst_base
        ncmp += compare_boxes(sub, recs)
        nbox += len(recs)
        got_lines = res.label_lines[v]
        assert len(got_lines) == len(lines)
        for a, b in zip(got_lines, lines):
            fa, fb = a.split(), b.split()
            assert fa[:3] == fb[:3] and fa[4:8] == fb[4:8]
            np.testing.assert_allclose(np.array(fa[8:], float), np.array(fb[8:], float), atol=1e-3)
        inst_base += n_inst
    print('config 4: 6 views, %d cloud rows, %d boxes (%d compared to 1e-3 m / 1e-3 rad, the rest near-ties)'
          % (cloud.shape[0], nbox, ncmp))
    assert nbox >= 6 and ncmp >= nbox // 2


# ----------------------------------------------------------------------------- a16: uniform_down_sample + radius
@pytest.mark.parametrize('every_k', [2, 4, 8, 16])
def test_uniform_down_sample_then_radius_matches_oracle(every_k):
    rng = np.random.default_rng(100 + every_k)
    segs = [rng.normal(size=(40000, 3)) * np.array([4.0, 3.0, 0.6]), rng.uniform(-6, 6, size=(9001, 3)),
            rng.normal(size=(every_k + 1, 3)), rng.normal(size=(1, 3))]
    pts = np.concatenate(segs)
    so = np.cumsum([0] + [s.shape[0] for s in segs])
    nb, r = 10, 0.6            # the Pedestrian schedule of my_loader_KittiPandasetWaymo.py:806-821
    got = ops.radius_filter(pts, nb, r, seg_off=so, every_k=every_k)
    want = np.zeros(pts.shape[0], bool)
    for s, seg in enumerate(segs):
        pc = fakes.FakePointCloud()
        pc.points = seg
        ds = pc.uniform_down_sample(every_k)
        _, ind = ds.remove_radius_outlier(nb_points=nb, radius=r)
        want[so[s] + np.arange(0, seg.shape[0], every_k)[np.asarray(ind, np.int64)]] = True
    assert np.array_equal(got, want), (int(got.sum()), int(want.sum()))
    assert 0 < want.sum() < pts.shape[0] // every_k + 4
