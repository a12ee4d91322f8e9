"""GPU parity of S1 (projection + label inheritance) and S2 (densification + angular dedup) through the
C ABI, against the oracle restatement on the same seeded inputs."""
import numpy as np
import pytest
import torch

from dfu3d_b200 import synth, pipeline, _lib
from tests import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('seed,shape,scale', [(3, 'nuscenes', 0.25), (4, 'nuscenes', 0.25), (11, 'kitti', 0.5)])
def test_project_and_densify_match_oracle(seed, shape, scale):
    fr = helpers.small_frame(seed, shape, scale=scale)
    inp, (fov, allp, lines, recs, dbg) = helpers.oracle_view(fr)
    pipe = pipeline.FusionPipeline(0)
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, cloud_dtype=torch.float64, instances=False)
    N = fr['lidar'].shape[0]
    pix, flags, label = helpers.emit_project(pipe, N)
    # --- S1: FOV flag, plane flag, pixel indices, labels: bit-exact
    assert np.array_equal((flags & 1) != 0, fov)
    nF = int(fov.sum())
    assert res.counts_project[0, 0] == nF
    calib = __import__('oracle.restate', fromlist=['x']).Calib(inp['P2'], inp['R0'], inp['V2C'])
    lid_fov = inp['lidar'][fov]
    from oracle import restate
    above = restate.above_plane(lid_fov[:, :3], inp['plane'])
    assert np.array_equal(((flags[fov] & 2) != 0), above)
    assert res.counts_project[0, 1] == int(above.sum())
    sel = np.nonzero(fov)[0][above]
    assert np.array_equal(pix[sel, 0], dbg['px'])
    assert np.array_equal(pix[sel, 1], dbg['py'])
    assert np.array_equal((flags[sel] & 4) != 0, dbg['inb'])
    n_inst = inp['masks'].shape[0]
    lab_gpu = label[sel][dbg['inb']]
    for j in range(n_inst):
        assert np.array_equal(((lab_gpu >> j) & 1).astype(bool), dbg['lab'][j]), 'instance %d labels' % j
    # --- S2: cloud
    cloud = res.cloud.cpu().numpy()
    assert cloud.shape == allp.shape, (cloud.shape, allp.shape, res.counts_densify)
    assert np.array_equal(cloud[:nF], allp[:nF])
    nv = int(res.counts_densify[0, 2])
    rz, ry, xz, xy = helpers.emit_voxels(pipe, nv)
    exact_pix = float((rz == dbg['rep_z_pix']).mean())
    exact_pix_y = float((ry == dbg['rep_y_pix']).mean())
    exact_xyz = float((cloud[nF:] == allp[nF:]).all(1).mean())
    print('voxels', nv, 'rep-z pixel exact', exact_pix, 'rep-y', exact_pix_y, 'rows bit-exact', exact_xyz)
    # representatives: index-exact (integer work)
    assert np.array_equal(rz, dbg['rep_z_pix'])
    assert np.array_equal(ry, dbg['rep_y_pix'])
    # coordinates: 1e-5 relative (north star); colours / flags exact
    np.testing.assert_allclose(cloud[nF:, :3], allp[nF:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(cloud[nF:, 3:], allp[nF:, 3:])


def test_kitti_preset_matches_oracle():
    """The 'kitti' literal set (my_loader_KittiPandasetWaymo.py:470,709): plane-gate range (-30, 80)^2, pixel bounds
    1216x352, no GenerateAnns; a sweep that reaches beyond x = 54 m and behind x = -30 m makes the ranges matter."""
    fr = helpers.small_frame(12, 'kitti', scale=0.5)
    inp = synth.view_as_reference_inputs(fr, 0)
    from oracle import restate
    fov, allp, lines, recs, dbg = restate.run_view(inp, return_debug=True, preset='kitti')
    fov_n, allp_n, _, _ = restate.run_view(inp, preset='nuscenes')
    pipe = pipeline.FusionPipeline(0, config=pipeline.FusionConfig.preset('kitti'))
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, cloud_dtype=torch.float64)
    assert res.boxes is None and lines == []
    N = fr['lidar'].shape[0]
    pix, flags, label = helpers.emit_project(pipe, N)
    lid_fov = inp['lidar'][fov]
    above = restate.above_plane(lid_fov[:, :3], inp['plane'], only_range=((-30, 80), (-30, 80)))
    assert np.array_equal(((flags[fov] & 2) != 0), above)
    assert not np.array_equal(above, restate.above_plane(lid_fov[:, :3], inp['plane'])), 'ranges must differ on this sweep'
    sel = np.nonzero(fov)[0][above]
    assert np.array_equal((flags[sel] & 4) != 0, dbg['inb'])
    cloud = res.cloud.cpu().numpy()
    assert cloud.shape == allp.shape
    nF = int(fov.sum())
    assert np.array_equal(cloud[:nF], allp[:nF])
    np.testing.assert_allclose(cloud[nF:, :3], allp[nF:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(cloud[nF:, 3:], allp[nF:, 3:])


@pytest.mark.parametrize('seed,maxv', [(3, 700), (4, 1), (5, 2500)])
def test_voxel_count_limit_keeps_the_first_voxels(seed, maxv):
    """The voxeliser's max_num_voxels (my_loader.py:74, 10^6 in the reference): a view keeps its first max_voxels angular
    voxels in first-appearance order, the points of the others are dropped from the cloud and from the instances' pseudo
    points.  A small limit on small frames; two views, so that the lists are re-packed across views."""
    fr = helpers.small_frame(seed, 'nuscenes', scale=0.25, n_cams=2)
    from oracle import restate
    free = pipeline.FusionPipeline(0).run(pipeline.batch_from_frames([fr], torch.device('cuda', 0)), cloud_dtype=torch.float64)
    pipe = pipeline.FusionPipeline(0, config=pipeline.FusionConfig(max_voxels=maxv))      # same context: run it last
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, cloud_dtype=torch.float64, instances=True)
    assert int(free.counts_densify[:, 2].max()) > maxv                      # the limit bites
    rz, ry, xz, xy = helpers.emit_voxels(pipe, int(res.counts_densify[:, 2].sum()))
    off = np.concatenate([[0], np.cumsum(res.counts_densify[:, 2])])
    inst0 = 0
    for v in range(2):
        inp = synth.view_as_reference_inputs(fr, v)
        fov, allp, lines, recs, dbg = restate.run_view(inp, return_debug=True, max_voxels=maxv)
        nv = int(res.counts_densify[v, 2])
        assert nv == min(maxv, int(free.counts_densify[v, 2])) == dbg['rep_z_pix'].shape[0]
        assert np.array_equal(rz[off[v]:off[v + 1]], dbg['rep_z_pix']) and np.array_equal(ry[off[v]:off[v + 1]], dbg['rep_y_pix'])
        cloud = res.cloud[int(res.row_off[v]):int(res.row_off[v + 1])].cpu().numpy()
        nF = int(fov.sum())
        assert cloud.shape == allp.shape
        np.testing.assert_allclose(cloud[nF:, :3], allp[nF:, :3], rtol=1e-5, atol=1e-9)
        for i, d in enumerate(dbg['inst']):
            assert tuple(res.seg_counts[inst0 + i][:2]) == (d['nL'], d['nP'])
        inst0 += len(dbg['inst'])


@pytest.mark.parametrize('seed,mpv', [(3, 12), (4, 10), (5, 16)])
def test_overflowing_voxels_keep_the_first_points_in_input_order(seed, mpv):
    """the reference's voxeliser keeps the first max_num_points_per_voxel points of an angular voxel in
    input order and la_sampling2 / 20 take the arg-min over those only (my_loader.py:69-83,253-257,268-272).  A small
    max_per_voxel makes thousands of voxels overflow on an ordinary view; representatives must stay index-exact."""
    fr = synth.make_frame(seed, 'nuscenes', n_cams=1, max_inst=4)        # full resolution: ~6 pixels per angular voxel
    inp = synth.view_as_reference_inputs(fr, 0)
    from oracle import restate
    fov, allp, lines, recs, dbg = restate.run_view(inp, return_debug=True, max_per_voxel=mpv)
    assert dbg['overflow'] > 50
    pipe = pipeline.FusionPipeline(0, config=pipeline.FusionConfig(max_per_voxel=mpv))
    vb = pipeline.batch_from_frames([fr], pipe.device)
    res = pipe.run(vb, cloud_dtype=torch.float64, instances=True)
    assert int(res.counts_densify[0, 3]) == dbg['overflow']
    nv = int(res.counts_densify[0, 2])
    rz, ry, xz, xy = helpers.emit_voxels(pipe, nv)
    assert np.array_equal(rz, dbg['rep_z_pix']) and np.array_equal(ry, dbg['rep_y_pix'])
    cloud = res.cloud.cpu().numpy()
    nF = int(fov.sum())
    assert cloud.shape == allp.shape
    np.testing.assert_allclose(cloud[nF:, :3], allp[nF:, :3], rtol=1e-5, atol=1e-9)
    for i, d in enumerate(dbg['inst']):
        assert tuple(res.seg_counts[i][:2]) == (d['nL'], d['nP'])
