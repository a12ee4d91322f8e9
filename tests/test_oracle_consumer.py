"""CPU: the vectorised voxeliser of the oracle (oracle/consumer.py over oracle/fakes.py:voxel_assign) against a literal
sequential restatement of the spconv v1 loop (points_to_voxel: dense coor_to_voxelidx, first-appearance numbering,
max_points / max_voxels cuts), and the modality split against the reference's own expressions."""
import numpy as np
import pytest

from oracle import consumer

RANGE = [0, -40, -3, 70.4, 40, 1]


def sequential_voxelize(points, voxel_size, coors_range, max_points, max_voxels):
    vs = np.asarray(voxel_size, np.float32)
    rg = np.asarray(coors_range, np.float32)
    grid = np.round((rg[3:] - rg[:3]) / vs).astype(np.int64)
    idx_of = {}
    voxels = np.zeros((max_voxels, max_points, points.shape[1]), points.dtype)
    coors = np.zeros((max_voxels, 3), np.int32)
    num = np.zeros(max_voxels, np.int32)
    nv = 0
    for p in points:
        c = np.floor((p[:3] - rg[:3]) / vs)
        if np.any(c < 0) or np.any(c >= grid):
            continue
        key = tuple(int(x) for x in c[::-1])          # (z, y, x)
        vid = idx_of.get(key, -1)
        if vid == -1:
            if nv >= max_voxels:
                continue
            vid = nv
            nv += 1
            idx_of[key] = vid
            coors[vid] = key
        if num[vid] < max_points:
            voxels[vid, num[vid]] = p
            num[vid] += 1
    return voxels[:nv], coors[:nv], num[:nv]


@pytest.mark.parametrize('seed,n,mp,mv', [(0, 4000, 5, 100000), (1, 4000, 3, 500), (2, 50, 1, 7)])
def test_vectorised_voxeliser_equals_sequential_loop(seed, n, mp, mv):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-2, 75, size=(n, 5)).astype(np.float32)
    pts[:, 1] = rng.uniform(-42, 42, n)
    pts[:, 2] = rng.uniform(-3.2, 1.2, n)
    pts[: n // 2, 0] = rng.uniform(10, 10.4, n // 2)       # a dense patch: several points per voxel
    pts[: n // 2, 1] = rng.uniform(0, 0.4, n // 2)
    pts[: n // 2, 2] = rng.uniform(-1, -0.6, n // 2)
    pts[::9, 0] = np.float32(10.0) + np.float32(0.05) * rng.integers(0, 8, len(pts[::9]))   # exactly on bin edges
    want = sequential_voxelize(pts, [0.05, 0.05, 0.1], RANGE, mp, mv)
    got = consumer.voxelize(pts, [0.05, 0.05, 0.1], RANGE, mp, mv)
    for a, b in zip(got, want):
        assert a.shape == b.shape and np.array_equal(a, b)
    assert got[2].max() == mp or n == 50


def test_split_is_the_reference_expression():
    rng = np.random.default_rng(3)
    c = rng.uniform(-5, 80, size=(5000, 8)).astype(np.float32)
    c[:, 1] = rng.uniform(-45, 45, 5000)
    c[:, -1] = rng.choice([1.0, 2.0, 0.0], size=5000)
    c[::5, 0] = np.float32(70.4)
    lidar, pseudo = consumer.split_cloud(c, RANGE)
    # dataset_virtual.py:272-273 then common_utils.py:78-81
    def ref(p):
        m = (p[:, 0] >= RANGE[0]) & (p[:, 0] <= RANGE[3]) & (p[:, 1] >= RANGE[1]) & (p[:, 1] <= RANGE[4])
        return p[m]
    assert np.array_equal(pseudo, ref(c[c[:, -1] == 1])) and np.array_equal(lidar, ref(c[c[:, -1] == 2]))
    assert lidar.shape[0] + pseudo.shape[0] < 5000
