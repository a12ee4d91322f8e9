"""GPU parity of the consumer-side row: modality split + range mask and hard voxelisation, against
oracle/consumer.py (first-party reference code for the split; restated spconv semantics for the voxeliser)."""
import numpy as np
import pytest

from dfu3d_b200 import ops
from oracle import consumer

pytestmark = pytest.mark.gpu
RANGE = [0, -40, -3, 70.4, 40, 1]


def cloud(seed, n, C=8):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-5, 80, size=(n, C)).astype(np.float32)
    c[:, 1] = rng.uniform(-45, 45, n)
    c[:, 2] = rng.uniform(-3.5, 1.5, n)
    c[:, -1] = rng.choice([1.0, 2.0, 0.0], size=n, p=[0.7, 0.28, 0.02])
    k = n // 3                                   # a dense patch: many points per voxel, some exactly on bin edges
    c[:k, 0] = rng.uniform(10, 12, k)
    c[:k, 1] = rng.uniform(-1, 1, k)
    c[:k:7, 0] = np.float32(10.0) + np.float32(0.05) * rng.integers(0, 40, len(c[:k:7]))
    c[5::11, 0] = np.float32(70.4)
    c[6::13, 1] = np.float32(-40.0)
    return c


@pytest.mark.parametrize('seed,n', [(0, 200000), (1, 1), (2, 777)])
def test_split_matches_reference_code(seed, n):
    c = cloud(seed, n)
    for rg in (None, RANGE):
        wl, wp = consumer.split_cloud(c, rg)
        gl, gp = ops.split_cloud(c, rg)
        assert np.array_equal(gl, wl) and np.array_equal(gp, wp)
    e = ops.split_cloud(c[:0])
    assert e[0].shape == (0, 8) and e[1].shape == (0, 8)


@pytest.mark.parametrize('seed,n,vs,mp,mv', [(3, 150000, [0.05, 0.05, 0.1], 5, 40000), (4, 60000, [0.16, 0.16, 4], 32, 16000),
                                             (5, 90000, [0.05, 0.05, 0.1], 5, 3000), (6, 3, [0.05, 0.05, 0.1], 5, 100)])
def test_voxelize_matches_oracle(seed, n, vs, mp, mv):
    pts = cloud(seed, n, C=4 if seed == 4 else 8)
    wv, wc, wn = consumer.voxelize(pts, vs, RANGE, mp, mv)
    gv, gc, gn = ops.voxelize(pts, vs, RANGE, mp, mv)
    assert gv.shape == wv.shape and gv.shape[0] > 0 or n == 3
    assert np.array_equal(gn, wn) and np.array_equal(gc, wc) and np.array_equal(gv, wv)
    if seed == 5:
        assert gv.shape[0] == mv                 # the max_voxels cap was hit
    # properties that hold at any size: every kept point lies in its voxel; counts bounded
    assert gn.max(initial=0) <= mp and gn.min(initial=1) >= 1


def test_voxelize_empty():
    v, c, k = ops.voxelize(np.zeros((0, 4), np.float32), [0.05, 0.05, 0.1], RANGE, 5, 100)
    assert v.shape == (0, 5, 4) and c.shape == (0, 3) and k.shape == (0,)
    out = np.full((10, 4), 500.0, np.float32)   # all outside the range
    v, c, k = ops.voxelize(out, [0.05, 0.05, 0.1], RANGE, 5, 100)
    assert v.shape[0] == 0
