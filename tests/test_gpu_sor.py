"""GPU parity of the statistical outlier filter (S4) through the C ABI against oracle/fakes.py
(Open3D remove_statistical_outlier semantics; parity UNPINNED for Open3D itself)."""
import numpy as np
import pytest

from dfu3d_b200 import ops
from oracle import fakes

pytestmark = pytest.mark.gpu

BOUNDARY = 1e-6      # north star: points within 1e-6 m of the threshold are reported separately


def clouds(seed, n):
    rng = np.random.default_rng(seed)
    # a depth-map-like surface cloud (spacing grows with range), a dense blob, and salt outliers far from anything
    u = rng.uniform(-0.6, 0.6, size=n); v = rng.uniform(-0.3, 0.1, size=n)
    d = rng.uniform(2.0, 70.0, size=n) ** 1.0
    surf = np.stack([d, d * u, d * v + 0.02 * rng.normal(size=n)], 1)
    blob = rng.normal(size=(n // 4, 3)) * 0.05 + np.array([5.0, 1.0, -1.0])
    salt = rng.uniform(-80, 80, size=(max(n // 100, 3), 3))
    pts = np.concatenate([surf, blob, salt])
    return pts[rng.permutation(pts.shape[0])]


def check(pts, seg_off, k, ratio):
    keep, mean, stats = ops.sor_filter(pts, k, ratio, seg_off=seg_off, return_stats=True)
    so = np.array([0, pts.shape[0]]) if seg_off is None else np.asarray(seg_off)
    nb = 0
    for s in range(so.shape[0] - 1):
        a, b = int(so[s]), int(so[s + 1])
        wk, wm, mu, sd = fakes.statistical_outlier_mask(pts[a:b], k, ratio, return_stats=True)
        np.testing.assert_allclose(mean[a:b], wm, rtol=1e-13, atol=0)
        exact = float((mean[a:b] == wm).mean()) if b > a else 1.0
        assert exact > 0.999, exact                      # same neighbours, same summation order
        assert abs(stats[s, 0] - mu) <= 1e-12 * max(mu, 1e-300) + 1e-300
        assert abs(stats[s, 1] - sd) <= 1e-10 * max(sd, 1e-300) + 1e-300
        near = np.abs(wm - (mu + ratio * sd)) < BOUNDARY
        nb += int(near.sum())
        assert np.array_equal(keep[a:b][~near], wk[~near])
    return nb


@pytest.mark.parametrize('seed,n,k,ratio', [(0, 20000, 20, 2.0), (1, 3000, 30, 0.3), (2, 60000, 20, 2.0), (3, 500, 5, 1.0),
                                            (4, 2000, 64, 1.5), (5, 4000, 1, 2.0)])
def test_sor_matches_oracle(seed, n, k, ratio):
    pts = clouds(seed, n)
    near = check(pts, None, k, ratio)
    print('points', pts.shape[0], 'within %.0e of the threshold (excluded):' % BOUNDARY, near)


def test_sor_segments_and_edge_cases():
    rng = np.random.default_rng(9)
    segs = [clouds(10, 1500), rng.normal(size=(7, 3)), np.zeros((0, 3)), rng.normal(size=(1, 3)),
            np.repeat(rng.normal(size=(3, 3)), 40, axis=0),        # duplicates: mean distance 0 -> dropped
            rng.normal(size=(2, 3))]
    pts = np.concatenate(segs)
    so = np.cumsum([0] + [s.shape[0] for s in segs])
    check(pts, so, 20, 2.0)
    keep = ops.sor_filter(np.zeros((0, 3)), 20, 2.0)
    assert keep.shape == (0,)


def test_sor_rejects_bad_k():
    with pytest.raises(RuntimeError):
        ops.sor_filter(np.zeros((10, 3)), 65, 2.0)
    with pytest.raises(RuntimeError):
        ops.sor_filter(np.full((10, 3), np.nan), 5, 2.0, return_stats=True)


def test_sor_idempotent_structure_full_size_property():
    """Size-independent property at a larger size: the mask is invariant under a permutation of the input."""
    pts = clouds(21, 300000)
    keep, mean, st = ops.sor_filter(pts, 20, 2.0, return_stats=True)
    perm = np.random.default_rng(1).permutation(pts.shape[0])
    keep2, mean2, st2 = ops.sor_filter(pts[perm], 20, 2.0, return_stats=True)
    assert np.array_equal(mean[perm], mean2)
    near = np.abs(mean2 - (st2[0, 0] + 2.0 * st2[0, 1])) < BOUNDARY
    assert np.array_equal(keep[perm][~near], keep2[~near])
    assert 0.5 < keep.mean() < 1.0
