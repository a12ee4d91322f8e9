"""GPU parity of the stand-alone back-projection (S2) against the oracle: same rows, same order, float64 coordinates
within 1e-5 relative (north star; observed bit-exact because the operation order is the reference's)."""
import numpy as np
import pytest

from dfu3d_b200 import ops
from oracle import restate
from tests import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('seed,shape,scale', [(3, 'nuscenes', 0.25), (11, 'kitti', 0.5)])
def test_backproject_matches_oracle(seed, shape, scale):
    fr = helpers.small_frame(seed, shape, scale=scale)
    vw = fr['views'][0]
    depth = vw['depth'].cpu().numpy()
    calib = restate.Calib(vw['P2'], vw['R0'], vw['V2C'])
    d0 = depth.copy()
    rows, cols, want = restate.backproject(d0, calib)
    xyz, pix, counts = ops.backproject(depth, [(vw['P2'], vw['R0'], vw['V2C'])])
    assert counts.tolist() == [want.shape[0]]
    assert np.array_equal(pix, rows * depth.shape[1] + cols)
    np.testing.assert_allclose(xyz, want, rtol=1e-5, atol=1e-9)
    print('rows', want.shape[0], 'bit-exact fraction', float((xyz == want).all(1).mean()))
    # height gate (my_loader.py:540)
    xyz2, pix2, _ = ops.backproject(depth, [(vw['P2'], vw['R0'], vw['V2C'])], max_z=1.0)
    low = want[:, 2] < 1.0
    assert np.array_equal(pix2, pix[low]) and np.array_equal(xyz2, xyz[low])


def test_backproject_empty_and_batch():
    fr = helpers.small_frame(4, 'nuscenes', scale=0.25)
    vw = fr['views'][0]
    c = (vw['P2'], vw['R0'], vw['V2C'])
    depth = vw['depth'].cpu().numpy()
    z = np.zeros_like(depth)
    xyz, pix, counts = ops.backproject(np.stack([z, depth, z]), [c, c, c])
    one, pix1, _ = ops.backproject(depth, [c])
    assert counts.tolist() == [0, one.shape[0], 0] and np.array_equal(xyz, one) and np.array_equal(pix, pix1)
