"""CPU: the travelling oracle (oracle/restate.py) against the committed golden fixtures, which
tests/golden/make_golden.py produced by EXECUTING the reference (tools/PENet/dataloaders/my_loader.py:502-617,
vis_utils.py:152-154) on seeded synthetic views."""
import glob
import os

import numpy as np
import pytest

from oracle import restate

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', '[nk]*_[a-z].npz')))


def load_golden(path):
    d = np.load(path, allow_pickle=False)
    inp = {k: d[k] for k in ('lidar', 'P2', 'R0', 'V2C', 'depth', 'classes', 'scores', 'boxes2d', 'plane')}
    inp['image'] = d['image'].astype(np.int32)
    inp['masks'] = d['masks'].astype(np.float32)
    inp['thing_classes'] = [str(s) for s in d['thing_classes']]
    return inp, d['ref_fov'], d['ref_all_points'], str(d['ref_label_text'])


def test_fixtures_present():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize('path', GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_restatement_reproduces_reference_outputs(path):
    inp, ref_fov, ref_all, ref_text = load_golden(path)
    fov, allp, lines, recs = restate.run_view(inp)
    assert np.array_equal(fov, ref_fov)                       # vis_utils.py:152-154
    assert allp.shape == ref_all.shape
    nF = int(fov.sum())
    assert np.array_equal(allp[:nF], ref_all[:nF])            # LiDAR rows: bit-exact
    # pseudo rows: same voxel representatives in the same (first-appearance) order; coordinates 1e-5 relative
    np.testing.assert_allclose(allp[nF:, :3], ref_all[nF:, :3], rtol=1e-5, atol=1e-9)
    assert np.array_equal(allp[nF:, 3:], ref_all[nF:, 3:])
    ref_lines = ref_text.splitlines(True)
    assert len(lines) == len(ref_lines)
    for a, b in zip(lines, ref_lines):
        fa, fb = a.split(), b.split()
        assert fa[:3] == fb[:3] and fa[4:8] == fb[4:8]        # class, truncation/occlusion, 2-D box text
        got = np.array([float(x) for x in fa[8:14]])
        want = np.array([float(x) for x in fb[8:14]])
        assert np.all(np.abs(got - want) <= 1e-3), (a, b)     # h w l x y z   (1e-3 m)
        assert abs(float(fa[14]) - float(fb[14])) <= 1e-3 and abs(float(fa[3]) - float(fb[3])) <= 1e-3   # ry, alpha
