"""TEST INFRASTRUCTURE ONLY — functional stand-ins for the third-party calls on the path.

The reference (tools/PENet/dataloaders/my_loader.py:12,25-33,569-599) calls into Open3D and
spconv, neither of which is installable here.  Their arithmetic is restated below from the
libraries' published semantics; parity for these pieces is therefore UNPINNED by the
reference.

Open3D (version not pinned by the reference: absent from requirements.txt / setup.py):
  * ``PointCloud.remove_radius_outlier(nb_points, radius)``  — KD-tree radius search on squared
    distance with strict ``<``; the result includes the query point; keep iff count > nb_points.
  * ``PointCloud.remove_statistical_outlier(nb_neighbors, std_ratio)`` — KNN(k) including self,
    per-point mean of sqrt(d2); cloud mean over points with mean > 0 divided by the number of
    points that returned any neighbour; Bessel-corrected std; keep iff 0 < mean_i < mu + ratio*std.
  * ``PointCloud.uniform_down_sample(every_k_points)`` — rows 0, k, 2k, ...
spconv v1 ``VoxelGeneratorV2`` / ``VoxelGenerator`` (first choices of my_loader.py:25-29):
  voxel_size / point_cloud_range are stored as float32; grid = round((hi-lo)/vsize);
  c_j = floor((p_j - lo_j) / vsize_j) evaluated in float64 on the float32 constants; a point
  with any c_j outside [0, grid_j) is dropped; voxels are numbered in first-appearance order;
  at most ``max_num_points`` points per voxel are kept in input order; a point that would open
  voxel number ``max_voxels`` or later is dropped.  Returned coordinates are (z, y, x).
"""
import numpy as np
from scipy.spatial import cKDTree


# --------------------------------------------------------------------------- Open3D
def radius_outlier_mask(points, nb_points, radius):
    """bool[n]: keep iff #{q : |p-q|^2 < r^2} > nb_points (self included), float64.

    Evaluated through the (nb_points+1)-th nearest neighbour: the count exceeds nb_points iff that neighbour's
    squared distance is below r^2.  The squared distance is recomputed as ((a-b)**2).sum(-1); queries whose
    value lies within 1e-9 (relative) of r^2 are re-done by brute force so that the KD-tree's own rounding can
    never decide a case."""
    pts = np.asarray(points, dtype=np.float64).reshape(-1, 3)
    n = pts.shape[0]
    if n == 0:
        return np.zeros(0, dtype=bool)
    r2 = float(radius) * float(radius)
    k = int(nb_points) + 1
    if k > n:
        return np.zeros(n, dtype=bool)
    tree = cKDTree(pts)
    _, idx = tree.query(pts, k=k, workers=-1)
    idx = idx.reshape(n, k)[:, -1]
    d2 = ((pts[idx] - pts) ** 2).sum(-1)
    keep = d2 < r2
    amb = np.nonzero(np.abs(d2 - r2) <= 1e-9 * r2)[0]
    for i in amb:
        dd = ((pts - pts[i]) ** 2).sum(-1)
        keep[i] = int((dd < r2).sum()) > nb_points
    return keep


def knn_mean_distance(points, k):
    """float64[n]: mean of the min(k, n) smallest distances (self, 0.0, included).

    Open3D (PointCloud::RemoveStatisticalOutliers) takes the squared distances of KDTreeFlann::SearchKNN, which
    arrive in ascending order, replaces each by its square root and adds them with std::accumulate, i.e.
    sequentially.  The neighbours come from the KD-tree; the squared distances are recomputed as
    ((a - b) ** 2).sum(-1) so that the tree's own rounding never enters the result."""
    pts = np.asarray(points, dtype=np.float64).reshape(-1, 3)
    n = pts.shape[0]
    if n == 0:
        return np.zeros(0)
    kk = min(int(k), n)
    tree = cKDTree(pts)
    out = np.empty(n)
    for a in range(0, n, 1 << 18):
        q = pts[a:a + (1 << 18)]
        _, idx = tree.query(q, k=kk, workers=-1)
        idx = idx.reshape(q.shape[0], kk)
        d = np.sqrt(((pts[idx] - q[:, None, :]) ** 2).sum(-1))
        d.sort(axis=1)
        acc = np.zeros(q.shape[0])
        for j in range(kk):
            acc = acc + d[:, j]
        out[a:a + q.shape[0]] = acc / kk
    return out


def statistical_outlier_mask(points, nb_neighbors, std_ratio, return_stats=False):
    mean = knn_mean_distance(points, nb_neighbors)
    n = mean.shape[0]
    if n == 0:
        return (np.zeros(0, dtype=bool), mean, 0.0, 0.0) if return_stats else np.zeros(0, dtype=bool)
    pos = mean > 0
    mu = mean[pos].sum() / n
    sq = ((mean[pos] - mu) ** 2).sum()
    std = np.sqrt(sq / (n - 1)) if n > 1 else 0.0
    thr = mu + std_ratio * std
    keep = pos & (mean < thr)
    if return_stats:
        return keep, mean, mu, std
    return keep


class _Vector3dVector(object):
    def __init__(self, arr):
        a = arr
        if hasattr(a, 'detach'):
            a = a.detach().cpu().numpy()
        self.arr = np.array(a, dtype=np.float64).reshape(-1, 3)

    def __array__(self, dtype=None, copy=None):
        return self.arr if dtype is None else self.arr.astype(dtype)

    def __len__(self):
        return self.arr.shape[0]


class FakePointCloud(object):
    def __init__(self):
        self._pts = np.zeros((0, 3), dtype=np.float64)

    @property
    def points(self):
        return _Vector3dVector(self._pts)

    @points.setter
    def points(self, v):
        self._pts = np.asarray(v, dtype=np.float64).reshape(-1, 3)

    def _sub(self, idx):
        out = FakePointCloud()
        out._pts = self._pts[idx]
        return out

    def uniform_down_sample(self, every_k_points):
        return self._sub(np.arange(0, self._pts.shape[0], int(every_k_points)))

    def remove_radius_outlier(self, nb_points, radius):
        keep = radius_outlier_mask(self._pts, nb_points, radius)
        ind = np.nonzero(keep)[0]
        return self._sub(ind), list(ind)

    def remove_statistical_outlier(self, nb_neighbors, std_ratio):
        keep = statistical_outlier_mask(self._pts, nb_neighbors, std_ratio)
        ind = np.nonzero(keep)[0]
        return self._sub(ind), list(ind)


# --------------------------------------------------------------------------- spconv v1
def voxel_assign(points, voxel_size, coors_range, max_points, max_voxels):
    """Returns (voxel_id[n] int64 (-1 = dropped), rank_in_voxel[n], n_voxels).

    Restates spconv v1 ``points_to_voxel`` (float32 constants, float64 point arithmetic).
    """
    pts = np.asarray(points)
    vs = np.asarray(voxel_size, dtype=np.float32)
    rg = np.asarray(coors_range, dtype=np.float32)
    grid = np.round((rg[3:] - rg[:3]) / vs).astype(np.int64)
    c = np.floor((pts[:, :3] - rg[:3]) / vs)          # float64 on float32 constants
    ok = np.all((c >= 0) & (c < grid), axis=1)
    c = np.where(ok[:, None], c, 0).astype(np.int64)
    lin = (c[:, 0] * grid[1] + c[:, 1]) * grid[2] + c[:, 2]
    n = pts.shape[0]
    voxel_id = np.full(n, -1, dtype=np.int64)
    rank = np.zeros(n, dtype=np.int64)
    idx_ok = np.nonzero(ok)[0]
    if idx_ok.size == 0:
        return voxel_id, rank, 0
    lin_ok = lin[idx_ok]
    uniq, first_pos, inv = np.unique(lin_ok, return_index=True, return_inverse=True)
    # number voxels by first appearance
    order = np.argsort(first_pos, kind='stable')
    vid_of_uniq = np.empty_like(order)
    vid_of_uniq[order] = np.arange(order.size)
    vid = vid_of_uniq[inv]
    # rank of each point inside its voxel, in input order
    srt = np.argsort(vid, kind='stable')
    vs_sorted = vid[srt]
    start = np.r_[0, np.nonzero(np.diff(vs_sorted))[0] + 1]
    seg_start = np.repeat(start, np.diff(np.r_[start, vs_sorted.size]))
    rk = np.empty(vid.size, dtype=np.int64)
    rk[srt] = np.arange(vid.size) - seg_start
    good = (vid < max_voxels) & (rk < max_points)
    voxel_id[idx_ok[good]] = vid[good]
    rank[idx_ok[good]] = rk[good]
    return voxel_id, rank, int(min(order.size, max_voxels))


class FakeVoxelGeneratorV2(object):
    """Stand-in for ``spconv.utils.VoxelGeneratorV2`` (my_loader.py:25-43)."""

    def __init__(self, voxel_size, point_cloud_range, max_num_points, max_voxels=20000, **kw):
        self.voxel_size = voxel_size
        self.point_cloud_range = point_cloud_range
        self.max_num_points = max_num_points
        self.max_voxels = max_voxels

    def generate(self, points, max_voxels=None):
        pts = np.asarray(points)
        vid, rank, nv = voxel_assign(pts, self.voxel_size, self.point_cloud_range,
                                     self.max_num_points, self.max_voxels)
        sel = np.nonzero(vid >= 0)[0]
        counts = np.bincount(vid[sel], minlength=nv).astype(np.int32)
        maxc = int(counts.max()) if nv else 0
        # dense [V, maxc, F] (the real generator allocates [V, max_num_points, F]; the reference
        # only ever reads voxel[:num_points], so the trailing zero rows are not needed)
        voxels = np.zeros((nv, max(maxc, 1), pts.shape[1]), dtype=pts.dtype)
        voxels[vid[sel], rank[sel]] = pts[sel]
        coords = np.zeros((nv, 3), dtype=np.int32)
        return voxels, coords, counts


class FakeBoxes(object):
    """detectron2 ``Boxes`` stand-in: only ``boxes[i].tensor[0]`` is read (my_loader.py:639)."""

    def __init__(self, arr):
        import torch
        self.tensor = torch.as_tensor(np.asarray(arr, dtype=np.float32)).reshape(-1, 4)

    def __getitem__(self, i):
        return FakeBoxes(self.tensor[i:i + 1].numpy())

    def __len__(self):
        return self.tensor.shape[0]
