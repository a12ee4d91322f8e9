"""TEST INFRASTRUCTURE ONLY — numpy restatement of the consumer side of the 8-column cloud.

  split_cloud   pcdet/datasets/dataset_virtual.py:270-296 (`points[:, -1] == 1` pseudo rows, `== 2` LiDAR rows) and
                pcdet/datasets/processor/data_processor_virtual.py:69-77 with pcdet/utils/common_utils.py:78-81
                (mask_points_by_range: x, y only, both ends inclusive) — first-party code, executed as written.
  voxelize      data_processor_virtual.py:133-185 -> VoxelGeneratorWrapper.generate (:16-58) -> spconv
                (third party, absent: semantics restated in oracle/fakes.py:voxel_assign; parity UNPINNED).
Never imported by the product (dfu3d_b200/)."""
import numpy as np

from . import fakes


def split_cloud(cloud, point_cloud_range=None):
    cloud = np.asarray(cloud)
    pseudo = cloud[cloud[:, -1] == 1]
    lidar = cloud[cloud[:, -1] == 2]
    if point_cloud_range is not None:
        r = point_cloud_range
        out = []
        for p in (lidar, pseudo):
            mask = (p[:, 0] >= r[0]) & (p[:, 0] <= r[3]) & (p[:, 1] >= r[1]) & (p[:, 1] <= r[4])
            out.append(p[mask])
        lidar, pseudo = out
    return lidar, pseudo


def voxelize(points, voxel_size, point_cloud_range, max_points, max_voxels):
    """-> voxels [V, max_points, C] (zero padded), coordinates [V, 3] (z, y, x) int32, num_points [V] int32."""
    pts = np.asarray(points)
    vid, rank, nv = fakes.voxel_assign(pts, voxel_size, point_cloud_range, max_points, max_voxels)
    vs = np.asarray(voxel_size, dtype=np.float32)
    rg = np.asarray(point_cloud_range, dtype=np.float32)
    sel = np.nonzero(vid >= 0)[0]
    voxels = np.zeros((nv, max_points, pts.shape[1]), dtype=pts.dtype)
    voxels[vid[sel], rank[sel]] = pts[sel]
    num = np.bincount(vid[sel], minlength=nv).astype(np.int32)
    coords = np.zeros((nv, 3), dtype=np.int32)
    first = sel[rank[sel] == 0]
    c = np.floor((pts[first, :3] - rg[:3]) / vs).astype(np.int32)
    coords[vid[first]] = c[:, ::-1]
    return voxels, coords, num
