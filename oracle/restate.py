"""TEST INFRASTRUCTURE ONLY — numpy/scipy restatement of the DFU3D fusion / pseudo-box path.

Travels to the GPU box (``/root/reference`` does not).  Every function cites the reference
lines it restates (paths relative to /root/reference/tools/PENet).  Pinned against the
reference's own code by tests/test_oracle_vs_reference.py (run where the reference exists)
and against tests/golden/*.npz (generated from the reference by tests/golden/make_golden.py).
Third-party pieces (Open3D, spconv, PIL) follow oracle/fakes.py — parity UNPINNED for those.

Never imported by dfu3d_b200.
"""
import numpy as np
from scipy.spatial import cKDTree
from scipy.sparse import coo_matrix
from scipy.sparse.csgraph import connected_components

from . import fakes

NUSC_VEHICLES = ('Car', 'Truck', 'Construction_vehicle', 'Bus', 'Trailer')                # my_loader.py:581
NUSC_SMALL = ('Pedestrian', 'Bicycle', 'Motorcycle', 'Barrier', 'Traffic_cone')          # my_loader.py:591
NUSC_NAMES = ['Car', 'Truck', 'Construction_vehicle', 'Bus', 'Trailer', 'Barrier', 'Motorcycle', 'Bicycle',
              'Pedestrian', 'Traffic_cone']                                              # my_loader.py:634


# ----------------------------------------------------------------------------- calibration
def _sgemm_fma(h, M):
    """float32 ``np.dot(h, M)`` as OpenBLAS evaluates it for K<=4: one accumulator per output,
    k ascending, fused multiply-add, first term a plain rounded product.  Written out so that the
    oracle does not depend on the BLAS build of the machine it runs on (checked bit-equal to
    np.dot on the reference container by tests/test_oracle_vs_reference.py).  The fused multiply-add
    is emulated as float32(float64(acc) + float64(a) * float64(b)): the float64 product of two float32
    values is exact, the float64 sum is rounded once and then again to float32 (double rounding, a
    difference from a true fmaf in about 2^-29 of the cases; none on the tested inputs)."""
    h = np.asarray(h, np.float32)
    M = np.asarray(M, np.float32)
    out = np.empty((h.shape[0], M.shape[1]), np.float32)
    h64 = h.astype(np.float64)
    for j in range(M.shape[1]):
        acc = (h64[:, 0] * np.float64(M[0, j])).astype(np.float32)
        for k in range(1, M.shape[0]):
            acc = (h64[:, k] * np.float64(M[k, j]) + acc.astype(np.float64)).astype(np.float32)
        out[:, j] = acc
    return out


class Calib(object):
    """calibration_kitti.py:62-79 plus the derived constant matrices of :89-123."""

    def __init__(self, P2, R0, V2C):
        self.P2 = np.asarray(P2, np.float32).reshape(3, 4)
        self.R0 = np.asarray(R0, np.float32).reshape(3, 3)
        self.V2C = np.asarray(V2C, np.float32).reshape(3, 4)
        self.cu, self.cv = self.P2[0, 2], self.P2[1, 2]
        self.fu, self.fv = self.P2[0, 0], self.P2[1, 1]
        self.tx = self.P2[0, 3] / (-self.fu)
        self.ty = self.P2[1, 3] / (-self.fv)
        self.M43 = _sgemm_fma(np.ascontiguousarray(self.V2C.T), np.ascontiguousarray(self.R0.T))  # :110
        self.P2T = np.ascontiguousarray(self.P2.T)
        # :95-101  (float32 product and float32 inverse)
        R0e = np.zeros((4, 4), np.float32); R0e[:3, :3] = self.R0; R0e[3, 3] = 1
        V2Ce = np.zeros((4, 4), np.float32); V2Ce[:3, :] = self.V2C; V2Ce[3, 3] = 1
        self.inv44 = np.linalg.inv(np.dot(R0e, V2Ce).T)

    def lidar_to_rect(self, pts):                      # :104-112
        pts = np.asarray(pts, np.float32)
        hom = np.hstack((pts, np.ones((pts.shape[0], 1), np.float32)))
        return _sgemm_fma(hom, self.M43)

    def rect_to_img(self, rect):                       # :114-123
        rect = np.asarray(rect, np.float32)
        hom = np.hstack((rect, np.ones((rect.shape[0], 1), np.float32)))
        h2 = _sgemm_fma(hom, self.P2T)
        img = (h2[:, 0:2].T / hom[:, 2]).T
        depth = h2[:, 2] - self.P2T[3, 2]
        return img, depth

    def lidar_to_img(self, pts):                       # :125-132
        return self.rect_to_img(self.lidar_to_rect(pts))

    def img_to_rect(self, u, v, d):                    # :134-144 (int64 - f32 scalar -> float64)
        x = ((u - self.cu) * d) / self.fu + self.tx
        y = ((v - self.cv) * d) / self.fv + self.ty
        return np.concatenate((x.reshape(-1, 1), y.reshape(-1, 1), d.reshape(-1, 1)), axis=1)

    def rect_to_lidar(self, rect):                     # :89-102
        hom = np.hstack((rect, np.ones((rect.shape[0], 1), np.float32)))
        return np.dot(hom, self.inv44)[:, 0:3]


def get_fov_flag(rect, img_shape, calib):              # vis_utils.py:108-123 == my_loader.py:85-100
    img, depth = calib.rect_to_img(rect)
    f = (img[:, 0] >= 0) & (img[:, 0] < img_shape[1]) & (img[:, 1] >= 0) & (img[:, 1] < img_shape[0])
    return f & (depth >= 0)


def above_plane(ptc, plane, offset=0.05, only_range=((-54, 54), (-54, 54))):   # my_loader.py:471-484
    plane = np.asarray(plane, np.float64)
    d = ptc @ plane[:3] + plane[3]
    d = d / np.sqrt((plane[:3] ** 2).sum())
    mask = d < offset
    if only_range is not None:
        rm = (ptc[:, 0] < only_range[0][1]) & (ptc[:, 0] > only_range[0][0]) & \
             (ptc[:, 1] < only_range[1][1]) & (ptc[:, 1] > only_range[1][0])
        mask = mask & rm
    return ~mask


def load_depth_input(calib, image, points):            # my_loader.py:102-122
    rect = calib.lidar_to_rect(points[:, 0:3])
    points = points[get_fov_flag(rect, image.shape, calib)]
    rect = calib.lidar_to_rect(points[:, 0:3])
    img, depth = calib.rect_to_img(rect)
    val = (img[:, 0] >= 0) & (img[:, 1] >= 0) & (img[:, 0] < image.shape[1]) & (img[:, 1] < image.shape[0])
    px = img[val].astype(np.int32)
    new_im = np.zeros(shape=image.shape[0:2])
    new_im[px[:, 1], px[:, 0]] = depth[val]
    return np.array(image, dtype='uint8'), np.expand_dims(new_im, -1)


# ----------------------------------------------------------------------------- label inheritance
def inherit_labels(lidar_mask, calib, masks, bounds=(1600, 900)):
    """my_loader.py:518-530 including quirk Q1 (labels of in-bounds points are index-aligned with the
    FIRST len(list) LiDAR rows).  ``masks`` [n,Hm,Wm] {0,1}; PIL getpixel((x, y)) == mask[int(y), int(x)].
    Returns (lidar_mask possibly truncated, label bool [n, len], pix_x, pix_y int64 (all points), inb)."""
    img, _ = calib.lidar_to_img(lidar_mask[:, :3])
    pr = np.round(img)
    inb = (pr[:, 0] >= 0) & (pr[:, 0] < bounds[0]) & (pr[:, 1] >= 0) & (pr[:, 1] < bounds[1])
    px = pr[inb, 0].astype(np.int64)
    py = pr[inb, 1].astype(np.int64)
    n_keep = min(int(inb.sum()), lidar_mask.shape[0])
    masks = np.asarray(masks)
    if masks.shape[0] > 0:
        lidar_mask = lidar_mask[:n_keep]
        lab = masks[:, py, px] > 0
    else:
        lab = np.zeros((0, lidar_mask.shape[0]), dtype=bool)
    return lidar_mask, lab, pr[:, 0].astype(np.int64), pr[:, 1].astype(np.int64), inb


# ----------------------------------------------------------------------------- densification + dedup
def backproject(depth, calib):
    """my_loader.py:507-509,535-536.  depth [H,W,1] or [H,W] float32 (mutated like the reference, Q10).
    Returns (uv rows, uv cols, p_lidar [M,3] float64)."""
    d3 = depth.reshape(depth.shape[0], depth.shape[1], 1)
    d3[d3 < 0.001] = 0
    uv = d3.nonzero()
    dv = d3[d3 > 0]
    rect = calib.img_to_rect(uv[1], uv[0], dv)
    return uv[0], uv[1], calib.rect_to_lidar(rect)


def sphere_voxels(xyz, vsize=0.002):
    """to_sphere_coords (my_loader.py:166-180) + the spconv voxel coordinates of the generators at
    my_loader.py:69-83.  Returns (sel index into xyz of points that enter a voxel, key int64)."""
    with np.errstate(all='ignore'):
        r = np.linalg.norm(xyz[:, 0:3], ord=2, axis=-1)
        theta = np.arccos(xyz[:, 2] / r)
        fan = np.arctan(xyz[:, 1] / xyz[:, 0])
    m1 = theta > 1.5
    vs = np.float32(vsize)
    lo = np.float32(-5.0)
    with np.errstate(all='ignore'):
        cr = np.floor((r - np.float32(-100.0)) / np.float32(200.0))
        ct = np.floor((theta - lo) / vs)
        cf = np.floor((fan - lo) / vs)
    ok = m1 & (cr >= 0) & (cr < 1) & (ct >= 0) & (ct < 5000) & (cf >= 0) & (cf < 5000)
    sel = np.nonzero(ok)[0]
    key = ct[sel].astype(np.int64) * 5000 + cf[sel].astype(np.int64)
    return sel, key


def angular_dedup(xyz, max_per_voxel=100, max_voxels=1000000):
    """la_sampling2 / la_sampling20 (my_loader.py:247-275) evaluated ONCE: per angular voxel, in
    voxel first-appearance order, the index (into xyz) of the min-z point (la_sampling2, column 10 of
    11) and of the min-y point (la_sampling20, column 10 of 12 — quirk Q2), among the first
    ``max_per_voxel`` points of the voxel; np.argmin takes the first minimum."""
    sel, key = sphere_voxels(xyz)
    if sel.size == 0:
        z = np.zeros(0, np.int64)
        return z, z, 0
    uniq, first_pos, inv = np.unique(key, return_index=True, return_inverse=True)
    order = np.argsort(first_pos, kind='stable')
    vid_of = np.empty_like(order); vid_of[order] = np.arange(order.size)
    vid = vid_of[inv]
    srt = np.argsort(vid, kind='stable')
    vs = vid[srt]
    start = np.r_[0, np.nonzero(np.diff(vs))[0] + 1]
    cnt = np.diff(np.r_[start, vs.size])
    rank = np.arange(vs.size) - np.repeat(start, cnt)
    good = (rank < max_per_voxel) & (vs < max_voxels)
    overflow = int((cnt > max_per_voxel).sum())
    pts_idx = sel[srt]                       # grouped by voxel, input order inside
    nv = int(min(order.size, max_voxels))

    def arg_first_min(col):
        val = np.where(good, xyz[pts_idx, col], np.inf)
        # first minimum per segment: lexsort by (rank, value) inside the voxel
        o = np.lexsort((rank, val, vs))
        firsts = o[np.r_[0, np.nonzero(np.diff(vs[o]))[0] + 1]]
        return pts_idx[firsts][:nv]

    return arg_first_min(2), arg_first_min(1), overflow


# ----------------------------------------------------------------------------- clustering + fitting
def cluster_adaptive(ox, oy, R0=3.0, Rd=0.001):
    """rectangle_fitting.py:161-191 with the wall-clock abort removed (Q6).  The set-merging fixpoint
    equals the connected components of the graph  i~j  iff  d(i,j) <= R_i or d(i,j) <= R_j,
    R_i = R0 + Rd*|p_i|; components are listed by their minimum point index (the merged set keeps the
    position of the lower index, rectangle_fitting.py:180).  Returns int labels [n] (0..C-1)."""
    ox = np.asarray(ox, np.float64); oy = np.asarray(oy, np.float64)
    n = ox.shape[0]
    if n == 0:
        return np.zeros(0, np.int64), 0
    P = np.stack([ox, oy], 1)
    R = R0 + Rd * np.sqrt(ox ** 2 + oy ** 2)
    rmax = float(R.max())
    side = R0 / np.sqrt(2.0) * (1 - 1e-9)
    cell = np.floor((P - P.min(0)) / side).astype(np.int64)
    ckey = cell[:, 0] * (cell[:, 1].max() + 1) + cell[:, 1]
    uniq, cfirst, cinv = np.unique(ckey, return_index=True, return_inverse=True)
    # every point is within R0 of all points of its own cell (cell diagonal < R0) -> star edges
    rows = [np.arange(n)]; cols = [cfirst[cinv]]
    # one representative test per pair of nearby cells is NOT enough (need "exists a pair")
    if n <= 3000:
        tree = cKDTree(P)
        pairs = tree.query_pairs(rmax * (1 + 1e-12) + 1e-12, output_type='ndarray')
        i, j = pairs[:, 0], pairs[:, 1]
        d = np.sqrt((ox[i] - ox[j]) ** 2 + (oy[i] - oy[j]) ** 2)
        okp = (d <= R[i]) | (d <= R[j])
        rows.append(i[okp]); cols.append(j[okp])
    else:
        # large inputs: cell pair by cell pair, "does ANY pair connect the two cells" with early exit, skipping cell
        # pairs that are already in one component (same graph, far fewer distance evaluations than all pairs)
        order = np.argsort(cinv, kind='stable')
        cstart = np.r_[0, np.cumsum(np.bincount(cinv, minlength=uniq.size))]
        ccoord = cell[cfirst]
        parent = np.arange(uniq.size)

        def find(a):
            while parent[a] != a:
                parent[a] = parent[parent[a]]
                a = parent[a]
            return a
        reach = int(np.ceil(rmax / side)) + 1
        ctree = cKDTree(ccoord.astype(np.float64))
        for ca, cb in ctree.query_pairs(reach * np.sqrt(2.0) + 1e-9, output_type='ndarray'):
            ra, rb = find(ca), find(cb)
            if ra == rb:
                continue
            ia = order[cstart[ca]:cstart[ca + 1]]
            ib = order[cstart[cb]:cstart[cb + 1]]
            hit = None
            for s0 in range(0, ia.size, 256):
                a_ = ia[s0:s0 + 256]
                d = np.sqrt((ox[a_][:, None] - ox[ib][None, :]) ** 2 + (oy[a_][:, None] - oy[ib][None, :]) ** 2)
                okp = (d <= R[a_][:, None]) | (d <= R[ib][None, :])
                if okp.any():
                    u, w = np.argwhere(okp)[0]
                    hit = (a_[u], ib[w])
                    break
            if hit is not None:
                parent[ra] = rb
                rows.append(np.array([hit[0]])); cols.append(np.array([hit[1]]))
    rr = np.concatenate(rows); cc = np.concatenate(cols)
    g = coo_matrix((np.ones(rr.size, np.int8), (rr, cc)), shape=(n, n))
    nc, lab = connected_components(g, directed=False)
    # renumber by minimum index
    first = np.full(nc, n, np.int64)
    np.minimum.at(first, lab, np.arange(n))
    order = np.argsort(first)
    ren = np.empty(nc, np.int64); ren[order] = np.arange(nc)
    return ren[lab], nc


def search_angles():
    dtheta = np.deg2rad(1.0)                                             # rectangle_fitting.py:117
    return np.arange(0.0, np.pi / 2.0 - dtheta, dtheta)                  # :119


def rectangle_search(x, y, return_costs=False):
    """rectangle_fitting.py:113-159 with the VARIANCE criterion (:83-111, :32)."""
    X = np.array([x, y]).T
    best = (-float('inf'), None)
    costs = []
    for theta in search_angles():
        e1 = np.array([np.cos(theta), np.sin(theta)])
        e2 = np.array([-np.sin(theta), np.cos(theta)])
        c1 = X @ e1.T
        c2 = X @ e2.T
        D1 = np.minimum(np.abs(c1.max() - c1), np.abs(c1 - c1.min()))
        D2 = np.minimum(np.abs(c2.max() - c2), np.abs(c2 - c2.min()))
        sel = D1 < D2
        E1 = D1[sel]; E2 = D2[~sel]
        V1 = -np.var(E1) if E1.size else 0.0
        V2 = -np.var(E2) if E2.size else 0.0
        cost = V1 + V2
        costs.append(cost)
        if best[0] < cost:
            best = (cost, theta)
    s, c = np.sin(best[1]), np.cos(best[1])
    c1s = X @ np.array([c, s]).T
    c2s = X @ np.array([-s, c]).T
    a = [c, -s, c, -s]
    b = [s, c, s, c]
    cc = [c1s.min(), c2s.min(), c1s.max(), c2s.max()]
    if return_costs:
        return a, b, cc, best[1], np.asarray(costs)
    return a, b, cc, best[1]


def _cross(a, b, c):                                                      # my_loader.py:699-702
    x = (b[0] * -c[1] - b[1] * -c[0]) / (a[0] * b[1] - a[1] * b[0])
    y = (a[1] * -c[0] - a[0] * -c[1]) / (a[0] * b[1] - a[1] * b[0])
    return x, y


def rect_contour(a, b, c):                                                # my_loader.py:686-697
    xs, ys = [None] * 4, [None] * 4
    xs[0], ys[0] = _cross(a[0:2], b[0:2], c[0:2])
    xs[1], ys[1] = _cross(a[1:3], b[1:3], c[1:3])
    xs[2], ys[2] = _cross(a[2:4], b[2:4], c[2:4])
    xs[3], ys[3] = _cross([a[3], a[0]], [b[3], b[0]], [c[3], c[0]])
    return xs, ys


def generate_anns(cls_name, pts, cls_idx, box2d, calib, names=NUSC_NAMES, legacy=False):
    """GenerateAnns (my_loader.py:633-684), nuScenes preset (center_type=None); ``legacy``: my_loader0.py:811-930
    (center_type 'prior_linecorner': height floor, aspect rule for every class, centre shift of short rectangles, every
    box 3.9 x 1.6, boxes closer than 10 m dropped, the class NAME is written as given).  Returns a list of
    records dict(alpha, box2d, h, w, l, x, y, z, ry, cluster, n, near_tie) and the label lines."""
    recs, lines = [], []
    if pts.shape[0] == 0:
        return recs, lines
    lab, nc = cluster_adaptive(pts[:, 0], pts[:, 1])
    zmax = pts[:, 2].max()
    x_min, y_min, x_max, y_max = [np.float32(v) for v in box2d]
    for k in range(nc):
        sel = lab == k
        a, b, c, th, costs = rectangle_search(list(pts[sel, 0]), list(pts[sel, 1]), return_costs=True)
        srt = np.sort(costs)[::-1]
        near_tie = bool(srt.size > 1 and (srt[0] - srt[1]) <= 1e-9 * max(1.0, abs(srt[0])))
        rx, ry_ = rect_contour(a, b, c)
        cx = (rx[0] + rx[2]) / 2
        cy = (ry_[0] + ry_[2]) / 2
        cz = zmax / 2 - 1.5
        height = zmax
        if legacy and height < 1.0:
            height = 1.56                                                     # my_loader0.py:832
        l1 = np.sqrt((rx[0] - rx[3]) ** 2 + (ry_[0] - ry_[3]) ** 2)
        l2 = np.sqrt((rx[0] - rx[1]) ** 2 + (ry_[0] - ry_[1]) ** 2)
        with np.errstate(all='ignore'):
            if (cls_name == 'Car' or legacy) and (l1 / l2 > 5 or l2 / l1 > 5):
                continue
        rot3 = np.arctan((ry_[3] - ry_[0]) / (rx[3] - rx[0] + 1e-8))
        rot1 = np.arctan((ry_[1] - ry_[0]) / (rx[1] - rx[0] + 1e-8))
        if l1 >= l2:
            length, width = l1, l2
            rot = rot3
        else:
            length, width = l2, l1
            rot = rot1
        if legacy:                                                            # my_loader0.py:848-857,889-898,876,916,918
            if width < 1.5 and length < 2:
                rot = rot1 if l1 >= l2 else rot3
                cx = (3.9 - length) / 2 * np.cos(rot) + cx
                cy = (3.9 - length) / 2 * np.sin(rot) + cy
            length, width = 3.9, 1.6
            if np.sqrt(cx ** 2 + cy ** 2) < 10:
                continue
        rot = -rot - np.pi / 2
        theta = np.arctan((-cx / (cy + 1e-8)))
        alpha = rot - theta
        hom = np.hstack((np.array([[cx, cy, cz]]), np.ones((1, 1), np.float32)))
        (crx, cry, crz), = np.dot(hom, calib.M43)
        lines.append((cls_name if legacy else names[cls_idx]) + ' 0 0 ' + str(alpha) + ' ' + str(x_min) + ' ' + str(y_min) + ' ' + str(x_max) +
                     ' ' + str(y_max) + ' ' + str(height) + ' ' + str(width) + ' ' + str(length) + ' ' + str(crx) +
                     ' ' + str(cry) + ' ' + str(crz) + ' ' + str(rot) + '\n')
        recs.append(dict(alpha=alpha, box2d=(x_min, y_min, x_max, y_max), h=height, w=width, l=length, x=crx, y=cry,
                         z=crz, ry=rot, cluster=k, n=int(sel.sum()), near_tie=near_tie, theta_fit=th))
    return recs, lines


# ----------------------------------------------------------------------------- per-instance filters
def instance_filter_params(cls_name):
    """my_loader.py:581-599 — returns ((lidar_enabled, nb, r), (nb, r)) or raises for Q9 classes."""
    if cls_name in NUSC_VEHICLES:
        return (True, 1, 3.0), (1, 3.0)
    if cls_name in NUSC_SMALL:
        return (False, 1, 0.6), (1, 0.6)          # Q5: the LiDAR filter runs on an EMPTY cloud
    raise AttributeError('class %r reaches BallQuery with numpy arrays in the reference (Q9)' % cls_name)


def ball_gate(pseudo, lidar, C=0.1):
    """BallQuery (my_loader.py:489-494) with Q4: keep iff min_j |p - l_j| < C (float64 sqrt of sum of squares).
    Nearest neighbour through a KD-tree, distance recomputed; queries within 1e-9 of C re-done by brute force."""
    pseudo = np.asarray(pseudo, np.float64).reshape(-1, 3)
    lidar = np.asarray(lidar, np.float64).reshape(-1, 3)
    if pseudo.shape[0] == 0 or lidar.shape[0] == 0:
        return np.zeros(pseudo.shape[0], bool)
    tree = cKDTree(lidar)
    _, idx = tree.query(pseudo, k=1, workers=-1)
    d = np.sqrt(((pseudo - lidar[idx]) ** 2).sum(-1))
    keep = d < C
    for i in np.nonzero(np.abs(d - C) <= 1e-9 * C)[0]:
        keep[i] = np.sqrt(((pseudo[i][None, :] - lidar) ** 2).sum(-1)).min() < C
    return keep


def legacy_filter_params(nL, nP):
    """my_loader0.py:748-766 ('multi-modal'): LiDAR (nb 20, r = 2 if n > 50 else 3); pseudo points
    uniform_down_sample(8 if n > 500, 4 if n > 100) then (nb 20, r 3).  Returns ((nb, r), (every_k, nb, r))."""
    return (20, 2.0 if nL > 50 else 3.0), (8 if nP > 500 else (4 if nP > 100 else 1), 20, 3.0)


PRESETS = {   # SURVEY.md 8(a) "Variants": literals of my_loader.py vs my_loader_KittiPandasetWaymo.py
    'nuscenes': dict(bounds=(1600, 900), only_range=((-54, 54), (-54, 54)), instances=True),       # my_loader.py:471,526
    'kitti': dict(bounds=(1216, 352), only_range=((-30, 80), (-30, 80)), instances=False),         # ..Waymo.py:470,709; GenerateAnns call commented out (:831)
    # my_loader0.py: no bounds test in the label loop (:680; PIL raises beyond the 1216 x 352 mask, so in-bounds inputs only),
    # KITTI plane range (:470), count-dependent filters (:748-766), GenerateAnns of :811-930
    'kitti_legacy': dict(bounds=(1216, 352), only_range=((-30, 80), (-30, 80)), instances=True, legacy=True),
}


def depth2pointsrgbpm(depth, image, calib, lidar, thing_classes, masks, classes, boxes2d, plane,
                      bounds=(1600, 900), return_debug=False, only_range=((-54, 54), (-54, 54)), instances=True,
                      legacy=False, max_per_voxel=100, max_voxels=1000000):
    """Restatement of my_loader.py:502-617 (nuScenes preset).  ``lidar`` is already FOV culled
    (vis_utils.py:152-154); ``plane`` replaces the unseeded RANSAC of :511 (Q8).  Returns
    (all_points [N+M',8] float64, label lines, records, debug dict)."""
    lidar = np.asarray(lidar, np.float32)
    masks = np.asarray(masks)
    n_inst = masks.shape[0]
    rows, cols, p_lidar = backproject(depth, calib)
    lidar_mask = lidar[above_plane(lidar[:, :3], plane, only_range=only_range)]
    lidar_mask, lab, px, py, inb = inherit_labels(lidar_mask, calib, masks, bounds)
    M = rows.shape[0]
    new_p = np.zeros((M, 8))
    new_p[:, 0:3] = p_lidar
    new_p[:, 4:7] = image[rows, cols] / 3
    low = new_p[:, 2] < 1.
    new_p0 = new_p[low]
    pix0 = (rows * depth.shape[1] + cols)[low]
    rep_z, rep_y, overflow = angular_dedup(new_p0, max_per_voxel=max_per_voxel, max_voxels=max_voxels)
    new_p1 = new_p0[rep_z].copy()
    new_p1[:, -1] = 1
    lines, recs = [], []
    dbg = dict(pix0=pix0, rep_z_pix=pix0[rep_z], rep_y_pix=pix0[rep_y], overflow=overflow, inst=[],
               lidar_mask=lidar_mask, lab=lab, px=px, py=py, inb=inb)
    for i in range(n_inst if instances else 0):
        cls_name = thing_classes[int(classes[i])]
        mrow = masks[i][rows, cols][low]
        P = new_p0[rep_y][mrow[rep_y] > 0][:, :3]
        L = lidar_mask[lab[i]][:, :3].astype(np.float64)
        if legacy:
            (l_nb, l_r), (p_k, p_nb, p_r) = legacy_filter_params(L.shape[0], P.shape[0])
            L1 = L[fakes.radius_outlier_mask(L, l_nb, l_r)]
            Pd = P[::p_k]
            P1 = Pd[fakes.radius_outlier_mask(Pd, p_nb, p_r)]
        else:
            (l_on, l_nb, l_r), (p_nb, p_r) = instance_filter_params(cls_name)
            L1 = L[fakes.radius_outlier_mask(L, l_nb, l_r)] if l_on else np.zeros((0, 3))
            P1 = P[fakes.radius_outlier_mask(P, p_nb, p_r)]
        if L1.shape[0] > 0 and P1.shape[0] > 0:
            P2 = P1[ball_gate(P1, L1, 0.1)]
        else:
            P2 = P1
        pts = np.concatenate([L1, P2], 0)
        r, ln = generate_anns(cls_name, pts, int(classes[i]), boxes2d[i], calib, legacy=legacy)
        for rec in r:
            rec['inst'] = i
        recs += r
        lines += ln
        dbg['inst'].append(dict(nL=L.shape[0], nP=P.shape[0], nL1=L1.shape[0], nP1=P1.shape[0], nP2=P2.shape[0], pts=pts))
    new_lidar = np.zeros((lidar.shape[0], 8))
    new_lidar[:, 0:4] = lidar[:, 0:4]
    new_lidar[:, 3] *= 10
    new_lidar[:, -1] = 2
    all_points = np.concatenate([new_lidar, new_p1], 0)
    if return_debug:
        return all_points, lines, recs, dbg
    return all_points, lines, recs


def run_view(inp, plane=None, return_debug=False, preset='nuscenes', max_per_voxel=100, max_voxels=1000000):
    """vis_utils.save_depth_as_points (vis_utils.py:142-166) minus file I/O, on the dict layout of
    dfu3d_b200.synth.view_as_reference_inputs."""
    calib = Calib(inp['P2'], inp['R0'], inp['V2C'])
    H, W = inp['depth'].shape[:2]
    lidar = np.asarray(inp['lidar'], np.float32)
    image = np.asarray(inp['image'], np.int32)[:H, :W]
    fov = get_fov_flag(calib.lidar_to_rect(lidar[:, :3]), image.shape, calib)
    lidar = lidar[fov]
    depth = np.array(inp['depth'], np.float32).reshape(H, W, 1)
    pl = inp['plane'] if plane is None else plane
    out = depth2pointsrgbpm(depth, image, calib, lidar, inp['thing_classes'], inp['masks'], inp['classes'],
                            inp['boxes2d'], pl, return_debug=return_debug, max_per_voxel=max_per_voxel, max_voxels=max_voxels, **PRESETS[preset])
    return (fov,) + tuple(out)


# ----------------------------------------------------------------------------- earlier variants (my_loader.py:142-164,415-442)
def depth2points(depth, calib):
    """my_loader.py:142-150 (threshold 0.1)."""
    d3 = depth.reshape(depth.shape[0], depth.shape[1], 1)
    d3[d3 < 0.1] = 0
    uv = d3.nonzero()
    return calib.rect_to_lidar(calib.img_to_rect(uv[1], uv[0], d3[d3 > 0]))


def depth2pointsrgb(depth, image, calib):
    """my_loader.py:152-164."""
    d3 = depth.reshape(depth.shape[0], depth.shape[1], 1)
    d3[d3 < 0.1] = 0
    uv = d3.nonzero()
    new_p = np.zeros((uv[0].shape[0], 6))
    new_p[:, 0:3] = calib.rect_to_lidar(calib.img_to_rect(uv[1], uv[0], d3[d3 > 0]))
    new_p[:, 3:] = image[uv[0], uv[1]]
    return new_p


def depth2pointsrgbp(depth, image, calib, lidar):
    """my_loader.py:415-442: threshold 0.01, height gate, la_sampling2 (restated spconv voxeliser), LiDAR rows x 10."""
    d3 = depth.reshape(depth.shape[0], depth.shape[1], 1)
    d3[d3 < 0.01] = 0
    uv = d3.nonzero()
    new_p = np.zeros((uv[0].shape[0], 8))
    new_p[:, 0:3] = calib.rect_to_lidar(calib.img_to_rect(uv[1], uv[0], d3[d3 > 0]))
    new_p[:, 4:7] = image[uv[0], uv[1]] / 3
    new_p = new_p[new_p[:, 2] < 1.]
    rep_z, _, _ = angular_dedup(new_p)
    new_p = new_p[rep_z].copy()
    new_p[:, -1] = 1
    lidar = np.asarray(lidar)
    new_lidar = np.zeros((lidar.shape[0], 8))
    new_lidar[:, 0:4] = lidar[:, 0:4]
    new_lidar[:, 3] *= 10
    new_lidar[:, -1] = 2
    return np.concatenate([new_lidar, new_p], 0)


def corners3d_to_img_boxes(P2, corners3d):
    """calibration_kitti.py:146-164."""
    n = corners3d.shape[0]
    hom = np.concatenate((corners3d, np.ones((n, 8, 1))), axis=2)
    img = np.matmul(hom, np.asarray(P2, np.float32).T)
    x, y = img[:, :, 0] / img[:, :, 2], img[:, :, 1] / img[:, :, 2]
    boxes = np.stack([x.min(1), y.min(1), x.max(1), y.max(1)], 1)
    return boxes, np.stack([x, y], 2)
