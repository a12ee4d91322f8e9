"""Seeded synthetic frames of the shapes BASELINE.json names.

SEEM and DepthAnything are outside the path (reference: tools/PENet/vis_utils.py:150,
tools/PENet/main.py:253-270); their outputs are replaced by an analytic ray-cast of a random
scene: a ground plane plus oriented boxes with the class size priors of
tools/PENet/dataloaders/my_loader_KittiPandasetWaymo.py:964-966.

A *frame* is one (multi-sweep) LiDAR cloud plus ``n_cams`` camera *views*; each view is what the
reference's ``save_depth_as_points`` (vis_utils.py:136) consumes for one image: calib, depth map,
RGB image, instance masks / classes / scores / 2-D boxes.

Everything is torch so that bench.py can build large batches on the GPU; with ``device='cpu'``
and a fixed seed the output is deterministic (tests and golden fixtures).
"""
import math

import numpy as np
import torch

THING_CLASSES_NUSC = ['Car', 'Truck', 'Construction_vehicle', 'Bus', 'Trailer', 'Barrier',
                      'Motorcycle', 'Bicycle', 'Pedestrian', 'Traffic_cone']   # vis_utils.py:182 / my_loader.py:634
THING_CLASSES_KITTI = ['Car', 'Pedestrian', 'Person ride a bike']              # vis_utils.py:181

# (class index in THING_CLASSES_NUSC, l, w, h)
_PRIORS = [(0, 3.9, 1.6, 1.56), (8, 0.8, 0.6, 1.73), (7, 1.76, 0.6, 1.73), (1, 6.5, 2.4, 2.8), (3, 10.5, 2.8, 3.2)]
_PRIOR_P = [0.55, 0.2, 0.15, 0.06, 0.04]

SHAPES = {
    # name: beams, azimuths, elev range (deg), lidar height, image (Hm, Wm), fx, cu, cv
    'nuscenes': dict(beams=32, az=1085, elev=(-30.0, 10.0), h=1.84, img=(900, 1600), f=1266.4, cu=816.3, cv=491.5),
    'kitti': dict(beams=64, az=1875, elev=(-24.8, 2.0), h=1.73, img=(375, 1242), f=721.5377, cu=609.5593, cv=172.854),
}

# tools/PENet/dataloaders/calibration_kitti.py:29-31 (default R0) and the P2 / Tr rows of the KITTI
# calibration the reference ships (tools/PENet/dataloaders/calib_cam_to_cam.txt:26 for P_rect_02)
_KITTI_R0 = np.array([[0.99992624, 0.00965411, -0.0072371], [-0.00968531, 0.99994343, -0.00433077],
                      [0.00719491, 0.00440054, 0.99996366]], dtype=np.float32)
_KITTI_P2 = np.array([[721.5377, 0.0, 609.5593, 44.85728], [0.0, 721.5377, 172.854, 0.2163791],
                      [0.0, 0.0, 1.0, 0.002745884]], dtype=np.float32)
_KITTI_V2C = np.array([[7.533745e-03, -9.999714e-01, -6.166020e-04, -4.069766e-03],
                       [1.480249e-02, 7.280733e-04, -9.998902e-01, -7.631618e-02],
                       [9.998621e-01, 7.523790e-03, 1.480755e-02, -2.717806e-01]], dtype=np.float32)
_NUSC_CAM_YAWS_DEG = [0.0, 55.0, -55.0, 110.0, -110.0, 180.0]


def make_scene(rng, n_obj=None, h=1.84, yaw_centers=(0.0,), fov=math.radians(70.0)):
    """Random boxes on the ground.  Returns float64 array [n, 8]: cx, cy, cz, l, w, hh, yaw, cls."""
    if n_obj is None:
        n_obj = int(rng.integers(4, 25))
    out = []
    for _ in range(n_obj):
        k = int(rng.choice(len(_PRIORS), p=_PRIOR_P))
        cls, l, w, hh = _PRIORS[k]
        rngm = float(rng.uniform(5.0, 60.0))
        yc = float(yaw_centers[int(rng.integers(0, len(yaw_centers)))])
        az = yc + float(rng.uniform(-0.5 * fov, 0.5 * fov)) * 0.9
        out.append([rngm * math.cos(az), rngm * math.sin(az), -h + hh / 2, l, w, hh,
                    float(rng.uniform(-math.pi, math.pi)), cls])
    return np.asarray(out, dtype=np.float64).reshape(-1, 8)


def _raycast(o, d, boxes, ground_z, device, chunk=1 << 20):
    """t along d (not normalised) of the first hit of ground or a box, and the box index (-1 = ground,
    -2 = miss).  o [3] or [R,3], d [R,3]; float64."""
    R = d.shape[0]
    t_out = torch.empty(R, dtype=torch.float64, device=device)
    id_out = torch.empty(R, dtype=torch.int64, device=device)
    B = torch.as_tensor(boxes, dtype=torch.float64, device=device)
    nb = B.shape[0]
    cy, sy = torch.cos(B[:, 6]), torch.sin(B[:, 6])
    half = B[:, 3:6] * 0.5
    for s in range(0, R, chunk):
        dd = d[s:s + chunk]
        oo = o[s:s + chunk] if o.dim() == 2 else o[None, :].expand(dd.shape[0], 3)
        inf = torch.full((dd.shape[0],), float('inf'), dtype=torch.float64, device=device)
        tg = torch.where(dd[:, 2] < 0, (ground_z - oo[:, 2]) / dd[:, 2], inf)
        tg = torch.where(tg > 0, tg, inf)
        best_t, best_id = tg, torch.where(torch.isinf(tg), -2, -1).to(torch.int64)
        for b in range(nb):
            rel = oo - B[b, 0:3]
            ox = rel[:, 0] * cy[b] + rel[:, 1] * sy[b]
            oy = -rel[:, 0] * sy[b] + rel[:, 1] * cy[b]
            oz = rel[:, 2]
            dx = dd[:, 0] * cy[b] + dd[:, 1] * sy[b]
            dy = -dd[:, 0] * sy[b] + dd[:, 1] * cy[b]
            dz = dd[:, 2]
            tn = torch.full_like(ox, -float('inf'))
            tf = torch.full_like(ox, float('inf'))
            for (oc, dc, hc) in ((ox, dx, half[b, 0]), (oy, dy, half[b, 1]), (oz, dz, half[b, 2])):
                inv = 1.0 / torch.where(dc == 0, torch.full_like(dc, 1e-30), dc)
                t1, t2 = (-hc - oc) * inv, (hc - oc) * inv
                tn = torch.maximum(tn, torch.minimum(t1, t2))
                tf = torch.minimum(tf, torch.maximum(t1, t2))
            hit = (tn <= tf) & (tn > 0) & (tn < best_t)
            best_t = torch.where(hit, tn, best_t)
            best_id = torch.where(hit, torch.full_like(best_id, b), best_id)
        t_out[s:s + chunk] = best_t
        id_out[s:s + chunk] = best_id
    return t_out, id_out


def make_lidar(shape, boxes, seed, sweeps=1, device='cpu'):
    """Spinning-LiDAR sweep(s): [N,4] float32 x,y,z,intensity in the KITTI velodyne convention
    (x forward, y left, z up; tools/nuscenes2kitti.py:234-236)."""
    cfg = SHAPES[shape]
    g = torch.Generator(device='cpu').manual_seed(int(seed) * 7919 + 13)
    el = torch.deg2rad(torch.linspace(cfg['elev'][0], cfg['elev'][1], cfg['beams'], dtype=torch.float64))
    az = torch.linspace(-math.pi, math.pi, cfg['az'] + 1, dtype=torch.float64)[:-1]
    E, A = torch.meshgrid(el, az, indexing='ij')
    d = torch.stack([torch.cos(E) * torch.cos(A), torch.cos(E) * torch.sin(A), torch.sin(E)], -1).reshape(-1, 3).to(device)
    pts = []
    for s in range(sweeps):
        shift = float(torch.rand(1, generator=g, dtype=torch.float64)) * 0.5 if s > 0 else 0.0
        o = torch.tensor([-shift * s, 0.0, 0.0], dtype=torch.float64, device=device)
        t, _ = _raycast(o, d, boxes, -cfg['h'], device)
        noise = torch.randn(d.shape[0], generator=g, dtype=torch.float64).to(device) * 0.02
        drop = torch.rand(d.shape[0], generator=g, dtype=torch.float64).to(device) < 0.02
        ok = torch.isfinite(t) & (t < 100.0) & (~drop)
        r = (t + noise)
        p = o[None, :] + d * r[:, None]
        inten = torch.rand(d.shape[0], generator=g, dtype=torch.float64).to(device)
        p4 = torch.cat([p, inten[:, None]], 1)[ok]
        pts.append(p4.to(torch.float32))
    return torch.cat(pts, 0).contiguous()


def make_calib(shape, cam=0):
    """float32 P2 [3,4], R0 [3,3], Tr_velo_to_cam [3,4]."""
    if shape == 'kitti':
        return _KITTI_P2.copy(), _KITTI_R0.copy(), _KITTI_V2C.copy()
    cfg = SHAPES[shape]
    P2 = np.array([[cfg['f'], 0, cfg['cu'], 0], [0, cfg['f'], cfg['cv'], 0], [0, 0, 1, 0]], dtype=np.float32)
    R0 = np.eye(3, dtype=np.float32)
    yaw = math.radians(_NUSC_CAM_YAWS_DEG[cam % 6])
    # small fixed mis-alignment so the rotation only *rounds* to the KITTI axis permutation
    # (the assertion of tools/nuscenes2kitti.py:234-236)
    a, b, c = 0.004, -0.007, 0.011
    Rx = np.array([[1, 0, 0], [0, math.cos(a), -math.sin(a)], [0, math.sin(a), math.cos(a)]])
    Ry = np.array([[math.cos(b), 0, math.sin(b)], [0, 1, 0], [-math.sin(b), 0, math.cos(b)]])
    Rz = np.array([[math.cos(c), -math.sin(c), 0], [math.sin(c), math.cos(c), 0], [0, 0, 1]])
    perm = np.array([[0, -1, 0], [0, 0, -1], [1, 0, 0]], dtype=np.float64)
    Ryaw = np.array([[math.cos(-yaw), -math.sin(-yaw), 0], [math.sin(-yaw), math.cos(-yaw), 0], [0, 0, 1]])
    R = perm @ Rx @ Ry @ Rz @ Ryaw
    t = np.array([0.012, -0.32, -0.61])
    V2C = np.concatenate([R, t[:, None]], 1).astype(np.float32)
    return P2, R0, V2C


def make_view(shape, boxes, seed, cam=0, crop=None, scale=1.0, device='cpu', max_inst=32, min_pixels=24,
              thing_classes=None, backdrop=None):
    """One camera view.  ``scale`` < 1 shrinks the image (and intrinsics) for fast CPU tests.
    ``crop=(H, W)``: the depth map covers image[:H, :W] (reference: vis_utils.py:147,161).
    ``backdrop`` (metres): rays that hit nothing within 100 m land on a sphere of that radius around the camera
    instead of getting depth 0 ("sky-free" view: every pixel back-projects, BASELINE.json configs[2] ~8.6 M points)."""
    cfg = SHAPES[shape]
    P2, R0, V2C = make_calib(shape, cam)
    Hm, Wm = cfg['img']
    if scale != 1.0:
        Hm, Wm = int(round(Hm * scale)), int(round(Wm * scale))
        P2 = P2.copy()
        P2[0:2, :] = (P2[0:2, :].astype(np.float64) * scale).astype(np.float32)
    H, W = (Hm, Wm) if crop is None else crop
    g = torch.Generator(device='cpu').manual_seed(int(seed) * 104729 + 1000 * cam + 7)
    fu, fv, cu, cv = float(P2[0, 0]), float(P2[1, 1]), float(P2[0, 2]), float(P2[1, 2])
    tx, ty = float(P2[0, 3]) / (-fu), float(P2[1, 3]) / (-fv)
    Rt = (R0.astype(np.float64) @ V2C[:, :3].astype(np.float64))
    tt = (R0.astype(np.float64) @ V2C[:, 3].astype(np.float64))
    RtT = torch.as_tensor(Rt.T.copy(), device=device)
    vv, uu = torch.meshgrid(torch.arange(Hm, dtype=torch.float64, device=device),
                            torch.arange(Wm, dtype=torch.float64, device=device), indexing='ij')
    d_c = torch.stack([(uu - cu) / fu, (vv - cv) / fv, torch.ones_like(uu)], -1).reshape(-1, 3)
    o_c = torch.tensor([tx, ty, 0.0], dtype=torch.float64, device=device)
    # rect -> lidar: p_l = R^T (p_r - t)
    o_l = (o_c - torch.as_tensor(tt, device=device)) @ RtT.T
    d_l = d_c @ RtT.T
    t, bid = _raycast(o_l, d_l, boxes, -cfg['h'], device)
    far = torch.zeros_like(t) if backdrop is None else float(backdrop) / d_c.norm(dim=1)
    depth = torch.where(torch.isfinite(t) & (t < 100.0), t, far)
    n = depth.shape[0]
    depth = depth * (1.0 + 0.02 * torch.randn(n, generator=g, dtype=torch.float64).to(device))
    salt = torch.rand(n, generator=g, dtype=torch.float64).to(device) < 0.01
    saltv = 1.0 + 79.0 * torch.rand(n, generator=g, dtype=torch.float64).to(device)
    depth = torch.where(salt, saltv, depth).clamp_min(0.0)
    depth = depth.reshape(Hm, Wm).to(torch.float32)
    bid = bid.reshape(Hm, Wm)
    masks, classes, boxes2d, obj_ids = [], [], [], []
    pad = max(1, int(round(3 * max(scale, 1.0 / 3))))
    for b in range(boxes.shape[0]):
        m = (bid == b)
        if int(m.sum()) < min_pixels or len(masks) >= max_inst:
            continue
        md = torch.nn.functional.max_pool2d(m[None, None].to(torch.float32), 2 * pad + 1, 1, pad)[0, 0] > 0
        ys, xs = torch.nonzero(md, as_tuple=True)
        masks.append(md.to(torch.uint8))
        classes.append(int(boxes[b, 7]))
        boxes2d.append([float(xs.min()), float(ys.min()), float(xs.max()), float(ys.max())])
        obj_ids.append(b)
    ninst = len(masks)
    masks_t = torch.stack(masks, 0) if ninst else torch.zeros((0, Hm, Wm), dtype=torch.uint8, device=device)
    image = torch.randint(0, 256, (Hm, Wm, 3), generator=g, dtype=torch.int32).to(device).to(torch.uint8)
    scores = 0.7 + 0.3 * torch.rand(max(ninst, 1), generator=g, dtype=torch.float64)[:ninst].to(torch.float32)
    return {
        'P2': P2, 'R0': R0, 'V2C': V2C,
        'depth': depth[:H, :W].contiguous(), 'image': image, 'masks': masks_t,
        'classes': np.asarray(classes, dtype=np.int64), 'scores': scores.numpy(),
        'boxes2d': np.asarray(boxes2d, dtype=np.float32).reshape(-1, 4), 'obj_ids': obj_ids,
        'thing_classes': list(thing_classes or THING_CLASSES_NUSC), 'H': H, 'W': W, 'Hm': Hm, 'Wm': Wm,
        'plane': np.array([0.0, 0.0, 1.0, cfg['h']], dtype=np.float64),   # ground z = -h  ->  n.p + h = 0
    }


def make_frame(seed, shape='nuscenes', n_cams=1, sweeps=1, scale=1.0, crop=None, device='cpu', n_obj=None, max_inst=32,
               only_cam=None, backdrop=None):
    """``dict(lidar [N,4] f32, views [n_cams], boxes_gt)``; seeded by ``seed`` alone.  ``only_cam`` keeps a
    single camera of the ``n_cams``-camera frame (same scene, same sweep)."""
    rng = np.random.default_rng(int(seed))
    cfg = SHAPES[shape]
    yaws = [math.radians(_NUSC_CAM_YAWS_DEG[c % 6]) for c in range(n_cams)] if shape == 'nuscenes' else [0.0]
    boxes = make_scene(rng, n_obj=n_obj, h=cfg['h'], yaw_centers=yaws)
    lidar = make_lidar(shape, boxes, seed, sweeps=sweeps, device=device)
    cams = range(n_cams) if only_cam is None else [only_cam]
    views = [make_view(shape, boxes, seed, cam=c, crop=crop, scale=scale, device=device, max_inst=max_inst,
                       backdrop=backdrop) for c in cams]
    return {'lidar': lidar, 'views': views, 'boxes_gt': boxes, 'shape': shape}


def view_as_reference_inputs(frame, v=0):
    """numpy dict in the layout oracle/ref_harness.run_depth2pointsrgbpm expects."""
    view = frame['views'][v]
    return {
        'lidar': frame['lidar'].cpu().numpy(), 'P2': view['P2'], 'R0': view['R0'], 'V2C': view['V2C'],
        'depth': view['depth'].cpu().numpy(), 'image': view['image'].cpu().numpy().astype(np.int32),
        'masks': view['masks'].cpu().numpy().astype(np.float32), 'classes': view['classes'],
        'scores': view['scores'], 'boxes2d': view['boxes2d'], 'thing_classes': view['thing_classes'],
        'plane': view['plane'],
    }
