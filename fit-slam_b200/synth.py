"""Seeded synthetic inputs of the BASELINE.json shapes: nav2-style costmaps, ORB-SLAM3-like
map-point clouds and candidate poses.  Pure numpy, byte-deterministic for a given (config, seed).

Costmap value set follows the reference's producers: 255 unknown, 0 free, 254 lethal, optionally an
inflation ring 253..1 (TRAV threshold_occupancy.cpp:31-45 + nav2 inflation layer).
"""
import math

import numpy as np

BASE_SEED = 20261104
MARGIN = 64  # unknown border: DEP and RE neighbourhood code agree away from the map edge


def _disc(a, cx, cy, r, value, only=None):
    H, W = a.shape
    x0, x1 = max(0, cx - r), min(W, cx + r + 1)
    y0, y1 = max(0, cy - r), min(H, cy + r + 1)
    if x0 >= x1 or y0 >= y1:
        return
    yy, xx = np.ogrid[y0:y1, x0:x1]
    m = (xx - cx) ** 2 + (yy - cy) ** 2 <= r * r
    sub = a[y0:y1, x0:x1]
    if only is not None:
        m = m & (sub == only)
    sub[m] = value


def make_costmap(W, H, seed, n_disc, r_lo=30, r_hi=120, spread=None, n_rooms=0, n_holes=0, hole_r=(5, 25), n_walls=None, wall_len=(8, 50),
                 porous=0.0, inflation=0, res=0.05, grid_radius=0, grid_step=110):
    """Explored region = a random-walk chain of sensor discs (so it is one reachable component) plus rooms;
    thin lethal wall segments inside; optional porous lethal cells on the free/unknown edge (fragmenting the
    frontier into small clusters, config C3) and an inflation ring.  Returns (map[H,W] uint8, meta dict)."""
    rng = np.random.default_rng(seed)
    a = np.full((H, W), 255, dtype=np.uint8)
    cx0, cy0 = W // 2, H // 2
    spread = spread if spread is not None else min(W, H) // 2 - MARGIN - r_hi
    lo_x, hi_x = max(MARGIN + r_hi, cx0 - spread), min(W - MARGIN - r_hi - 1, cx0 + spread)
    lo_y, hi_y = max(MARGIN + r_hi, cy0 - spread), min(H - MARGIN - r_hi - 1, cy0 + spread)
    centres = [(cx0, cy0, (r_lo + r_hi) // 2)]
    _disc(a, cx0, cy0, centres[0][2], 0)
    if grid_radius:
        # a jittered grid of overlapping discs inside a circle (config C4: a large explored area around the robot);
        # n_disc is ignored, neighbouring discs overlap (r_lo + r_lo > grid_step), so the region is one component
        g = int(grid_radius // grid_step)
        for gy in range(-g, g + 1):
            for gx in range(-g, g + 1):
                if (gx or gy) and gx * gx + gy * gy <= g * g:
                    cx = int(min(hi_x, max(lo_x, cx0 + gx * grid_step + rng.integers(-8, 9))))
                    cy = int(min(hi_y, max(lo_y, cy0 + gy * grid_step + rng.integers(-8, 9))))
                    r = int(rng.integers(r_lo, r_hi + 1))
                    _disc(a, cx, cy, r, 0)
                    centres.append((cx, cy, r))
        n_disc = 1
    for _ in range(n_disc - 1):
        # grow mostly from recent discs so the chain spreads out instead of piling up at the centre
        lo = max(0, len(centres) - 8) if rng.random() < 0.7 else 0
        px, py, pr = centres[int(rng.integers(lo, len(centres)))]
        r = int(rng.integers(r_lo, r_hi + 1))
        ang = float(rng.uniform(0, 2 * math.pi))
        d = float(rng.uniform(0.6, 0.98)) * (pr + r)
        cx = int(min(hi_x, max(lo_x, px + d * math.cos(ang))))
        cy = int(min(hi_y, max(lo_y, py + d * math.sin(ang))))
        _disc(a, cx, cy, r, 0)
        centres.append((cx, cy, r))
    for _ in range(n_rooms):
        px, py, pr = centres[int(rng.integers(0, len(centres)))]
        w, h = int(rng.integers(40, 200)), int(rng.integers(40, 200))
        x0, y0 = max(MARGIN, px - w // 2), max(MARGIN, py - h // 2)
        a[y0:min(H - MARGIN, y0 + h), x0:min(W - MARGIN, x0 + w)] = 0
    # unexplored pockets inside the explored region (each one adds a closed frontier loop)
    for _ in range(n_holes):
        px, py, pr = centres[int(rng.integers(0, len(centres)))]
        r = int(rng.integers(hole_r[0], hole_r[1] + 1))
        x = int(px + rng.integers(-pr, pr + 1))
        y = int(py + rng.integers(-pr, pr + 1))
        if abs(x - cx0) + abs(y - cy0) > r + 12:
            _disc(a, x, y, r, 255)
    # lethal wall segments, 1-3 cells thick, only over explored cells
    n_walls = n_walls if n_walls is not None else n_disc
    for _ in range(n_walls):
        px, py, pr = centres[int(rng.integers(0, len(centres)))]
        L = int(rng.integers(wall_len[0], wall_len[1] + 1))
        t = int(rng.integers(1, 4))
        x = int(px + rng.integers(-pr, pr + 1))
        y = int(py + rng.integers(-pr, pr + 1))
        if rng.integers(0, 2):
            sub = a[max(0, y):y + t, max(0, x):x + L]
        else:
            sub = a[max(0, y):y + L, max(0, x):x + t]
        sub[sub == 0] = 254
    if porous > 0.0:
        free = a == 0
        unk = a == 255
        edge = np.zeros_like(free)
        edge[:, 1:] |= unk[:, :-1]
        edge[:, :-1] |= unk[:, 1:]
        edge[1:, :] |= unk[:-1, :]
        edge[:-1, :] |= unk[1:, :]
        edge &= free
        pick = rng.random(size=int(edge.sum())) < porous
        ys, xs = np.nonzero(edge)
        a[ys[pick], xs[pick]] = 254
    _disc(a, cx0, cy0, 6, 0, only=254)  # the robot stands on free space
    a[cy0 - 2:cy0 + 3, cx0 - 2:cx0 + 3] = 0
    if inflation > 0:
        cur = (a == 254)
        for d in range(1, inflation + 1):
            g = cur.copy()
            g[:, 1:] |= cur[:, :-1]
            g[:, :-1] |= cur[:, 1:]
            h = g.copy()
            h[1:, :] |= g[:-1, :]
            h[:-1, :] |= g[1:, :]
            ring = h & ~cur & (a == 0)
            a[ring] = max(1, 253 - (d - 1) * (252 // max(1, inflation)))
            cur = h
        a[cy0, cx0] = 0
    ox, oy = -W * res / 2.0, -H * res / 2.0
    meta = dict(W=W, H=H, res=res, ox=ox, oy=oy, robot_x=ox + (cx0 + 0.5) * res, robot_y=oy + (cy0 + 0.5) * res,
                poly=[ox, oy, ox + W * res, oy + H * res], centres=centres)
    return a, meta


def explored_bbox(a, meta):
    ys, xs = np.nonzero(a != 255) if a.size <= (1 << 26) else (None, None)
    if ys is None:
        c = np.array([(x, y, r) for x, y, r in meta["centres"]])
        x0, x1 = (c[:, 0] - c[:, 2]).min(), (c[:, 0] + c[:, 2]).max()
        y0, y1 = (c[:, 1] - c[:, 2]).min(), (c[:, 1] + c[:, 2]).max()
    else:
        x0, x1, y0, y1 = xs.min(), xs.max(), ys.min(), ys.max()
    r = meta["res"]
    return meta["ox"] + x0 * r, meta["oy"] + y0 * r, meta["ox"] + (x1 + 1) * r, meta["oy"] + (y1 + 1) * r


def make_points(P, bbox, seed, z_range=(0.0, 3.0)):
    """ORB-SLAM3-like landmarks in the ROS map frame: x, y uniform over the explored bounding box, z in [0, 3]."""
    rng = np.random.default_rng(seed ^ 0x5EED)
    x0, y0, x1, y1 = bbox
    pts = np.empty((P, 3), dtype=np.float32)
    pts[:, 0] = rng.uniform(x0, x1, P).astype(np.float32)
    pts[:, 1] = rng.uniform(y0, y1, P).astype(np.float32)
    pts[:, 2] = rng.uniform(z_range[0], z_range[1], P).astype(np.float32)
    return pts


def yaw_shape(R, Y):
    """delta_theta / camera_fov that give exactly R rays and Y yaw windows: k = R - Y + 1."""
    dt = 2 * math.pi / (R - 0.5)
    n = 0
    th = 0.0
    while th <= 2 * math.pi:
        n += 1
        th += dt
    assert n == R, (n, R)
    k = R - Y + 1
    fov = (k + 0.5) * dt
    assert int(fov / dt) == k
    return dt, fov


# name -> generator arguments + parameter overrides.  n_disc tuned so searchFrom yields ~ the named F.
CONFIGS = {
    "C1": dict(W=1000, H=1000, n_disc=14, r_lo=30, r_hi=90, n_holes=25, P=20_000, R=64, Y=8, max_frontier_distance=50.0),
    "C2": dict(W=4096, H=4096, n_disc=450, r_lo=30, r_hi=120, n_holes=1000, P=200_000, R=64, Y=16, max_frontier_distance=150.0),
    "C3": dict(W=2048, H=2048, n_disc=300, r_lo=25, r_hi=60, n_holes=1500, hole_r=(2, 6), n_walls=0, porous=0.3, P=50_000, R=64, Y=16,
               max_frontier_distance=80.0),
    "C4": dict(W=16384, H=16384, n_disc=1, r_lo=60, r_hi=120, grid_radius=2600, n_holes=9000, hole_r=(4, 14), n_walls=600, P=1_000_000,
               R=64, Y=32, max_frontier_distance=150.0),
}


def make_config(name, seed=None, reference_rays=False, lib_params=None):
    """Returns (map, meta, points, overrides) for a BASELINE.json config; `overrides` are the fitslam_params
    fields that differ from the defaults (ray/yaw shape unless reference_rays, search distance)."""
    cfg = dict(CONFIGS[name])
    idx = int(name[1:])
    seed = BASE_SEED + idx if seed is None else seed
    P, R, Y = cfg.pop("P"), cfg.pop("R"), cfg.pop("Y")
    mfd = cfg.pop("max_frontier_distance")
    a, meta = make_costmap(seed=seed, **cfg)
    pts = make_points(P, explored_bbox(a, meta), seed)
    ov = dict(max_frontier_distance=mfd)
    if not reference_rays:
        dt, fov = yaw_shape(R, Y)
        ov.update(delta_theta=dt, camera_fov=fov)
    return a, meta, pts, ov


def apply_overrides(params, ov):
    for k, v in ov.items():
        setattr(params, k, v)
    return params


def reveal_step(a, meta, cx, cy, rng, r_cells=40, lethal_frac=0.05):
    """Rapid-mode replay (C5): reveal a disc around (cx, cy); returns the dirty rectangle (x0, y0, w, h)."""
    H, W = a.shape
    x0, x1 = max(0, cx - r_cells), min(W, cx + r_cells + 1)
    y0, y1 = max(0, cy - r_cells), min(H, cy + r_cells + 1)
    yy, xx = np.ogrid[y0:y1, x0:x1]
    sub = a[y0:y1, x0:x1]
    m = ((xx - cx) ** 2 + (yy - cy) ** 2 <= r_cells * r_cells) & (sub == 255)
    leth = m & (rng.random(size=sub.shape) < lethal_frac)
    sub[m] = 0
    sub[leth] = 254
    sub2 = a[cy - 2:cy + 3, cx - 2:cx + 3]
    sub2[sub2 == 254] = 0
    return (x0, y0, x1 - x0, y1 - y0)


def frontier_poses(frontiers, theta, z, orientation_fn=None):
    """Candidate poses at frontier goals with yaw theta (setGoalOrientation, DEP/src/Frontier.cpp:50-56)."""
    from .capi import POSE_DTYPE
    n = len(frontiers)
    p = np.zeros(n, dtype=POSE_DTYPE)
    p["x"], p["y"], p["z"] = frontiers["goal_x"], frontiers["goal_y"], z
    th = np.broadcast_to(np.asarray(theta, dtype=np.float64), (n,))
    p["qz"] = np.sin(th * 0.5)
    p["qw"] = np.cos(th * 0.5)
    return p


def make_keyframes(pts, n_kf, seed, max_range=8.0, fov=1.2, per_kf=None):
    """A synthetic ORB-SLAM3 atlas over the map points `pts` (ROS map frame): n_kf keyframes along a random walk through the
    cloud's bounding box, each listing the points in front of it within `max_range` and a horizontal field of view `fov`
    (so neighbouring keyframes share many points, as covisible keyframes do).  Returns (KEYFRAME_DTYPE records with
    GetPoseInverse() in the ORB world frame, concatenated point indices)."""
    from .capi import KEYFRAME_DTYPE
    rng = np.random.default_rng(seed)
    lo, hi = pts[:, :2].min(axis=0), pts[:, :2].max(axis=0)
    pos = rng.uniform(lo, hi)
    yaw = rng.uniform(-np.pi, np.pi)
    perm = np.array([[0, -1, 0], [0, 0, -1], [1, 0, 0]], dtype=np.float64)  # ROS map axes -> ORB world axes
    kfs = np.zeros(n_kf, dtype=KEYFRAME_DTYPE)
    lists = []
    at = 0
    for k in range(n_kf):
        yaw += rng.normal(0, 0.5)
        pos = np.clip(pos + 0.8 * np.array([np.cos(yaw), np.sin(yaw)]) + rng.normal(0, 0.1, 2), lo, hi)
        z = 0.3
        c, s_ = np.cos(yaw), np.sin(yaw)
        R_ros = np.array([[c, -s_, 0], [s_, c, 0], [0, 0, 1]])
        kfs[k]["Rwc"] = (perm @ R_ros @ perm.T).astype(np.float32).ravel()
        kfs[k]["twc"] = (perm @ np.array([pos[0], pos[1], z])).astype(np.float32)
        d = pts[:, :2] - pos
        r = np.hypot(d[:, 0], d[:, 1])
        ang = np.abs((np.arctan2(d[:, 1], d[:, 0]) - yaw + np.pi) % (2 * np.pi) - np.pi)
        idx = np.nonzero((r < max_range) & (ang < fov / 2))[0]
        if per_kf is not None and len(idx) > per_kf:
            idx = rng.choice(idx, per_kf, replace=False)
        idx = rng.permutation(idx).astype(np.uint32)
        if len(idx) > 4:  # a map point listed twice by one keyframe must be taken once as well
            idx = np.concatenate([idx, idx[:2]])
        kfs[k]["first"], kfs[k]["count"] = at, len(idx)
        at += len(idx)
        lists.append(idx)
    return kfs, (np.concatenate(lists) if lists else np.zeros(0, np.uint32)).astype(np.uint32)
