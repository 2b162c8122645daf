"""Seeded synthetic organized clouds for the five BASELINE.json configs.

All generators are plain numpy arithmetic (analytic ray/plane/box/cylinder
casting, no graphics API).  Clouds are returned in the reference's boundary
layout (ps/cloud.h:37-41,74; ps/multichannel_cloud_test.cc:209-218):
column-major grid, index = col * rows + row, AoS float32 xyz, NaN = invalid.
Scene definitions follow SURVEY.md section 8(d).
"""
import math

import numpy as np

NAN = np.float32(np.nan)


def _cam_rotation(yaw, pitch_down):
    """Camera (x right, y down, z forward) -> world (x fwd, y left, z up)."""
    cp, sp = math.cos(pitch_down), math.sin(pitch_down)
    cy, sy = math.cos(yaw), math.sin(yaw)
    fwd = np.array([cp * cy, cp * sy, -sp])
    right = np.array([sy, -cy, 0.0])
    down = np.cross(fwd, right)
    return np.stack([right, down, fwd], axis=1)  # columns = camera axes in world


def _ray_boxes(o, w, boxes):
    """Nearest positive hit parameter of rays o + t*w with AABBs [lo, hi]."""
    t_best = np.full(w.shape[0], np.inf)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / w
        for lo, hi in boxes:
            t0 = (np.asarray(lo) - o) * inv
            t1 = (np.asarray(hi) - o) * inv
            tn = np.minimum(t0, t1).max(axis=1)
            tf = np.maximum(t0, t1).min(axis=1)
            hit = (tn <= tf) & (tf > 1e-9)
            t = np.where(tn > 1e-9, tn, tf)  # inside the box -> exit point
            t_best = np.where(hit & (t < t_best), t, t_best)
    return t_best


def _ray_floor(o, w, z=0.0):
    with np.errstate(divide="ignore", invalid="ignore"):
        t = (z - o[2]) / w[:, 2]
    return np.where((t > 1e-9) & np.isfinite(t), t, np.inf)


def _finish_depth_cloud(dirs_cam, t, rows, cols, rng_noise, rng_drop, sigma_rel,
                        drop_frac, max_range):
    g = rng_noise.standard_normal(t.shape[0])
    t = t * (1.0 + sigma_rel * g)
    p = dirs_cam * t[:, None]
    rng2 = np.einsum("ij,ij->i", p, p)
    bad = ~np.isfinite(t) | (rng2 > max_range * max_range)
    bad |= rng_drop.random(t.shape[0]) < drop_frac
    p = p.astype(np.float32)
    p[bad] = NAN
    return np.ascontiguousarray(p.reshape(rows * cols, 3))


def _pinhole_dirs(rows, cols, fx, fy, cx, cy):
    # linear index = col * rows + row  (column-major)
    u = np.repeat(np.arange(cols, dtype=np.float64), rows)
    v = np.tile(np.arange(rows, dtype=np.float64), cols)
    return np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u)], axis=1)


C1_BOXES = [((2.0, -0.8), (0.5, 0.5, 0.5)),
            ((3.0, 0.5), (0.4, 0.6, 0.8)),
            ((2.5, 1.2), (0.3, 0.3, 1.0))]


def depth_frame(rows=480, cols=640, seed=1234, yaw_deg=0.0, box_jitter=None,
                fx=525.0, fy=525.0, sigma_rel=0.002, drop_frac=0.01,
                max_range=8.0, cam_height=1.2, pitch_deg=30.0):
    """C1 scene: floor z=0 + 3 boxes, pinhole depth camera, sensor-frame xyz."""
    cx, cy = (cols - 1) / 2.0, (rows - 1) / 2.0
    d = _pinhole_dirs(rows, cols, fx * cols / 640.0, fy * rows / 480.0, cx, cy)
    R = _cam_rotation(math.radians(yaw_deg), math.radians(pitch_deg))
    w = d @ R.T
    o = np.array([0.0, 0.0, cam_height])
    boxes = []
    for k, ((bx, by), (sx, sy, sz)) in enumerate(C1_BOXES):
        jx, jy = (box_jitter[k] if box_jitter is not None else (0.0, 0.0))
        boxes.append(((bx + jx - sx / 2, by + jy - sy / 2, 0.0),
                      (bx + jx + sx / 2, by + jy + sy / 2, sz)))
    t = np.minimum(_ray_floor(o, w), _ray_boxes(o, w, boxes))
    return _finish_depth_cloud(d, t, rows, cols, np.random.default_rng(seed),
                               np.random.default_rng(seed + 1), sigma_rel,
                               drop_frac, max_range)


def c1(rows=480, cols=640):
    return depth_frame(rows, cols, seed=1234)


def c2_frame(i, n_frames=256, rows=480, cols=640, replica=0):
    """Frame i of the C2 batch.  replica > 0: the same camera sweep and box
    jitter with other noise / dropout seeds (the batches of the other GPUs in a weak-scaling
    run: same workload per GPU, different data)."""
    jit = np.random.default_rng(99).uniform(-0.2, 0.2, size=(n_frames, 3, 2))
    i = i % n_frames
    yaw = -20.0 + 40.0 * i / max(n_frames - 1, 1)
    return depth_frame(rows, cols, seed=1234 + 2 * i + 100000 * replica, yaw_deg=yaw,
                       box_jitter=jit[i])


def c2_batch(n_frames=256, rows=480, cols=640, distinct=None):
    """[n_frames, rows*cols, 3].  distinct < n_frames tiles the first frames."""
    distinct = n_frames if distinct is None else min(distinct, n_frames)
    base = [c2_frame(i, n_frames, rows, cols) for i in range(distinct)]
    out = np.empty((n_frames, rows * cols, 3), np.float32)
    for i in range(n_frames):
        out[i] = base[i % distinct]
    return out


def c3(rows=720, cols=1280, seed=2234):
    """Indoor room, many planes (walls, floor, ceiling, 6 table tops, 12 boxes)."""
    cx, cy = (cols - 1) / 2.0, (rows - 1) / 2.0
    d = _pinhole_dirs(rows, cols, 900.0 * cols / 1280.0, 900.0 * rows / 720.0, cx, cy)
    R = _cam_rotation(math.radians(8.0), math.radians(18.0))
    w = d @ R.T
    o = np.array([0.4, 0.0, 1.5])
    room = [((0.0, -2.5, 0.0), (6.0, 2.5, 2.6))]
    rng = np.random.default_rng(seed + 10)
    objs = []
    for k in range(6):  # table tops
        x = 1.6 + 0.7 * k
        y = -1.8 + 0.7 * k
        h = 0.45 if k % 2 == 0 else 0.75
        objs.append(((x - 0.4, y - 0.3, h - 0.04), (x + 0.4, y + 0.3, h)))
    for k in range(12):  # boxes
        x = rng.uniform(1.5, 5.5)
        y = rng.uniform(-2.2, 2.2)
        s = rng.uniform(0.15, 0.45, size=3)
        objs.append(((x - s[0], y - s[1], 0.0), (x + s[0], y + s[1], 2 * s[2])))
    t = np.minimum(_ray_boxes(o, w, room), _ray_boxes(o, w, objs))
    return _finish_depth_cloud(d, t, rows, cols, np.random.default_rng(seed),
                               np.random.default_rng(seed + 1), 0.002, 0.01, 12.0)


def c4(rows=128, cols=16384, seed=3234, n_objects=200):
    """Organized lidar range image; returns (xyz, labels) with ground masked
    (kMaskedOut = -2 where height above ground < 0.05 m)."""
    el = np.radians(np.linspace(15.0, -25.0, rows))
    az = np.linspace(-math.pi, math.pi, cols, endpoint=False)
    e = np.tile(el, cols)
    a = np.repeat(az, rows)
    w = np.stack([np.cos(e) * np.cos(a), np.cos(e) * np.sin(a), np.sin(e)], axis=1)
    o = np.array([0.0, 0.0, 1.8])
    rng = np.random.default_rng(seed + 10)
    t = _ray_floor(o, w)
    boxes = []
    for _ in range(n_objects):
        r = rng.uniform(4.0, 40.0)
        th = rng.uniform(-math.pi, math.pi)
        s = rng.uniform(0.3, 1.5, size=3)
        x, y = r * math.cos(th), r * math.sin(th)
        boxes.append(((x - s[0], y - s[1], 0.0), (x + s[0], y + s[1], 2 * s[2])))
    # a box can only be hit by the azimuth columns its footprint spans (+ a margin): same arithmetic per
    # (ray, box) pair and same box order as testing every ray against every box, 50x less work
    daz = 2.0 * math.pi / cols
    for lo, hi in boxes:
        ang = [math.atan2(y, x) for x in (lo[0], hi[0]) for y in (lo[1], hi[1])]
        a0, a1 = min(ang), max(ang)
        if a1 - a0 > math.pi or (lo[0] <= 0.0 <= hi[0] and lo[1] <= 0.0 <= hi[1]):   # wraps around +-pi (or contains the sensor)
            c0, c1 = 0, cols
        else:
            c0 = max(int((a0 + math.pi) / daz) - 2, 0)
            c1 = min(int((a1 + math.pi) / daz) + 3, cols)
        sl = slice(c0 * rows, c1 * rows)
        t[sl] = np.minimum(t[sl], _ray_boxes(o, w[sl], [(lo, hi)]))
    g = np.random.default_rng(seed).standard_normal(t.shape[0])
    t = t + 0.01 * g
    p = w * t[:, None]
    bad = ~np.isfinite(t) | (t > 60.0) | (t <= 0.0)
    bad |= np.random.default_rng(seed + 1).random(t.shape[0]) < 0.05
    height = p[:, 2] + o[2]
    p = p.astype(np.float32)
    p[bad] = NAN
    labels = np.full(rows * cols, -1, np.int32)
    labels[(height < 0.05) & ~bad] = -2
    return np.ascontiguousarray(p), labels


def c5(rows=2500, cols=4000, seed=4234, n_blobs=2000, pitch=0.02, nan_frac=0.10, tiles=None):
    """Stereo-like grid: Gaussian blobs (sigma 0.15 m) on a ground sheet.
    tiles = (tile_rows, tile_cols, gap): the sheet is cut into tiles separated by `gap` NaN rows /
    columns (wider than the growth window), so that it is not ONE closure under the 0.2 m neighbour
    rule of ps/mean_shift_segmentation.h:152-199 and the sort + suppression loop sees many regions."""
    rng = np.random.default_rng(seed + 10)
    z = np.zeros((cols, rows), np.float64)  # [col, row] == column-major
    sig = 0.15
    half = int(math.ceil(4 * sig / pitch))
    bc = rng.uniform(0, cols, size=n_blobs)
    br = rng.uniform(0, rows, size=n_blobs)
    amp = rng.uniform(0.5, 1.5, size=n_blobs)
    for k in range(n_blobs):
        c0, r0 = int(bc[k]), int(br[k])
        cs = slice(max(c0 - half, 0), min(c0 + half + 1, cols))
        rs = slice(max(r0 - half, 0), min(r0 + half + 1, rows))
        dc = (np.arange(cs.start, cs.stop) - bc[k]) * pitch
        dr = (np.arange(rs.start, rs.stop) - br[k]) * pitch
        z[cs, rs] += amp[k] * np.exp(-(dc[:, None] ** 2 + dr[None, :] ** 2) / (2 * sig * sig))
    x = np.repeat(np.arange(cols) * pitch, rows)
    y = np.tile(np.arange(rows) * pitch, cols)
    p = np.stack([x, y, z.reshape(-1)], axis=1).astype(np.float32)
    p[:, 2] += (0.003 * np.random.default_rng(seed).standard_normal(rows * cols)).astype(np.float32)
    p[np.random.default_rng(seed + 1).random(rows * cols) < nan_frac] = NAN
    if tiles:
        tr, tc, gap = tiles
        rr = np.tile(np.arange(rows), cols)
        cc = np.repeat(np.arange(cols), rows)
        p[((rr % tr) >= tr - gap) | ((cc % tc) >= tc - gap)] = NAN
    return np.ascontiguousarray(p)


def c5_full(rows=2500, cols=4000):
    """Config C5 at full size: 10 M points, 2 000 blobs, cut into 20 x 20 tiles of
    125 x 200 cells (2.5 m x 4 m) by 8-cell NaN gaps => several hundred mean-shift regions."""
    return c5(rows, cols, tiles=(max(rows // 20, 16), max(cols // 20, 16), 8))


def terraces(rows, cols, tr, tc, pitch, step, seed, nan_frac=0.03):
    """Tiles of tr x tc cells at heights step * ((ti + 2 tj) % 4), no gaps: the regions TOUCH (their
    windows overlap) without merging, and cells within 1 m of a neighbour's mode cross the steps --
    the case in which regions grown at the same time interact (meanshift_wave.cuh)."""
    rng = np.random.default_rng(seed)
    cc, rr = np.meshgrid(np.arange(cols), np.arange(rows), indexing="ij")
    z = step * (((rr // tr) + 2 * (cc // tc)) % 4) + 0.002 * rng.standard_normal((cols, rows))
    p = np.stack([cc * pitch, rr * pitch, z], axis=-1).reshape(-1, 3).astype(np.float32)
    p[rng.random(rows * cols) < nan_frac] = np.nan
    return np.ascontiguousarray(p)


def speckle_cloud(rows, cols, seed, pitch=0.05):
    """A sheet with random height steps, dropouts and isolated speckles: hundreds of connected components
    of all sizes, regions that touch one another, candidates that are rejected inside a component
    (quirk Q5) and components too small to yield a region."""
    rng = np.random.default_rng(seed)
    cc, rr = np.meshgrid(np.arange(cols), np.arange(rows), indexing="ij")
    z = 0.35 * rng.integers(0, 4, size=(cols // 16 + 1, rows // 12 + 1))[cc // 16, rr // 12].astype(np.float64)
    z += 0.02 * rng.standard_normal((cols, rows))
    spk = rng.random((cols, rows)) < 0.04
    z[spk] += rng.uniform(0.15, 0.6, size=int(spk.sum()))
    p = np.stack([cc * pitch, rr * pitch, z], axis=-1).reshape(-1, 3).astype(np.float32)
    p[rng.random(rows * cols) < 0.05] = np.nan
    return np.ascontiguousarray(p)


def clustering_test_cloud():
    """ps/clustering_test.cc:53-113 PointCloudData(): 10x10 golden cloud,
    returns (xyz[100,3], seeds[100]) with seed[99 - idx] = idx."""
    k = 10
    pts = np.zeros((k * k, 3), np.float32)
    seeds = np.zeros(k * k, np.int32)
    f = np.float32
    for col in range(k):
        for row in range(k):
            idx = col * k + row
            seeds[k * k - 1 - idx] = idx
            if col < k // 2 - 1:
                if row < k // 2 - 1:
                    p = (f(0.05) * f(col) + f(2.0), f(0.05) * f(row) + f(2.0), f(2.0))
                elif row > k // 2 - 1:
                    p = (f(0.05) * f(col), f(0.05) * f(row) + f(2.0), f(0.0))
                else:
                    p = (f(0.05) * f(col), f(0.05) * f(row) + f(2.0), f(1.0))
            elif col == k // 2 - 1:
                p = (f(0.1) * f(col), f(0.0), f(0.1) * f(row))
            elif row + col == k or row + col == k - 1:
                p = (f(0.05) * f(col) + f(5.0), f(-0.05) * f(row) - f(2.0), f(2.0) - f(0.05) * f(row))
            elif row == k - 2 and col == k - 2:
                p = (f(0.05) * f(col) + f(2.0), f(-0.05) * f(row) - f(2.0), f(4.0))
            elif (row == k - 4 or row == k - 3) and col == k - 2:
                p = (f(6.0), f(8.0), f(2.0))
            else:
                p = (f(0.05) * f(col) + f(2.0), f(-0.05) * f(row) - f(2.0), f(2.0) - f(0.05) * f(row))
            pts[idx] = p
    return pts, seeds
