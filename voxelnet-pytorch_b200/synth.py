"""Synthetic KITTI-shaped LiDAR frames (SURVEY.md section 8d "Synthetic inputs").

A 64-beam spinning sensor at the origin (Velodyne frame: x forward, y left, z up) casts
`beams x azimuth_steps` rays against a ground plane at z = -1.73 m, a ring of far walls and a few
dozen random axis-aligned boxes (cars 3.9 x 1.6 x 1.56 m and larger clutter).  Range noise is
sigma = 2 cm, reflectance ~ U[0,1), rays without a hit inside `max_range` are dropped.  The
result is a float32 (N, 4) array [x, y, z, r] with N ~= 120k for the default settings and roughly
half the points inside the Car crop - the shape of a KITTI velodyne `.bin`
(reference loader: voxelnet/dataset.py:51).

Uniform-random points would give ~115k voxels for 120k points; real scans give ~13k
(SURVEY section 8), which is why the generator ray-casts instead.
"""
from __future__ import annotations

import numpy as np


def lidar_frame(
    seed: int = 0,
    beams: int = 64,
    azimuth_steps: int = 1900,
    n_boxes: int = 40,
    foliage: float = 0.08,
    max_range: float = 120.0,
    clutter: float = 0.0,
) -> np.ndarray:
    """One synthetic frame.  `clutter` > 0 adds that fraction of extra near-field scatter
    points (used by the dense-scene stress config to push the voxel count up)."""
    rng = np.random.default_rng(seed)
    elev = np.deg2rad(np.linspace(2.0, -24.8, beams, dtype=np.float64))
    azim = np.linspace(-np.pi, np.pi, azimuth_steps, endpoint=False, dtype=np.float64)
    azim = azim + rng.uniform(0, 2 * np.pi / azimuth_steps)
    ce, se = np.cos(elev)[:, None], np.sin(elev)[:, None]
    dx = (ce * np.cos(azim)[None, :]).ravel()
    dy = (ce * np.sin(azim)[None, :]).ravel()
    dz = np.broadcast_to(se, (beams, azimuth_steps)).ravel()
    n = dx.size

    t_hit = np.full(n, np.inf)
    # ground plane z = -1.73 (+ gentle undulation so voxels are not all in one z-slab)
    ground_z = -1.73
    down = dz < -1e-6
    t_ground = np.where(down, ground_z / np.where(down, dz, -1.0), np.inf)
    t_hit = np.minimum(t_hit, t_ground)

    # boxes: (cx, cy, cz, lx, ly, lz); cars in front, larger walls/buildings around
    boxes = []
    for _ in range(n_boxes):
        if rng.random() < 0.7:
            lx, ly, lz = (3.9, 1.6, 1.56) if rng.random() < 0.5 else (1.6, 3.9, 1.56)
            cx, cy = rng.uniform(4, 65), rng.uniform(-30, 30)
        else:
            lx, ly, lz = rng.uniform(4, 20), rng.uniform(4, 20), rng.uniform(2.5, 6)
            r, a = rng.uniform(15, 60), rng.uniform(-np.pi, np.pi)
            cx, cy = r * np.cos(a), r * np.sin(a)
        boxes.append((cx, cy, ground_z + lz / 2, lx, ly, lz))
    # far walls closing the scene so that most upward rays still return
    for (cx, cy, lx, ly) in ((80, 0, 1, 200), (-80, 0, 1, 200), (0, 80, 200, 1), (0, -80, 200, 1)):
        boxes.append((cx, cy, 4.0, lx, ly, 14.0))
    b = np.asarray(boxes, dtype=np.float64)
    lo = b[:, :3] - b[:, 3:] / 2
    hi = b[:, :3] + b[:, 3:] / 2
    d = np.stack([dx, dy, dz], axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / d
    for j in range(len(b)):  # slab test, ray origin = 0
        t0 = lo[j][None, :] * inv
        t1 = hi[j][None, :] * inv
        tmin = np.nanmax(np.minimum(t0, t1), axis=1)
        tmax = np.nanmin(np.maximum(t0, t1), axis=1)
        ok = (tmax >= tmin) & (tmax > 0.5)
        t_hit = np.where(ok & (tmin > 0.5), np.minimum(t_hit, tmin), t_hit)

    # vegetation / semi-transparent returns: a fraction of rays comes back early, which scatters
    # points through the volume the way trees and poles do in real scans
    early = rng.random(n) < foliage
    t_hit = np.where(early, t_hit * rng.uniform(0.25, 1.0, n), t_hit)

    keep = np.isfinite(t_hit) & (t_hit < max_range)
    t = t_hit[keep] + rng.normal(0.0, 0.02, int(keep.sum()))
    pts = np.stack([dx[keep] * t, dy[keep] * t, dz[keep] * t, rng.random(t.size)], axis=1)

    if clutter > 0:
        m = int(clutter * len(pts))
        c = np.stack(
            [rng.uniform(0, 70.4, m), rng.uniform(-40, 40, m), rng.uniform(-2.6, 0.6, m), rng.random(m)], axis=1
        )
        pts = np.concatenate([pts, c], axis=0)
        pts = pts[rng.permutation(len(pts))]
    return np.ascontiguousarray(pts.astype(np.float32))


def lidar_batch(batch: int, seed0: int = 0, **kw) -> list:
    """`batch` frames with seeds seed0, seed0+1, ..."""
    return [lidar_frame(seed0 + i, **kw) for i in range(batch)]
