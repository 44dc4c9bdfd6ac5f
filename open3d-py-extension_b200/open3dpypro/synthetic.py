"""Synthetic LiDAR-like scans.  The reference ships no generator; this is the
definition of the benchmark inputs: a ground plane with 1/r^2 density, vertical rectangles,
small spheres/cylinders and uniform clutter, cast to float32 and ordered by azimuth."""
import numpy as np

SEED_C2, SEED_C3, SEED_C5 = 1234, 1235, 1236


def frame_seed(f):
    return 2000 + int(f)


def lidar_like(n, seed=SEED_C2, dtype=np.float64):
    """(n,3) points, float32-representable, returned as `dtype`."""
    n = int(n)
    rng = np.random.Generator(np.random.Philox(int(seed)))
    n_ground = int(0.55 * n)
    n_wall = int(0.30 * n)
    n_obj = int(0.10 * n)
    n_clut = n - n_ground - n_wall - n_obj
    parts = []
    # ground: range log-uniform in [2,100] m, azimuth uniform, z = 0.01 x + N(0, 0.02)
    r = np.exp(rng.uniform(np.log(2.0), np.log(100.0), n_ground))
    a = rng.uniform(-np.pi, np.pi, n_ground)
    x, y = r * np.cos(a), r * np.sin(a)
    parts.append(np.stack([x, y, 0.01 * x + rng.normal(0.0, 0.02, n_ground)], 1))
    # 40 vertical rectangles
    nw = 40
    wc_r = np.exp(rng.uniform(np.log(2.0), np.log(100.0), nw))
    wc_a = rng.uniform(-np.pi, np.pi, nw)
    w_dir = rng.uniform(-np.pi, np.pi, nw)
    w_w = rng.uniform(5.0, 30.0, nw)
    w_h = rng.uniform(2.0, 8.0, nw)
    which = rng.integers(0, nw, n_wall)
    u = rng.uniform(-0.5, 0.5, n_wall) * w_w[which]
    h = rng.uniform(0.0, 1.0, n_wall) * w_h[which]
    off = rng.normal(0.0, 0.02, n_wall)
    cx, cy = wc_r[which] * np.cos(wc_a[which]), wc_r[which] * np.sin(wc_a[which])
    dx, dy = np.cos(w_dir[which]), np.sin(w_dir[which])
    parts.append(np.stack([cx + u * dx - off * dy, cy + u * dy + off * dx, h], 1))
    # 200 spheres / cylinders of radius 0.3 .. 2 m
    no = 200
    oc_r = np.exp(rng.uniform(np.log(2.0), np.log(100.0), no))
    oc_a = rng.uniform(-np.pi, np.pi, no)
    o_rad = rng.uniform(0.3, 2.0, no)
    o_cyl = rng.integers(0, 2, no)
    which = rng.integers(0, no, n_obj)
    th = rng.uniform(-np.pi, np.pi, n_obj)
    ph = np.arccos(rng.uniform(-1.0, 1.0, n_obj))
    rad = o_rad[which]
    cx, cy = oc_r[which] * np.cos(oc_a[which]), oc_r[which] * np.sin(oc_a[which])
    cyl = o_cyl[which] == 1
    sx = np.where(cyl, rad * np.cos(th), rad * np.sin(ph) * np.cos(th))
    sy = np.where(cyl, rad * np.sin(th), rad * np.sin(ph) * np.sin(th))
    sz = np.where(cyl, rng.uniform(0.0, 3.0, n_obj), rad + rad * np.cos(ph))
    parts.append(np.stack([cx + sx, cy + sy, sz], 1))
    # clutter
    parts.append(np.stack([rng.uniform(-100, 100, n_clut), rng.uniform(-100, 100, n_clut), rng.uniform(0, 10, n_clut)], 1))
    p = np.concatenate(parts, 0).astype(np.float32)
    order = np.argsort(np.arctan2(p[:, 1], p[:, 0]), kind="stable")
    return np.ascontiguousarray(p[order].astype(dtype))


def frames(n_frames, n_points, first=0, dtype=np.float32):
    """(F,N,3) batch for the frame pipeline (config 4: 256 x 131072)."""
    return np.stack([lidar_like(n_points, frame_seed(first + f), dtype) for f in range(n_frames)], 0)
