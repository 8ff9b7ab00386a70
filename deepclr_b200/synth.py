"""Seeded synthetic registration pairs shaped like the reference's two datasets.

numpy.random.Generator(PCG64(seed)), fp32.  KITTI-shaped: an HDL-64-like scan (64 beams, ground plane 1.73 m
below the sensor, 30 % obstacle returns) + intensity; the source cloud is an independently re-sampled scan of the
same model moved by a small rigid motion (translation U(-1,1) m, Euler U(-1,1) deg) with N(0, 0.01 m) noise, the
recipe of the reference's configs/training/kitti_pairs.yaml:22-30.  ModelNet40-shaped: points on a random union
of boxes/cylinders scaled into the unit sphere; source = moved (U(-0.1,0.1), U(-5,5) deg) noisy (N(0,0.02)) copy,
configs/training/modelnet40.yaml:11-21; `partial=True` additionally crops the source with a random plane keeping
70 % and re-samples to N (BASELINE.json configs[3]).
"""
import numpy as np


def _euler_to_mat(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def random_motion(rng, trans_scale, rot_deg_scale):
    t = rng.uniform(-trans_scale, trans_scale, 3)
    e = np.deg2rad(rng.uniform(-rot_deg_scale, rot_deg_scale, 3))
    m = np.eye(4)
    m[:3, :3] = _euler_to_mat(*e)
    m[:3, 3] = t
    return m


def _lidar_scan(rng, n):
    theta = rng.uniform(0.0, 2.0 * np.pi, n)
    beams = np.deg2rad(np.linspace(-24.8, 2.0, 64))
    phi = beams[rng.integers(0, 64, n)]
    down = phi < 0
    rng_m = np.where(down, np.clip(1.73 / np.sin(np.where(down, -phi, 1.0)), 2.0, 80.0), rng.uniform(5.0, 80.0, n))
    obstacle = rng.uniform(0.0, 1.0, n) < 0.3
    rng_m = np.where(obstacle, rng.uniform(3.0, 40.0, n), rng_m)
    xyz = np.stack([rng_m * np.cos(phi) * np.cos(theta), rng_m * np.cos(phi) * np.sin(theta), rng_m * np.sin(phi)], 1)
    return xyz, rng.uniform(0.0, 1.0, n)


def kitti_pair(seed: int, n: int = 16384):
    """-> template [n,4], source [n,4] float32, motion 4x4 float64 (source = motion applied to a re-sampled scan)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    t_xyz, t_i = _lidar_scan(rng, n)
    s_xyz, s_i = _lidar_scan(rng, n)
    m = random_motion(rng, 1.0, 1.0)
    s_xyz = s_xyz @ m[:3, :3].T + m[:3, 3]
    t_xyz = t_xyz + rng.normal(0.0, 0.01, t_xyz.shape)
    s_xyz = s_xyz + rng.normal(0.0, 0.01, s_xyz.shape)
    tmpl = np.concatenate([t_xyz, t_i[:, None]], 1).astype(np.float32)
    src = np.concatenate([s_xyz, s_i[:, None]], 1).astype(np.float32)
    return tmpl, src, m


def _shape_points(rng, n):
    parts = int(rng.integers(3, 7))
    per = np.full(parts, n // parts)
    per[: n - per.sum()] += 1
    pts = []
    for c in per:
        centre = rng.uniform(-0.5, 0.5, 3)
        size = rng.uniform(0.1, 0.5, 3)
        if rng.uniform() < 0.5:                      # box surface
            p = rng.uniform(-1.0, 1.0, (c, 3))
            ax = rng.integers(0, 3, c)
            p[np.arange(c), ax] = np.sign(p[np.arange(c), ax])
        else:                                        # cylinder side
            a = rng.uniform(0.0, 2.0 * np.pi, c)
            p = np.stack([np.cos(a), np.sin(a), rng.uniform(-1.0, 1.0, c)], 1)
        pts.append(p * size + centre)
    p = np.concatenate(pts, 0)
    p = p - p.mean(0)
    return p / np.linalg.norm(p, axis=1).max()


def modelnet_pair(seed: int, n: int = 1024, partial: bool = False, noise: float = 0.02):
    """-> template [n,3], source [n,3] float32, motion 4x4."""
    rng = np.random.Generator(np.random.PCG64(seed))
    t_xyz = _shape_points(rng, n)
    m = random_motion(rng, 0.1, 5.0)
    s_xyz = t_xyz[rng.permutation(n)] @ m[:3, :3].T + m[:3, 3]
    if partial:
        normal = rng.normal(size=3)
        normal /= np.linalg.norm(normal)
        d = s_xyz @ normal
        keep = s_xyz[d <= np.quantile(d, 0.7)]
        s_xyz = keep[rng.integers(0, keep.shape[0], n)]
    t_xyz = t_xyz + rng.normal(0.0, noise, t_xyz.shape)
    s_xyz = s_xyz + rng.normal(0.0, noise, s_xyz.shape)
    return t_xyz.astype(np.float32), s_xyz.astype(np.float32), m


def batch(kind: str, num_pairs: int, seed0: int = 0, n=None, **kw) -> np.ndarray:
    """Reference batch layout (deepclr/data/build.py:82, base.py:122-136): [2B,N,C], templates first."""
    gen = kitti_pair if kind == 'kitti' else modelnet_pair
    args = {} if n is None else {'n': n}
    pairs = [gen(seed0 + i, **args, **kw) for i in range(num_pairs)]
    return np.ascontiguousarray(np.stack([p[0] for p in pairs] + [p[1] for p in pairs], 0))
