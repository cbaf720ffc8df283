"""Synthetic Ouster-OS1-64-shaped LiDAR keyframes (64 x 1024 beams) for benchmarks and parity tests.

No ROS, no datasets: ranges come from analytic ray casting against a closed room with vertical
cylinders ("trunks").  Shapes follow the reference's inputs:
  * beam pattern: 64 elevations uniform in [-22.5, +22.5] deg x 1024 azimuths
    (reference cfg/fusion_portable/canteen.yaml:28, analysis/renderer_lidar.py:39-60)
  * LidarScan fields: ray_directions [3,n], distances [n]  (reference src/common/sensors.py:57-231)
  * world cube = bounding cube of the trajectory + ray range, `scale_factor`, `shift`
    (reference src/common/pose_utils.py:23-57,222-314)
Pure torch (runs on CPU or CUDA tensors); deterministic for a given seed.
"""
import math
from dataclasses import dataclass

import torch

OS1_ROWS, OS1_COLS = 64, 1024


def os1_directions(rows=OS1_ROWS, cols=OS1_COLS, fov_deg=(-22.5, 22.5)):
    """[3, rows*cols] unit vectors, column-major over azimuth like a spinning scan."""
    elev = torch.deg2rad(torch.linspace(fov_deg[0], fov_deg[1], rows))
    azim = torch.linspace(0, 2 * math.pi, cols + 1)[:-1]
    az, el = torch.meshgrid(azim, elev, indexing="ij")          # [cols, rows]
    d = torch.stack([torch.cos(el) * torch.cos(az), torch.cos(el) * torch.sin(az), torch.sin(el)], 0)
    return d.reshape(3, -1).float()


@dataclass
class Scene:
    half_extent: tuple = (18.0, 14.0)      # room half sizes in x, y (metres)
    floor: float = -1.8
    ceiling: float = 9.0
    trunks: tuple = ((5.0, 3.0, 0.4), (-6.0, -4.0, 0.55), (9.0, -7.0, 0.35), (-3.0, 8.0, 0.5),
                     (12.0, 6.0, 0.45), (-11.0, 2.0, 0.6))          # (x, y, radius)


def cast(scene, origin, dirs):
    """Analytic ranges for rays origin[3] + t*dirs[3,n]; returns t [n] (inf where nothing is hit)."""
    ox, oy, oz = origin[0], origin[1], origin[2]
    dx, dy, dz = dirs[0], dirs[1], dirs[2]
    inf = torch.full_like(dx, float("inf"))
    eps = 1e-9

    def plane(o, d, c):
        t = (c - o) / torch.where(d.abs() < eps, torch.full_like(d, eps), d)
        return torch.where(t > 1e-4, t, inf)

    hx, hy = scene.half_extent
    t = torch.minimum(plane(ox, dx, hx), plane(ox, dx, -hx))
    t = torch.minimum(t, torch.minimum(plane(oy, dy, hy), plane(oy, dy, -hy)))
    t = torch.minimum(t, torch.minimum(plane(oz, dz, scene.floor), plane(oz, dz, scene.ceiling)))
    for (cx, cy, r) in scene.trunks:
        a = dx * dx + dy * dy
        b = 2 * ((ox - cx) * dx + (oy - cy) * dy)
        c = (ox - cx) ** 2 + (oy - cy) ** 2 - r * r
        disc = b * b - 4 * a * c
        ok = (disc > 0) & (a > eps)
        sq = torch.sqrt(disc.clamp(min=0))
        t0 = (-b - sq) / (2 * a.clamp(min=eps))
        tc = torch.where(ok & (t0 > 1e-4), t0, inf)
        zc = oz + tc * dz
        tc = torch.where((zc > scene.floor) & (zc < scene.ceiling), tc, inf)
        t = torch.minimum(t, tc)
    return t


def arc_poses(n_kf, spacing=1.0, radius=12.0):
    """n_kf pose 6-vectors [t, axis-angle] on a smooth arc, ~`spacing` metres apart, yawing along it."""
    poses = []
    for k in range(n_kf):
        th = (k - (n_kf - 1) / 2) * spacing / radius
        x, y = radius * math.sin(th), radius * (1 - math.cos(th)) - 2.0
        poses.append(torch.tensor([x, y, 0.0, 0.0, 0.0, th], dtype=torch.float32))
    return poses


@dataclass
class SynthKeyFrame:
    directions: torch.Tensor      # [3, n] sensor-frame unit vectors of returns
    distances: torch.Tensor       # [n] metres
    sky_directions: torch.Tensor  # [3, m] beams with no return inside ray_range
    pose: torch.Tensor            # [6]  (t, axis-angle), world frame


def make_keyframes(n_kf, ray_range=(1.0, 50.0), seed=0, scene=None, noise_std=0.02, poses=None):
    """n_kf synthetic 64x1024 scans. Returns (list[SynthKeyFrame], scale_factor, shift)."""
    scene = scene or Scene()
    g = torch.Generator().manual_seed(seed)
    dirs = os1_directions()
    poses = poses if poses is not None else arc_poses(n_kf)
    kfs = []
    for p in poses:
        yaw = float(p[5])
        c, s = math.cos(yaw), math.sin(yaw)
        R = torch.tensor([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])
        t = cast(scene, p[:3], R @ dirs)
        t = t + noise_std * torch.randn(t.shape, generator=g)
        hit = torch.isfinite(t) & (t > ray_range[0]) & (t < ray_range[1])
        kfs.append(SynthKeyFrame(dirs[:, hit].contiguous(), t[hit].float().contiguous(),
                                 dirs[:, ~hit].contiguous(), p.clone()))
    # world cube: bounding cube of trajectory +- max range, padded (reference pose_utils.py:222-314)
    ts = torch.stack([p[:3] for p in poses])
    lo, hi = ts.min(0)[0] - ray_range[1], ts.max(0)[0] + ray_range[1]
    centre = 0.5 * (lo + hi)
    scale = float(0.5 * (hi - lo).max() * 1.3)
    return kfs, torch.tensor(scale), (-centre).float()
