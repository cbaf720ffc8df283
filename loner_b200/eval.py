"""Eval / render drivers as batched device loops.

Mirrors the per-frame render helpers of the reference's analysis scripts on top of `Model.forward(testing=True)`:
  * render_frame            analysis/renderer.py:181-215 (render_dataset_frame): colour, depth and the
                            peak-depth-consistency |z[argmax w] - depth| images, one chunk loop on the device
  * render_lidar_depth      analysis/renderer_lidar.py:71-93: depth / variance of a LiDAR-shaped ray bundle
  * l1_depth_error          fdt_optimize_implicit_map_utils.py:260-282: mean |depth - depth_gt| over rays with a return
  * build_lidar_scan        analysis/mesher.py:29-50 / renderer_lidar.py:39-60: the synthetic 360-degree scan pattern
  * Mesher.get_grid_uniform / Mesher.splat_volume   analysis/mesher.py:70-180: the occupancy-like volume marching
                            cubes runs on (max compositing weight per voxel over virtual scans from the keyframe
                            poses); lb_voxel_splat_max does the bucketise + max on the device.  The marching-cubes
                            call itself (skimage, absent from this image) and the Open3D mesh object are not built.
The reference runs these in chunks of 512-1024 rays with a host round trip per chunk; here a chunk is
`cfg.render.chunk` rays (16384) and nothing leaves the device until the images are complete.
"""
import numpy as np
import torch

from . import ops


@torch.no_grad()
def render_frame(model, ray_sampler, rays, scale_factor, camera=True, render_strategy='default'):
    """rays [N,13] (one image or scan, any N) -> dict(rgb [N,C] or None, depth [N] (metres),
    peak_depth_consistency [N] (metres), variance [N], opacity [N])."""
    keep = model.cfg.render.retraw
    model.cfg["render"]["retraw"] = True              # samples_fine is needed for the peak metric
    try:
        res = model(rays, ray_sampler, scale_factor, testing=True, camera=camera, return_variance=True,
                    render_strategy=render_strategy)
    finally:
        model.cfg["render"]["retraw"] = keep
    depth = res['depth_fine']
    s_peaks = torch.gather(res['samples_fine'], 1, res['weights_fine'].argmax(dim=1, keepdim=True))[:, 0]
    out = dict(depth=depth * scale_factor, peak_depth_consistency=(s_peaks - depth).abs() * scale_factor,
               variance=res['variance'], opacity=res['opacity_fine'],
               rgb=res['rgb_fine'].clamp(0, 1) if camera else None)
    return out


@torch.no_grad()
def render_lidar_depth(model, ray_sampler, rays, scale_factor, render_strategy='default'):
    """Depth (metres) and variance of LiDAR rays, sigma head only."""
    r = render_frame(model, ray_sampler, rays, scale_factor, camera=False, render_strategy=render_strategy)
    return r["depth"], r["variance"]


@torch.no_grad()
def l1_depth_error(model, ray_sampler, rays, depths_gt, scale_factor, render_strategy='default'):
    """Mean absolute depth error in metres over rays with a valid return (depth_gt > 0 and inside the ray range)."""
    depth, _ = render_lidar_depth(model, ray_sampler, rays, scale_factor, render_strategy)
    gt = depths_gt * scale_factor
    ok = (depths_gt > 0) & (depths_gt <= rays[:, -1])
    return ((depth - gt).abs() * ok).sum() / ok.sum().clamp_min(1)


def build_lidar_scan(lidar_intrinsics, device="cuda"):
    """reference analysis/mesher.py:29-50: unit directions [3, n] of a spinning LiDAR with the given vertical FOV
    (degrees) and angular resolutions; n = n_elevations * n_azimuths, azimuth fastest."""
    vert_fov = lidar_intrinsics["vertical_fov"]
    phi = torch.arange(vert_fov[0], vert_fov[1], lidar_intrinsics["vertical_resolution"]).deg2rad()
    theta = torch.arange(0, 360, lidar_intrinsics["horizontal_resolution"]).deg2rad()
    phi_grid, theta_grid = torch.meshgrid(phi, theta, indexing="ij")
    phi_grid = torch.pi / 2 - phi_grid.reshape(-1, 1)
    theta_grid = theta_grid.reshape(-1, 1)
    xyz = torch.hstack((torch.cos(theta_grid) * torch.sin(phi_grid), torch.sin(theta_grid) * torch.sin(phi_grid),
                        torch.cos(phi_grid)))
    return xyz.T.contiguous().to(device)


class Mesher:
    """reference analysis/mesher.py:53-180 up to the volume handed to marching cubes."""

    def __init__(self, model, world_cube, ray_range, resolution=0.2, marching_cubes_bound=((-40, 20), (0, 20), (-3, 15)),
                 lidar_vertical_fov=(-22.5, 22.5), chunk_size=512):
        self.marching_cubes_bound = np.array(marching_cubes_bound, dtype=np.float64)
        self.world_cube = world_cube
        self.world_cube_shift = torch.as_tensor(world_cube.shift).detach().cpu().numpy()
        self.world_cube_scale_factor = float(world_cube.scale_host)
        self.model = model
        self.resolution = resolution
        self.ray_range = tuple(float(v) for v in ray_range)
        self.lidar_vertical_fov = lidar_vertical_fov
        self.chunk_size = chunk_size

    def get_grid_uniform(self, resolution=None):
        """mesher.py:70-96: axis coordinates of the marching-cubes grid in cube units (fp64)."""
        resolution = self.resolution if resolution is None else resolution
        bound = (self.marching_cubes_bound + np.expand_dims(self.world_cube_shift, 1)) / self.world_cube_scale_factor
        length = self.marching_cubes_bound[:, 1] - self.marching_cubes_bound[:, 0]
        num = (length / resolution).astype(int)
        return [np.linspace(bound[a][0], bound[a][1], num[a]) for a in range(3)]

    @torch.no_grad()
    def splat_volume(self, ray_sampler, lidar_poses, var_threshold=None, lidar_intrinsics=None):
        """mesher.py:100-180: virtual scans from `lidar_poses` (pose 6-vectors), Model.forward in training mode
        (N_samples_train, perturbation and noise on, as the reference calls it), rays kept when depth <
        ray_range[1] - 0.25 (and variance < var_threshold), max weight per voxel.  Returns (volume [nx, ny, nz] fp32
        on the device -- the array the reference passes to marching cubes -- and the three axis coordinate arrays)."""
        from .mapping import build_lidar_rays
        dev = next(self.model.parameters()).device
        xyz = self.get_grid_uniform()
        bx, by, bz = (torch.from_numpy(a).contiguous().to(dev) for a in xyz)
        nx, ny, nz = len(xyz[0]), len(xyz[1]), len(xyz[2])
        volume = torch.zeros(ny, nx, nz, device=dev, dtype=torch.float32)
        intr = lidar_intrinsics or {"vertical_fov": self.lidar_vertical_fov, "vertical_resolution": 0.25,
                                    "horizontal_resolution": 0.25}
        directions = build_lidar_scan(intr, dev)
        keep = self.model.cfg.render.retraw
        self.model.cfg["render"]["retraw"] = True
        try:
            for pose6 in lidar_poses:
                rays, _ = build_lidar_rays(directions, None, pose6.to(dev).float(), self.world_cube, self.ray_range,
                                           ignore_world_cube=False, const_depth=1.0)   # fetch_chunk_rays keeps the validity filter
                res = self.model(rays, ray_sampler, self.world_cube_scale_factor, testing=False, camera=False,
                                 return_variance=True)
                ok = res["depth_fine"].view(-1) < self.ray_range[1] - 0.25
                if var_threshold is not None:
                    ok = ok & (res["variance"].view(-1) < var_threshold)
                ops.voxel_splat_max(rays, res["samples_fine"], res["weights_fine"], bx, by, bz, volume, ray_mask=ok)
        finally:
            self.model.cfg["render"]["retraw"] = keep
        return volume.permute(1, 0, 2), xyz
