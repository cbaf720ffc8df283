"""Eval / render drivers as batched device loops.

Mirrors the per-frame render helpers of the reference's analysis scripts on top of `Model.forward(testing=True)`:
  * render_frame            analysis/renderer.py:181-215 (render_dataset_frame): colour, depth and the
                            peak-depth-consistency |z[argmax w] - depth| images, one chunk loop on the device
  * render_lidar_depth      analysis/renderer_lidar.py:71-93: depth / variance of a LiDAR-shaped ray bundle
  * l1_depth_error          fdt_optimize_implicit_map_utils.py:260-282: mean |depth - depth_gt| over rays with a return
The reference runs these in chunks of 512-1024 rays with a host round trip per chunk; here a chunk is
`cfg.render.chunk` rays (16384) and nothing leaves the device until the images are complete.
"""
import torch


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
