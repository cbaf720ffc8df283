"""Host-side mirror of the reference's model layer (src/models/*) on top of the sm_100a kernels.

Same class / function names, argument meaning, result-dict keys and error behaviour as
  src/models/nerf_tcnn.py       DecoupledNeRF
  src/models/model_tcnn.py      Model, OccupancyGridModel
  src/models/ray_sampling.py    UniformRaySampler, OccGridRaySampler
  src/models/rendering_tcnn.py  render_rays, raw2outputs, raw2outputs_adjusted, inference, sample_pdf
  src/models/losses.py          get_weights_gt, get_logits_grad
so a caller of the reference can switch imports.  Every tensor op of the reference on this path is
one kernel call here; the chunk loop and the three RNG draws per chunk (u1, u2, noise -- same shapes,
same order, same torch generator) are kept so a seeded run consumes the same random stream.
"""
from collections import defaultdict

import torch
import torch.nn as nn

from . import ops, tcnn
from ._lib import LonerB200Error
from .functional import Composite, CompositeRGB, SigmaFromRays


class DecoupledNeRF(nn.Module):
    """src/models/nerf_tcnn.py:19-95.  Parameter names match the reference state_dict:
    _model_sigma.params, _pos_encoding.params, _dir_encoding.params, _model_intensity.params."""

    def __init__(self, cfg, num_colors=3):
        super().__init__()
        self._num_colors = num_colors
        self.cfg = cfg
        self._enable_view_dependence = cfg["enable_view_dependence"]
        self._model_sigma = tcnn.NetworkWithInputEncoding(
            n_input_dims=3, n_output_dims=1, encoding_config=cfg["pos_encoding_sigma"],
            network_config=cfg["sigma_network"])
        self._pos_encoding = tcnn.Encoding(3, cfg["pos_encoding_intensity"])
        if self._enable_view_dependence:
            self._dir_encoding = tcnn.Encoding(3, cfg["dir_encoding_intensity"])
            n_in = self._pos_encoding.n_output_dims + self._dir_encoding.n_output_dims
        else:
            self._dir_encoding = None
            n_in = self._pos_encoding.n_output_dims
        self._model_intensity = tcnn.Network(n_input_dims=n_in, n_output_dims=num_colors,
                                             network_config=cfg["intensity_network"])
        self._max_float = torch.finfo(self._model_intensity.dtype).max
        self._min_float = torch.finfo(self._model_intensity.dtype).min

    def sigma_from_rays(self, rays, z, sink=None):
        """Fused sample-point construction + sigma head: [N,13], [N,S] -> [N,S].
        (rendering_tcnn.py:390-392 + nerf_tcnn.py:63,71-72; the intensity hash grid that the reference
        evaluates and discards on sigma-only calls, nerf_tcnn.py:64, is skipped.)"""
        return SigmaFromRays.apply(self._model_sigma.plan, self._model_sigma.params, rays, z, sink)

    def forward(self, pos, dir, sigma_only=False, detach_sigma=True):
        pos = (pos + 1) / 2
        if detach_sigma and not sigma_only:
            with torch.no_grad():
                sigma = self._model_sigma(pos)[..., [0]]
        else:
            sigma = self._model_sigma(pos)[..., [0]]
        # non-finite outputs are clamped like nerf_tcnn.py:74-78, without the host sync
        sigma = sigma.nan_to_num(posinf=self._max_float, neginf=self._min_float)
        if sigma_only:
            return sigma
        # intensity head (nerf_tcnn.py:64,83-95): 2nd hash grid (+ SH-4 of the view direction) -> fused MLP; the
        # torch.sigmoid of :93 runs in the network kernel's output epilogue
        h = self._pos_encoding(pos)
        if self._enable_view_dependence:
            h = torch.cat([h, self._dir_encoding((dir + 1) / 2)], dim=-1)
        color = self._model_intensity(h, out_act=ops.OUT_ACT["Sigmoid"])
        return torch.cat([color, sigma], dim=-1)


class OccupancyGridModel(nn.Module):
    """src/models/model_tcnn.py:111-134: D^3 log-odds parameter + trilinear lookup."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        v = cfg.voxel_size if hasattr(cfg, "voxel_size") else cfg["voxel_size"]
        self.occupancy_grid = nn.Parameter(torch.zeros(1, 1, v, v, v))

    def forward(self):
        return self.occupancy_grid

    @staticmethod
    def interpolate(occupancy_grid, ray_bin_centers, mode='bilinear'):
        n_rays, n_bins, _ = ray_bin_centers.shape
        grid_values = ray_bin_centers.reshape(1, 1, n_rays, n_bins, 3)
        return nn.functional.grid_sample(occupancy_grid, grid_values, mode=mode,
                                         align_corners=False).reshape(n_rays, n_bins)


class UniformRaySampler:
    """src/models/ray_sampling.py:18-43."""

    def __init__(self):
        pass

    def get_samples(self, rays, N_samples, perturb, rows=None):
        """rows = (lo, hi): the random draw covers all of `rays` (same generator consumption as a single-GPU run),
        the depths are produced for rays[lo:hi] only (RayDataParallel.shard_batch)."""
        r = rays.detach().contiguous()
        u = torch.rand((r.shape[0], N_samples), device=r.device) if perturb > 0 else None
        if rows is not None:
            r = r[rows[0]:rows[1]].contiguous()
            u = u[rows[0]:rows[1]].contiguous() if u is not None else None
        return ops.sample_uniform(r, u, N_samples, perturb)


class OccGridRaySampler:
    """src/models/ray_sampling.py:46-92: S/2 stratified + S/2 occupancy-PDF samples, sorted; no grad."""

    def __init__(self):
        self._occ_gamma = None

    def update_occ_grid(self, occ_gamma):
        self._occ_gamma = occ_gamma

    def get_samples(self, rays, N_samples, perturb, debug=False, rows=None):
        """rows = (lo, hi): see UniformRaySampler.get_samples."""
        r = rays.detach().contiguous()
        s2 = N_samples // 2
        u1 = torch.rand((r.shape[0], s2), device=r.device) if perturb > 0 else None   # ray_sampling.py:71-72
        u2 = torch.rand(r.shape[0], s2, device=r.device)                                # rendering_tcnn.py:49
        if rows is not None:
            r = r[rows[0]:rows[1]].contiguous()
            u1 = u1[rows[0]:rows[1]].contiguous() if u1 is not None else None
            u2 = u2[rows[0]:rows[1]].contiguous()
        return ops.sample_occ(r, self._occ_gamma.detach().contiguous(), u1, u2, N_samples, perturb, debug=debug)


def get_weights_gt(sampled_depth, gt_depth, eps, norm=True):
    """src/models/losses.py:29-51 (kept in torch for callers outside the fused loss)."""
    sigma = eps / 9
    t = (sampled_depth - gt_depth) / sigma
    w = 0.3989422804014327 * torch.exp(-0.5 * t * t) / sigma          # Phi(9)-Phi(-9) == 1 in fp32
    w = w * ((sampled_depth > gt_depth - eps) & (sampled_depth < gt_depth + eps)).to(w.dtype)
    if norm:
        w = w / (w.sum(dim=1, keepdim=True) + 1e-6)
    return w


def get_logits_grad(z_vals, depth, eps=2, l_free=0.25, l_occ=2.5):
    """src/models/losses.py:54-62."""
    x = z_vals - depth
    return l_free * (x < -eps).to(x.dtype) - l_occ * ((x > -eps) & (x < eps)).to(x.dtype)


def sample_pdf(bins, weights, N_importance, det=False, eps=1e-5):
    """src/models/rendering_tcnn.py:19-68 is fused into the sampling kernel (lb_sample_occ); a
    stand-alone inverse-CDF draw is not part of the mapping path."""
    raise LonerB200Error("sample_pdf is fused into OccGridRaySampler.get_samples (lb_sample_occ)")


def raw2outputs(raw, z_vals, rays_d, raw_noise_std=0, white_bkgd=False, sigma_only=False, num_colors=3,
                softplus=False, far=None, ret_var=False, rays=None):
    """src/models/rendering_tcnn.py:219-295 for sigma-only raw.  `rays` [N,13] may be passed instead of
    (rays_d, far) to avoid re-packing; gradients flow to sigma, rays_d (|d|) and far."""
    if softplus:
        raise LonerB200Error("raw2outputs: softplus=True is never used by the reference path")
    sigma = raw[..., 0] if sigma_only else raw[..., num_colors]
    if rays is None:
        n = z_vals.shape[0]
        pad = torch.zeros(n, 3, device=z_vals.device, dtype=torch.float32)
        far_col = far if far is not None else torch.zeros(n, 1, device=z_vals.device)
        rays = torch.cat([pad, rays_d, pad, pad[:, :2], pad[:, :1], far_col], 1)
    noise = torch.randn(sigma.shape, device=sigma.device) * raw_noise_std if raw_noise_std > 0 else None
    w, depth, opac, var = Composite.apply(sigma, z_vals, rays, noise)
    if far is None and rays_d is not None:
        depth = torch.sum(w * z_vals, -1)
    rgb = torch.tensor([-1.]) if sigma_only else CompositeRGB.apply(w, raw[..., :num_colors], bool(white_bkgd))
    return rgb, depth, w, opac, (var if ret_var else None)


def raw2outputs_adjusted(raw, z_vals, rays_o, rays_d, raw_noise_std=0, white_bkgd=False, sigma_only=False,
                         num_colors=3, softplus=False, far=None, ret_var=True, rays=None):
    """src/models/rendering_tcnn.py:70-214, effective semantics: noise forced to 0 (:104), depth = z at
    the sample where transmittance crosses 0.5 (:130-132) else 0; the peak/prominence write-backs
    (:196-199) hit a temporary copy in the reference and change nothing.  Inference only."""
    sigma = (raw[..., 0] if sigma_only else raw[..., num_colors]).detach().contiguous()
    if rays is None:
        n = z_vals.shape[0]
        pad = torch.zeros(n, 3, device=z_vals.device, dtype=torch.float32)
        rays = torch.cat([pad, rays_d, pad, pad[:, :2], pad[:, :2]], 1)
    if raw_noise_std > 0:      # the reference still draws (and discards) the noise: keep the RNG stream aligned
        torch.randn(sigma.shape, device=sigma.device)
    w, depth, opac, var = ops.composite_fwd(sigma, z_vals.detach().contiguous(), rays.detach().contiguous(), None,
                                            "adjusted", True)
    rgb = torch.tensor([-1.]) if sigma_only else CompositeRGB.apply(w, raw[..., :num_colors], bool(white_bkgd))
    return rgb, depth, w, opac, (var if ret_var else None)


def inference(model, xyz_, dir_, sigma_only=False, netchunk=32768, detach_sigma=True, meshing=False):
    """src/models/rendering_tcnn.py:297-335 (netchunk > 0 re-runs the full batch per chunk in the
    reference, :327-330; only netchunk == 0 is meaningful and supported)."""
    if netchunk != 0:
        raise LonerB200Error("inference: only netchunk == 0 is supported (reference bug at rendering_tcnn.py:327-330)")
    n_rays, n_samples = xyz_.shape[0:2]
    xyz = xyz_.reshape(-1, 3).contiguous()
    d = None if sigma_only else torch.repeat_interleave(dir_, repeats=n_samples, dim=0).contiguous()
    out = model(xyz, d, sigma_only, detach_sigma)
    return out if meshing else out.view(n_rays, n_samples, -1)


_EVAL_COLOUR_BLOCK_POINTS = 1 << 23       # sample points per block of the colour head in inference (render_rays)


def render_rays(rays, ray_sampler, nerf_model, ray_range, scale_factor, N_samples=64, retraw=False, perturb=0,
                white_bkgd=False, raw_noise_std=0., netchunk=32768, num_colors=3, sigma_only=False, DEBUG=True,
                detach_sigma=True, return_variance=False, render_strategy='default'):
    """src/models/rendering_tcnn.py:340-425.  Three kernels per chunk: sample -> fused field -> composite.
    The reference's NaN/inf prints (:419-424, eight host syncs per chunk) are dropped."""
    if render_strategy not in ('default', 'adjusted'):
        raise ValueError(f"Unknown render strategy: {render_strategy}")
    if netchunk != 0:
        raise LonerB200Error("render_rays: only netchunk == 0 is supported (reference bug at rendering_tcnn.py:327-330)")
    rays = rays.float()
    z_vals = ray_sampler.get_samples(rays, N_samples, perturb)
    if sigma_only:
        raw = nerf_model.sigma_from_rays(rays, z_vals).unsqueeze(-1)
    else:
        # camera rays (rendering_tcnn.py:387-392): colour + sigma per sample point through DecoupledNeRF.forward
        block = _EVAL_COLOUR_BLOCK_POINTS // max(1, N_samples)
        if torch.is_grad_enabled() or rays.shape[0] <= block:
            xyz = rays[:, None, 0:3] + rays[:, None, 3:6] * z_vals.unsqueeze(2)
            raw = inference(nerf_model, xyz, rays[:, 6:9], sigma_only=False, netchunk=0, detach_sigma=detach_sigma)
        else:
            # inference (eval renders: 16384 rays x 2048 samples per chunk): the per-point intermediates of the colour
            # head (positions, both encodings, their concatenation: ~450 B per point) are produced for blocks of rays
            # instead of the whole chunk -- same numbers, 15 GB of temporaries per chunk become < 4 GB
            raw = torch.empty(rays.shape[0], N_samples, num_colors + 1, device=rays.device, dtype=torch.float32)
            for b0 in range(0, rays.shape[0], block):
                r_b, z_b = rays[b0:b0 + block], z_vals[b0:b0 + block]
                xyz = r_b[:, None, 0:3] + r_b[:, None, 3:6] * z_b.unsqueeze(2)
                raw[b0:b0 + block] = inference(nerf_model, xyz, r_b[:, 6:9], sigma_only=False, netchunk=0, detach_sigma=detach_sigma)
    if render_strategy == 'default':
        rgb, depth, weights, opacity, variance = raw2outputs(
            raw, z_vals, None, raw_noise_std, white_bkgd, sigma_only=sigma_only, num_colors=num_colors,
            ret_var=return_variance, rays=rays)
    else:
        rgb, depth, weights, opacity, variance = raw2outputs_adjusted(
            raw, z_vals, None, None, raw_noise_std, white_bkgd, sigma_only=sigma_only, num_colors=num_colors,
            ret_var=return_variance, rays=rays)
    result = {'rgb_fine': rgb, 'depth_fine': depth, 'weights_fine': weights, 'opacity_fine': opacity}
    if return_variance:
        result["variance"] = variance
    if retraw:
        result['samples_fine'] = z_vals
        result['points_fine'] = rays[:, None, 0:3] + rays[:, None, 3:6] * z_vals.unsqueeze(2)
    if DEBUG:
        result['raw_fine'] = raw
    return result


class Model(nn.Module):
    """src/models/model_tcnn.py:22-108: holder of all trainables + chunked ray rendering."""

    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        if cfg.model_type == 'nerf_decoupled':
            self.nerf_model = DecoupledNeRF(cfg.nerf_config, cfg.num_colors)
        else:
            raise NotImplementedError()

    def get_rgb_parameters(self, ignore_requires_grad=False):
        n = self.nerf_model
        ps = list(n._model_intensity.parameters()) + list(n._pos_encoding.parameters()) + \
            ([] if n._dir_encoding is None else list(n._dir_encoding.parameters()))
        return ps if ignore_requires_grad else [p for p in ps if p.requires_grad]

    def get_rgb_mlp_parameters(self):
        return list(self.nerf_model._model_intensity.parameters())

    def get_rgb_feature_parameters(self):
        n = self.nerf_model
        ps = list(n._pos_encoding.parameters()) + ([] if n._dir_encoding is None else list(n._dir_encoding.parameters()))
        return [p for p in ps if p.requires_grad]

    def get_sigma_parameters(self, ignore_requires_grad=False):
        ps = list(self.nerf_model._model_sigma.parameters())
        return ps if ignore_requires_grad else [p for p in ps if p.requires_grad]

    def freeze_sigma_head(self, should_freeze=True):
        for p in self.get_sigma_parameters(True):
            p.requires_grad = not should_freeze

    def freeze_rgb_head(self, should_freeze=True):
        for p in self.get_rgb_parameters(True):
            p.requires_grad = not should_freeze

    def inference_points(self, xyz_, dir_, sigma_only):
        return inference(self.nerf_model, xyz_, dir_, netchunk=0, sigma_only=sigma_only, meshing=True)

    def forward(self, rays, ray_sampler, scale_factor, testing=False, camera=True, detach_sigma=True,
                return_variance=False, render_strategy='default'):
        if testing:
            n_samples, perturb = self.cfg.render.N_samples_test, 0.
        else:
            n_samples, perturb = self.cfg.render.N_samples_train, self.cfg.render.perturb
        results = defaultdict(list)
        for i in range(0, rays.shape[0], self.cfg.render.chunk):
            out = render_rays(rays[i:i + self.cfg.render.chunk, :], ray_sampler, self.nerf_model, self.cfg.ray_range,
                              scale_factor, N_samples=n_samples, retraw=self.cfg.render.retraw, perturb=perturb,
                              white_bkgd=True, raw_noise_std=self.cfg.render.raw_noise_std,
                              netchunk=self.cfg.render.netchunk, num_colors=self.cfg.num_colors,
                              sigma_only=(not camera), DEBUG=True, detach_sigma=detach_sigma,
                              return_variance=return_variance, render_strategy=render_strategy)
            for k, v in out.items():
                results[k] += [v]
        return {k: torch.cat(v, 0) for k, v in results.items()}
