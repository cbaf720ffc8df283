"""ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product path (loner_b200/).

Self-contained CPU restatement (torch-CPU, fp32 unless noted) of the reference's mapping hot path.
It travels to the GPU box (unlike /root/reference) and is what `tests/ -m gpu` compare the CUDA
kernels against.  Every function cites the reference lines it follows (paths relative to
/root/reference).  Random numbers are *inputs* (u1, u2, noise) so CPU and CUDA consume identical
streams; oracle/make_golden.py records the tensors the unmodified reference drew.

Two places deliberately fix an arithmetic ORDER that torch leaves unspecified, so that a parallel
GPU implementation can be bit-exact (both are at least as accurate as the reference's own):
  * `sum(weights)` in sample_pdf (rendering_tcnn.py:36): computed exactly (fp64 accumulation of fp32
    terms is exact for <= 2^10 terms in [1e-5, 1+1e-5]) then rounded once to fp32.  torch-CPU uses a
    vectorised cascade sum whose association depends on the host ISA; the results agree to <= 1 ulp.
  * `exp(-logit)` in the occupancy sigmoid (ray_sampling.py:79): correctly rounded fp32 (fp64 exp
    rounded to fp32).  torch-CPU uses Sleef's 1-ulp expf; ~1 % of values differ by one ulp.
`cumsum` needs no such choice: torch-CPU accumulates fp32 cumsum in fp64 (acc_type) which is exact
here, so any scan order reproduces it bit-for-bit.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

from oracle import tinycudann as tcnn_shim

# --------------------------------------------------------------------------------------------
# R2: sampling  (src/models/ray_sampling.py:18-92, src/models/rendering_tcnn.py:19-68,
#                src/models/model_tcnn.py:126-134)
# --------------------------------------------------------------------------------------------


def t_steps(n):
    """linspace(0,1,n) exactly as torch-CPU produces it (ray_sampling.py:59); kernels take it as input."""
    return torch.linspace(0, 1, n)


def stratified_depths(near, far, n, perturb, u):
    """ray_sampling.py:59-73 / :27-41.  near, far [N,1]; u [N,n] uniform in [0,1)."""
    z_steps = t_steps(n)
    z = near * (1 - z_steps) + far * z_steps
    z = z.expand(near.shape[0], n)
    if perturb > 0:
        mid = 0.5 * (z[:, :-1] + z[:, 1:])
        upper = torch.cat([mid, z[:, -1:]], -1)
        lower = torch.cat([z[:, :1], mid], -1)
        z = lower + (upper - lower) * (perturb * u)
    return z


def occ_interpolate(occ, pts):
    """model_tcnn.py:126-134: trilinear grid_sample(align_corners=False, zeros padding).
    occ [1,1,D,H,W]; pts [N,B,3] (x->W, y->H, z->D).  torch-CPU's kernel accumulates the 8 corners
    sequentially (tnw,tne,tsw,tse,bnw,bne,bsw,bse) with separate mul/add; the CUDA kernel copies
    that order, verified bit-exact in tests/test_oracle.py::test_grid_sample_restatement."""
    n, b, _ = pts.shape
    return F.grid_sample(occ, pts.reshape(1, 1, n, b, 3), mode="bilinear", align_corners=False).reshape(n, b)


def occ_interpolate_explicit(occ, pts):
    """numpy restatement of ATen grid_sampler_3d_cpu (bilinear, zeros, align_corners=False): the exact
    sequence of fp32 operations the CUDA sampler reproduces."""
    f32 = np.float32
    g = occ[0, 0].numpy()
    D, H, W = g.shape
    p = pts.numpy().reshape(-1, 3).astype(f32)

    def unnorm(c, size):
        return ((c + f32(1)) * f32(size) - f32(1)) / f32(2)

    ix, iy, iz = unnorm(p[:, 0], W), unnorm(p[:, 1], H), unnorm(p[:, 2], D)
    x0, y0, z0 = np.floor(ix), np.floor(iy), np.floor(iz)
    x1, y1, z1 = x0 + f32(1), y0 + f32(1), z0 + f32(1)
    wx1, wx0 = ix - x0, x1 - ix
    wy1, wy0 = iy - y0, y1 - iy
    wz1, wz0 = iz - z0, z1 - iz
    out = np.zeros(len(p), dtype=f32)
    corners = [(z0, y0, x0, wz0, wy0, wx0), (z0, y0, x1, wz0, wy0, wx1), (z0, y1, x0, wz0, wy1, wx0),
               (z0, y1, x1, wz0, wy1, wx1), (z1, y0, x0, wz1, wy0, wx0), (z1, y0, x1, wz1, wy0, wx1),
               (z1, y1, x0, wz1, wy1, wx0), (z1, y1, x1, wz1, wy1, wx1)]
    for (z, y, x, wz, wy, wx) in corners:
        ok = (x >= 0) & (x < W) & (y >= 0) & (y < H) & (z >= 0) & (z < D)
        xi = np.clip(x, 0, W - 1).astype(np.int64)
        yi = np.clip(y, 0, H - 1).astype(np.int64)
        zi = np.clip(z, 0, D - 1).astype(np.int64)
        w = ((wx * wy).astype(f32) * wz).astype(f32)
        out = np.where(ok, (out + (g[zi, yi, xi] * w).astype(f32)).astype(f32), out)
    return torch.from_numpy(out.reshape(pts.shape[0], pts.shape[1]))


def occ_probs(logits):
    """ray_sampling.py:79-80 with a correctly-rounded fp32 exp (see module docstring)."""
    e = torch.exp(-(logits.double())).float()
    p = 1.0 / (1 + e)
    return 2 * (p.clamp(min=0.5, max=1.0) - 0.5)


def sample_pdf(bins, weights, u, eps=1e-5):
    """rendering_tcnn.py:19-68 with det=False and `u` supplied.  Returns (samples, inds)."""
    n_rays, n_w = weights.shape
    w = weights + eps
    wsum = w.double().sum(-1, keepdim=True).float()          # exact sum, one rounding
    pdf = w / wsum
    cdf = torch.cumsum(pdf.double(), -1).float()              # == torch-CPU fp32 cumsum (fp64 acc)
    cdf = torch.cat([torch.zeros_like(cdf[:, :1]), cdf], -1)
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp_min(inds - 1, 0)
    above = torch.clamp_max(inds, n_w)
    cdf_b, cdf_a = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    bins_b, bins_a = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = cdf_a - cdf_b
    denom = torch.where(denom < eps, torch.ones_like(denom), denom)
    samples = bins_b + (u - cdf_b) / denom * (bins_a - bins_b)
    return samples, inds


def sample_occ(rays, occ, u1, u2, n_samples, perturb):
    """OccGridRaySampler.get_samples (ray_sampling.py:53-92).  Returns dict with the sorted depths
    `z` [N,S] plus the two bit-exact intermediates north_star names: `occ_bits` (prob > 0) and
    `inds` (searchsorted, int64)."""
    s2 = n_samples // 2
    o, d = rays[:, 0:3], rays[:, 3:6]
    near, far = rays[:, -2:-1], rays[:, -1:]
    zc = stratified_depths(near, far, s2, perturb, u1)
    pts = o.unsqueeze(1) + d.unsqueeze(1) * zc.unsqueeze(2)
    logits = occ_interpolate(occ, pts)
    probs = occ_probs(logits)
    mid = 0.5 * (zc[:, :-1] + zc[:, 1:])
    zf, inds = sample_pdf(mid, probs[:, 1:-1], u2)
    z, _ = torch.sort(torch.cat([zc, zf], -1), -1)
    return dict(z=z, zc=zc, zf=zf, probs=probs, logits=logits, occ_bits=probs > 0, inds=inds)


def sample_uniform(rays, u, n_samples, perturb):
    """UniformRaySampler.get_samples (ray_sampling.py:22-43)."""
    return stratified_depths(rays[:, -2:-1], rays[:, -1:], n_samples, perturb, u)


# --------------------------------------------------------------------------------------------
# R4: sigma field  (src/models/nerf_tcnn.py:59-81 + tiny-cuda-nn, see oracle/tinycudann)
# --------------------------------------------------------------------------------------------

SIGMA_GRID_CFG = dict(otype="HashGrid", n_levels=16, n_features_per_level=2, log2_hashmap_size=18,
                      base_resolution=16)                       # cfg/nerf_config/default_nerf_hash.yaml:20-25
SIGMA_MLP_CFG = dict(otype="FullyFusedMLP", activation="ReLU", output_activation="None",
                     n_neurons=64, n_hidden_layers=1)           # default_nerf_hash.yaml:26-31
MLP_PARAMS = 64 * 32 + 16 * 64                                  # W1 [64,32] then W2 [16,64] (row 0 used)


def grid_cfg(log2_hashmap_size=18):
    c = dict(SIGMA_GRID_CFG)
    c["log2_hashmap_size"] = log2_hashmap_size
    return c


def init_sigma_params(log2_hashmap_size=18, seed=1337):
    """Flat fp32 parameter vector with the tcnn layout: MLP (3072) then tables."""
    m = tcnn_shim.NetworkWithInputEncoding(3, 1, grid_cfg(log2_hashmap_size), SIGMA_MLP_CFG, seed=seed)
    return m.params.detach().clone()


def hash_encode(x01, table_flat, log2_hashmap_size=18):
    """[P,3] in [0,1] -> [P,32] features (differentiable wrt table and x01)."""
    return tcnn_shim._HashGridFn(grid_cfg(log2_hashmap_size))(x01, table_flat)


def sigma_mlp(e, mlp_flat):
    """h = relu(W1 e); sigma = (W2 h)[0]   (no bias, no output activation)."""
    w1 = mlp_flat[:2048].view(64, 32)
    w2 = mlp_flat[2048:3072].view(16, 64)
    h = torch.relu(e @ w1.t())
    return (h @ w2.t())[:, 0]


def sigma_field(xyz, params, log2_hashmap_size=18):
    """DecoupledNeRF.forward(pos, None, sigma_only=True) (nerf_tcnn.py:63,71-72): xyz in [-1,1]."""
    x01 = (xyz + 1) / 2
    e = hash_encode(x01, params[MLP_PARAMS:], log2_hashmap_size)
    return sigma_mlp(e, params[:MLP_PARAMS])


# --------------------------------------------------------------------------------------------
# R5/R6: compositing  (src/models/rendering_tcnn.py:219-295, :70-214)
# --------------------------------------------------------------------------------------------


def composite(sigma, z, rays_d, far, noise):
    """raw2outputs sigma-only (rendering_tcnn.py:241-293).  sigma,z,noise [N,S]; far [N,1].
    `noise` is the already-scaled N(0, raw_noise_std^2) draw (or None)."""
    deltas = z[:, 1:] - z[:, :-1]
    deltas = torch.cat([deltas, 1e10 * torch.ones_like(deltas[:, :1])], -1)
    deltas = deltas * torch.norm(rays_d.unsqueeze(1), dim=-1)
    s = sigma if noise is None else sigma + noise
    alphas = 1 - torch.exp(-deltas * torch.relu(s))
    shifted = torch.cat([torch.ones_like(alphas[:, :1]), 1. - alphas + 1e-10], -1)
    trans = torch.cumprod(shifted, -1)[:, :-1]
    weights = alphas * trans
    opacity = weights.sum(-1)
    depth = torch.sum(torch.cat([weights, 1 - weights.sum(1, keepdim=True)], 1) * torch.cat([z, far], -1), -1)
    variance = (weights * (depth.view(-1, 1) - z) ** 2).sum(1)
    return dict(weights=weights, opacity=opacity, depth=depth, variance=variance, trans=trans, alphas=alphas)


def composite_adjusted(sigma, z, rays_d):
    """raw2outputs_adjusted *effective* semantics (rendering_tcnn.py:70-214): noise forced to 0 (:104);
    depth = z at the first sample where transmittance crosses 0.5 (:130-132), else 0; the
    peak/prominence write-backs (:196-199) assign into a temporary copy and are no-ops."""
    out = composite(sigma, z, rays_d, torch.zeros_like(z[:, :1]), None)
    T = out["trans"]
    T_prev = torch.cat([torch.ones_like(T[:, :1]), T[:, :-1]], 1)
    hit = (~(T > 0.5)) & (T_prev > 0.5)
    depth = (hit.float() * z).sum(1)      # at most one crossing per ray (T is non-increasing)
    variance = (out["weights"] * (depth.view(-1, 1) - z) ** 2).sum(1)
    out.update(depth=depth, variance=variance)
    return out


# --------------------------------------------------------------------------------------------
# R4/R5 colour branch  (src/models/nerf_tcnn.py:40-52,64,83-95; src/models/rendering_tcnn.py:284-289;
#                       src/mapping/optimizer.py:861-894)
# --------------------------------------------------------------------------------------------

INTENSITY_MLP_CFG = dict(otype="FullyFusedMLP", activation="ReLU", output_activation="None",
                         n_neurons=64, n_hidden_layers=4)       # default_nerf_hash.yaml:8-13


def sh4(d01):
    """tcnn SphericalHarmonics degree 4 on 2*d01-1 (nerf_tcnn.py:43,86)."""
    return tcnn_shim.sh4(d01)


def init_intensity_params(log2_hashmap_size=19, n_colors=3, seed=1337):
    """(table_flat, mlp_flat) of the intensity head with the reference's module seeds (tcnn default 1337 each)."""
    enc = tcnn_shim.Encoding(3, grid_cfg(log2_hashmap_size), seed=seed)
    net = tcnn_shim.Network(48, n_colors, INTENSITY_MLP_CFG, seed=seed)
    return enc.params.detach().clone(), net.params.detach().clone()


def intensity_mlp(x, mlp_flat, n_in=48, n_out=3, cfg=None):
    """tcnn.Network: n_in -> 64 x n_hidden -> n_out, ReLU, no bias, no output activation."""
    return tcnn_shim._MLP(n_in, n_out, cfg or INTENSITY_MLP_CFG)(x, mlp_flat)


def color_field(xyz, viewdirs, table_flat, mlp_flat, log2_hashmap_size=19, n_colors=3):
    """Colour half of DecoupledNeRF.forward (nerf_tcnn.py:63-64,83-93): xyz [P,3] in [-1,1], viewdirs [P,3]."""
    h_x = hash_encode((xyz + 1) / 2, table_flat, log2_hashmap_size)
    h_d = sh4((viewdirs + 1) / 2)
    return torch.sigmoid(intensity_mlp(torch.cat([h_x, h_d], -1), mlp_flat, 48, n_colors))


def composite_rgb(weights, colors, white_bkgd=True):
    """rendering_tcnn.py:284-289."""
    rgb = torch.sum(weights.unsqueeze(-1) * colors, -2)
    return rgb + 1 - weights.sum(-1, keepdim=True) if white_bkgd else rgb


def camera_forward(rays, intensities, sigma_params, table_int, mlp_int, z, noise, log2_T_sigma, log2_T_int,
                   n_colors=3, white_bkgd=True):
    """Model.forward(camera=True, detach_sigma=True) + compute_loss_camera (model_tcnn.py:70-108,
    rendering_tcnn.py:380-425, optimizer.py:861-894) at given sample depths z [N,S] and noise [N,S]."""
    n, s = z.shape
    xyz = rays[:, None, 0:3] + rays[:, None, 3:6] * z.unsqueeze(2)
    with torch.no_grad():                                        # nerf_tcnn.py:66-69
        sigma = sigma_field(xyz.reshape(-1, 3), sigma_params, log2_T_sigma).view(n, s)
    dirs = torch.repeat_interleave(rays[:, 6:9], repeats=s, dim=0)
    col = color_field(xyz.reshape(-1, 3), dirs, table_int, mlp_int, log2_T_int, n_colors).view(n, s, n_colors)
    out = composite(sigma, z, rays[:, 3:6], rays[:, -1:], noise)
    rgb = composite_rgb(out["weights"], col, white_bkgd)
    loss = F.l1_loss(rgb.reshape(-1, 1), intensities.reshape(-1, 1))
    out.update(rgb=rgb, colors=col, sigma=sigma, loss=loss)
    return out


# --------------------------------------------------------------------------------------------
# L1-L4: JS dynamic-margin depth loss  (src/mapping/optimizer.py:701-859, src/models/losses.py:29-51)
# --------------------------------------------------------------------------------------------

DEFAULT_LOSS_CFG = dict(loss_selection="L1_JS", min_js_score=1.0, max_js_score=10.0, alpha=1.0,
                        los_lambda=1000.0, min_depth_eps=0.5, depthloss_lambda=0.005,
                        depth_eps=3.0)   # cfg/model_config/default_model_config.yaml:39-58


def kl_gauss(m1, s1, m2, s2):
    """optimizer.py:913-920."""
    return torch.log(s2 / s1) + (s1 * s1 + (m1 - m2) ** 2) / (2 * s2 * s2) - 0.5


def js_gauss(m1, s1, m2, s2):
    """optimizer.py:922-925."""
    mm = 0.5 * (m1 + m2)
    sm = 0.5 * torch.sqrt(s1 ** 2 + s2 ** 2)
    return 0.5 * kl_gauss(m1, s1, mm, sm) + 0.5 * kl_gauss(m2, s2, mm, sm)


def weights_gt(zm, g, eps):
    """get_weights_gt (losses.py:29-51), norm=True.  zm [N,S]; g [N,1]; eps [N,1] or float."""
    sig = eps / 9
    clip_a = (g - eps - g) / sig
    clip_b = (g + eps - g) / sig
    pdf = 1. / np.sqrt(2 * np.pi) * torch.exp(-0.5 * (((zm - g) / sig) ** 2))
    cdf = lambda x: 0.5 * (1 + torch.erf(x / np.sqrt(2)))
    w = pdf / sig / (cdf(torch.as_tensor(clip_b)) - cdf(torch.as_tensor(clip_a)))
    zero = torch.zeros_like(zm)
    w = torch.heaviside(zm - (g - eps), zero) * torch.heaviside((g + eps) - zm, zero) * w
    return w / (w.sum(1, keepdim=True) + 1e-6)


def depth_loss(weights, depth, opacity, z, far, depth_gt, scale, cfg=None, los_lambda=None, depth_eps=None):
    """Optimizer.compute_loss lidar branch (optimizer.py:719-848).  depth_gt [N] in cube units.
    Returns dict(loss, js, eps, w_gt, opaque, mean, std)."""
    cfg = dict(DEFAULT_LOSS_CFG, **(cfg or {}))
    los_lambda = cfg["los_lambda"] if los_lambda is None else los_lambda
    dg = depth_gt.reshape(-1, 1)
    transparent = (dg > far.reshape(-1, 1))[:, 0]
    opaque = (dg > 0)[:, 0] & ~transparent
    zm = z * scale
    g = dg * scale
    wsum = weights.sum(1)
    mean = (zm * weights).sum(1) / (wsum + 1e-10)
    var = ((zm - mean.unsqueeze(-1)) ** 2 * weights).sum(1) / (wsum + 1e-10) + 1e-10
    std = torch.sqrt(var)
    eps_min = cfg["min_depth_eps"]
    js = js_gauss(g, eps_min / 3., mean.unsqueeze(1), std.unsqueeze(1)).squeeze(-1)
    loss = cfg["depthloss_lambda"] * F.mse_loss((depth * scale)[opaque], g[opaque, 0])
    sel = cfg["loss_selection"]
    if sel in ("L1_JS", "L2_JS"):
        jsc = js.detach().clone()
        jsc[jsc < cfg["min_js_score"]] = 0
        jsc[jsc > cfg["max_js_score"]] = cfg["max_js_score"]
        eps = (eps_min * (1 + cfg["alpha"] * jsc)).unsqueeze(-1)
    else:
        eps = cfg["depth_eps"] if depth_eps is None else depth_eps
        jsc = js.detach()
    w_gt = weights_gt(zm, g, eps)
    w_gt = w_gt * opaque.unsqueeze(1).float()
    if sel in ("L1_JS", "L1_LOS"):
        los = F.l1_loss(weights, w_gt)
    else:
        los = F.mse_loss(weights, w_gt)
    loss = loss + los_lambda * los
    loss = loss + torch.abs(opacity[opaque] - 1).mean()
    eps_t = eps if torch.is_tensor(eps) else torch.full_like(g, float(eps))
    return dict(loss=loss, js=jsc, js_raw=js.detach(), eps=eps_t.reshape(-1), w_gt=w_gt, opaque=opaque,
                mean=mean.detach(), std=std.detach())


# --------------------------------------------------------------------------------------------
# Whole forward (R1-R5 + L1-L4) with pre-drawn randoms; autograd gives L5.
# --------------------------------------------------------------------------------------------


def mapping_forward(rays, depth_gt, params, occ, u1, u2, noise, scale, n_samples,
                    log2_hashmap_size=18, perturb=1.0, cfg=None, sampler="OGM"):
    """rays [N,13] (may require grad), params flat (may require grad).  Mirrors
    Optimizer.compute_loss -> Model.forward -> render_rays with chunk >= N."""
    with torch.no_grad():
        if sampler == "OGM":
            smp = sample_occ(rays.detach(), occ, u1, u2, n_samples, perturb)
            z = smp["z"]
        else:
            smp = {}
            z = sample_uniform(rays.detach(), u1, n_samples, perturb)
    o, d, far = rays[:, 0:3], rays[:, 3:6], rays[:, -1:]
    xyz = o.unsqueeze(1) + d.unsqueeze(1) * z.unsqueeze(2)
    sigma = sigma_field(xyz.reshape(-1, 3), params, log2_hashmap_size).view(z.shape)
    comp = composite(sigma, z, d, far, noise)
    ls = depth_loss(comp["weights"], comp["depth"], comp["opacity"], z, far, depth_gt, scale, cfg)
    return dict(z=z, xyz=xyz, sigma=sigma, **{k: comp[k] for k in ("weights", "opacity", "depth", "variance")},
                **ls, sample=smp)


# --------------------------------------------------------------------------------------------
# O1/O2: optimiser step and occupancy update
# --------------------------------------------------------------------------------------------


def adam_step(p, g, m, v, step, lr, b1=0.9, b2=0.999, eps=1e-8):
    """torch.optim.Adam defaults, no weight decay (optimizer.py:261-265); `step` is 1-based. In place."""
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


def logits_grad(zm, g, eps=2.0, l_free=0.25, l_occ=2.5):
    """get_logits_grad (losses.py:54-62); heaviside(.,0) is strict."""
    x = zm - g
    return l_free * (x < -eps).float() - l_occ * ((x > -eps) & (x < eps)).float()


def occ_update(occ, xyz, zm, g, lr=1e-4):
    """_step_occupancy_grid (optimizer.py:897-908): logits -= lr * trilinear-scatter(logits_grad)."""
    occ = occ.detach().clone().requires_grad_(True)
    logits = occ_interpolate(occ, xyz.detach())
    logits.backward(gradient=logits_grad(zm, g))
    return (occ - lr * occ.grad).detach()


# --------------------------------------------------------------------------------------------
# R0: ray construction  (src/common/ray_utils.py:31-60,269-322; src/common/pose_utils.py:354-368)
# --------------------------------------------------------------------------------------------


def axis_angle_to_matrix(aa):
    """pytorch3d.transforms.axis_angle_to_matrix (via quaternion), for the pose 6-vector chain."""
    ang = torch.norm(aa, p=2, dim=-1, keepdim=True)
    half = ang * 0.5
    small = ang.abs() < 1e-6
    s = torch.where(small, 0.5 - ang * ang / 48, torch.sin(half) / torch.where(small, torch.ones_like(ang), ang))
    q = torch.cat([torch.cos(half), aa * s], -1)
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    m = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return m.reshape(q.shape[:-1] + (3, 3))


def far_val(o, d):
    """get_far_val(no_nan=True) (ray_utils.py:31-60)."""
    d = d + 1e-15
    dirs = torch.tensor([[-1.], [1.]])
    t = (dirs[..., None] - o[:, [0, 1, 2]]) / d[:, [0, 1, 2]]
    return t.clamp(min=0).max(dim=0)[0].min(dim=1)[0].unsqueeze(1)


def build_lidar_rays(directions, distances, pose6, shift, scale, ray_range):
    """LidarRayDirections.build_lidar_rays (ray_utils.py:269-322) for pose 6-vector [t, axis-angle].
    directions [3,n]; returns (rays [m,13], depths [m]) after the validity filter (:321-322)."""
    R = axis_angle_to_matrix(pose6[3:].unsqueeze(0))[0]
    o = ((pose6[:3] + shift) / scale).tile(directions.shape[1], 1)
    dirs = (R @ directions.to(R.dtype)).T
    dirs = dirs / torch.norm(dirs, dim=1, keepdim=True)
    near = ray_range[0] / scale * torch.ones_like(o[:, :1])
    far_r = ray_range[1] / scale * torch.ones_like(o[:, :1])
    far = torch.minimum(far_r, far_val(o, dirs))
    rays = torch.cat([o, dirs, -dirs, torch.zeros_like(o[:, :2]), near, far], 1)
    valid = (far > (near + 1. / scale))[..., 0]
    return rays[valid], (distances / scale)[valid]


# --------------------------------------------------------------------------------------------
# f2: camera rays  (src/common/ray_utils.py:62-124,168-206; kornia undistort_points restated in
#     oracle/stubs/kornia/geometry/calibration.py -- parity unpinned for distorted intrinsics)
# --------------------------------------------------------------------------------------------


def camera_directions(H, W, new_k, dist=None, K=None):
    """get_ray_directions (ray_utils.py:62-124), sppd = 1: (directions [H*W,3], pix_i [H*W,1], pix_j [H*W,1])."""
    xs = torch.linspace(0, W - 1., W)
    ys = torch.linspace(0, H - 1., H)
    gx, gy = torch.meshgrid([xs, ys], indexing="ij")
    gx = gx.permute(1, 0).reshape(-1, 1)
    gy = gy.permute(1, 0).reshape(-1, 1)
    nx, ny = gx, gy
    if dist is not None:
        import importlib.util
        import os
        spec = importlib.util.spec_from_file_location(
            "_oracle_kornia_calibration",
            os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs", "kornia", "geometry", "calibration.py"))
        cal = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(cal)
        und = cal.undistort_points(torch.cat([gx, gy], -1).unsqueeze(0), K.reshape(1, 3, 3), dist.reshape(1, -1),
                                   new_k.reshape(1, 3, 3))
        nx, ny = und[0, :, 0:1], und[0, :, 1:2]
    d = torch.cat([(nx - new_k[0, 2]) / new_k[0, 0], (ny - new_k[1, 2]) / new_k[1, 1], torch.ones_like(gx)], -1)
    return d, gx, gy


def build_camera_rays(directions, pix_i, pix_j, indices, T, shift, scale, near_m, image=None):
    """CameraRayDirections.build_rays (ray_utils.py:168-206): T = camera-to-world 4x4; returns
    (rays [m,13] = [o, d, -d, pix_i, pix_j, near, cube exit], intensities [m,C] | None)."""
    d = directions[indices]
    R = T[:3, :3].float()
    o = ((T[:3, 3] + shift) / scale).float()
    rd = d @ R.T
    rd = rd / torch.norm(rd, dim=-1, keepdim=True)
    ro = o.reshape(1, 3).expand(rd.shape[0], 3)
    near = near_m / scale * torch.ones_like(ro[:, :1])
    far = far_val(ro, rd)
    rays = torch.cat([ro, rd, -rd, pix_i[indices], pix_j[indices], near, far], 1).float()
    inten = None if image is None else image.reshape(-1, image.shape[-1])[indices]
    return rays, inten
