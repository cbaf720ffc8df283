"""Tensor-level wrappers over the C-ABI (loner_b200/_lib.py).  PyTorch is used only for device
memory and streams: every function takes CUDA fp32 tensors, passes raw pointers to
libloner_b200.so on the current stream, and returns torch tensors.  No autograd here (see
loner_b200/functional.py) and no fallback: a CPU tensor is an error.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import GridConfig, LossConfig, LOSS_KINDS, LonerB200Error

MLP_PARAMS = 64 * 32 + 16 * 64


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t, dtype=torch.float32, name="tensor"):
    """Raw device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise LonerB200Error(f"{name} must be a CUDA tensor: loner_b200 has no CPU path")
    if t.dtype != dtype:
        raise LonerB200Error(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise LonerB200Error(f"{name} must be contiguous")
    return C.c_void_p(t.data_ptr())


class Plan:
    """Immutable hash-grid level table + MLP dims (lb_plan). Mirrors the tcnn config dict keys
    (reference cfg/nerf_config/default_nerf_hash.yaml:14-31)."""

    def __init__(self, n_levels=16, n_features_per_level=2, log2_hashmap_size=18, base_resolution=16,
                 per_level_scale=2.0, mlp_width=64, mlp_hidden_layers=1, l2_bytes=None):
        self.cfg = GridConfig(n_levels, n_features_per_level, log2_hashmap_size, base_resolution,
                              per_level_scale, mlp_width, mlp_hidden_layers)
        self._h = C.c_void_p()
        _lib.call("lb_plan_create", C.byref(self.cfg), C.byref(self._h))
        if l2_bytes is not None:       # level-group budget override (tests force group passes on small tables)
            _lib.call("lb_plan_set_l2_bytes", self._h, int(l2_bytes))
        a, b = C.c_int64(), C.c_int64()
        _lib.call("lb_plan_param_counts", self._h, C.byref(a), C.byref(b))
        self.n_mlp_params, self.n_table_params = a.value, b.value
        self.n_params = a.value + b.value
        self.n_levels = n_levels
        self.n_enc = n_levels * n_features_per_level
        self.enc_optional = bool(_lib.load().lb_sigma_fwd_enc_optional(self._h))
        # forward-mode input derivative (lb_sigma_fwd_jvp) available: the backward then needs no second table gather
        self.jvp_supported = bool(_lib.load().lb_sigma_fwd_jvp_supported(self._h))

    @classmethod
    def from_tcnn_config(cls, encoding_config, network_config=None):
        if encoding_config.get("otype") != "HashGrid":
            raise LonerB200Error(f"unsupported encoding {encoding_config.get('otype')}")
        kw = dict(n_levels=int(encoding_config["n_levels"]),
                  n_features_per_level=int(encoding_config["n_features_per_level"]),
                  log2_hashmap_size=int(encoding_config["log2_hashmap_size"]),
                  base_resolution=int(encoding_config["base_resolution"]),
                  per_level_scale=float(encoding_config.get("per_level_scale", 2.0)))
        if network_config is not None:
            kw.update(mlp_width=int(network_config["n_neurons"]),
                      mlp_hidden_layers=int(network_config["n_hidden_layers"]))
        return cls(**kw)

    def level(self, l):
        s, r, o, n, d = C.c_float(), C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_int32()
        _lib.call("lb_plan_level", self._h, l, C.byref(s), C.byref(r), C.byref(o), C.byref(n), C.byref(d))
        return dict(scale=s.value, res=r.value, offset=o.value, size=n.value, dense=bool(d.value))

    def __del__(self):
        try:
            if self._h:
                _lib.load().lb_plan_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass


_T_STEPS = {}


def pose_to_rot_origin(pose6, shift, scale):
    """pose 6-vectors [K,6] -> (rot [K,9] row-major, origin [K,3] = (t + shift) / scale)."""
    K = pose6.shape[0]
    rot = torch.empty(K, 9, device=pose6.device, dtype=torch.float32)
    org = torch.empty(K, 3, device=pose6.device, dtype=torch.float32)
    _lib.call("lb_pose_to_rot_origin", _p(pose6, name="pose6"), _p(shift, name="shift"), float(scale), K, _p(rot), _p(org), _stream())
    return rot, org


def pose_to_rot_origin_bwd(pose6, scale, d_rot, d_origin):
    K = pose6.shape[0]
    d_pose = torch.empty(K, 6, device=pose6.device, dtype=torch.float32)
    _lib.call("lb_pose_to_rot_origin_bwd", _p(pose6, name="pose6"), float(scale), K, _p(d_rot, name="d_rot"),
              _p(d_origin, name="d_origin"), _p(d_pose), _stream())
    return d_pose


def build_lidar_rays(directions, distances, indices, rot, origin, near, far_max, inv_scale, const_depth=0.0,
                     out_rays=None, out_depths=None, mask_span=0.0):
    """R0 forward.  directions [3,n]; distances [n] or None (constant depth); indices [m] int64 or None.
    Writes into out_rays [m,13] / out_depths [m] when given (slices of a batch buffer).  mask_span > 0: rays with
    far <= near + mask_span (the ones the reference drops) get depth NaN = masked (see include/loner_b200.h)."""
    n = directions.shape[1]
    m = n if indices is None else indices.shape[0]
    dev = directions.device
    rays = out_rays if out_rays is not None else torch.empty(m, 13, device=dev, dtype=torch.float32)
    depths = out_depths if out_depths is not None else torch.empty(m, device=dev, dtype=torch.float32)
    _lib.call("lb_build_lidar_rays", _p(directions, name="directions"), _p(distances, name="distances"),
              _p(indices, torch.int64, name="indices"), n, m, _p(rot, name="rot"), _p(origin, name="origin"),
              float(near), float(far_max), float(inv_scale), float(const_depth), float(mask_span), _p(rays), _p(depths),
              _stream())
    return rays, depths


def build_camera_rays(directions, pix_i, pix_j, indices, rot, origin, near, image=None):
    """Camera rays.  directions [n,3]; pix_i / pix_j [n] or None; image [n,C] or None; indices [m] int64 or None.
    Returns (rays [m,13], intensities [m,C] or None)."""
    n = directions.shape[0]
    m = n if indices is None else indices.shape[0]
    dev = directions.device
    rays = torch.empty(m, 13, device=dev, dtype=torch.float32)
    c = 0 if image is None else image.shape[1]
    inten = None if image is None else torch.empty(m, c, device=dev, dtype=torch.float32)
    _lib.call("lb_build_camera_rays", _p(directions, name="directions"), _p(pix_i, name="pix_i"), _p(pix_j, name="pix_j"),
              _p(image, name="image"), c, _p(indices, torch.int64, name="indices"), n, m, _p(rot, name="rot"),
              _p(origin, name="origin"), float(near), _p(rays), _p(inten), _stream())
    return rays, inten


def build_lidar_rays_bwd(directions, indices, rot, origin, far_max, d_rays, d_rot, d_origin):
    """R0 backward: accumulates into d_rot [9] and d_origin [3]."""
    n = directions.shape[1]
    m = d_rays.shape[0]
    _lib.call("lb_build_lidar_rays_bwd", _p(directions, name="directions"), _p(indices, torch.int64, name="indices"),
              n, m, _p(rot, name="rot"), _p(origin, name="origin"), float(far_max), _p(d_rays, name="d_rays"),
              _p(d_rot), _p(d_origin), _stream())


def t_steps(n, device):
    """linspace(0,1,n) computed on the CPU (so every device consumes identical bits), cached."""
    key = (n, str(device))
    if key not in _T_STEPS:
        _T_STEPS[key] = torch.linspace(0, 1, n).to(device)
    return _T_STEPS[key]


def sample_occ(rays, occ, u1, u2, n_samples, perturb, debug=False):
    """OccGridRaySampler.get_samples. rays [N,13], occ [..,D,D,D], u1/u2 [N,S/2] -> z [N,S] sorted.
    debug=True also returns (occ_bits uint8 [N,S/2], inds int64 [N,S/2])."""
    n = rays.shape[0]
    s2 = n_samples // 2
    D = occ.shape[-1]
    z = torch.empty(n, n_samples, device=rays.device, dtype=torch.float32)
    bits = torch.empty(n, s2, device=rays.device, dtype=torch.uint8) if debug else None
    inds = torch.empty(n, s2, device=rays.device, dtype=torch.int64) if debug else None
    _lib.call("lb_sample_occ", _p(rays, name="rays"), _p(occ, name="occ"), D, _p(t_steps(s2, rays.device)),
              _p(u1, name="u1"), _p(u2, name="u2"), n, n_samples, float(perturb), _p(z),
              _p(bits, torch.uint8), _p(inds, torch.int64), _stream())
    return (z, bits, inds) if debug else z


def sample_uniform(rays, u, n_samples, perturb):
    n = rays.shape[0]
    z = torch.empty(n, n_samples, device=rays.device, dtype=torch.float32)
    _lib.call("lb_sample_uniform", _p(rays, name="rays"), _p(t_steps(n_samples, rays.device)), _p(u, name="u"),
              n, n_samples, float(perturb), _p(z), _stream())
    return z


def encode_fwd(plan, x01, table):
    P = x01.shape[0]
    enc = torch.empty(P, plan.n_enc, device=x01.device, dtype=torch.float32)
    _lib.call("lb_encode_fwd", plan._h, _p(x01, name="x01"), _p(table, name="table"), P, _p(enc), _stream())
    return enc


def encode_bwd(plan, x01, table, d_enc, grad_table, need_dx):
    P = x01.shape[0]
    dx = torch.empty(P, 3, device=x01.device, dtype=torch.float32) if need_dx else None
    _lib.call("lb_encode_bwd", plan._h, _p(x01, name="x01"), _p(table, name="table"), _p(d_enc, name="d_enc"), P,
              _p(grad_table, name="grad_table"), _p(dx), _stream())
    return dx


def sigma_fwd(plan, params, rays=None, z=None, x01=None, save_enc=True, tensor_cores=True):
    """Fused xyz -> hash grid -> MLP. Returns (sigma [N,S] or [P], saved) where saved = (enc, act) for
    sigma_bwd (None when save_enc is False).  tensor_cores=False selects the fp32 CUDA-core MLP."""
    if rays is not None:
        n, s = z.shape
        shape = (n, s)
    else:
        n, s = x01.shape[0], 1
        shape = (n,)
    if not tensor_cores and not _lib.test_switches():
        raise LonerB200Error("sigma_fwd(tensor_cores=False): the fp32 CUDA-core MLP exists only in a test build "
                             "(python -m loner_b200.build --test-switches)")
    sigma = torch.empty(shape, device=params.device, dtype=torch.float32)
    # the encodings are the encode -> MLP hand-off of the two-kernel path; the fused gather + MLP kernel needs
    # neither buffer when nothing is kept for lb_sigma_bwd
    inference = (not save_enc) and tensor_cores and plan.enc_optional
    enc = None if inference else torch.empty(n * s, plan.n_enc, device=params.device, dtype=torch.float32)
    act = torch.empty(n * s, device=params.device, dtype=torch.int64) if (tensor_cores and not inference) else None
    _lib.call("lb_sigma_fwd", plan._h, _p(rays, name="rays"), _p(z, name="z"), _p(x01, name="x01"),
              _p(params, name="params"), n, s, _p(sigma), _p(enc), _p(act, torch.int64), _stream())
    return sigma, ((enc, act) if save_enc else None)


def sigma_fwd_jvp(plan, params, rays=None, z=None, x01=None, save_enc=True):
    """lb_sigma_fwd_jvp: sigma plus d sigma / d x01 [P,3] from ONE forward kernel (plans with plan.jvp_supported).
    Returns (sigma, saved, dsigma_dx); saved = (enc, act) for sigma_bwd (None when save_enc is False)."""
    if rays is not None:
        n, s = z.shape
        shape = (n, s)
    else:
        n, s = x01.shape[0], 1
        shape = (n,)
    dev = params.device
    sigma = torch.empty(shape, device=dev, dtype=torch.float32)
    dsdx = torch.empty(n * s, 3, device=dev, dtype=torch.float32)
    enc = torch.empty(n * s, plan.n_enc, device=dev, dtype=torch.float32) if save_enc else None
    act = torch.empty(n * s, device=dev, dtype=torch.int64) if save_enc else None
    _lib.call("lb_sigma_fwd_jvp", plan._h, _p(rays, name="rays"), _p(z, name="z"), _p(x01, name="x01"),
              _p(params, name="params"), n, s, _p(sigma), _p(enc), _p(act, torch.int64), _p(dsdx), _stream())
    return sigma, ((enc, act) if save_enc else None), dsdx


def rays_grad_from_jvp(z, dsigma_dx, d_sigma, d_rays):
    """d_rays[:, 0:3] += sum_s d_sigma g / 2, d_rays[:, 3:6] += sum_s d_sigma z g / 2 (g = dsigma_dx of sigma_fwd_jvp)."""
    n, s = z.shape
    _lib.call("lb_rays_grad_from_jvp", _p(z, name="z"), _p(dsigma_dx, name="dsigma_dx"), _p(d_sigma, name="d_sigma"),
              n, s, _p(d_rays, name="d_rays"), _stream())
    return d_rays


NOALLOC_MAX_POINTS = 65536       # LB_NOALLOC_MAX_POINTS (include/loner_b200.h)


def sigma_fwd_noalloc(plan, params, sigma_out, rays=None, z=None, x01=None):
    """lb_sigma_fwd_noalloc: sigma-head inference into the caller's `sigma_out` with the library's built-in scratch
    (at most NOALLOC_MAX_POINTS sample points).  Allocates nothing, neither here nor in the library; calls that overlap
    on the GPU race on the shared scratch (see include/loner_b200.h).  No counterpart in the reference."""
    if rays is not None:
        n, s = z.shape
    else:
        n, s = x01.shape[0], 1
    if sigma_out.numel() != n * s:
        raise LonerB200Error(f"sigma_out must hold {n * s} values, got {sigma_out.numel()}")
    _lib.call("lb_sigma_fwd_noalloc", plan._h, _p(rays, name="rays"), _p(z, name="z"), _p(x01, name="x01"),
              _p(params, name="params"), n, s, _p(sigma_out, name="sigma_out"), _stream())
    return sigma_out


def sigma_bwd(plan, params, saved, d_sigma, rays=None, z=None, x01=None, grad_params=None, d_rays=None, need_dx=False):
    """Accumulates into grad_params / d_rays (caller zero-fills); returns d_x01 when asked.
    `saved` = (enc, act) from sigma_fwd; enc is consumed (overwritten in place with dL/d_enc)."""
    enc, act = saved
    if rays is not None:
        n, s = z.shape
    else:
        n, s = x01.shape[0], 1
    dx = torch.empty(n, 3, device=params.device, dtype=torch.float32) if (need_dx and rays is None) else None
    _lib.call("lb_sigma_bwd", plan._h, _p(rays, name="rays"), _p(z, name="z"), _p(x01, name="x01"),
              _p(params, name="params"), _p(enc, name="enc"), _p(act, torch.int64, name="act"),
              _p(d_sigma, name="d_sigma"), n, s,
              _p(grad_params, name="grad_params"), _p(d_rays, name="d_rays"), _p(dx), _stream())
    return dx


def composite_fwd(sigma, z, rays, noise=None, strategy="default", want_variance=True):
    n, s = z.shape
    dev = z.device
    w = torch.empty(n, s, device=dev, dtype=torch.float32)
    depth = torch.empty(n, device=dev, dtype=torch.float32)
    opac = torch.empty(n, device=dev, dtype=torch.float32)
    var = torch.empty(n, device=dev, dtype=torch.float32) if want_variance else None
    strat = {"default": 0, "adjusted": 1}[strategy]
    _lib.call("lb_composite_fwd", _p(sigma, name="sigma"), _p(z, name="z"), _p(rays, name="rays"), _p(noise, name="noise"),
              n, s, strat, _p(w), _p(depth), _p(opac), _p(var), _stream())
    return w, depth, opac, var


def composite_bwd(sigma, z, rays, noise, g_w, g_depth, g_opacity, g_variance, d_rays=None):
    n, s = z.shape
    ds = torch.empty(n, s, device=z.device, dtype=torch.float32)
    _lib.call("lb_composite_bwd", _p(sigma, name="sigma"), _p(z, name="z"), _p(rays, name="rays"), _p(noise, name="noise"),
              _p(g_w, name="g_w"), _p(g_depth, name="g_depth"), _p(g_opacity, name="g_opacity"),
              _p(g_variance, name="g_variance"), n, s, _p(ds), _p(d_rays, name="d_rays"), _stream())
    return ds


def count_opaque(rays, depth_gt):
    """-> device float[2] = {opaque rays, rays} of this batch: the global loss normalisers once summed over ranks."""
    c = torch.zeros(2, device=rays.device, dtype=torch.float32)
    _lib.call("lb_count_opaque", _p(rays, name="rays"), _p(depth_gt, name="depth_gt"), rays.shape[0], _p(c), _stream())
    return c


def make_loss_config(kind, scale_factor, n_opaque, n_total_samples, min_js_score=1.0, max_js_score=10.0,
                     js_alpha=1.0, min_depth_eps=0.5, depth_eps=3.0, los_lambda=1000.0, depthloss_lambda=0.005):
    return LossConfig(LOSS_KINDS[kind], float(scale_factor), float(min_js_score), float(max_js_score),
                      float(js_alpha), float(min_depth_eps), float(depth_eps), float(los_lambda),
                      float(depthloss_lambda), float(n_opaque), float(n_total_samples))


def composite_loss(sigma, z, rays, noise, depth_gt, cfg, loss_parts, want_grad=True, n_opaque_dev=None):
    """Fused compositing + loss (+ analytic gradient). loss_parts [4] is accumulated into.  n_opaque_dev = device
    float[2] {opaque-ray count, ray count} from count_opaque (global normalisers), or None to use cfg's.
    Returns dict(weights, depth, opacity, variance, js, eps, d_sigma, d_rays)."""
    if n_opaque_dev is not None and n_opaque_dev.numel() != 2:
        raise LonerB200Error("composite_loss: the normaliser tensor must hold {n_opaque, n_rays} (ops.count_opaque)")
    n, s = z.shape
    dev = z.device
    f = lambda *shape: torch.empty(*shape, device=dev, dtype=torch.float32)
    out = dict(weights=f(n, s), depth=f(n), opacity=f(n), variance=f(n), js=f(n), eps=f(n),
               d_sigma=f(n, s) if want_grad else None, d_rays=f(n, 13) if want_grad else None)
    _lib.call("lb_composite_loss", _p(sigma, name="sigma"), _p(z, name="z"), _p(rays, name="rays"), _p(noise, name="noise"),
              _p(depth_gt, name="depth_gt"), C.byref(cfg), _p(n_opaque_dev, name="n_opaque_dev"), n, s, _p(out["weights"]), _p(out["depth"]),
              _p(out["opacity"]), _p(out["variance"]), _p(out["js"]), _p(out["eps"]), _p(loss_parts, name="loss_parts"),
              _p(out["d_sigma"]), _p(out["d_rays"]), _stream())
    return out


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=1.0, zero_grad=False):
    _lib.call("lb_adam_step", _p(p, name="p"), _p(g, name="g"), _p(m, name="m"), _p(v, name="v"), p.numel(),
              float(lr), float(beta1), float(beta2), float(eps), int(step), float(grad_scale), int(zero_grad), _stream())


def sgd_step(p, g, lr, zero_grad=False):
    _lib.call("lb_sgd_step", _p(p, name="p"), _p(g, name="g"), p.numel(), float(lr), int(zero_grad), _stream())


def occ_grad(rays, z, depth_gt, scale_factor, occ_dim, out=None):
    n, s = z.shape
    if out is None:
        out = torch.zeros(occ_dim, occ_dim, occ_dim, device=z.device, dtype=torch.float32)
    _lib.call("lb_occ_grad", _p(rays, name="rays"), _p(z, name="z"), _p(depth_gt, name="depth_gt"), float(scale_factor),
              n, s, occ_dim, _p(out, name="occ_grad"), _stream())
    return out


# ---- colour branch ----------------------------------------------------------------
def sh4_fwd(d01):
    P = d01.shape[0]
    out = torch.empty(P, 16, device=d01.device, dtype=torch.float32)
    _lib.call("lb_sh4_fwd", _p(d01, name="d01"), P, _p(out), _stream())
    return out


def sh4_bwd(d01, d_out):
    P = d01.shape[0]
    dx = torch.empty(P, 3, device=d01.device, dtype=torch.float32)
    _lib.call("lb_sh4_bwd", _p(d01, name="d01"), _p(d_out, name="d_out"), P, _p(dx), _stream())
    return dx


OUT_ACT = {"None": 0, "none": 0, None: 0, "Sigmoid": 1, "sigmoid": 1}


def net_fwd(x, params, n_hidden, n_out, out_act=0, save_acts=False):
    """tcnn.Network forward: x [P, n_in] (n_in multiple of 16) -> out [P, n_out] (+ activations for the backward)."""
    P, n_in = x.shape
    out = torch.empty(P, n_out, device=x.device, dtype=torch.float32)
    # activations in the kernels' tile-transposed layout [layer][tile of 128 points][16-byte chunk][row]
    acts = torch.empty(n_hidden, (P + 127) // 128 * 128, 64, device=x.device, dtype=torch.float32) if save_acts else None
    _lib.call("lb_net_fwd", _p(x, name="x"), P, n_in, n_hidden, n_out, out_act, _p(params, name="params"), _p(out),
              _p(acts), _stream())
    return out, acts


def net_bwd(x, params, acts, out, d_out, n_hidden, n_out, out_act=0, grad_params=None, need_dx=False):
    P, n_in = x.shape
    dx = torch.empty(P, n_in, device=x.device, dtype=torch.float32) if need_dx else None
    _lib.call("lb_net_bwd", _p(x, name="x"), P, n_in, n_hidden, n_out, out_act, _p(params, name="params"),
              _p(acts, name="acts"), _p(out, name="out"), _p(d_out, name="d_out"), _p(grad_params), _p(dx), _stream())
    return dx


def composite_rgb_fwd(weights, colors, white_bkgd):
    N, S, Cn = colors.shape
    rgb = torch.empty(N, Cn, device=colors.device, dtype=torch.float32)
    _lib.call("lb_composite_rgb_fwd", _p(weights, name="weights"), _p(colors, name="colors"), N, S, Cn,
              1 if white_bkgd else 0, _p(rgb), _stream())
    return rgb


def composite_rgb_bwd(weights, colors, g_rgb, white_bkgd, need_dw=True, need_dc=True):
    N, S, Cn = colors.shape
    dw = torch.empty(N, S, device=colors.device, dtype=torch.float32) if need_dw else None
    dc = torch.empty(N, S, Cn, device=colors.device, dtype=torch.float32) if need_dc else None
    _lib.call("lb_composite_rgb_bwd", _p(weights, name="weights"), _p(colors, name="colors"), _p(g_rgb, name="g_rgb"),
              N, S, Cn, 1 if white_bkgd else 0, _p(dw), _p(dc), _stream())
    return dw, dc


def voxel_splat_max(rays, z, weights, x_bounds, y_bounds, z_bounds, volume, ray_mask=None):
    """f3: max-weight voxel splat into volume [ny, nx, nz] (fp32, accumulated in place).  Bounds are fp64 device
    tensors (the axis grid coordinates); ray_mask [N] uint8 / bool or None."""
    n, s_ = z.shape
    if ray_mask is not None and ray_mask.dtype == torch.bool:
        ray_mask = ray_mask.to(torch.uint8)
    _lib.call("lb_voxel_splat_max", _p(rays, name="rays"), _p(z, name="z"), _p(weights, name="weights"),
              _p(ray_mask, torch.uint8, name="ray_mask"), n, s_, _p(x_bounds, torch.float64, name="x_bounds"),
              x_bounds.numel(), _p(y_bounds, torch.float64, name="y_bounds"), y_bounds.numel(),
              _p(z_bounds, torch.float64, name="z_bounds"), z_bounds.numel(), _p(volume, name="volume"), _stream())
    return volume
