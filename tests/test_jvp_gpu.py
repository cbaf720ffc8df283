"""lb_sigma_fwd_jvp / lb_rays_grad_from_jvp (include/loner_b200.h): the forward kernel that also returns d sigma / d x01,
used for pose refinement (reference src/mapping/optimizer.py:701-712,849: the lidar pose is a leaf, autograd reaches the
ray origins / directions through tcnn's input gradient).  Checker: the reverse-mode route of the same library
(lb_sigma_bwd with d_x01 / d_rays), which tests/test_reference_chain_gpu.py pins to the reference's recorded ray
gradients -- and, through SigmaFromRays (which now takes the forward-mode route), those reference tests themselves.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def make_inputs(n, s, seed, plan, table_scale=1.0):
    gen = torch.Generator().manual_seed(seed)
    params = (torch.rand(plan.n_params, generator=gen) - 0.5) * 0.6
    params[3072:] *= table_scale
    rays = torch.zeros(n, 13)
    rays[:, 0:3] = (torch.rand(n, 3, generator=gen) - 0.5) * 0.4
    d = torch.randn(n, 3, generator=gen)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    z = torch.rand(n, s, generator=gen).sort(dim=1).values * 0.7
    d_sigma = torch.randn(n, s, generator=gen)
    d_sigma[torch.rand(n, s, generator=gen) < 0.4] = 0.0          # ReLU-dead / masked samples
    return params.to(DEV), rays.to(DEV), z.contiguous().to(DEV), d_sigma.to(DEV)


@pytest.mark.parametrize("log2_T,n,s", [(15, 300, 129), (15, 37, 64), (19, 2048, 512), (14, 1, 1)])
def test_jvp_forward_equals_plain_forward_and_reverse_mode_gradient(log2_T, n, s):
    from loner_b200 import ops
    plan = ops.Plan(log2_hashmap_size=log2_T)
    assert plan.jvp_supported
    params, rays, z, d_sigma = make_inputs(n, s, 7 + n, plan)
    sigma0, (enc0, act0) = ops.sigma_fwd(plan, params, rays=rays, z=z)
    sigma1, (enc1, act1), dsdx = ops.sigma_fwd_jvp(plan, params, rays=rays, z=z)
    # same arithmetic for the primal outputs: bit-identical
    assert torch.equal(sigma0, sigma1) and torch.equal(enc0, enc1) and torch.equal(act0, act1)
    assert torch.isfinite(dsdx).all() and dsdx.shape == (n * s, 3)
    # (1) d sigma / d x01 against the reverse-mode input gradient for d_sigma = 1 (x01 point source)
    x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3) + 1) / 2
    x01 = x01.contiguous()
    _, saved = ops.sigma_fwd(plan, params, x01=x01)
    want = ops.sigma_bwd(plan, params, saved, torch.ones(n * s, device=DEV), x01=x01, need_dx=True)
    assert rel_err(dsdx, want) < 1e-4
    # the x01 source of the jvp kernel: same derivative
    sigma2, _, dsdx2 = ops.sigma_fwd_jvp(plan, params, x01=x01, save_enc=False)
    assert rel_err(sigma2, sigma0.reshape(-1)) < 1e-6 and rel_err(dsdx2, want) < 1e-4
    # (2) ray gradients: lb_rays_grad_from_jvp against lb_sigma_bwd's d_rays on a sparse d_sigma
    dr_want = torch.zeros_like(rays)
    gp_want = torch.zeros_like(params)
    ops.sigma_bwd(plan, params, (enc0, act0), d_sigma, rays=rays, z=z, grad_params=gp_want, d_rays=dr_want)
    dr = torch.zeros_like(rays)
    gp = torch.zeros_like(params)
    ops.sigma_bwd(plan, params, (enc1, act1), d_sigma, rays=rays, z=z, grad_params=gp)
    ops.rays_grad_from_jvp(z, dsdx, d_sigma, dr)
    assert rel_err(dr[:, :6], dr_want[:, :6]) < 1e-4
    assert (dr[:, 6:] == 0).all()
    assert rel_err(gp, gp_want) < 1e-5          # atomics order only


def test_jvp_against_finite_differences_of_the_forward():
    """A coarse sanity check that is independent of the reverse-mode kernels: central differences of the fp32 forward on
    a coarse-only field (fine-level tables zeroed).  Coarse because sigma is piecewise linear in x with a kink wherever
    one of the 64 hidden units changes sign or a cell boundary is crossed, and most +-h intervals contain one: the
    direction and the bulk magnitude must agree, not the last digits (those are pinned by the reference's recorded ray
    gradients in tests/test_reference_chain_gpu.py, which run through this kernel)."""
    from loner_b200 import ops
    plan = ops.Plan(log2_hashmap_size=15)
    params, rays, z, _ = make_inputs(64, 32, 3, plan)
    off = 3072
    for l in range(3, 16):
        lv = plan.level(l)
        params[off + 2 * lv["offset"]: off + 2 * (lv["offset"] + lv["size"])] = 0.0
    x01 = (((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3) + 1) / 2).contiguous()
    _, _, dsdx = ops.sigma_fwd_jvp(plan, params, x01=x01, save_enc=False)
    h = 2e-4
    fd = torch.zeros_like(dsdx)
    for d in range(3):
        e = torch.zeros(3, device=DEV)
        e[d] = h
        sp, _ = ops.sigma_fwd(plan, params, x01=(x01 + e).contiguous(), save_enc=False)
        sm, _ = ops.sigma_fwd(plan, params, x01=(x01 - e).contiguous(), save_enc=False)
        fd[:, d] = (sp.double() - sm.double()).float() / (2 * h)
    cos = torch.nn.functional.cosine_similarity(dsdx.reshape(1, -1).double(), fd.reshape(1, -1).double()).item()
    err = (dsdx - fd).abs().max(dim=1).values / fd.abs().max().clamp_min(1e-12)
    print("cosine", cos, "median err", err.median().item())
    assert cos > 0.99
    assert err.median().item() < 5e-2


def test_autograd_nodes_take_the_forward_mode_route():
    """SigmaFromRays / SigmaFromPoints with a differentiable input: gradients equal the reverse-mode kernels', and the
    backward launches no input-gradient scatter (lb_rays_grad_from_jvp / one elementwise product instead)."""
    from loner_b200 import ops, _lib
    from loner_b200 import functional as F
    plan = ops.Plan(log2_hashmap_size=15)
    params, rays, z, d_sigma = make_inputs(200, 96, 11, plan)
    p = params.clone().requires_grad_(True)
    r = rays.clone().requires_grad_(True)
    _lib.reset_counters()
    sigma = F.SigmaFromRays.apply(plan, p, r, z)
    (sigma * d_sigma).sum().backward()
    assert _lib.LAUNCHES.get("lb_sigma_fwd_jvp", 0) == 1 and _lib.LAUNCHES.get("lb_rays_grad_from_jvp", 0) == 1
    _, saved = ops.sigma_fwd(plan, params, rays=rays, z=z)
    dr_want, gp_want = torch.zeros_like(rays), torch.zeros_like(params)
    ops.sigma_bwd(plan, params, saved, d_sigma, rays=rays, z=z, grad_params=gp_want, d_rays=dr_want)
    assert rel_err(r.grad, dr_want) < 1e-4 and rel_err(p.grad, gp_want) < 1e-5
    # input gradient only (frozen field): no encodings are kept, no table scatter runs
    r2 = rays.clone().requires_grad_(True)
    _lib.reset_counters()
    (F.SigmaFromRays.apply(plan, params, r2, z) * d_sigma).sum().backward()
    assert "lb_sigma_bwd" not in _lib.LAUNCHES
    assert rel_err(r2.grad, dr_want) < 1e-4
    # x01 source (tcnn.NetworkWithInputEncoding with a differentiable input)
    x01 = (((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3) + 1) / 2).contiguous()
    x = x01.clone().requires_grad_(True)
    (F.SigmaFromPoints.apply(plan, params, x) * d_sigma.reshape(-1)).sum().backward()
    _, saved = ops.sigma_fwd(plan, params, x01=x01)
    dx_want = ops.sigma_bwd(plan, params, saved, d_sigma.reshape(-1).contiguous(), x01=x01, need_dx=True)
    assert rel_err(x.grad, dx_want) < 1e-4


def test_jvp_rejects_plans_beyond_the_one_kernel_path():
    from loner_b200 import ops
    from loner_b200._lib import LonerB200Error
    plan = ops.Plan(log2_hashmap_size=15, l2_bytes=700 * 1024)       # forces level groups
    assert not plan.jvp_supported
    params, rays, z, _ = make_inputs(8, 16, 5, plan)
    with pytest.raises(LonerB200Error, match="lb_sigma_fwd_jvp_supported"):
        ops.sigma_fwd_jvp(plan, params, rays=rays, z=z)
    # the autograd node falls back to the reverse-mode input gradient there
    from loner_b200 import functional as F
    r = rays.clone().requires_grad_(True)
    F.SigmaFromRays.apply(plan, params, r, z).sum().backward()
    assert torch.isfinite(r.grad).all() and r.grad[:, :6].abs().sum() > 0
