"""GPU parity tests: every C-ABI kernel (through loner_b200.ops) against the CPU oracle
(oracle/loner_oracle.py) and the golden vectors produced by the reference's own Python.

Bars (BASELINE.json north_star): occupancy bits and sample indices bit-exact; everything
floating-point within 1e-3 relative (we assert much tighter where fp32 allows).
"""
import numpy as np
import pytest
import torch

from conftest import load_golden, golden_params

pytestmark = pytest.mark.gpu

OGM_CASES = ["c1_ogm_l1js", "s128_ogm_l2js"]
ALL_CASES = OGM_CASES + ["c1_uniform_l2los"]
REL = 1e-3  # north_star tolerance


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def ops():
    from loner_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def O():
    from oracle import loner_oracle
    return loner_oracle


# ------------------------------------------------------------------------------------------ sampling
@pytest.mark.parametrize("name", OGM_CASES)
def test_sample_occ_bit_exact_golden(name, dev, ops, O):
    g = load_golden(name)
    ref = O.sample_occ(g["rays"], g["occ"], g["u1"], g["u2"], g["n_samples"], 1.0)
    z, bits, inds = ops.sample_occ(g["rays"].to(dev), g["occ"].to(dev).contiguous(), g["u1"].to(dev), g["u2"].to(dev),
                                   g["n_samples"], 1.0, debug=True)
    assert torch.equal(bits.cpu().bool(), ref["occ_bits"]), "occupancy bits differ from oracle"
    assert torch.equal(inds.cpu(), ref["inds"]), "searchsorted indices differ from oracle"
    assert torch.equal(z.cpu(), ref["z"]), "sorted depths differ from oracle"
    # and against the reference's own run: bit-exact (the benchmark-shaped fixture, where 5 of 524288 draws sit on a
    # cdf knot, is covered in test_reference_chain_gpu.py)
    assert torch.equal(bits.cpu().bool(), g["occ_bits"])
    assert torch.equal(inds.cpu(), g["inds"])


@pytest.mark.parametrize("S,N,D,perturb", [(512, 1024, 100, 1.0), (64, 777, 33, 1.0), (256, 300, 50, 0.0), (2048, 64, 100, 0.0)])
def test_sample_occ_bit_exact_seeded(S, N, D, perturb, dev, ops, O):
    gen = torch.Generator().manual_seed(S + N)
    o = (torch.rand(N, 3, generator=gen) - 0.5) * 0.6
    d = torch.nn.functional.normalize(torch.randn(N, 3, generator=gen), dim=1)
    near = torch.full((N, 1), 0.02)
    far = near + 0.05 + 0.3 * torch.rand(N, 1, generator=gen)
    rays = torch.cat([o, d, -d, torch.zeros(N, 2), near, far], 1)
    occ = torch.randn(1, 1, D, D, D, generator=gen) * 0.5
    occ[occ.abs() < 0.2] = 0.0        # exact-zero logits exercise the prob == 0 branch
    u1 = torch.rand(N, S // 2, generator=gen)
    u2 = torch.rand(N, S // 2, generator=gen)
    u2[0, :4] = torch.tensor([0.0, 1.0 - 2 ** -24, 0.5, 2 ** -30])
    ref = O.sample_occ(rays, occ, u1, u2, S, perturb)
    z, bits, inds = ops.sample_occ(rays.to(dev), occ.to(dev), u1.to(dev) if perturb > 0 else None, u2.to(dev), S, perturb, debug=True)
    assert torch.equal(bits.cpu().bool(), ref["occ_bits"])
    assert torch.equal(inds.cpu(), ref["inds"])
    assert torch.equal(z.cpu(), ref["z"])
    assert (z[:, 1:] >= z[:, :-1]).all()


def test_sample_uniform_exact(dev, ops, O):
    g = load_golden("c1_uniform_l2los")
    z = ops.sample_uniform(g["rays"].to(dev), g["u1"].to(dev), g["n_samples"], 1.0)
    assert torch.equal(z.cpu(), g["z"])
    z0 = ops.sample_uniform(g["rays"].to(dev), None, 96, 0.0)
    assert torch.equal(z0.cpu(), O.sample_uniform(g["rays"], None, 96, 0.0))


# ------------------------------------------------------------------------------------------ field
@pytest.mark.parametrize("log2T", [14, 19])
def test_encode_fwd_bwd(log2T, dev, ops, O):
    gen = torch.Generator().manual_seed(log2T)
    plan = ops.Plan(log2_hashmap_size=log2T)
    P = 5000
    x = torch.rand(P, 3, generator=gen)
    x[:8] = torch.tensor([[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [1, 0, 0], [0, 1, 0.25], [0.999999, 0.5, 0], [1e-7, 1e-7, 1e-7], [0.25, 0.75, 1]])
    table = (torch.rand(plan.n_table_params, generator=gen) * 2 - 1) * 0.3
    de = torch.randn(P, 32, generator=gen)
    xr, tr = x.clone().requires_grad_(True), table.clone().requires_grad_(True)
    ref = O.hash_encode(xr, tr, log2T)
    ref.backward(de)
    enc = ops.encode_fwd(plan, x.to(dev), table.to(dev))
    assert rel_err(enc, ref) < 1e-5
    gt = torch.zeros_like(table, device=dev)
    dx = ops.encode_bwd(plan, x.to(dev), table.to(dev), de.to(dev), gt, need_dx=True)
    assert rel_err(gt, tr.grad) < 1e-4
    # positions exactly on cell faces have one-sided derivatives; compare the generic points
    assert rel_err(dx[8:], xr.grad[8:]) < 1e-4


def test_encode_bwd_skips_zero_rows(dev, ops, O):
    """Rows of dL/d_enc that are exactly zero (ReLU-inactive samples) issue no atomics; whole zero warps are
    skipped, and dL/dx of such points is written as zero."""
    gen = torch.Generator().manual_seed(5)
    plan = ops.Plan(log2_hashmap_size=15)
    P = 4099
    x = torch.rand(P, 3, generator=gen)
    table = (torch.rand(plan.n_table_params, generator=gen) * 2 - 1) * 0.3
    de = torch.randn(P, 32, generator=gen)
    de[torch.rand(P, generator=gen) < 0.5] = 0          # scattered zero rows
    de[1024:2048] = 0                                   # whole warps / tiles of zero rows
    de[7] = -0.0
    xr, tr = x.clone().requires_grad_(True), table.clone().requires_grad_(True)
    O.hash_encode(xr, tr, 15).backward(de)
    gt = torch.zeros_like(table, device=dev)
    dx = ops.encode_bwd(plan, x.to(dev), table.to(dev), de.to(dev), gt, need_dx=True)
    assert rel_err(gt, tr.grad) < 1e-4
    assert rel_err(dx, xr.grad) < 1e-4
    assert (dx[1024:2048] == 0).all()
    gt2 = torch.zeros_like(gt)
    ops.encode_bwd(plan, x.to(dev), table.to(dev), de.to(dev), gt2, need_dx=False)
    assert rel_err(gt2, tr.grad) < 1e-4


def _need_test_build(tc):
    from loner_b200 import _lib
    if not tc and not _lib.test_switches():
        pytest.skip("fp32 CUDA-core MLP cross-check: test build only (python -m loner_b200.build --test-switches)")


@pytest.mark.parametrize("tc", [True, False], ids=["tcgen05", "simt"])
@pytest.mark.parametrize("name", ALL_CASES)
def test_sigma_fwd_bwd_golden(name, tc, dev, ops, O):
    _need_test_build(tc)
    g = load_golden(name)
    p = golden_params(g)
    plan = ops.Plan(log2_hashmap_size=g["log2_T"])
    assert plan.n_params == p.numel()
    rays, z = g["rays"].to(dev), g["z"].to(dev)
    sigma, enc = ops.sigma_fwd(plan, p.to(dev), rays=rays, z=z, tensor_cores=tc)
    assert (enc[1] is not None) == tc
    assert rel_err(sigma, g["sigma"]) < 1e-4
    # backward against oracle autograd for a random upstream gradient
    gen = torch.Generator().manual_seed(1)
    ds = torch.randn(z.shape, generator=gen)
    # A hidden unit whose pre-activation is a rounding tie (|h| < 1e-6 sum|W1||e|) has no well-defined ReLU
    # derivative in floating point: which side an evaluation lands on depends on its summation order (tcnn
    # itself evaluates in fp16).  The upstream gradient is zeroed at such points (one in 2 M units here) so
    # that every other contribution can be compared tightly.
    pr, rr = p.clone().requires_grad_(True), g["rays"].clone().requires_grad_(True)
    xyz = rr[:, None, 0:3] + rr[:, None, 3:6] * g["z"][..., None]
    with torch.no_grad():
        e0 = O.hash_encode((xyz.reshape(-1, 3) + 1) / 2, p[3072:], g["log2_T"]).double()
        w1 = p[:2048].view(64, 32).double()
        tie = ((e0 @ w1.t()).abs() < 1e-6 * (e0.abs() @ w1.abs().t())).any(1)
        ds.view(-1)[tie] = 0
    O.sigma_field(xyz.reshape(-1, 3), pr, g["log2_T"]).view(z.shape).backward(ds)
    gp = torch.zeros_like(p, device=dev)
    dr = torch.zeros(rays.shape, device=dev)
    ops.sigma_bwd(plan, p.to(dev), enc, ds.to(dev), rays=rays, z=z, grad_params=gp, d_rays=dr)
    assert rel_err(gp[:3072], pr.grad[:3072]) < 1e-4
    assert rel_err(gp[3072:], pr.grad[3072:]) < 1e-4
    assert rel_err(dr[:, :6], rr.grad[:, :6]) < 1e-3
    assert (dr[:, 6:] == 0).all()
    # x01 point source (tcnn module path) gives the same numbers
    x01 = ((xyz.detach().reshape(-1, 3) + 1) / 2).to(dev).contiguous()
    s2, enc2 = ops.sigma_fwd(plan, p.to(dev), x01=x01, tensor_cores=tc)
    assert rel_err(s2.view(z.shape), g["sigma"]) < 1e-4
    gp2 = torch.zeros_like(gp)
    dx = ops.sigma_bwd(plan, p.to(dev), enc2, ds.to(dev).reshape(-1), x01=x01, grad_params=gp2, need_dx=True)
    assert rel_err(gp2, pr.grad) < 1e-4
    assert dx.shape == (x01.shape[0], 3)
    # frozen sigma head (tracking phase): ray gradients only
    dr3 = torch.zeros(rays.shape, device=dev)
    _, enc3 = ops.sigma_fwd(plan, p.to(dev), rays=rays, z=z, tensor_cores=tc)      # sigma_bwd consumes its enc buffer
    ops.sigma_bwd(plan, p.to(dev), enc3, ds.to(dev), rays=rays, z=z, grad_params=None, d_rays=dr3)
    assert rel_err(dr3, dr) < 1e-5


@pytest.mark.parametrize("n,s", [(1, 1), (3, 7), (5, 128), (300, 129), (4099, 64)])
def test_fused_sigma_forward_matches_two_kernel_path(n, s, dev, ops):
    """The one-kernel gather + MLP forward (rays, ring hand-off in shared memory) against the two-kernel path (x01
    point source): identical encodings, sigma and masks; ragged tiles, many tiles per CTA; and the inference call
    that saves nothing returns the same sigma."""
    gen = torch.Generator().manual_seed(n * 1000 + s)
    plan = ops.Plan(log2_hashmap_size=15)
    assert plan.enc_optional
    params = ((torch.rand(plan.n_params, generator=gen) - 0.5) * 0.6).to(dev)
    rays = torch.zeros(n, 13)
    rays[:, 0:3] = (torch.rand(n, 3, generator=gen) - 0.5) * 0.4
    d = torch.randn(n, 3, generator=gen)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    z = torch.rand(n, s, generator=gen).sort(dim=1).values * 0.7
    rays, z = rays.to(dev), z.to(dev)
    sig, (enc, act) = ops.sigma_fwd(plan, params, rays=rays, z=z)
    x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3) + 1) / 2
    enc2 = ops.encode_fwd(plan, x01.contiguous(), params[plan.n_mlp_params:])
    assert rel_err(enc, enc2) < 1e-5                 # xyz is formed with an fma inside the kernel
    sig2, (_, act2) = ops.sigma_fwd(plan, params, x01=x01.contiguous())
    assert rel_err(sig.view(-1), sig2) < 1e-4
    sig3, saved = ops.sigma_fwd(plan, params, rays=rays, z=z, save_enc=False)
    assert saved is None
    assert torch.equal(sig3, sig)
    # masks agree except at rounding ties of the hidden pre-activation (the two point sources round xyz differently)
    x = act ^ act2
    flips = sum(int(((x >> b) & 1).sum()) for b in range(64))
    assert flips <= max(1, int(1e-4 * 64 * n * s))


@pytest.mark.parametrize("P", [1, 127, 128, 129, 5000, 128 * 700 + 5])
def test_mlp_tensor_cores_match_fp32(P, dev, ops, O):
    """tcgen05 MLP (bf16-split operands, fp32 TMEM accumulators) against the oracle's fp32 MLP and the fp32
    CUDA-core kernels: ragged tile tails, many tiles per CTA, large activations with mixed signs."""
    gen = torch.Generator().manual_seed(P)
    log2T = 14
    plan = ops.Plan(log2_hashmap_size=log2T)
    p = torch.cat([(torch.rand(3072, generator=gen) - 0.5) * 1.5, (torch.rand(plan.n_table_params, generator=gen) - 0.5) * 4.0])
    x = torch.rand(P, 3, generator=gen)
    ds = torch.randn(P, generator=gen)
    with torch.no_grad():     # no upstream gradient at ReLU rounding ties (see test_sigma_fwd_bwd_golden)
        e0 = O.hash_encode(x, p[3072:], log2T).double()
        w1 = p[:2048].view(64, 32).double()
        ds[((e0 @ w1.t()).abs() < 1e-6 * (e0.abs() @ w1.abs().t())).any(1)] = 0
    pr, xr = p.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ref = O.sigma_mlp(O.hash_encode(xr, pr[3072:], log2T), pr[:3072])
    ref.backward(ds)
    out = {}
    from loner_b200 import _lib
    for tc in ((True, False) if _lib.test_switches() else (True,)):
        sigma, saved = ops.sigma_fwd(plan, p.to(dev), x01=x.to(dev), tensor_cores=tc)
        gp = torch.zeros_like(p, device=dev)
        dx = ops.sigma_bwd(plan, p.to(dev), saved, ds.to(dev), x01=x.to(dev), grad_params=gp, need_dx=True)
        out[tc] = (sigma, gp, dx)
        assert rel_err(sigma, ref) < 2e-5, tc
        assert rel_err(gp[:2048], pr.grad[:2048]) < 1e-4, tc
        assert rel_err(gp[2048:2112], pr.grad[2048:2112]) < 1e-4, tc
        assert (gp[2112:3072] == 0).all()
        assert rel_err(gp[3072:], pr.grad[3072:]) < 1e-4, tc
        assert rel_err(dx, xr.grad) < 1e-3, tc
    if False in out:
        assert rel_err(out[True][0], out[False][0]) < 2e-5


# ------------------------------------------------------------------------------------------ compositing
@pytest.mark.parametrize("name", ALL_CASES)
def test_composite_fwd_golden(name, dev, ops, O):
    g = load_golden(name)
    w, depth, opac, var = ops.composite_fwd(g["sigma"].to(dev), g["z"].to(dev), g["rays"].to(dev), g["noise"].to(dev))
    assert rel_err(w, g["weights"]) < 1e-5
    assert rel_err(depth, g["depth"]) < 1e-5
    assert rel_err(opac, g["opacity"]) < 1e-5
    assert rel_err(var, g["variance"]) < 1e-4
    if "adj_depth" in g:
        w2, d2, _, v2 = ops.composite_fwd(g["sigma"].to(dev), g["z"].to(dev), g["rays"].to(dev), g["noise"].to(dev), strategy="adjusted")
        assert (d2.cpu() != g["adj_depth"]).float().mean().item() < 0.01     # a crossing within 1 ulp of 0.5 may move one sample
        assert rel_err(w2, g["adj_weights"]) < 1e-5
        same = d2.cpu() == g["adj_depth"]
        assert rel_err(v2.cpu()[same], g["adj_variance"][same]) < 1e-4


@pytest.mark.parametrize("S,N", [(64, 130), (512, 257), (100, 33), (2048, 9)])
def test_composite_bwd_matches_autograd(S, N, dev, ops, O):
    gen = torch.Generator().manual_seed(S * 7 + N)
    z = torch.sort(torch.rand(N, S, generator=gen) * 0.4 + 0.01, dim=1).values
    sigma = torch.randn(N, S, generator=gen) * 30
    noise = torch.randn(N, S, generator=gen)
    d = torch.randn(N, 3, generator=gen)
    far = torch.rand(N, 1, generator=gen) * 0.3 + 0.45
    rays = torch.cat([torch.zeros(N, 3), d, -d, torch.zeros(N, 2), torch.full((N, 1), 0.01), far], 1)
    gw, gd, go, gv = (torch.randn(N, S, generator=gen), torch.randn(N, generator=gen),
                      torch.randn(N, generator=gen), torch.randn(N, generator=gen))
    sr, rr = sigma.clone().requires_grad_(True), rays.clone().requires_grad_(True)
    c = O.composite(sr, z, rr[:, 3:6], rr[:, -1:], noise)
    ((c["weights"] * gw).sum() + (c["depth"] * gd).sum() + (c["opacity"] * go).sum() + (c["variance"] * gv).sum()).backward()
    w, depth, opac, var = ops.composite_fwd(sigma.to(dev), z.to(dev), rays.to(dev), noise.to(dev))
    assert rel_err(w, c["weights"]) < 1e-5 and rel_err(depth, c["depth"]) < 1e-5 and rel_err(var, c["variance"]) < 1e-4
    dr = torch.zeros(N, 13, device=dev)
    ds = ops.composite_bwd(sigma.to(dev), z.to(dev), rays.to(dev), noise.to(dev), gw.to(dev), gd.to(dev), go.to(dev), gv.to(dev), d_rays=dr)
    assert rel_err(ds, sr.grad) < 1e-4
    assert rel_err(dr, rr.grad) < 1e-4


# ------------------------------------------------------------------------------------------ loss
@pytest.mark.parametrize("name", ALL_CASES)
def test_composite_loss_golden(name, dev, ops, O):
    g = load_golden(name)
    rays, z, sig, noise, dgt = (g[k].to(dev) for k in ("rays", "z", "sigma", "noise", "depths"))
    n_op = ops.count_opaque(rays, dgt)
    far = g["rays"][:, -1]
    assert n_op.tolist() == [float(((g["depths"] > 0) & ~(g["depths"] > far)).sum()), float(rays.shape[0])]
    cfg = ops.make_loss_config(g["loss_selection"], float(g["scale"]), n_op[0].item(), z.numel(), min_js_score=g["min_js"])
    parts = torch.zeros(4, device=dev)
    out = ops.composite_loss(sig, z, rays, noise, dgt, cfg, parts)
    pc = parts.cpu()
    loss = 0.005 * pc[0] / pc[3] + 1000.0 * pc[1] / z.numel() + pc[2] / pc[3]
    assert abs(loss.item() - g["loss"].item()) < 1e-4 * abs(g["loss"].item())
    assert rel_err(out["weights"], g["weights"]) < 1e-5
    assert abs(out["eps"].mean().item() - g["depth_eps"]) < 1e-4 * g["depth_eps"]
    # gradients: oracle autograd from sigma on the same samples
    sr, rr = g["sigma"].clone().requires_grad_(True), g["rays"].clone().requires_grad_(True)
    c = O.composite(sr, g["z"], rr[:, 3:6], rr[:, -1:], g["noise"])
    ls = O.depth_loss(c["weights"], c["depth"], c["opacity"], g["z"], rr[:, -1:], g["depths"], g["scale"],
                      dict(loss_selection=g["loss_selection"], min_js_score=g["min_js"]))
    ls["loss"].backward()
    assert rel_err(out["js"], ls["js"]) < 1e-3
    assert rel_err(out["eps"], ls["eps"]) < 1e-4
    assert rel_err(out["d_sigma"], sr.grad) < 1e-3
    assert rel_err(out["d_rays"], rr.grad) < 1e-3


# ------------------------------------------------------------------------------------------ optimiser / occupancy
def test_adam_and_sgd(dev, ops, O):
    gen = torch.Generator().manual_seed(0)
    n = 100003
    p = torch.randn(n, generator=gen); m = torch.zeros(n); v = torch.zeros(n)
    pd, md, vd = p.to(dev), m.to(dev), v.to(dev)
    for step in range(1, 5):
        gr = torch.randn(n, generator=gen) * 0.1
        gr[::3] = 0
        O.adam_step(p, gr, m, v, step, 0.01)
        gd = gr.to(dev)
        ops.adam_step(pd, gd, md, vd, step, 0.01, zero_grad=(step == 4))
    assert rel_err(pd, p) < 1e-5 and rel_err(md, m) < 1e-5 and rel_err(vd, v) < 1e-5
    assert (gd == 0).all()
    q = torch.randn(1000, generator=gen); gq = torch.randn(1000, generator=gen)
    qd, gqd = q.to(dev), gq.to(dev)
    ops.sgd_step(qd, gqd, 1e-2)
    assert rel_err(qd, q - 1e-2 * gq) < 1e-6


@pytest.mark.parametrize("name", OGM_CASES)
def test_occ_update_golden(name, dev, ops, O):
    g = load_golden(name)
    D = g["occ"].shape[-1]
    grad = ops.occ_grad(g["rays"].to(dev), g["z"].to(dev), g["depths"].to(dev), float(g["scale"]), D)
    occ = g["occ"].to(dev).clone().contiguous()
    ops.sgd_step(occ.view(-1), grad.view(-1), 1e-4)
    assert rel_err(occ - g["occ"].to(dev), g["occ_after"] - g["occ"]) < 1e-4
    assert rel_err(occ, g["occ_after"]) < 1e-6


# ------------------------------------------------------------------------------------------ ray construction
@pytest.mark.parametrize("far_clip", [False, True])
def test_build_lidar_rays_matches_oracle(far_clip, dev, O):
    """lb_build_lidar_rays (+ backward to the pose 6-vector) against the oracle's restatement of
    ray_utils.py:269-322; far_clip places the sensor near a cube face so that far = cube exit for most rays."""
    from loner_b200 import mapping as M
    gen = torch.Generator().manual_seed(7)
    n = 3000
    dirs = torch.randn(3, n, generator=gen)
    dirs = dirs / dirs.norm(dim=0, keepdim=True)
    dist = torch.rand(n, generator=gen) * 40 + 1
    idx = torch.randint(n, (777,), generator=gen)
    pose = torch.tensor([20.0 if far_clip else 1.0, -2.0, 0.5, 0.3, -0.2, 0.9])
    scale, shift = torch.tensor(60.0 if not far_clip else 25.0), torch.tensor([0.5, 1.0, -0.25])
    rr = torch.tensor([1.0, 50.0])
    pr = pose.clone().requires_grad_(True)
    rays_ref, depths_ref = O.build_lidar_rays(dirs[:, idx], dist[idx], pr, shift, scale, rr)
    g = torch.randn(rays_ref.shape, generator=gen)
    rays_ref.backward(g)
    pd = pose.to(dev).requires_grad_(True)
    wc = M.WorldCube(scale.to(dev), shift.to(dev))
    rays, depths = M.build_lidar_rays(dirs.to(dev), dist.to(dev), pd, wc, (1.0, 50.0), ignore_world_cube=False,
                                      indices=idx.to(dev))
    assert rays.shape == rays_ref.shape
    if far_clip:
        assert (rays_ref[:, 12] < 50.0 / 25.0 - 1e-6).float().mean() > 0.3
    assert rel_err(rays, rays_ref) < 1e-6 and rel_err(depths, depths_ref) < 1e-6
    rays.backward(g.to(dev))
    assert rel_err(pd.grad, pr.grad) < 1e-4
    # sky rays: constant depth, detached pose
    srays, sdepths = M.build_lidar_rays(dirs.to(dev), None, pd.detach(), wc, (1.0, 50.0), False, indices=idx.to(dev),
                                        const_depth=51.0)
    assert torch.equal(srays, rays.detach()) and torch.allclose(sdepths, torch.full_like(sdepths, 51.0 / float(scale)))


# ---- colour branch ---------------------------------------------------------------
def test_sh4_matches_oracle():
    from loner_b200 import ops
    from oracle import loner_oracle as O
    torch.manual_seed(5)
    d01 = torch.rand(4099, 3)
    ref_in = d01.clone().requires_grad_(True)
    ref = O.sh4(ref_in)
    out = ops.sh4_fwd(d01.to("cuda:0"))
    assert rel_err(out, ref) < 1e-6
    g = torch.randn(4099, 16)
    ref.backward(g)
    dx = ops.sh4_bwd(d01.to("cuda:0"), g.to("cuda:0"))
    assert rel_err(dx, ref_in.grad) < 1e-5


def _mlp_ref(x, params, n_in, n_hidden, n_out, sigmoid):
    """fp64 evaluation of tcnn.Network (layout as oracle/tinycudann._MLP).  Also returns, per point, the smallest
    |pre-activation| relative to sum|w||h_prev| over all hidden units: points below ~1e-5 sit on a ReLU tie that
    fp32 / tf32-split arithmetic may resolve the other way, which changes that point's gradient by O(1)."""
    off, h = 0, x.double()
    shapes = [(64, n_in)] + [(64, 64)] * (n_hidden - 1) + [(16, 64)]
    margin = torch.full((x.shape[0],), float("inf"), dtype=torch.float64)
    for li, (o, i) in enumerate(shapes):
        W = params[off:off + o * i].view(o, i).double()
        off += o * i
        pre = h @ W.t()
        if li < len(shapes) - 1:
            scale = (h.abs() @ W.abs().t()).clamp_min(1e-30)
            margin = torch.minimum(margin, (pre.abs() / scale).min(dim=1).values.detach())
            h = torch.relu(pre)
        else:
            h = pre
    h = h[:, :n_out]
    return (torch.sigmoid(h) if sigmoid else h), margin


@pytest.mark.parametrize("P,n_in,n_hidden,n_out,act", [(1, 48, 4, 3, 1), (127, 48, 4, 3, 1), (129, 32, 1, 1, 0),
                                                       (5000, 16, 2, 16, 0), (33000, 64, 4, 4, 1), (777, 32, 6, 2, 0), (70001, 48, 4, 3, 0)])
def test_fused_network_matches_fp64(P, n_in, n_hidden, n_out, act):
    """tcnn.Network forward (tcgen05, tf32 hi/lo 3-pass) and backward (fp32 tiles) vs an fp64 evaluation."""
    from loner_b200 import ops
    torch.manual_seed(P)
    n_params = 64 * n_in + (n_hidden - 1) * 64 * 64 + 16 * 64
    params = (torch.rand(n_params) * 2 - 1) * 0.3
    x = torch.randn(P, n_in) * 0.5
    pr = params.clone().requires_grad_(True)
    xr = x.clone().requires_grad_(True)
    ref, margin = _mlp_ref(xr, pr, n_in, n_hidden, n_out, act == 1)
    out, acts = ops.net_fwd(x.to("cuda:0"), params.to("cuda:0"), n_hidden, n_out, act, save_acts=True)
    assert rel_err(out, ref) < 1e-5          # tf32 hi/lo 3-pass: ~2^-21 per layer
    out_inf, none = ops.net_fwd(x.to("cuda:0"), params.to("cuda:0"), n_hidden, n_out, act, save_acts=False)
    assert none is None and torch.equal(out_inf, out)
    g = torch.randn(P, n_out) * (margin > 1e-5).float()[:, None]      # no upstream gradient at ReLU ties
    ref.backward(g.double())
    gp = torch.zeros(n_params, device="cuda:0")
    dx = ops.net_bwd(x.to("cuda:0"), params.to("cuda:0"), acts, out, g.to("cuda:0"), n_hidden, n_out, act, grad_params=gp, need_dx=True)
    assert rel_err(gp, pr.grad) < 1e-4
    assert rel_err(dx, xr.grad) < 1e-4
    used_out = n_out * 64
    out_off = n_params - 16 * 64
    assert (gp[out_off + used_out:] == 0).all()          # padded output rows get no gradient


def test_composite_rgb_matches_oracle():
    from loner_b200 import ops
    from oracle import loner_oracle as O
    torch.manual_seed(9)
    N, S = 301, 64
    w = (torch.rand(N, S) / S).requires_grad_(True)
    col = torch.rand(N, S, 3).requires_grad_(True)
    for white in (True, False):
        ref = O.composite_rgb(w, col, white)
        out = ops.composite_rgb_fwd(w.detach().to("cuda:0"), col.detach().to("cuda:0"), white)
        assert rel_err(out, ref) < 1e-5
        g = torch.randn(N, 3)
        w.grad = col.grad = None
        ref.backward(g)
        dw, dc = ops.composite_rgb_bwd(w.detach().to("cuda:0"), col.detach().to("cuda:0"), g.to("cuda:0"), white)
        assert rel_err(dw, w.grad) < 1e-5 and rel_err(dc, col.grad) < 1e-5


def test_level_group_passes_match_single_pass(monkeypatch):
    """Tables larger than the L2 budget are walked in level groups (one launch per group, lb_encode.cu); forcing a
    tiny budget on a small table must reproduce the single-launch results: forward bit for bit, gradients to
    rounding (atomics reorder)."""
    from loner_b200 import ops
    dev = "cuda:0"
    torch.manual_seed(11)
    N, S, T = 96, 64, 15
    plan1 = ops.Plan(log2_hashmap_size=T)
    planN = ops.Plan(log2_hashmap_size=T, l2_bytes=700 * 1024)     # ~385 KB budget: groups of 2 hashed levels
    params = torch.cat([(torch.rand(3072) - 0.5) * 0.5, (torch.rand(plan1.n_table_params) - 0.5) * 0.4]).to(dev)
    rays = torch.zeros(N, 13, device=dev)
    rays[:, 0:3] = (torch.rand(N, 3, device=dev) - 0.5) * 0.2
    d = torch.randn(N, 3, device=dev)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    z = torch.sort(torch.rand(N, S, device=dev) * 0.8, dim=1).values.contiguous()
    x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None] + 1) / 2).reshape(-1, 3).contiguous()
    table = params[3072:].contiguous()
    e1, eN = ops.encode_fwd(plan1, x01, table), ops.encode_fwd(planN, x01, table)
    assert torch.equal(e1, eN)
    s1, sv1 = ops.sigma_fwd(plan1, params, rays=rays, z=z)
    sN, svN = ops.sigma_fwd(planN, params, rays=rays, z=z)
    assert torch.equal(s1, sN)
    d_enc = torch.randn(N * S, 32, device=dev)
    g1, gN = torch.zeros_like(table), torch.zeros_like(table)
    dx1 = ops.encode_bwd(plan1, x01, table, d_enc, g1, True)
    dxN = ops.encode_bwd(planN, x01, table, d_enc, gN, True)
    assert rel_err(gN, g1) < 1e-5 and rel_err(dxN, dx1) < 1e-5
    ds = torch.randn(N, S, device=dev)
    gp1, gpN = torch.zeros_like(params), torch.zeros_like(params)
    dr1, drN = torch.zeros_like(rays), torch.zeros_like(rays)
    ops.sigma_bwd(plan1, params, sv1, ds, rays=rays, z=z, grad_params=gp1, d_rays=dr1)
    ops.sigma_bwd(planN, params, svN, ds, rays=rays, z=z, grad_params=gpN, d_rays=drN)
    assert rel_err(gpN, gp1) < 1e-5 and rel_err(drN, dr1) < 1e-4


@pytest.mark.parametrize("n,s,dims", [(1, 1, (3, 2, 4)), (37, 65, (17, 9, 23)), (2000, 128, (60, 40, 31))])
def test_voxel_splat_max_matches_bucketize_amax(n, s, dims, dev, ops):
    """lb_voxel_splat_max (mesher.py:137-180) against torch.bucketize + scatter amax on the same points: bound filter,
    fp64 boundaries, voxel index layout [y][x][z], exact maxima, ray mask; accumulation over two calls."""
    gen = torch.Generator().manual_seed(n + s)
    rays = torch.zeros(n, 13)
    rays[:, 0:3] = (torch.rand(n, 3, generator=gen) - 0.5) * 0.6
    d = torch.randn(n, 3, generator=gen)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    z = torch.rand(n, s, generator=gen).sort(dim=1).values * 1.2
    w = torch.rand(n, s, generator=gen) * (torch.rand(n, s, generator=gen) < 0.7)
    mask = torch.rand(n, generator=gen) < 0.8
    nx, ny, nz = dims
    xb = torch.from_numpy(np.linspace(-0.61, 0.52, nx))
    yb = torch.from_numpy(np.linspace(-0.3, 0.45, ny))
    zb = torch.from_numpy(np.linspace(-0.2, 0.7, nz))
    # put a few points exactly on boundaries (bucketize right=False edge) and on the bound limits
    if n > 1:
        rays[0, 0:3] = torch.tensor([xb[1], yb[0], zb[nz - 1]]).float()
        z[0, 0] = 0.0
        w[0, 0] = 0.5
        mask[0] = True

    def reference(vol, m):
        pts = rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]
        pts, ww = pts[m].reshape(-1, 3), w[m].reshape(-1)
        good = torch.ones(pts.shape[0], dtype=torch.bool)
        for a, b in enumerate((xb, yb, zb)):
            good &= (pts[:, a].double() >= b[0]) & (pts[:, a].double() <= b[-1])
        pts, ww = pts[good], ww[good]
        idx = torch.bucketize(pts[:, 0].double().contiguous(), xb) * nz + \
            torch.bucketize(pts[:, 1].double().contiguous(), yb) * nx * nz + torch.bucketize(pts[:, 2].double().contiguous(), zb)
        ok = idx < vol.numel()
        return vol.view(-1).scatter_reduce(0, idx[ok], ww[ok], "amax", include_self=True).view(ny, nx, nz)

    vol = torch.zeros(ny, nx, nz, device=dev)
    ops.voxel_splat_max(rays.to(dev), z.to(dev), w.to(dev), xb.to(dev), yb.to(dev), zb.to(dev), vol, ray_mask=mask.to(dev))
    ref = reference(torch.zeros(ny, nx, nz), mask)
    assert torch.equal(vol.cpu(), ref)
    ops.voxel_splat_max(rays.to(dev), z.to(dev), (w * 0.5).to(dev), xb.to(dev), yb.to(dev), zb.to(dev), vol)    # accumulate, no mask
    w_keep = w
    w = w * 0.5
    ref2 = reference(ref.clone(), torch.ones(n, dtype=torch.bool))
    w = w_keep
    assert torch.equal(vol.cpu(), ref2)
