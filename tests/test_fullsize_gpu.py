"""GPU tests at BASELINE.json's full sizes (configs[1]/[2]: 32768 rays x 512 samples, T = 2^19; configs[3]: T = 2^22;
configs[4]: 2048 samples per ray) through size-independent properties -- the oracle cannot run these sizes in
seconds -- plus oracle spot checks on random subsets, and the empty / ragged inputs of every entry point.

Properties used:
  * sampler: rows ascending, inside [near, far], deterministic, and a random subset of rows bit-exact vs. the oracle;
  * hash grid: the encoding is linear in the table, so <E(theta; x), g> == <theta, E^T(g; x)> (adjoint identity), for
    the single-launch path (T = 2^19) and the level-group passes (T = 2^22);
  * sigma head: fused gather + tensor-core MLP vs. a plain fp32 torch MLP on the stand-alone encodings; backward
    linear in d_sigma and equal to torch autograd through the same plain MLP;
  * compositing + loss: weights >= 0, opacity = sum w <= 1, depth inside [near, far], loss parts and d_sigma equal to
    torch autograd through a plain restatement at full size;
  * Adam: equal to torch.optim.Adam on 14.2 M and 105 M parameters.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

N_FULL, S_FULL = 32768, 512


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


@pytest.fixture(scope="module")
def scene(dev):
    """32768 synthetic OS1-64 rays from 8 keyframes (the bench workload), a warmed occupancy grid."""
    from loner_b200 import mapping as M, synth
    torch.manual_seed(7)
    kfs, scale, shift = synth.make_keyframes(8, ray_range=(1.0, 50.0), seed=3)
    wc = M.WorldCube(scale.to(dev), shift.to(dev))
    rays, depths = [], []
    for k in kfs:
        idx = torch.randint(k.distances.shape[0], (N_FULL // 8,), device=dev)
        r, d = M.build_lidar_rays(k.directions.to(dev), k.distances.to(dev), k.pose.to(dev), wc, (1.0, 50.0), True, indices=idx)
        rays.append(r); depths.append(d)
    rays, depths = torch.vstack(rays).contiguous(), torch.cat(depths).contiguous()
    occ = (torch.randn(100, 100, 100, device=dev) * 0.7).contiguous()
    return dict(rays=rays, depths=depths, occ=occ, scale=float(scale))


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


# ------------------------------------------------------------------------------------------ sampler
@pytest.mark.parametrize("S", [512, 2048])
def test_sampler_full_size_properties(S, scene, dev, ops, O):
    n = N_FULL if S == 512 else 4096
    rays, occ = scene["rays"][:n].contiguous(), scene["occ"]
    g = torch.Generator(device=dev).manual_seed(11)
    u1 = torch.rand(n, S // 2, device=dev, generator=g)
    u2 = torch.rand(n, S // 2, device=dev, generator=g)
    z, bits, inds = ops.sample_occ(rays, occ, u1, u2, S, 1.0, debug=True)
    assert z.shape == (n, S) and torch.isfinite(z).all()
    assert (z[:, 1:] >= z[:, :-1]).all(), "depths must be ascending along every ray"
    assert (z >= rays[:, 11:12]).all() and (z <= rays[:, 12:13]).all(), "depths must stay inside [near, far]"
    assert (inds >= 0).all() and (inds <= S // 2 - 1).all()
    z2 = ops.sample_occ(rays, occ, u1, u2, S, 1.0)
    assert torch.equal(z, z2), "sampler must be deterministic"
    # oracle spot check: a random subset of rays, bit-exact (occupancy bits, searchsorted indices, depths)
    sel = torch.randperm(n, generator=torch.Generator().manual_seed(5))[:48]
    sd = sel.to(dev)
    ref = O.sample_occ(rays[sd].cpu(), occ.cpu()[None, None], u1[sd].cpu(), u2[sd].cpu(), S, 1.0)
    assert torch.equal(bits[sd].cpu().bool(), ref["occ_bits"])
    assert torch.equal(inds[sd].cpu(), ref["inds"])
    assert torch.equal(z[sd].cpu(), ref["z"])


def test_sampler_unsorted_coarse_depths_take_the_full_sort(dev, ops, O):
    """near > far makes the coarse depths descending: the merge path must notice and fall back to the full sort."""
    n, S = 64, 128
    torch.manual_seed(2)
    rays = torch.zeros(n, 13)
    rays[:, 0:3] = (torch.rand(n, 3) - 0.5) * 0.2
    d = torch.randn(n, 3)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    rays[:, 6:9] = -rays[:, 3:6]
    rays[:, 11], rays[:, 12] = 0.6, 0.05
    occ = torch.randn(40, 40, 40) * 0.5
    u1, u2 = torch.rand(n, S // 2), torch.rand(n, S // 2)
    z = ops.sample_occ(rays.to(dev), occ.to(dev), u1.to(dev), u2.to(dev), S, 1.0)
    zo = O.sample_occ(rays, occ[None, None], u1, u2, S, 1.0)["z"]
    assert torch.equal(z.cpu(), zo)


# ------------------------------------------------------------------------------------------ hash grid
@pytest.mark.parametrize("log2T,P", [(19, N_FULL // 2 * S_FULL), (22, 1 << 21)])
def test_encode_adjoint_identity_full_size(log2T, P, scene, dev, ops):
    plan = ops.Plan(log2_hashmap_size=log2T)
    g = torch.Generator(device=dev).manual_seed(log2T)
    x01 = torch.rand(P, 3, device=dev, generator=g)
    table = (torch.rand(plan.n_table_params, device=dev, generator=g) - 0.5)
    enc = ops.encode_fwd(plan, x01, table)
    d_enc = torch.randn(P, 32, device=dev, generator=g)
    lhs = (enc.double() * d_enc.double()).sum().item()
    grad = torch.zeros_like(table)
    ops.encode_bwd(plan, x01, table, d_enc.clone(), grad, False)
    rhs = (table.double() * grad.double()).sum().item()
    scale = (enc.double().abs() * d_enc.double().abs()).sum().item()
    assert abs(lhs - rhs) <= 1e-6 * scale, (lhs, rhs, scale)
    # every sample adds trilinear weights that sum to one per level and feature: column sums of the gradient
    # (per level) equal column sums of d_enc
    off = 0
    for l in range(16):
        n_l = plan.level(l)["size"]
        gl = grad[2 * off:2 * (off + n_l)].view(-1, 2).double().sum(0)
        dl = d_enc[:, 2 * l:2 * l + 2].double().sum(0)
        assert torch.allclose(gl, dl, rtol=0, atol=2e-5 * d_enc[:, 2 * l:2 * l + 2].double().abs().sum().item() / 1e3), (l, gl, dl)
        off += n_l


def test_sigma_head_full_chunk_matches_plain_fp32(scene, dev, ops):
    """One bench chunk (16384 x 512 = 8.4 M points): fused forward vs. stand-alone gather + torch fp32 MLP; backward
    vs. torch autograd through that MLP and the stand-alone scatter."""
    rays = scene["rays"][:16384].contiguous()
    g = torch.Generator(device=dev).manual_seed(3)
    z = ops.sample_occ(rays, scene["occ"], torch.rand(16384, 256, device=dev, generator=g),
                       torch.rand(16384, 256, device=dev, generator=g), 512, 1.0)
    plan = ops.Plan(log2_hashmap_size=19)
    params = torch.cat([(torch.rand(3072, device=dev, generator=g) - 0.5) * 0.6,
                        (torch.rand(plan.n_table_params, device=dev, generator=g) - 0.5) * 0.5])
    sigma, saved = ops.sigma_fwd(plan, params, rays=rays, z=z, save_enc=True)
    x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None] + 1) / 2).reshape(-1, 3).contiguous()
    enc = ops.encode_fwd(plan, x01, params[3072:].contiguous())
    assert torch.equal(enc, saved[0]), "the fused forward must save the stand-alone kernel's encodings bit for bit"
    W1, w2 = params[:2048].view(64, 32), params[2048:2048 + 64]
    h = enc @ W1.t()
    ref = torch.relu(h) @ w2
    assert _rel(sigma.reshape(-1), ref) < 1e-5
    # the backward is defined by the ReLU mask the forward saved; a unit with |h| ~ 1e-7 may be on one side in the
    # tensor-core evaluation and on the other in torch's (a handful of the 537 M units), and with 8.4 M terms per
    # weight such a flip is visible at the 1e-3 level, so the reference gradient uses the saved mask
    mask = ((saved[1][:, None] >> torch.arange(64, device=dev)) & 1).float()
    assert (mask.bool() != (h > 0)).float().mean().item() < 1e-5
    d_sigma = torch.randn(16384, 512, device=dev, generator=g) * (torch.rand(16384, 512, device=dev, generator=g) < 0.6)
    ds = d_sigma.reshape(-1, 1)
    A = mask * (ds * w2[None, :])                       # dL/dh
    dW1_ref, dw2_ref, dE_ref = A.t() @ enc, (mask * h * ds).sum(0), A @ W1
    del A, mask, h
    grad = torch.zeros_like(params)
    ops.sigma_bwd(plan, params, saved, d_sigma, rays=rays, z=z, grad_params=grad)
    assert _rel(grad[:2048].view(64, 32), dW1_ref) < 1e-4
    assert _rel(grad[2048:2048 + 64], dw2_ref) < 1e-4
    gt = torch.zeros(plan.n_table_params, device=dev)
    ops.encode_bwd(plan, x01, params[3072:].contiguous(), dE_ref.contiguous(), gt, False)
    assert _rel(grad[3072:], gt) < 1e-4
    # linearity in d_sigma (scaling by a power of two is exact, only the atomics' order differs)
    grad2 = torch.zeros_like(params)
    ops.sigma_bwd(plan, params, saved, d_sigma * 4.0, rays=rays, z=z, grad_params=grad2)
    assert _rel(grad2, grad * 4.0) < 1e-5


# ------------------------------------------------------------------------------------------ compositing + loss
def test_composite_loss_full_size_vs_torch_autograd(scene, dev, ops):
    n, S = N_FULL, S_FULL
    rays, depths, s = scene["rays"], scene["depths"], scene["scale"]
    g = torch.Generator(device=dev).manual_seed(9)
    z = ops.sample_occ(rays, scene["occ"], torch.rand(n, S // 2, device=dev, generator=g),
                       torch.rand(n, S // 2, device=dev, generator=g), S, 1.0)
    sigma = (torch.randn(n, S, device=dev, generator=g) * 2.0 + 0.3).requires_grad_(True)
    noise = torch.randn(n, S, device=dev, generator=g)
    n_op = ops.count_opaque(rays, depths)
    cfg = ops.make_loss_config("L1_LOS", s, 0.0, float(n * S), depth_eps=3.0)
    parts = torch.zeros(4, device=dev)
    out = ops.composite_loss(sigma.detach(), z, rays, noise, depths, cfg, parts, want_grad=True, n_opaque_dev=n_op)
    w, depth, opac = out["weights"], out["depth"], out["opacity"]
    assert (w >= 0).all() and (opac <= 1.0 + 1e-4).all()
    assert _rel(w.sum(1), opac) < 1e-5
    assert (depth >= rays[:, 11] - 1e-5).all() and (depth <= rays[:, 12] + 1e-5).all()
    # plain torch restatement (reference rendering_tcnn.py:219-295, optimizer.py:780-848, losses.py:29-51)
    dn = rays[:, 3:6].norm(dim=1, keepdim=True)
    dist = torch.cat([z[:, 1:] - z[:, :-1], torch.full((n, 1), 1e10, device=dev)], 1) * dn
    alpha = 1 - torch.exp(-dist * torch.relu(sigma + noise))
    T = torch.cumprod(torch.cat([torch.ones(n, 1, device=dev), 1 - alpha + 1e-10], 1), 1)[:, :-1]
    w_t = alpha * T
    op_t = w_t.sum(1)
    depth_t = (w_t * z).sum(1) + (1 - op_t) * rays[:, 12]
    assert _rel(w, w_t) < 1e-4 and _rel(depth, depth_t) < 1e-4
    opaque = (depths > 0) & ~(depths > rays[:, 12])
    zm, gm, eps = z * s, (depths * s)[:, None], 3.0
    sig_t = eps / 9.0
    pdf = torch.exp(-0.5 * ((zm - gm) / sig_t) ** 2) / (sig_t * (2 * torch.pi) ** 0.5)
    norm_c = torch.erf(torch.tensor(9.0 / 2 ** 0.5)).item()          # Phi(9) - Phi(-9)
    wg = pdf / norm_c * ((zm > gm - eps) & (zm < gm + eps))
    wg = wg / (wg.sum(1, keepdim=True) + 1e-6)
    wg = wg * opaque[:, None]
    l_depth = ((depth_t * s - depths * s) ** 2)[opaque].sum()
    l_los = (w_t - wg).abs().sum()
    l_op = (op_t - 1).abs()[opaque].sum()
    assert abs(parts[0].item() - l_depth.item()) <= 1e-4 * abs(l_depth.item()) + 1e-6
    assert abs(parts[1].item() - l_los.item()) <= 1e-4 * abs(l_los.item())
    assert abs(parts[2].item() - l_op.item()) <= 1e-4 * abs(l_op.item()) + 1e-6
    loss = 0.005 * l_depth / n_op[0] + 1000.0 * l_los / (n * S) + l_op / n_op[0]
    loss.backward()
    assert _rel(out["d_sigma"], sigma.grad) < 1e-3


# ------------------------------------------------------------------------------------------ Adam
@pytest.mark.parametrize("log2T", [19, 22])
def test_adam_full_size_matches_torch(log2T, dev, ops):
    plan = ops.Plan(log2_hashmap_size=log2T)
    n = 3072 + plan.n_table_params
    g = torch.Generator(device=dev).manual_seed(1)
    p = torch.randn(n, device=dev, generator=g) * 1e-2
    ref = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=0.01)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step in range(1, 4):
        grad = torch.randn(n, device=dev, generator=g) * 1e-3 * (torch.rand(n, device=dev, generator=g) < 0.3)
        ref.grad = grad.clone()
        opt.step()
        ops.adam_step(p, grad, m, v, step, 0.01)
    assert (p - ref.detach()).abs().max().item() <= 2e-7
    del opt


# ------------------------------------------------------------------------------------------ empty / ragged inputs
def test_empty_inputs_are_no_ops(dev, ops):
    plan = ops.Plan(log2_hashmap_size=14)
    params = torch.zeros(3072 + plan.n_table_params, device=dev)
    rays = torch.zeros(0, 13, device=dev)
    occ = torch.zeros(20, 20, 20, device=dev)
    z = ops.sample_occ(rays, occ, torch.zeros(0, 32, device=dev), torch.zeros(0, 32, device=dev), 64, 1.0)
    assert z.shape == (0, 64)
    assert ops.sample_uniform(rays, torch.zeros(0, 64, device=dev), 64, 1.0).shape == (0, 64)
    sigma, saved = ops.sigma_fwd(plan, params, rays=rays, z=z, save_enc=True)
    assert sigma.shape == (0, 64)
    grad = torch.zeros_like(params)
    ops.sigma_bwd(plan, params, saved, torch.zeros(0, 64, device=dev), rays=rays, z=z, grad_params=grad)
    assert not grad.any()
    assert ops.encode_fwd(plan, torch.zeros(0, 3, device=dev), params[3072:].contiguous()).shape == (0, 32)
    parts = torch.zeros(4, device=dev)
    cfg = ops.make_loss_config("L1_JS", 10.0, 0.0, 1.0)
    out = ops.composite_loss(sigma, z, rays, None, torch.zeros(0, device=dev), cfg, parts, n_opaque_dev=torch.ones(2, device=dev))
    assert out["weights"].shape == (0, 64) and not parts.any()
    assert not ops.count_opaque(rays, torch.zeros(0, device=dev)).any()
    assert not ops.occ_grad(rays, z, torch.zeros(0, device=dev), 10.0, 20).any()
    ops.adam_step(torch.zeros(0, device=dev), torch.zeros(0, device=dev), torch.zeros(0, device=dev), torch.zeros(0, device=dev), 1, 0.01)
    torch.cuda.synchronize()


@pytest.mark.parametrize("n,S", [(1, 64), (7, 66), (129, 130), (1000, 98)])
def test_ragged_sizes_match_oracle(n, S, dev, ops, O):
    """Ray counts / sample counts that fill no warp, tile or power of two: sampler bit-exact, field within 1e-3."""
    torch.manual_seed(n * 1000 + S)
    rays = torch.zeros(n, 13)
    rays[:, 0:3] = (torch.rand(n, 3) - 0.5) * 0.4
    d = torch.randn(n, 3)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    rays[:, 6:9] = -rays[:, 3:6]
    rays[:, 11] = 0.02
    rays[:, 12] = 0.3 + 0.2 * torch.rand(n)
    occ = torch.randn(30, 30, 30) * 0.8
    u1, u2 = torch.rand(n, S // 2), torch.rand(n, S // 2)
    z = ops.sample_occ(rays.to(dev), occ.to(dev), u1.to(dev), u2.to(dev), S, 1.0)
    zo = O.sample_occ(rays, occ[None, None], u1, u2, S, 1.0)["z"]
    assert torch.equal(z.cpu(), zo)
    params = O.init_sigma_params(14, seed=4)
    params[3072:] *= 3e3
    plan = ops.Plan(log2_hashmap_size=14)
    sigma, _ = ops.sigma_fwd(plan, params.to(dev), rays=rays.to(dev), z=z, save_enc=True)
    xyz = rays[:, None, 0:3] + rays[:, None, 3:6] * zo[..., None]
    ref = O.sigma_field(xyz.reshape(-1, 3), params, 14).reshape(n, S)
    assert _rel(sigma.cpu(), ref) < 1e-3
