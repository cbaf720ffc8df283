"""GPU parity against the REFERENCE'S OWN RUN at north_star's 1e-3 ("next round" item 1).

(a) Gradient chain on the reference's recorded sample depths: the golden z replaces the sampler (a replay sampler
    object in the Optimizer's `_ray_sampler` slot), the recorded noise replaces torch.randn, and the product's
    Optimizer.compute_loss -> backward runs sigma_fwd -> lb_composite_loss -> sigma_bwd (-> lb_build_lidar_rays_bwd).
    With bit-identical sample positions every gradient the reference recorded is reproduced to <= 1e-3:
    MLP, hash tables, rays, pose.
(b) The benchmark-shaped fixture tests/golden/c2_shape_ogm_l1js.pt (2048 rays x 512 samples, T = 2^19, warmed 100^3
    occupancy grid, two chunks, tables x1 and x2000; oracle/make_golden_c2.py): full pipeline with the product's own
    sampler -- occupancy bits, searchsorted indices, depths, per-ray outputs, loss, gradients, occupancy update.
(c) MASK ray selection against the reference's rule (src/mapping/optimizer.py:367-379) on a replayed randperm.
"""
import contextlib

import numpy as np
import pytest
import torch

from conftest import load_golden, golden_params, c2_shape_random, c2_shape_params

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REL = 1e-3          # north_star tolerance


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


class ReplaySampler:
    """Hands out recorded sample depths chunk by chunk (duck-types OccGridRaySampler / UniformRaySampler)."""

    def __init__(self, z):
        self.z, self.pos = z, 0

    def update_occ_grid(self, occ):
        pass

    def get_samples(self, rays, N_samples, perturb):
        n = rays.shape[0]
        out = self.z[self.pos:self.pos + n].contiguous()
        self.pos += n
        assert out.shape == (n, N_samples)
        return out


@contextlib.contextmanager
def replay_randn(noise_chunks):
    orig = torch.randn
    q = list(noise_chunks)

    def fake(*shape, **kw):
        t = q.pop(0)
        return t.to(kw.get("device", "cpu"))
    torch.randn = fake
    try:
        yield
    finally:
        torch.randn = orig


def make_optimizer(g, params, chunk=16384, loss_selection=None):
    from loner_b200 import mapping as M
    s = M.default_settings(n_samples_train=g["n_samples"], log2_hashmap_size=g["log2_T"], ray_range=tuple(g["ray_range"]),
                           voxel_size=g["voxel"], loss_selection=loss_selection or g["loss_selection"],
                           strategy=g["sampler"], chunk=chunk)
    s["model_config"]["loss"]["JS_loss"]["min_js_score"] = g["min_js"]
    s["model_config"]["loss"]["decay_depth_eps"] = False
    opt = M.Optimizer(s, None, M.WorldCube(g["scale"], g["shift"]), DEV)
    with torch.no_grad():
        opt._model.nerf_model._model_sigma.params.copy_(params.to(DEV))
        if g["sampler"] == "OGM":
            opt._occupancy_grid.copy_(g["occ"].to(DEV))
    return opt, M


def check_gradients(opt, M, g, st, rays, l1_sign_flips=False):
    """MLP / table / ray gradients of the product vs the reference's recorded ones, then the pose gradient through
    lb_build_lidar_rays_bwd fed with the product's own ray gradient.

    l1_sign_flips: the L1 line-of-sight term's gradient is sign(w - w_gt).  w_gt depends on the per-ray margin eps, which
    is a function of (g - mean)^2 with g ~ mean ~ 20 m: a cancellation that leaves ~1e-5 relative fp32 summation-order
    noise in eps (any two fp32 evaluations differ there, e.g. the reference on the CPU and on a GPU).  At 512 samples
    per ray a few dozen of the 2048 rays have one sample with |w - w_gt| below that noise; its sign, and with it the
    ray's gradient, differs by up to ~1e-2 of the maximum.  Sums over all rays (MLP, tables: <= 1e-3) do not notice; per-ray
    quantities are then asserted on all but those rays, and the L2_JS state (no sign) asserts every ray."""
    gp = opt._model.nerf_model._model_sigma.params.grad
    tg = gp[3072:]
    assert rel_err(gp[:3072], st["mlp_grad"]) < REL
    assert rel_err(tg[st["table_grad_idx"].to(DEV)], st["table_grad_val"]) < REL
    assert abs(tg.double().norm().item() - st["table_grad_norm"]) < REL * st["table_grad_norm"]
    assert abs(tg.double().sum().item() - st["table_grad_sum"]) < REL * max(abs(st["table_grad_sum"]), st["table_grad_norm"])
    if l1_sign_flips:
        ref = st["rays_grad"].double()
        per_ray = (rays.grad.double().cpu() - ref).abs().amax(1) / ref.abs().max()
        assert (per_ray > REL).float().mean().item() < 0.03, (per_ray > REL).sum().item()      # measured: ~25 of 2048 rays
        assert per_ray.median().item() < 1e-5 and per_ray.max().item() < 3e-2
        return
    assert rel_err(rays.grad, st["rays_grad"]) < REL
    if "pose_grad" in st:
        pose = g["kf_pose"].to(DEV).requires_grad_(True)
        wc = M.WorldCube(g["scale"].to(DEV), g["shift"].to(DEV))
        rays_l, _ = M.build_lidar_rays(g["kf_directions"].to(DEV), g["kf_distances"].to(DEV), pose, wc,
                                       torch.tensor(g["ray_range"], device=DEV))
        n_l = rays_l.shape[0]
        assert rel_err(rays_l, g["rays"][:n_l]) < 1e-6
        rays_l.backward(rays.grad[:n_l])
        assert rel_err(pose.grad, st["pose_grad"]) < REL


@pytest.mark.parametrize("name", ["c1_ogm_l1js", "s128_ogm_l2js", "c1_uniform_l2los"])
def test_gradient_chain_on_reference_depths(name):
    g = load_golden(name)
    opt, M = make_optimizer(g, golden_params(g))
    opt._ray_sampler = ReplaySampler(g["z"].to(DEV))
    rays = g["rays"].to(DEV).requires_grad_(True)
    with replay_randn([g["noise"]]):
        loss = opt.compute_loss(None, (rays, g["depths"].to(DEV)), 0)
    loss.backward()
    res = opt._results_lidar
    assert abs(loss.item() - g["loss"].item()) < 1e-5 * abs(g["loss"].item())
    for k, gk in (("weights_fine", "weights"), ("depth_fine", "depth"), ("opacity_fine", "opacity"), ("variance", "variance")):
        assert rel_err(res[k], g[gk]) < 1e-4, k
    assert abs(opt._depth_eps - g["depth_eps"]) < 1e-4 * g["depth_eps"]
    check_gradients(opt, M, g, g, rays)


@pytest.mark.parametrize("state", ["l2js_2000", 2000, 1])
def test_c2_shape_gradient_chain_on_reference_depths(state):
    """(a) at the benchmark shape: 2048 rays x 512 samples, T = 2^19, two chunks of 1024 rays."""
    g = load_golden("c2_shape_ogm_l1js")
    st = g["states"][state]
    opt, M = make_optimizer(g, c2_shape_params(g, state), chunk=g["chunk"], loss_selection=st["loss_selection"])
    opt._ray_sampler = ReplaySampler(g["z"].to(DEV))
    rays = g["rays"].to(DEV).requires_grad_(True)
    with replay_randn([r[2] for r in c2_shape_random(g)]):
        loss = opt.compute_loss(None, (rays, g["depths"].to(DEV)), 0)
    loss.backward()
    res = opt._results_lidar
    k = g["row_stride"]
    assert abs(loss.item() - st["loss"].item()) < 1e-5 * abs(st["loss"].item())
    assert rel_err(res["depth_fine"], st["depth"]) < 1e-4
    assert rel_err(res["opacity_fine"], st["opacity"]) < 1e-4
    assert rel_err(res["variance"], st["variance"]) < 1e-4
    assert rel_err(res["weights_fine"][::k], st["weights_rows"]) < 1e-4
    assert abs(res["weights_fine"].double().sum().item() - st["weights_sum"]) < 1e-5 * st["weights_sum"]
    assert abs(opt._depth_eps - st["depth_eps"]) < 1e-4 * st["depth_eps"]
    check_gradients(opt, M, g, st, rays, l1_sign_flips=st["loss_selection"] == "L1_JS")
    # occupancy update on the reference's depths (optimizer.py:897-908) at 100^3
    occ0 = opt._occupancy_grid.detach().clone()
    opt._step_occupancy_grid()
    delta = (opt._occupancy_grid.detach() - occ0).flatten()
    assert rel_err(delta[st["occ_delta_idx"].to(DEV)], st["occ_delta_val"]) < REL
    assert abs(delta.double().norm().item() - st["occ_delta_norm"]) < REL * st["occ_delta_norm"]


@pytest.mark.parametrize("state", ["l2js_2000", 2000, 1])
def test_c2_shape_full_pipeline_against_reference(state):
    """(b) everything, with the product's own sampler, on the random streams the reference consumed."""
    from oracle import portable_rng as R
    g = load_golden("c2_shape_ogm_l1js")
    st = g["states"][state]
    opt, M = make_optimizer(g, c2_shape_params(g, state), chunk=g["chunk"], loss_selection=st["loss_selection"])
    rays = g["rays"].to(DEV).requires_grad_(True)
    # sampler intermediates chunk by chunk: occupancy bits and searchsorted indices
    bits, inds = [], []
    for c, (u1, u2, _) in enumerate(c2_shape_random(g)):
        r_c = g["rays"][c * g["chunk"]:(c + 1) * g["chunk"]].to(DEV)
        _, b, i = M.ops.sample_occ(r_c, opt._occupancy_grid.detach().contiguous(), u1.to(DEV), u2.to(DEV),
                                   g["n_samples"], 1.0, debug=True)
        bits.append(b.cpu().bool())
        inds.append(i.cpu())
    bits, inds = torch.cat(bits), torch.cat(inds)
    gold_bits = torch.from_numpy(np.unpackbits(g["occ_bits_packed"].numpy(), axis=1)).bool()
    assert torch.equal(bits, gold_bits), "occupancy bits differ from the reference run"
    # searchsorted indices: the reference's pdf uses torch-CPU's vectorised fp32 sum (association depends on the host
    # ISA), the kernel the exactly rounded sum; the cdf knots then differ by <= 1 ulp and a draw lying exactly between
    # the two knots lands in the neighbouring bin: measured 5 of 524288 at this shape (0 at S <= 128), never by more than 1
    mism = inds != g["inds"].long()
    assert mism.sum().item() <= 16 and (inds - g["inds"].long()).abs().max().item() <= 1
    with R.Patch(g["rng_base_stream"]):
        loss = opt.compute_loss(None, (rays, g["depths"].to(DEV)), 0)
    loss.backward()
    res = opt._results_lidar
    dz = (res["samples_fine"].cpu() - g["z"]).abs()
    assert (dz > 1e-6).float().mean().item() < 1e-3 and dz.max().item() < 5e-3
    k = g["row_stride"]
    assert abs(loss.item() - st["loss"].item()) < 1e-5 * abs(st["loss"].item())
    assert rel_err(res["depth_fine"], st["depth"]) < REL
    assert rel_err(res["opacity_fine"], st["opacity"]) < REL
    assert rel_err(res["variance"], st["variance"]) < REL
    assert abs(opt._depth_eps - st["depth_eps"]) < 1e-4 * st["depth_eps"]
    gp = opt._model.nerf_model._model_sigma.params.grad
    tg = gp[3072:]
    assert rel_err(gp[:3072], st["mlp_grad"]) < REL
    assert rel_err(tg[st["table_grad_idx"].to(DEV)], st["table_grad_val"]) < REL
    assert abs(tg.double().norm().item() - st["table_grad_norm"]) < REL * st["table_grad_norm"]
    # ray gradients run through the x-derivative of the finest hash levels and are discontinuous in z at the ulp
    # level: compared on the rays whose depths are bit-identical to the reference's
    same = (res["samples_fine"].cpu() == g["z"]).all(1)
    assert same.float().mean().item() > 0.02
    if st["loss_selection"] == "L2_JS":
        assert rel_err(rays.grad.cpu()[same], st["rays_grad"][same]) < REL


def test_mask_ray_selection_matches_reference_rule():
    """O3 MASK (reference optimizer.py:367-379): 75 % of the rays from the trunk-height band 0.5 m < z < 8 m, drawn
    without replacement by randperm, the rest from the complement.  torch.randperm is replayed so the index lists
    of the product and of the reference rule (restated below on the CPU) must be identical."""
    from loner_b200 import mapping as M, synth
    kfs, scale, shift = synth.make_keyframes(1, ray_range=(1.0, 30.0), seed=5)
    s = M.default_settings(n_samples_train=64, log2_hashmap_size=14, ray_range=(1.0, 30.0), lidar=1000, sky=0, voxel_size=32)
    s["rays_selection"]["strategy"] = "MASK"
    opt = M.Optimizer(s, None, M.WorldCube(scale, shift), DEV)
    kf = M.KeyFrame(M.LidarScanData(kfs[0].directions, kfs[0].distances, None), kfs[0].pose, device=DEV)
    n = kfs[0].distances.shape[0]
    perms, orig = [], torch.randperm

    def fake_randperm(m, **kw):
        p = orig(m, generator=torch.Generator().manual_seed(100 + len(perms)))
        perms.append(p)
        return p.to(kw.get("device", "cpu"))
    torch.randperm = fake_randperm
    try:
        idx, sky_idx = opt._select_rays(kf)
    finally:
        torch.randperm = orig
    # the reference rule on the CPU with the same two permutations
    xyz = kfs[0].directions * kfs[0].distances
    trunk_mask = (0.5 < xyz[2, :]) * (xyz[2, :] < 8)
    allidx = torch.arange(n)
    trunk, non_trunk = allidx[trunk_mask], allidx[~trunk_mask]
    assert len(perms) == 2 and len(perms[0]) == len(trunk) and len(perms[1]) == len(non_trunk)
    want = torch.cat((trunk[perms[0][0:750]], non_trunk[perms[1][0:250]]))
    assert sky_idx is None
    assert torch.equal(idx.cpu(), want)
    assert trunk_mask[idx.cpu()[:750]].all() and not trunk_mask[idx.cpu()[750:]].any()
    assert idx.cpu().unique().numel() == idx.numel() == 1000
    # and through the whole iteration: a MASK-selected batch trains
    s2 = M.default_settings(n_samples_train=64, log2_hashmap_size=14, ray_range=(1.0, 30.0), lidar=256, sky=0,
                            voxel_size=32, num_iterations=3)
    s2["rays_selection"]["strategy"] = "MASK"
    opt2 = M.Optimizer(s2, None, M.WorldCube(scale, shift), DEV)
    opt2.iterate_optimizer([kf])
    assert opt2._global_step == 3
