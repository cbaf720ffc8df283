"""The UNMODIFIED reference on top of the kernels.

`sys.modules["tinycudann"] = loner_b200.tcnn`, then the reference's own src/models/model_tcnn.Model,
src/models/nerf_tcnn.DecoupledNeRF (call sites nerf_tcnn.py:15,35-55,59-95: the three tcnn constructors, `.dtype` for
torch.finfo, `.n_output_dims`, `h[..., [0]]`), src/models/rendering_tcnn.render_rays and
src/mapping/optimizer.Optimizer.compute_loss run on CUDA from the verbatim copy oracle/_ref (oracle/make_ref_copy.py),
and the results are compared with the golden vectors the same code produced on the CPU behind the fp32 tcnn shim.
"""
import os
import sys

import pytest
import torch

from conftest import load_golden, golden_params, golden_rgb_params
from test_reference_chain_gpu import ReplaySampler, replay_randn, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REL = 1e-3


@pytest.fixture(scope="module")
def ref():
    from oracle.reference_import import import_reference, reference_available
    if not reference_available():
        pytest.skip("no reference tree: run `python -m oracle.make_ref_copy` in the build container (oracle/_ref travels)")
    import loner_b200.tcnn as tcnn
    ns = import_reference(tcnn)
    yield ns
    import_reference()          # rebind the CPU shim for whatever runs next in this process


def reference_optimizer(ns, g, sampler):
    from oracle import make_golden as G
    cfg = G.model_cfg(ns, g["n_samples"], g["log2_T"], g["voxel"], tuple(g["ray_range"]), g["loss_selection"], g["min_js"])
    opt = G.bare_optimizer(ns, cfg, g["scale"].to(DEV), sampler, device=DEV)
    sig = opt._model.nerf_model._model_sigma
    assert type(sig).__module__ == "loner_b200.tcnn" and sig.params.is_cuda
    with torch.no_grad():
        sig.params.copy_(golden_params(g).to(DEV))
        if sampler == "OGM":
            opt._occupancy_grid_model.occupancy_grid.copy_(g["occ"].to(DEV))
            opt._ray_sampler.update_occ_grid(opt._occupancy_grid.detach())
    return opt


@pytest.mark.parametrize("name", ["c1_ogm_l1js", "s128_ogm_l2js", "c1_uniform_l2los"])
def test_unmodified_reference_compute_loss_on_kernels(ref, name):
    g = load_golden(name)
    opt = reference_optimizer(ref, g, g["sampler"])
    # (1) the reference's own depths (replay sampler): every recorded output and gradient at 1e-3
    opt._ray_sampler = ReplaySampler(g["z"].to(DEV))
    rays = g["rays"].to(DEV).requires_grad_(True)
    with replay_randn([g["noise"]]):
        loss = opt.compute_loss(None, (rays, g["depths"].to(DEV)), 0)
    loss.backward()
    res = opt._results_lidar
    assert set(res) >= {"rgb_fine", "depth_fine", "weights_fine", "opacity_fine", "variance", "samples_fine", "points_fine", "raw_fine"}
    assert abs(loss.item() - g["loss"].item()) < 1e-5 * abs(g["loss"].item())
    assert rel_err(res["raw_fine"][..., 0], g["sigma"]) < 1e-4
    for k, gk in (("weights_fine", "weights"), ("depth_fine", "depth"), ("opacity_fine", "opacity"), ("variance", "variance")):
        assert rel_err(res[k], g[gk]) < 1e-4, k
    gp = opt._model.nerf_model._model_sigma.params.grad
    assert rel_err(gp[:3072], g["mlp_grad"]) < REL
    assert rel_err(gp[3072:][g["table_grad_idx"].to(DEV)], g["table_grad_val"]) < REL
    assert abs(gp[3072:].double().norm().item() - g["table_grad_norm"]) < REL * g["table_grad_norm"]
    assert rel_err(rays.grad, g["rays_grad"]) < REL
    # (2) the reference's own torch sampler on CUDA (its sums / exp differ from the CPU run at the ulp level)
    if g["sampler"] == "OGM":
        opt._ray_sampler = ref.ray_sampling.OccGridRaySampler()
        opt._ray_sampler.update_occ_grid(opt._occupancy_grid.detach())
        orig = torch.rand
        q = [g["u1"], g["u2"]]
        torch.rand = lambda *a, **k: q.pop(0).to(k.get("device", "cpu"))
        try:
            with replay_randn([g["noise"]]), torch.no_grad():
                loss2 = opt.compute_loss(None, (g["rays"].to(DEV), g["depths"].to(DEV)), 0)
        finally:
            torch.rand = orig
        assert abs(loss2.item() - g["loss"].item()) < 1e-4 * abs(g["loss"].item())
        assert rel_err(opt._results_lidar["depth_fine"], g["depth"]) < REL
        # and its occupancy update
        opt._step_occupancy_grid()
        assert rel_err(opt._occupancy_grid.detach().cpu() - g["occ"], g["occ_after"] - g["occ"]) < 5e-3


def test_unmodified_reference_camera_forward_on_kernels(ref):
    """Model.forward(camera=True) of the unmodified reference: both hash grids, SH-4 and the 4-layer colour MLP through
    loner_b200.tcnn (Encoding x2, Network, NetworkWithInputEncoding) vs tests/golden/rgb_camera.pt."""
    from oracle import make_golden as G
    g = load_golden("rgb_camera")
    p_sigma, table, mlp = golden_rgb_params(g)
    cfg = G.model_cfg(ref, g["n_samples"], g["log2_T"], 32, tuple(g["ray_range"]))
    cfg["model"]["nerf_config"]["pos_encoding_intensity"]["log2_hashmap_size"] = g["log2_T_int"]
    model = ref.model_tcnn.Model(cfg.model).to(DEV)
    nerf = model.nerf_model
    assert torch.finfo(nerf._model_intensity.dtype).max == nerf._max_float
    with torch.no_grad():
        nerf._model_sigma.params.copy_(p_sigma.to(DEV))
        nerf._pos_encoding.params.copy_(table.to(DEV))
        nerf._model_intensity.params.copy_(mlp.to(DEV))
    rays = g["rays"].to(DEV)
    sampler = ReplaySampler(g["z"].to(DEV))
    with replay_randn([g["noise"]]):
        out = model(rays, sampler, g["scale"].to(DEV), camera=True, return_variance=True)
    assert rel_err(out["raw_fine"][..., :3], g["raw"][..., :3]) < 1e-4
    assert rel_err(out["raw_fine"][..., 3], g["raw"][..., 3]) < 1e-4
    assert rel_err(out["rgb_fine"], g["rgb"]) < 1e-4
    loss = torch.nn.functional.l1_loss(out["rgb_fine"].reshape(-1, 1), g["intensities"].to(DEV).reshape(-1, 1))
    assert abs(loss.item() - g["loss"].item()) < 1e-5 * abs(g["loss"].item())
    loss.backward()
    assert rel_err(nerf._model_intensity.params.grad, g["mlp_int_grad"]) < REL
    tg = nerf._pos_encoding.params.grad
    assert rel_err(tg[g["table_int_grad_idx"].to(DEV)], g["table_int_grad_val"]) < REL
