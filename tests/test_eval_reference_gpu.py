"""Eval drivers against fixtures produced by the reference's OWN analysis code
(oracle/make_golden_eval.py: analysis/mesher.py Mesher.get_mesh up to the marching-cubes volume, analysis/renderer.py
render_dataset_frame), replayed on the same portable random streams with the same 512-ray chunking."""
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_model(g, n_train=64, n_test=128):
    from loner_b200 import mapping as M
    from oracle import loner_oracle as O
    s = M.default_settings(n_samples_train=n_train, n_samples_test=n_test, log2_hashmap_size=g["log2_T"],
                           ray_range=tuple(g["ray_range"]), strategy="UNIFORM", chunk=g["chunk"], voxel_size=32)
    s["model_config"]["model"]["nerf_config"]["pos_encoding_intensity"]["log2_hashmap_size"] = g.get("log2_T_int", 14)
    opt = M.Optimizer(s, None, M.WorldCube(g["scale"], g["shift"]), DEV, lidar_only=False)
    p = O.init_sigma_params(g["log2_T"], g["param_seed"])
    p[O.MLP_PARAMS:] *= g["table_scale"]
    table, mlp = O.init_intensity_params(g.get("log2_T_int", 14), 3, g["param_seed"])
    nerf = opt._model.nerf_model
    with torch.no_grad():
        nerf._model_sigma.params.copy_(p.to(DEV))
        nerf._pos_encoding.params.copy_((table * g["table_scale"]).to(DEV))
        nerf._model_intensity.params.copy_(mlp.to(DEV))
    return opt, M


def test_render_frame_matches_reference_render_dataset_frame():
    from loner_b200 import eval as E
    from oracle import portable_rng as R
    g = load_golden("eval_render")
    opt, M = make_model(g, n_test=g["n_samples_test"])
    rays = g["rays"].to(DEV)
    with R.Patch(0):
        out = E.render_frame(opt._model, M.UniformRaySampler(), rays, opt._scale_factor, camera=True)
    H, W = g["H"], g["W"]
    img = lambda t, c: t.reshape(1, H, W, c).permute(0, 3, 1, 2).cpu()
    rgb, depth, peak = img(out["rgb"], 3), img(out["depth"], 1), img(out["peak_depth_consistency"], 1)
    assert (rgb - g["rgb"]).abs().max().item() < 1e-3
    assert ((depth - g["depth"]).abs() / g["depth"].abs().clamp_min(1e-3)).max().item() < 1e-3
    # |z[argmax w] - depth|: the argmax may land on a neighbouring sample where two weights tie to rounding
    close = (peak - g["peak_depth_consistency"]).abs() < 1e-3 * g["depth"].abs().clamp_min(1.0)
    assert close.float().mean().item() > 0.995


def test_mesher_volume_matches_reference_get_mesh():
    from loner_b200 import eval as E
    from oracle import portable_rng as R
    g = load_golden("eval_mesher")
    opt, M = make_model(g, n_train=g["n_samples_train"])
    wc = M.WorldCube(g["scale"].to(DEV), g["shift"].to(DEV))
    mesher = E.Mesher(opt._model, wc, g["ray_range"], resolution=g["resolution"], marching_cubes_bound=g["bound"],
                      lidar_vertical_fov=tuple(g["fov"]))
    for a, b in zip(mesher.get_grid_uniform(), g["xyz"]):
        assert torch.equal(torch.from_numpy(a), b)
    with R.Patch(0):
        vol, xyz = mesher.splat_volume(M.UniformRaySampler(), g["poses"])
    vol, ref = vol.cpu().double(), g["volume"].double()
    assert vol.shape == ref.shape
    # rays are built on the device (ulp-level differences to the CPU rays): a sample within an ulp of a voxel face may
    # land in the neighbouring voxel, and the x2000-scaled finest hash levels turn ulps of position into percent-level
    # changes of single sigmas -- so: same support up to a handful of voxels, same maxima for the bulk
    support_diff = ((vol != 0) != (ref != 0)).sum().item()
    assert support_diff <= 0.002 * ref.numel(), support_diff
    both = (vol != 0) & (ref != 0)
    rel = ((vol - ref).abs() / ref.abs().clamp_min(1e-6))[both]
    assert rel.median().item() < 1e-3, rel.median().item()
    assert (rel < 5e-2).float().mean().item() > 0.99, (rel < 5e-2).float().mean().item()
    assert abs(vol.max().item() - g["vmax"]) < 2e-2 * g["vmax"]
