"""ORACLE / TEST INFRASTRUCTURE ONLY.  Container-only script (needs /root/reference).

Runs the reference's OWN Python, unmodified (src/models/*, src/mapping/optimizer.py::compute_loss and
::_step_occupancy_grid, src/common/ray_utils.py::build_lidar_rays, src/common/pose.py), on torch-CPU
behind the fp32 `tinycudann` shim, on seeded synthetic OS1-64 keyframes, and writes small golden
fixtures to tests/golden/*.pt.  The random tensors the reference drew (torch.rand / torch.randn) and
the `searchsorted` indices are captured by wrapping those torch functions for the duration of the
call -- the reference source is not touched.

    python -m oracle.make_golden            # regenerates tests/golden/

Fixture keys: inputs (rays, depths, occ, u1, u2, noise, scale, cfg), reference outputs (z, sigma,
weights, depth, opacity, variance, loss, depth_eps, inds, occ_logits, occ_bits), gradients
(mlp_grad, table_grad_idx/val, table_grad_norm, rays_grad, pose_grad) and occ grid after the update.
"""
import contextlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.reference_import import import_reference  # noqa: E402
from loner_b200 import synth  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
TABLE_SCALE = 2000.0   # tables are U(-1e-4,1e-4) at init; scale them to "trained" magnitudes (+-0.2)


def model_cfg(ns, n_samples, log2_T, voxel, ray_range, loss_selection="L1_JS", min_js=1.0, chunk=16384):
    A = ns.settings.Settings
    nerf = dict(enable_view_dependence=True,
                dir_encoding_intensity=dict(degree=4, otype="SphericalHarmonics"),
                intensity_network=dict(activation="ReLU", n_hidden_layers=4, n_neurons=64,
                                       otype="FullyFusedMLP", output_activation="None"),
                pos_encoding_intensity=dict(base_resolution=16, log2_hashmap_size=14, n_features_per_level=2,
                                            n_levels=16, otype="HashGrid"),
                pos_encoding_sigma=dict(base_resolution=16, log2_hashmap_size=log2_T, n_features_per_level=2,
                                        n_levels=16, otype="HashGrid"),
                sigma_network=dict(activation="ReLU", n_hidden_layers=1, n_neurons=64,
                                   otype="FullyFusedMLP", output_activation="None"))
    return A(dict(
        data=dict(ray_range=list(ray_range)),
        model=dict(num_colors=3, model_type="nerf_decoupled", nerf_config=nerf, ray_range=list(ray_range),
                   render=dict(N_samples_train=n_samples, N_samples_test=n_samples * 2, retraw=True, lindisp=False,
                               perturb=1.0, white_bkgd=False, raw_noise_std=1.0, chunk=chunk, netchunk=0),
                   occ_model=dict(voxel_size=voxel, lr=1e-4, N_iters_acc=10)),
        train=dict(lrate_sigma_mlp=0.01, lrate_rgb=0.01, lrate_pose=0.001, lrate_gamma=1.0),
        loss=dict(loss_selection=loss_selection,
                  JS_loss=dict(min_js_score=min_js, max_js_score=10.0, alpha=1.0),
                  decay_los_lambda=False, los_lambda=1000.0, min_los_lambda=10.0,
                  los_lambda_decay_rate=0.001, los_lambda_decay_steps=15000,
                  decay_depth_eps=False, depth_eps=3.0, min_depth_eps=0.5,
                  depth_eps_decay_rate=0.95, depth_eps_decay_steps=1,
                  depthloss_lambda=0.005, cam_lambda=1, std_lambda=0.0)))


def bare_optimizer(ns, cfg, scale, sampler="OGM", device="cpu"):
    """A reference Optimizer with only the state compute_loss/_step_occupancy_grid read
    (its __init__ needs calibration/settings plumbing that is outside the hot path)."""
    Opt = ns.optimizer.Optimizer
    o = Opt.__new__(Opt)
    dbg = {k: False for k in ("visualize_loss", "draw_samples", "draw_rays_eps", "store_ray",
                              "write_ray_point_clouds", "draw_comp_graph", "log_losses")}
    o._settings = ns.settings.Settings(dict(debug=dbg, samples_selection=dict(strategy=sampler)))
    o._model_config = cfg
    o._device = device
    o._scale_factor = scale
    o._optimization_settings = ns.optimizer.OptimizationSettings()
    o._global_step = 0
    o._keyframe_count = 0
    o._model = ns.model_tcnn.Model(cfg.model).to(device)
    if sampler == "OGM":
        o._occupancy_grid_model = ns.model_tcnn.OccupancyGridModel(cfg.model.occ_model).to(device)
        o._occupancy_grid = o._occupancy_grid_model()
        o._occupancy_grid_optimizer = torch.optim.SGD(o._occupancy_grid_model.parameters(), lr=cfg.model.occ_model.lr)
        o._ray_sampler = ns.ray_sampling.OccGridRaySampler()
        o._ray_sampler.update_occ_grid(o._occupancy_grid.detach())
    else:
        o._ray_sampler = ns.ray_sampling.UniformRaySampler()
    return o


@contextlib.contextmanager
def capture_torch():
    """Record every torch.rand / torch.randn / torch.searchsorted result while active."""
    log = dict(rand=[], randn=[], searchsorted=[], grid_sample=[])
    orig = (torch.rand, torch.randn, torch.searchsorted, torch.nn.functional.grid_sample)

    def wrap(name, fn):
        def f(*a, **k):
            r = fn(*a, **k)
            log[name].append(r.detach().clone())
            return r
        return f
    torch.rand, torch.randn = wrap("rand", orig[0]), wrap("randn", orig[1])
    torch.searchsorted = wrap("searchsorted", orig[2])
    torch.nn.functional.grid_sample = wrap("grid_sample", orig[3])
    try:
        yield log
    finally:
        torch.rand, torch.randn, torch.searchsorted, torch.nn.functional.grid_sample = orig


def reference_rays(ns, kf, scale, shift, ray_range, n_lidar, n_sky, gen, pose_requires_grad):
    """Rays through the reference's own KeyFrame-less path: LidarScan + Pose + LidarRayDirections
    (ray_utils.py:269-322), sky rays from the detached pose with depth r1+1 (keyframe.py:96-98)."""
    wc = ns.pose_utils.WorldCube(scale, shift)
    n = kf.distances.shape[0]
    scan = ns.sensors.LidarScan(kf.directions, kf.distances, torch.arange(n).float() * 1e-6)
    pose = ns.pose.Pose(pose_tensor=kf.pose.clone(), fixed=not pose_requires_grad)
    rr = torch.tensor(ray_range, dtype=torch.float32)
    idx = torch.randint(n, (n_lidar,), generator=gen)
    T = pose.get_transformation_matrix()
    rays, depths = ns.ray_utils.LidarRayDirections(scan).build_lidar_rays(idx, rr, wc, T)
    if n_sky > 0 and kf.sky_directions.shape[1] > 0:
        m = kf.sky_directions.shape[1]
        sky = ns.sensors.LidarScan(kf.sky_directions, torch.full((m,), ray_range[1] + 1.0), torch.arange(m).float() * 1e-6)
        sidx = torch.randint(m, (n_sky,), generator=gen)
        srays, sdepths = ns.ray_utils.LidarRayDirections(sky).build_lidar_rays(sidx, rr, wc, T.detach())
        rays, depths = torch.cat((rays, srays)), torch.cat((depths, sdepths))
    return rays, depths, pose, idx


def set_params(opt, log2_T, seed=1337):
    """Deterministic 'trained-like' parameters: seeded init, tables scaled by TABLE_SCALE."""
    from oracle import loner_oracle as O
    p = O.init_sigma_params(log2_T, seed)
    p[O.MLP_PARAMS:] *= TABLE_SCALE
    with torch.no_grad():
        opt._model.nerf_model._model_sigma.params.copy_(p)
    return p


def run_case(ns, name, n_samples, sampler, loss_selection, n_lidar, n_sky, log2_T=14, voxel=32,
             ray_range=(1.0, 30.0), warm_steps=12, pose_grad=True, min_js=1.0, seed=0):
    torch.manual_seed(seed)
    gen = torch.Generator().manual_seed(seed)
    kfs, scale, shift = synth.make_keyframes(2, ray_range=ray_range, seed=seed)
    cfg = model_cfg(ns, n_samples, log2_T, voxel, ray_range, loss_selection, min_js)
    opt = bare_optimizer(ns, cfg, scale, sampler)
    params0 = set_params(opt, log2_T)

    # warm the occupancy grid with the reference's own update rule (big lr so logits matter)
    if sampler == "OGM":
        for g in opt._occupancy_grid_optimizer.param_groups:
            g["lr"] = 0.05
        for w in range(warm_steps):
            rays, depths, _, _ = reference_rays(ns, kfs[w % 2], scale, shift, ray_range, n_lidar, n_sky, gen, False)
            with torch.no_grad():
                opt.compute_loss(None, (rays.float(), depths.float()), w)
            opt._step_occupancy_grid()
        for g in opt._occupancy_grid_optimizer.param_groups:
            g["lr"] = cfg.model.occ_model.lr
        occ_before = opt._occupancy_grid.detach().clone()
    else:
        occ_before = None

    # the golden iteration
    rays, depths, pose, idx = reference_rays(ns, kfs[1], scale, shift, ray_range, n_lidar, n_sky, gen, pose_grad)
    rays = rays.float()
    rays.retain_grad() if rays.requires_grad else None
    sigma_p = opt._model.nerf_model._model_sigma.params
    sigma_p.grad = None
    with capture_torch() as log:
        loss = opt.compute_loss(None, (rays, depths.float()), 0)
    loss.backward()
    res = opt._results_lidar
    out = dict(name=name, n_samples=n_samples, sampler=sampler, loss_selection=loss_selection, log2_T=log2_T,
               voxel=voxel, ray_range=list(ray_range), table_scale=TABLE_SCALE, param_seed=1337, min_js=min_js,
               scale=scale.clone(), shift=shift.clone(),
               rays=rays.detach().clone(), depths=depths.detach().float().clone(),
               kf_pose=kfs[1].pose.clone(), kf_idx=idx,
               kf_directions=kfs[1].directions[:, idx].clone(), kf_distances=kfs[1].distances[idx].clone(),
               z=res["samples_fine"].detach().clone(), sigma=res["raw_fine"].detach()[..., 0].clone(),
               weights=res["weights_fine"].detach().clone(), depth=res["depth_fine"].detach().clone(),
               opacity=res["opacity_fine"].detach().clone(), variance=res["variance"].detach().clone(),
               loss=loss.detach().clone(), depth_eps=float(opt._depth_eps),
               params_checksum=params0.double().sum().item(), params_probe=params0[::100003].clone())
    if sampler == "OGM":
        out.update(u1=log["rand"][0], u2=log["rand"][1], noise=log["randn"][0], inds=log["searchsorted"][0],
                   occ=occ_before, occ_logits=log["grid_sample"][0].reshape(rays.shape[0], n_samples // 2))
        lg = out["occ_logits"]
        probs = 1. / (1 + torch.exp(-lg))                      # ray_sampling.py:79-80, verbatim ops
        probs = 2 * (probs.clamp(min=0.5, max=1.0) - 0.5)
        out.update(occ_probs=probs, occ_bits=probs > 0)
    else:
        out.update(u1=log["rand"][0], noise=log["randn"][0])
    g = sigma_p.grad.detach()
    tg = g[3072:]
    top = torch.topk(tg.abs(), 4096).indices.sort().values
    out.update(mlp_grad=g[:3072].clone(), table_grad_idx=top, table_grad_val=tg[top].clone(),
               table_grad_norm=tg.double().norm().item(), table_grad_sum=tg.double().sum().item(),
               table_grad_nnz=int((tg != 0).sum()))
    if rays.grad is not None:
        out["rays_grad"] = rays.grad.detach().clone()
    if pose_grad:
        out["pose_grad"] = pose.get_pose_tensor().grad.detach().clone()
    if sampler == "OGM":
        opt._step_occupancy_grid()
        out["occ_after"] = opt._occupancy_grid.detach().clone()
    # 'adjusted' (peak) render strategy on the same inputs (rendering_tcnn.py:70-214)
    with torch.no_grad():
        raw = res["raw_fine"].detach()
        try:
            _, d_adj, w_adj, op_adj, var_adj = ns.rendering_tcnn.raw2outputs_adjusted(
                raw, out["z"], rays[:, 0:3].detach(), rays[:, 3:6].detach(), 1.0, True, sigma_only=True,
                far=rays[:, -1:].detach(), ret_var=True)
            out.update(adj_depth=d_adj.clone(), adj_weights=w_adj.clone(), adj_variance=var_adj.clone())
        except Exception as e:  # the reference's peak code indexes empty tensors on some inputs
            out["adj_error"] = repr(e)
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, f"{name}.pt")
    torch.save(out, path)
    print(f"{name}: N={rays.shape[0]} S={n_samples} loss={float(loss.detach()):.6f} eps={out['depth_eps']:.4f} "
          f"-> {path} ({os.path.getsize(path) / 1e6:.2f} MB)")
    return out


def main():
    ns = import_reference()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    run_case(ns, "c1_ogm_l1js", 64, "OGM", "L1_JS", 448, 64, log2_T=18, ray_range=(1.0, 16.0), min_js=0.05)
    run_case(ns, "c1_uniform_l2los", 64, "UNIFORM", "L2_LOS", 320, 0, seed=1)
    run_case(ns, "s128_ogm_l2js", 128, "OGM", "L2_JS", 192, 32, seed=2, min_js=1.0)


if __name__ == "__main__":
    main()
