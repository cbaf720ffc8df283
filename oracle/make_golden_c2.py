"""ORACLE / TEST INFRASTRUCTURE ONLY.  Container-only script (needs /root/reference).

A golden fixture at the BENCHMARK shape (BASELINE configs[1]): the reference's OWN, unmodified
Optimizer.compute_loss (src/mapping/optimizer.py:701-859) -> Model.forward chunk loop (src/models/model_tcnn.py:70-108)
-> loss.backward() -> _step_occupancy_grid (optimizer.py:897-908) on torch-CPU behind the fp32 `tinycudann` shim with

    2048 rays (1792 lidar + 256 sky) x 512 samples, 16-level hash grid T = 2^19, 100^3 occupancy grid warmed by the
    reference's own update rule, cfg.render.chunk = 1024 (two chunks, so the per-chunk RNG draws and the concatenation
    are exercised), pose gradient on; states: L1_JS with tables x1 (initialisation) and x2000 ("trained"), and L2_JS
    with tables x2000 (the same pipeline without the sign discontinuity of the L1 line-of-sight term).

The random tensors come from oracle/portable_rng.py (patched over torch.rand / torch.randn while the reference runs),
so the fixture stores none of them: tests regenerate the same bits.  Stored in full: rays, occupancy grid, sample
depths z, `searchsorted` indices (int16), occupancy bits (packed); per parameter state: per-ray outputs, every 8th
ray's sigma / weights rows, loss, eps, MLP gradient, top-4096 table-gradient entries + norms, ray and pose
gradients, and the top-16384 entries of the occupancy update.

    python -m oracle.make_golden_c2            # writes tests/golden/c2_shape_ogm_l1js.pt (~10 MB)
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.reference_import import import_reference  # noqa: E402
from oracle import make_golden as G  # noqa: E402
from oracle import portable_rng as R  # noqa: E402
from loner_b200 import synth  # noqa: E402

NAME = "c2_shape_ogm_l1js"
N_LIDAR, N_SKY, S, LOG2_T, VOXEL, CHUNK = 1792, 256, 512, 19, 100, 1024
RAY_RANGE = (1.0, 50.0)
ROW_STRIDE = 8          # sigma / weights rows stored for every 8th ray


def capture_searchsorted():
    log = []
    orig = torch.searchsorted

    def f(*a, **k):
        r = orig(*a, **k)
        log.append(r.detach().clone())
        return r
    torch.searchsorted = f
    return log, orig


def golden_iteration(ns, opt, rays_in, depths, pose, base_stream):
    """One compute_loss + backward of the reference with the portable RNG; returns (dict of outputs, rng log)."""
    sigma_p = opt._model.nerf_model._model_sigma.params
    sigma_p.grad = None
    if pose.get_pose_tensor().grad is not None:
        pose.get_pose_tensor().grad = None
    log_ss, orig_ss = capture_searchsorted()
    try:
        with R.Patch(base_stream) as rng:
            loss = opt.compute_loss(None, (rays_in, depths), 0)
    finally:
        torch.searchsorted = orig_ss
    loss.backward(retain_graph=True)
    res = opt._results_lidar
    g = sigma_p.grad.detach()
    tg = g[3072:]
    top = torch.topk(tg.abs(), 4096).indices.sort().values
    rows = slice(0, None, ROW_STRIDE)
    out = dict(loss=loss.detach().clone(), depth_eps=float(opt._depth_eps),
               depth=res["depth_fine"].detach().clone(), opacity=res["opacity_fine"].detach().clone(),
               variance=res["variance"].detach().clone(),
               sigma_rows=res["raw_fine"].detach()[rows, :, 0].clone(), weights_rows=res["weights_fine"].detach()[rows].clone(),
               weights_sum=res["weights_fine"].detach().double().sum().item(),
               sigma_abs_sum=res["raw_fine"].detach().double().abs().sum().item(),
               mlp_grad=g[:3072].clone(), table_grad_idx=top, table_grad_val=tg[top].clone(),
               table_grad_norm=tg.double().norm().item(), table_grad_sum=tg.double().sum().item(),
               table_grad_nnz=int((tg != 0).sum()),
               rays_grad=rays_in.grad.detach().clone(), pose_grad=pose.get_pose_tensor().grad.detach().clone())
    return out, rng, log_ss, res["samples_fine"].detach().clone()


def main():
    ns = import_reference()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    torch.manual_seed(0)
    gen = torch.Generator().manual_seed(7)
    # a long hall: beams along x leave the 50 m ray range, so the scan has sky rays (transparent-ray branch of the loss)
    kfs, scale, shift = synth.make_keyframes(2, ray_range=RAY_RANGE, seed=0, scene=synth.Scene(half_extent=(70.0, 14.0)))
    cfg = G.model_cfg(ns, S, LOG2_T, VOXEL, RAY_RANGE, "L1_JS", 1.0, chunk=CHUNK)
    opt = G.bare_optimizer(ns, cfg, scale, "OGM")
    G.set_params(opt, LOG2_T)

    # warm the 100^3 occupancy grid with the reference's own update rule (large lr so the logits matter)
    for g_ in opt._occupancy_grid_optimizer.param_groups:
        g_["lr"] = 0.05
    for w in range(10):
        rays, depths, _, _ = G.reference_rays(ns, kfs[w % 2], scale, shift, RAY_RANGE, N_LIDAR, N_SKY, gen, False)
        with torch.no_grad(), R.Patch(1000 + 10 * w):
            opt.compute_loss(None, (rays.float(), depths.float()), w)
        opt._step_occupancy_grid()
        print(f"warm step {w}: occ |logit| max {float(opt._occupancy_grid.detach().abs().max()):.3f}", flush=True)
    for g_ in opt._occupancy_grid_optimizer.param_groups:
        g_["lr"] = cfg.model.occ_model.lr
    occ_before = opt._occupancy_grid.detach().clone()

    rays, depths, pose, idx = G.reference_rays(ns, kfs[1], scale, shift, RAY_RANGE, N_LIDAR, N_SKY, gen, True)
    rays = rays.float()
    rays.retain_grad()
    depths = depths.float()
    assert rays.shape[0] == N_LIDAR + N_SKY, rays.shape
    out = dict(name=NAME, n_samples=S, sampler="OGM", loss_selection="L1_JS", log2_T=LOG2_T, voxel=VOXEL, chunk=CHUNK,
               ray_range=list(RAY_RANGE), param_seed=1337, min_js=1.0, rng_base_stream=0, row_stride=ROW_STRIDE,
               scale=scale.clone(), shift=shift.clone(), rays=rays.detach().clone(), depths=depths.detach().clone(),
               kf_pose=kfs[1].pose.clone(), kf_directions=kfs[1].directions[:, idx].clone(),
               kf_distances=kfs[1].distances[idx].clone(), occ=occ_before, states={})
    from oracle import loner_oracle as O
    # L1_JS at both parameter states; L2_JS (a loss without the sign discontinuity of |w - w_gt|) at the trained state
    for key, table_scale, loss_sel in ((2000, 2000.0, "L1_JS"), (1, 1.0, "L1_JS"), ("l2js_2000", 2000.0, "L2_JS")):
        opt._model_config["loss"]["loss_selection"] = loss_sel      # raw dict access: attribute reads return wrapped copies
        assert opt._model_config.loss.loss_selection == loss_sel
        p = O.init_sigma_params(LOG2_T, 1337)
        p[O.MLP_PARAMS:] *= table_scale
        with torch.no_grad():
            opt._model.nerf_model._model_sigma.params.copy_(p)
            opt._occupancy_grid_model.occupancy_grid.copy_(occ_before)      # the previous state's update is undone
        opt._occupancy_grid = opt._occupancy_grid_model()
        opt._ray_sampler.update_occ_grid(opt._occupancy_grid.detach())
        assert torch.equal(opt._occupancy_grid.detach(), occ_before)
        rays.grad = None
        st, rng, log_ss, z = golden_iteration(ns, opt, rays, depths, pose, 0)
        st.update(table_scale=table_scale, loss_selection=loss_sel, params_checksum=p.double().sum().item())
        # sampler outputs do not depend on the parameters: stored once, asserted equal for the second state
        inds = torch.cat(log_ss, 0)
        if "z" not in out:
            n_chunks = len(log_ss)
            assert n_chunks == 2 and len(rng.rand) == 4 and len(rng.randn) == 2
            assert int(inds.max()) < 32768
            out.update(z=z, inds=inds.to(torch.int16))
            # occupancy bits at the coarse depths, re-derived with the reference's own ops (ray_sampling.py:59-81)
            bits = []
            for c in range(n_chunks):
                r_c = rays.detach()[c * CHUNK:(c + 1) * CHUNK]
                t = torch.linspace(0, 1, S // 2)
                zc = r_c[:, -2:-1] * (1 - t) + r_c[:, -1:] * t
                mid = 0.5 * (zc[:, :-1] + zc[:, 1:])
                upper, lower = torch.cat([mid, zc[:, -1:]], -1), torch.cat([zc[:, :1], mid], -1)
                zc = lower + (upper - lower) * (1.0 * rng.rand[2 * c])
                p3 = r_c[:, None, 0:3] + r_c[:, None, 3:6] * zc.unsqueeze(2)
                lg = ns.model_tcnn.OccupancyGridModel.interpolate(occ_before, p3)
                pr = 1. / (1 + torch.exp(-lg))
                pr = 2 * (pr.clamp(min=0.5, max=1.0) - 0.5)
                bits.append(pr.reshape(r_c.shape[0], S // 2) > 0)
            bits = torch.cat(bits, 0)
            out.update(occ_bits_packed=torch.from_numpy(np.packbits(bits.numpy(), axis=1)),
                       occ_bits_count=int(bits.sum()))
        else:
            assert torch.equal(out["z"], z) and torch.equal(out["inds"], inds.to(torch.int16))
        opt._step_occupancy_grid()
        delta = (opt._occupancy_grid.detach() - occ_before).flatten()
        top = torch.topk(delta.abs(), 16384).indices.sort().values
        st.update(occ_delta_idx=top, occ_delta_val=delta[top].clone(), occ_delta_norm=delta.double().norm().item(),
                  occ_delta_nnz=int((delta != 0).sum()))
        out["states"][key] = st
        print(f"state {key} ({loss_sel}, tables x{table_scale:g}): loss {float(st['loss']):.6f} eps {st['depth_eps']:.4f} "
              f"|table grad| {st['table_grad_norm']:.4e} nnz {st['table_grad_nnz']}", flush=True)
    path = os.path.join(G.GOLDEN_DIR, f"{NAME}.pt")
    torch.save(out, path)
    print(f"{NAME}: N={rays.shape[0]} S={S} -> {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    main()
