"""Worker of tests/test_dp_nccl_gpu.py (one process per GPU, launched with torch.distributed.run): an N-GPU mapping
step with a sharded global batch must reproduce the 1-GPU step ("depth/loss equal to 1-GPU run")
and the reference's own recorded run (tests/golden/c2_shape_ogm_l1js.pt)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def run(dev, dp, g, state, M, R, params0):
    s = M.default_settings(n_samples_train=g["n_samples"], log2_hashmap_size=g["log2_T"], ray_range=tuple(g["ray_range"]),
                           voxel_size=g["voxel"], loss_selection=g["states"][state]["loss_selection"], strategy="OGM", chunk=g["chunk"])
    s["model_config"]["loss"]["decay_depth_eps"] = False
    opt = M.Optimizer(s, None, M.WorldCube(g["scale"], g["shift"]), dev, data_parallel=dp)
    p = opt._model.nerf_model._model_sigma.params
    with torch.no_grad():
        p.copy_(params0.to(dev))
        opt._occupancy_grid.copy_(g["occ"].to(dev))
    rays = g["rays"].to(dev).requires_grad_(True)
    with R.Patch(g["rng_base_stream"]):
        loss = opt.compute_loss(None, (rays, g["depths"].to(dev)), 0)
    loss.backward()
    res = opt._results_lidar
    out = dict(loss=loss.detach().clone(), grad=p.grad.detach().clone(), rays_grad=rays.grad.detach().clone(),
               depth=res["depth_fine"].detach().clone(), z=res["samples_fine"].detach().clone())
    adam = M.FusedAdam([{"params": [p], "lr": 0.01}], data_parallel=dp)
    adam.step()
    opt._step_occupancy_grid()
    out.update(params=p.detach().clone(), occ=opt._occupancy_grid.detach().clone())
    return out


def train(dev, dp, g, state, M, R, params0, overlap, n_steps=4):
    """n_steps of compute_loss -> backward -> sharded Adam on the same batch; with `overlap` the optimiser step (reduce-
    scatter, Adam, all-gather) runs on a side stream and compute_loss's wait_params() orders it before the next field
    evaluation."""
    s = M.default_settings(n_samples_train=g["n_samples"], log2_hashmap_size=g["log2_T"], ray_range=tuple(g["ray_range"]),
                           voxel_size=g["voxel"], loss_selection=g["states"][state]["loss_selection"], strategy="OGM", chunk=g["chunk"])
    opt = M.Optimizer(s, None, M.WorldCube(g["scale"], g["shift"]), dev, data_parallel=dp)
    p = opt._model.nerf_model._model_sigma.params
    with torch.no_grad():
        p.copy_(params0.to(dev))
        opt._occupancy_grid.copy_(g["occ"].to(dev))
    adam = M.FusedAdam([{"params": [p], "lr": 0.01}], data_parallel=dp, overlap_step=overlap)
    rays, depths = g["rays"].to(dev), g["depths"].to(dev)
    losses = []
    for i in range(n_steps):
        with R.Patch(g["rng_base_stream"] + i):
            loss = opt.compute_loss(None, (rays, depths), i)
        loss.backward()
        adam.step()
        adam.zero_grad(set_to_none=True)
        losses.append(loss.detach().clone())
    dp.wait_params()
    return p.detach().clone(), torch.stack(losses)


def main():
    from conftest import load_golden, c2_shape_params
    from loner_b200 import mapping as M
    from oracle import portable_rng as R
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    import datetime
    dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    g = load_golden("c2_shape_ogm_l1js")
    state = 2000
    st = g["states"][state]
    params0 = c2_shape_params(g, state)
    dp = M.RayDataParallel(shard_batch=True)
    assert dp.enabled and dp.world_size == world
    multi = run(dev, dp, g, state, M, R, params0)
    # assemble the N-GPU step: losses add up (per-rank partial sums over GLOBAL normalisers), gradients add up
    loss = multi["loss"].clone(); dist.all_reduce(loss)
    grad = multi["grad"].clone(); dist.all_reduce(grad)
    rays_grad = multi["rays_grad"].clone(); dist.all_reduce(rays_grad)      # each rank's rows are non-zero only in its slices
    report = {}

    # the 1-GPU step on every rank (a disabled DP context), for the comparison
    class Off(M.RayDataParallel):
        def __init__(self):
            self.enabled, self.world_size, self.rank, self.shard_batch, self.events, self.group = False, 1, 0, False, None, None
            self._param_work = []
    single = run(dev, Off(), g, state, M, R, params0)
    lo, hi = dp.rows(g["chunk"])
    rows = torch.cat([torch.arange(c + lo, min(c + hi, g["rays"].shape[0])) for c in range(0, g["rays"].shape[0], g["chunk"])]).to(dev)
    report["z_equal"] = bool(torch.equal(multi["z"], single["z"][rows]))
    report["depth_equal"] = rel(multi["depth"], single["depth"][rows])
    report["loss_vs_single"] = abs(loss.item() - single["loss"].item()) / abs(single["loss"].item())
    report["grad_vs_single"] = rel(grad, single["grad"])
    report["rays_grad_vs_single"] = rel(rays_grad, single["rays_grad"])
    report["params_after_adam_vs_single"] = rel(multi["params"] - params0.to(dev), single["params"] - params0.to(dev))
    report["occ_after_update_vs_single"] = rel(multi["occ"] - g["occ"].to(dev), single["occ"] - g["occ"].to(dev))
    report["loss_vs_reference"] = abs(loss.item() - st["loss"].item()) / abs(st["loss"].item())
    report["mlp_grad_vs_reference"] = rel(grad[:3072], st["mlp_grad"])
    report["table_grad_vs_reference"] = rel(grad[3072:][st["table_grad_idx"].to(dev)], st["table_grad_val"])
    # every rank must hold the same parameters after the sharded step
    chk = multi["params"].double().sum().reshape(1)
    lst = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(lst, chk)
    report["params_identical_across_ranks"] = bool(all(torch.equal(lst[0], t) for t in lst))
    # side-stream optimiser step (what Optimizer.iterate_optimizer runs) against the in-line one: a step that a field
    # evaluation did not wait for would leave the parameters one step stale and bend the loss curve.
    # Not bit-equal by construction: the table gradients are fp32 atomics, their order differs from run to run.
    p_sync, l_sync = train(dev, dp, g, state, M, R, params0, overlap=False)
    p_over, l_over = train(dev, dp, g, state, M, R, params0, overlap=True)
    p0 = params0.to(dev)
    report["overlap_losses"] = ((l_sync - l_over).abs() / l_sync.abs()).max().item()
    report["overlap_loss_curve"] = [round(x, 6) for x in l_over.tolist()]
    report["overlap_params_mean"] = ((p_sync - p_over).abs().mean() / (p_sync - p0).abs().mean()).item()
    if rank == 0:
        print("DP_EQUIVALENCE " + json.dumps(report))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
