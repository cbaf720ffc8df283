"""Per-call timeline of the bench step (BASELINE configs[1]: 8 x 4096 rays, two 16384-ray chunks): CUDA-event time of
every C-ABI call of the last steps in launch order, plus the gap to the previous call (PyTorch glue + idle).

    python profiles/prof_timeline.py [steps] [workload]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from loner_b200 import _lib, mapping as M  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
W = bench.WORKLOADS[sys.argv[2] if len(sys.argv) > 2 else "c2"]
bench.WORKLOAD = W
dev = torch.device("cuda:0")
torch.manual_seed(1234)
kfs, scale, shift = bench.make_scene(W["keyframes"], seed=0)
settings = M.default_settings(n_samples_train=W["n_samples"], log2_hashmap_size=W["log2_hashmap_size"], ray_range=W["ray_range"],
                              lidar=W["rays_per_keyframe"], sky=0, chunk=W["chunk"], voxel_size=W["voxel_size"])
opt = M.Optimizer(settings, None, M.WorldCube(scale, shift), dev)
window = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=dev) for k in kfs]
for i, kf in enumerate(window):
    free = bool(W["pose_grad"]) and i > 0
    kf.is_anchored = not free
    kf.get_lidar_pose().set_fixed(not free)
sigma_params = opt._model.get_sigma_parameters()
groups = [{"params": sigma_params, "lr": 0.01}]
pose_params = [kf.get_lidar_pose().get_pose_tensor() for kf in window if not kf.is_anchored]
if pose_params:
    groups.append({"params": pose_params, "lr": 0.001})
adam = M.FusedAdam(groups)
marks = []
for i in range(steps):
    if i == steps - 2:
        torch.cuda.synchronize()
        _lib.reset_counters()
        _lib.TIMING["on"] = True
    rays, depths = opt.build_batch(window, ignore_world_cube=True)
    loss = opt.compute_loss(None, (rays, depths), i)
    loss.backward()
    adam.step()
    adam.zero_grad(set_to_none=True)
    if i % 10 == 0:
        opt._step_occupancy_grid()
torch.cuda.synchronize()
ev = _lib.TIMING["events"]
t_first = ev[0][1]
prev_end = None
tot = 0.0
for name, e0, e1 in ev:
    ms = e0.elapsed_time(e1)
    gap = prev_end.elapsed_time(e0) if prev_end is not None else 0.0
    tot += ms
    print(f"{t_first.elapsed_time(e0):9.3f} ms  {name:22s} {ms:8.4f} ms   gap before {gap:7.4f} ms")
    prev_end = e1
print(f"span {t_first.elapsed_time(ev[-1][2]):.3f} ms over 2 steps, kernels {tot:.3f} ms")

# ---- the same forward / backward calls replayed back to back on the last step's inputs (context vs. data) ----
if len(sys.argv) > 3 and sys.argv[3] == "replay":
    from loner_b200 import ops
    _lib.TIMING["on"] = False
    nerf = opt._model.nerf_model
    r_c = rays[:W["chunk"]].detach().contiguous()
    z = opt._results_lidar["samples_fine"][:W["chunk"]].contiguous()
    params = sigma_params[0].detach()

    def timeit(name, fn, reps=5):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        print(f"{name:50s} {e0.elapsed_time(e1) / reps:8.4f} ms")

    plan = nerf._sigma_plan if hasattr(nerf, "_sigma_plan") else nerf._model_sigma.plan
    timeit("replay sigma_fwd save_enc, step inputs", lambda: ops.sigma_fwd(plan, params, rays=r_c, z=z, save_enc=True))
    big = torch.cat([params[:3072], (torch.rand(params.numel() - 3072, device=dev) - 0.5) * 0.4])
    timeit("replay sigma_fwd save_enc, large table values", lambda: ops.sigma_fwd(plan, big, rays=r_c, z=z, save_enc=True))
    zu = torch.sort(torch.rand_like(z) * (r_c[:, 12:13] - r_c[:, 11:12]) + r_c[:, 11:12], dim=1).values.contiguous()
    timeit("replay sigma_fwd save_enc, uniform z", lambda: ops.sigma_fwd(plan, params, rays=r_c, z=zu, save_enc=True))
    sig, saved = ops.sigma_fwd(plan, params, rays=r_c, z=z, save_enc=True)
    gp = torch.zeros_like(params)
    for frac in (1.0, 0.5):
        ds = torch.randn_like(z) * 1e-3
        ds = ds * (torch.rand_like(z) < frac)
        timeit(f"replay sigma_bwd params, {frac:.0%} non-zero d_sigma", lambda: ops.sigma_bwd(plan, params, saved, ds, rays=r_c, z=z, grad_params=gp, d_rays=None))
    print("z stats: mean spacing", float((z[:, 1:] - z[:, :-1]).mean()), "median", float((z[:, 1:] - z[:, :-1]).median()))
