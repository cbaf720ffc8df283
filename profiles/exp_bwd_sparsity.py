"""lb_sigma_bwd as a function of the fraction of non-zero d_sigma rows (0 % = the MLP-backward producer chain alone),
and the fraction of exactly-zero d_sigma entries in the bench step.  python profiles/exp_bwd_sparsity.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = [sys.argv[0]]
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench_levels_z.py")).read()
exec(src.split('print(f"P={P} T=2^{T}")')[0])
full = ops.Plan(log2_hashmap_size=T)
params = torch.cat([(torch.rand(3072) - 0.5) * 0.5, (torch.rand(full.n_table_params) - 0.5) * 0.4]).to(dev)
z = zs["surface"]
sig, saved = ops.sigma_fwd(full, params, rays=rays, z=z, save_enc=True)
gp = torch.zeros_like(params)
for frac in (1.0, 0.75, 0.5, 0.25, 0.1, 0.0):
    ds = torch.randn_like(z) * 1e-3 * (torch.rand_like(z) < frac)
    t = timeit(lambda: ops.sigma_bwd(full, params, saved, ds, rays=rays, z=z, grad_params=gp, d_rays=None), reps=5)
    print(f"non-zero rows {frac:4.0%}: sigma_bwd {t:7.3f} ms")
# coherent sparsity: the first 60 % of every ray (free space) has zero gradient
ds = torch.randn_like(z) * 1e-3
ds[:, : int(0.6 * S)] = 0
t = timeit(lambda: ops.sigma_bwd(full, params, saved, ds, rays=rays, z=z, grad_params=gp, d_rays=None), reps=5)
print(f"first 60 % of every ray zero: sigma_bwd {t:7.3f} ms")

import bench
from loner_b200 import functional as F
W = bench.WORKLOADS["c2"]; bench.WORKLOAD = W
kfs2, scale2, shift2 = bench.make_scene(8, seed=0)
settings = M.default_settings(n_samples_train=512, log2_hashmap_size=19, ray_range=W["ray_range"], lidar=4096, sky=0, chunk=16384, voxel_size=100)
opt = M.Optimizer(settings, None, M.WorldCube(scale2, shift2), dev)
window = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=dev) for k in kfs2]
for kf in window:
    kf.is_anchored = True; kf.get_lidar_pose().set_fixed(True)
adam = M.FusedAdam([{"params": opt._model.get_sigma_parameters(), "lr": 0.01}])
orig = ops.sigma_bwd
fr = []
def spy(plan, params, saved, d_sigma, **kw):
    fr.append((d_sigma == 0).float().mean().item())
    return orig(plan, params, saved, d_sigma, **kw)
ops.sigma_bwd = spy
F.ops.sigma_bwd = spy
for i in range(1, 46):
    r, d = opt.build_batch(window, ignore_world_cube=True)
    loss = opt.compute_loss(None, (r, d), i)
    loss.backward(); adam.step(); adam.zero_grad(set_to_none=True)
    if i % 10 == 0:
        opt._step_occupancy_grid()
print("zero fraction of d_sigma per sigma_bwd call, steps 1..45:", " ".join(f"{f:.2f}" for f in fr[::6]))
