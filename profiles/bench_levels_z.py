"""Per-level encode cost for two depth distributions along the same rays: "uniform" (flat occupancy: stratified +
uniform inverse-CDF draws) and "surface" (half of the samples within +-2 m of the LiDAR return, which is what the
occupancy grid produces after a few updates).  python profiles/bench_levels_z.py [log2_T]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from loner_b200 import mapping as M, ops, synth

T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
N, S = 16384, 512
dev = torch.device("cuda:0")
torch.manual_seed(0)
kfs, scale, shift = synth.make_keyframes(4, ray_range=(1.0, 50.0), seed=0)
wc = M.WorldCube(scale.to(dev), shift.to(dev))
rays, depths = [], []
for k in kfs:
    idx = torch.randint(k.distances.shape[0], (N // 4,), device=dev)
    r, d = M.build_lidar_rays(k.directions.to(dev), k.distances.to(dev), k.pose.to(dev), wc, (1.0, 50.0), True, indices=idx)
    rays.append(r); depths.append(d)
rays, depths = torch.vstack(rays).contiguous(), torch.cat(depths).contiguous()
near, far = rays[:, 11:12], rays[:, 12:13]
t = (torch.arange(S // 2, device=dev) + torch.rand(N, S // 2, device=dev)) / (S // 2)
coarse = near + (far - near) * t
fine_u = near + (far - near) * torch.rand(N, S // 2, device=dev)
fine_s = (depths[:, None] + (torch.rand(N, S // 2, device=dev) * 2 - 1) * 2.0 / float(scale)).clamp(min=near, max=far)
zs = {"uniform": torch.sort(torch.cat([coarse, fine_u], 1), 1).values.contiguous(),
      "surface": torch.sort(torch.cat([coarse, fine_s], 1), 1).values.contiguous()}
P = N * S


def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


print(f"P={P} T=2^{T}")
full = ops.Plan(log2_hashmap_size=T)
params = torch.cat([(torch.rand(3072) - 0.5) * 0.5, (torch.rand(full.n_table_params) - 0.5) * 0.4]).to(dev)
for name, z in zs.items():
    tf = timeit(lambda: ops.sigma_fwd(full, params, rays=rays, z=z, save_enc=True))
    sig, saved = ops.sigma_fwd(full, params, rays=rays, z=z, save_enc=True)
    ds = torch.randn_like(z) * 1e-3
    gp = torch.zeros_like(params)
    tb = timeit(lambda: ops.sigma_bwd(full, params, saved, ds, rays=rays, z=z, grad_params=gp, d_rays=None))
    print(f"{name:8s} sigma_fwd {tf:7.3f} ms   sigma_bwd {tb:7.3f} ms")
for first, n in [(0, 16), (0, 1), (1, 1), (2, 1), (3, 1), (4, 1), (5, 1), (6, 1), (7, 1), (8, 1), (10, 1), (15, 1), (0, 3), (3, 4), (7, 9)]:
    plan = ops.Plan(n_levels=n, log2_hashmap_size=T, base_resolution=16 * 2 ** first)
    table = (torch.rand(plan.n_table_params, device=dev) - 0.5) * 0.4
    d_enc = torch.randn(P, 2 * n, device=dev) * 1e-3
    gt = torch.zeros_like(table)
    row = f"levels {first:2d}..{first + n - 1:2d}  res {plan.level(0)['res']:7d}  dense={plan.level(0)['dense']}"
    for name, z in zs.items():
        x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None] + 1) / 2).reshape(-1, 3).contiguous()
        tf = timeit(lambda: ops.encode_fwd(plan, x01, table))
        tb = timeit(lambda: ops.encode_bwd(plan, x01, table, d_enc, gt, False))
        row += f"   {name}: fwd {tf:6.3f} bwd {tb:6.3f}"
    print(row)
