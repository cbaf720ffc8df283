"""Which hash-grid levels cost what: encode fwd / bwd timings for sub-ranges of the 16 levels (same sample
positions as bench_kernels.py).  python profiles/bench_levels.py [log2_T]"""
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
rays = []
for k in kfs:
    idx = torch.randint(k.distances.shape[0], (N // 4,), device=dev)
    r, d = M.build_lidar_rays(k.directions.to(dev), k.distances.to(dev), k.pose.to(dev), wc, (1.0, 50.0), True, indices=idx)
    rays.append(r)
rays = torch.vstack(rays).contiguous()
occ = (torch.randn(100, 100, 100) * 0.5).to(dev)
z = ops.sample_occ(rays, occ, torch.rand(N, S // 2, device=dev), torch.rand(N, S // 2, device=dev), S, 1.0)
x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None] + 1) / 2).reshape(-1, 3).contiguous()
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
for first, n in [(0, 16), (0, 1), (1, 1), (2, 1), (3, 1), (4, 1), (5, 1), (6, 1), (7, 1), (8, 1), (10, 1), (12, 1), (15, 1), (0, 3), (3, 3), (6, 10)]:
    plan = ops.Plan(n_levels=n, log2_hashmap_size=T, base_resolution=16 * 2 ** first)
    table = (torch.rand(plan.n_table_params, device=dev) - 0.5) * 0.4
    d_enc = torch.randn(P, 2 * n, device=dev) * 1e-3
    gt = torch.zeros_like(table)
    tf = timeit(lambda: ops.encode_fwd(plan, x01, table))
    tb = timeit(lambda: ops.encode_bwd(plan, x01, table, d_enc, gt, False))
    print(f"levels {first:2d}..{first + n - 1:2d}  res {plan.level(0)["res"]:7d}  dense={plan.level(0)["dense"]}  fwd {tf:7.3f} ms  bwd {tb:7.3f} ms")
