"""Per-kernel CUDA-event timings at the BASELINE configs[1] chunk size (16384 rays x 512 samples, T=2^19)
on realistic sample positions (synthetic OS1-64 rays, occupancy-guided depths).

    python profiles/bench_kernels.py [log2_T] [n_rays] [n_samples]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from loner_b200 import mapping as M, ops, synth  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
N = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
S = int(sys.argv[3]) if len(sys.argv) > 3 else 512
dev = torch.device("cuda:0")
torch.manual_seed(0)
kfs, scale, shift = synth.make_keyframes(4, ray_range=(1.0, 50.0), seed=0)
wc = M.WorldCube(scale.to(dev), shift.to(dev))
rr = torch.tensor([1.0, 50.0], device=dev)
rays, depths = [], []
for k in kfs:
    idx = torch.randint(k.distances.shape[0], (N // 4,))
    r, d = M.build_lidar_rays(k.directions[:, idx].to(dev), k.distances[idx].to(dev), k.pose.to(dev), wc, rr, True)
    rays.append(r); depths.append(d)
rays, depths = torch.vstack(rays).contiguous(), torch.cat(depths).contiguous()
plan = ops.Plan(log2_hashmap_size=T)
params = torch.cat([(torch.rand(3072) - 0.5) * 0.5, (torch.rand(plan.n_table_params) - 0.5) * 0.4]).to(dev)
occ = (torch.randn(100, 100, 100) * 0.5).to(dev)
u1, u2 = torch.rand(N, S // 2, device=dev), torch.rand(N, S // 2, device=dev)
noise = torch.randn(N, S, device=dev)
z = ops.sample_occ(rays, occ, u1, u2, S, 1.0)
x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None] + 1) / 2).reshape(-1, 3).contiguous()
P = N * S
sigma, enc = ops.sigma_fwd(plan, params, rays=rays, z=z)
d_enc = torch.randn(P, 32, device=dev) * 1e-3
d_sigma = torch.randn(N, S, device=dev) * 1e-3
gp = torch.zeros_like(params)
gt = gp[3072:]
dr = torch.zeros_like(rays)
table = params[3072:]
n_op = ops.count_opaque(rays, depths)
cfg = ops.make_loss_config("L1_JS", float(scale), 0.0, P)
parts = torch.zeros(4, device=dev)
m, v = torch.zeros_like(params), torch.zeros_like(params)


def timeit(name, fn, bytes_per_point=None, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    extra = f"  {bytes_per_point * P / ms / 1e6:8.0f} GB/s algorithmic" if bytes_per_point else ""
    print(f"{name:34s} {ms:9.4f} ms  {P / ms / 1e6:8.2f} Gpt/s{extra}")


print(f"N={N} S={S} P={P} T=2^{T}")
timeit("sample_occ", lambda: ops.sample_occ(rays, occ, u1, u2, S, 1.0), 12300 / S)
timeit("encode_fwd (x01)", lambda: ops.encode_fwd(plan, x01, table), 1024 + 12 + 128)
timeit("encode_bwd table only", lambda: ops.encode_bwd(plan, x01, table, d_enc, gt, False), 2048 + 128 + 12)
d_enc_half = d_enc * (torch.rand(P, 1, device=dev) < 0.5)
timeit("encode_bwd table only, 50% zero rows", lambda: ops.encode_bwd(plan, x01, table, d_enc_half, gt, False), 2048 + 128 + 12)
timeit("encode_bwd table + dx", lambda: ops.encode_bwd(plan, x01, table, d_enc, gt, True), 3072 + 128 + 24)
timeit("sigma_fwd (rays) + save enc", lambda: ops.sigma_fwd(plan, params, rays=rays, z=z), 1024 + 4 + 128 + 4)
timeit("sigma_fwd (rays) no save", lambda: ops.sigma_fwd(plan, params, rays=rays, z=z, save_enc=False), 1024 + 8)
timeit("sigma_bwd params", lambda: ops.sigma_bwd(plan, params, enc, d_sigma, rays=rays, z=z, grad_params=gp), 2048 + 136)
timeit("sigma_bwd params + rays", lambda: ops.sigma_bwd(plan, params, enc, d_sigma, rays=rays, z=z, grad_params=gp, d_rays=dr), 3072 + 136)
timeit("sigma_bwd rays only", lambda: ops.sigma_bwd(plan, params, enc, d_sigma, rays=rays, z=z, grad_params=None, d_rays=dr), 1024 + 136)
timeit("composite_fwd", lambda: ops.composite_fwd(sigma, z, rays, noise), 16)
timeit("composite_loss (fwd+grad)", lambda: ops.composite_loss(sigma, z, rays, noise, depths, cfg, parts, n_opaque_dev=n_op), 20)
timeit("adam_step", lambda: ops.adam_step(params, gp, m, v, 1, 0.0), None)
timeit("occ_grad", lambda: ops.occ_grad(rays, z, depths, float(scale), 100), 76)
timeit("torch randn [N,S]", lambda: torch.randn(N, S, device=dev))
timeit("torch zeros_like(params)", lambda: torch.zeros_like(params))
# colour branch: 48 -> 64 x 4 -> 3 network on the same point count
net_params = ((torch.rand(64 * 48 + 3 * 64 * 64 + 16 * 64) * 2 - 1) * 0.2).to(dev)
xin = torch.randn(P, 48, device=dev) * 0.3
d01 = torch.rand(P, 3, device=dev)
timeit("sh4_fwd", lambda: ops.sh4_fwd(d01), 12 + 64)
timeit("net_fwd 48-64x4-3 (inference)", lambda: ops.net_fwd(xin, net_params, 4, 3, 1), 192 + 12)
out_n, acts_n = ops.net_fwd(xin, net_params, 4, 3, 1, save_acts=True)
timeit("net_fwd 48-64x4-3 (+ save acts)", lambda: ops.net_fwd(xin, net_params, 4, 3, 1, save_acts=True), 192 + 12 + 1024)
g_out = torch.randn(P, 3, device=dev)
gnp = torch.zeros_like(net_params)
timeit("net_bwd 48-64x4-3 (params + dx)", lambda: ops.net_bwd(xin, net_params, acts_n, out_n, g_out, 4, 3, 1, grad_params=gnp, need_dx=True), 192 * 2 + 12 + 1024, reps=2)
