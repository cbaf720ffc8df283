import os, sys
sys.path.insert(0, "/root/repo")
sys.argv = ["x"]
exec(open("/root/repo/profiles/bench_levels_z.py").read().split("print(f\"P={P} T=2^{T}\")")[0])
print("marginal cost of the coarse levels (encode_bwd, table only)")
for first, n in [(0, 16), (1, 15), (2, 14), (3, 13), (4, 12), (7, 9)]:
    plan = ops.Plan(n_levels=n, log2_hashmap_size=T, base_resolution=16 * 2 ** first)
    table = (torch.rand(plan.n_table_params, device=dev) - 0.5) * 0.4
    d_enc = torch.randn(P, 2 * n, device=dev) * 1e-3
    gt = torch.zeros_like(table)
    row = f"levels {first:2d}..15"
    for name, z in zs.items():
        x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None] + 1) / 2).reshape(-1, 3).contiguous()
        tf = timeit(lambda: ops.encode_fwd(plan, x01, table))
        tb = timeit(lambda: ops.encode_bwd(plan, x01, table, d_enc, gt, False))
        row += f"   {name}: fwd {tf:6.3f} bwd {tb:6.3f}"
    print(row)
