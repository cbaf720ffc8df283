"""Debug helper: where does lb_sigma_fwd_jvp differ from lb_sigma_fwd?  python profiles/debug_jvp.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from loner_b200 import ops
from test_jvp_gpu import make_inputs

for log2_T, n, s in [(15, 2048, 512), (19, 300, 129), (19, 2048, 512), (19, 2048, 512)]:
    plan = ops.Plan(log2_hashmap_size=log2_T)
    params, rays, z, d_sigma = make_inputs(n, s, 7 + n, plan)
    s0, (e0, a0) = ops.sigma_fwd(plan, params, rays=rays, z=z)
    s0b, (e0b, a0b) = ops.sigma_fwd(plan, params, rays=rays, z=z)
    s1, (e1, a1), g = ops.sigma_fwd_jvp(plan, params, rays=rays, z=z)
    torch.cuda.synchronize()
    print(f"T=2^{log2_T} n={n} s={s} plain-vs-plain equal: {torch.equal(s0, s0b)} {torch.equal(e0, e0b)} {torch.equal(a0, a0b)}")
    ds = (s0.reshape(-1) != s1.reshape(-1)).nonzero().reshape(-1)
    de = (e0 != e1).any(dim=1).nonzero().reshape(-1)
    da = (a0 != a1).nonzero().reshape(-1)
    print("  mismatching points: sigma", ds.numel(), "enc", de.numel(), "act", da.numel(), "of", n * s)
    for name, idx in (("sigma", ds), ("enc", de)):
        if idx.numel():
            tile = idx // 128
            rnd = tile // 148
            print("   ", name, "first", idx[:12].tolist(), "rows", (idx % 128)[:12].tolist(), "rounds", torch.unique(rnd)[:20].tolist(),
                  "n_tiles_hit", torch.unique(tile).numel())
            if name == "sigma":
                print("    max abs diff", (s0.reshape(-1)[idx] - s1.reshape(-1)[idx]).abs().max().item(),
                      "median", (s0.reshape(-1)[idx] - s1.reshape(-1)[idx]).abs().median().item())
            else:
                cols = (e0[idx[:2000]] != e1[idx[:2000]]).float().mean(dim=0)
                print("    enc cols hit freq", [round(c, 2) for c in cols.tolist()])
                print("    enc max abs diff", (e0 - e1).abs().max().item())
