import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from conftest import load_golden, golden_params
from loner_b200 import ops
from oracle import loner_oracle as O
dev = torch.device("cuda:0")
for name in ["c1_uniform_l2los", "c1_ogm_l1js"]:
    g = load_golden(name); p = golden_params(g)
    plan = ops.Plan(log2_hashmap_size=g["log2_T"])
    rays, z = g["rays"].to(dev), g["z"].to(dev)
    sigma, (enc, act) = ops.sigma_fwd(plan, p.to(dev), rays=rays, z=z)
    xyz = g["rays"][:, None, 0:3] + g["rays"][:, None, 3:6] * g["z"][..., None]
    e_ref = O.hash_encode((xyz.reshape(-1, 3) + 1) / 2, p[3072:], g["log2_T"])
    W1 = p[:2048].view(64, 32)
    h32 = e_ref @ W1.t()
    h64 = e_ref.double() @ W1.double().t()
    hk = enc.cpu().double() @ W1.double().t()     # from the kernel's own encodings
    bits = torch.stack([(act.cpu() >> j) & 1 for j in range(64)], 1).bool()
    scale = (e_ref.abs().double() @ W1.abs().double().t())
    for nm, h in (("fp32 oracle", h32), ("fp64 on oracle enc", h64), ("fp64 on kernel enc", hk)):
        mm = bits != (h > 0)
        print(name, nm, "mismatches", mm.sum().item(), "max |h|/scale at mismatch", (h.abs().double() / scale)[mm].max().item() if mm.any() else 0)
    print("  enc rel err", ((enc.cpu() - e_ref).abs().max() / e_ref.abs().max()).item(), " zero-h units:", (h32 == 0).sum().item(), "exact zero enc rows:", (e_ref.abs().sum(1) == 0).sum().item())
