"""Debug driver: tcgen05 MLP forward/backward against an fp64 torch evaluation on the same encodings."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from loner_b200 import ops

dev = torch.device("cuda:0")
P = int(sys.argv[1]) if len(sys.argv) > 1 else 128 * 300
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 4.0
torch.manual_seed(0)
plan = ops.Plan(log2_hashmap_size=14)
p = torch.cat([(torch.rand(3072) - 0.5) * 1.5, (torch.rand(plan.n_table_params) - 0.5) * scale]).to(dev)
x = torch.rand(P, 3, device=dev)
ds = torch.randn(P, device=dev)
W1 = p[:2048].view(64, 32).double(); w2 = p[2048:2112].double()
rel = lambda a, b: ((a.double() - b.double()).abs().max() / b.double().abs().max()).item()
for tc in (True, False):
    sigma, (enc, act) = ops.sigma_fwd(plan, p, x01=x, tensor_cores=tc)
    E = enc.clone().double()
    h = E @ W1.t()
    mask = (h > 0)
    s_ref = (torch.relu(h) * w2).sum(1)
    gp = torch.zeros_like(p)
    ops.sigma_bwd(plan, p, (enc, act), ds, x01=x, grad_params=gp)
    dE = enc
    u = ds.double()[:, None] * mask
    dE_ref = (u * w2) @ W1
    dW1_ref = (u * w2).t() @ E
    dW2_ref = (ds.double()[:, None] * torch.relu(h)).sum(0)
    print(f"tc={tc}: sigma {rel(sigma, s_ref):.2e}  dE {rel(dE, dE_ref):.2e}  dW1 {rel(gp[:2048].view(64,32), dW1_ref):.2e}  dW2 {rel(gp[2048:2112], dW2_ref):.2e}")
    if tc:
        bits = torch.stack([(act >> j) & 1 for j in range(64)], 1).bool()
        print("   mask mismatches:", (bits != mask).sum().item(), "of", mask.numel())
        err = (dE.double() - dE_ref).abs()
        print("   dE err by column:", (err.max(0).values / dE_ref.abs().max()).cpu().numpy().round(6))
        rows = err.max(1).values.view(-1, 128) if P % 128 == 0 else None
        if rows is not None:
            print("   dE err by row-in-tile (max over tiles), first 16:", (rows.max(0).values[:16] / dE_ref.abs().max()).cpu().numpy().round(6))
        e1 = ((gp[:2048].view(64, 32).double() - dW1_ref).abs() / dW1_ref.abs().max())
        print("   dW1 err by j (max over k) first 16:", e1.max(1).values[:16].cpu().numpy().round(6))
        print("   dW1 err by k (max over j):", e1.max(0).values.cpu().numpy().round(6))
