import os, sys
sys.path.insert(0, "/root/repo")
import torch
from loner_b200 import ops
dev = torch.device("cuda:0")
torch.manual_seed(0)
P = 16384 * 512
plan = ops.Plan(log2_hashmap_size=22)
x01 = torch.rand(P, 3, device=dev)
table = (torch.rand(plan.n_table_params, device=dev) - 0.5) * 0.4
d_enc = torch.randn(P, 32, device=dev) * 1e-3
gt = torch.zeros_like(table)
for _ in range(2):
    enc = ops.encode_fwd(plan, x01, table)
    ops.encode_bwd(plan, x01, table, d_enc, gt, False)
torch.cuda.synchronize()
