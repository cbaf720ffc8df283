"""Times lb_sigma_fwd (encode + tcgen05 MLP) for one 8.39 M-point chunk; used with the kernel's debug switches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from loner_b200 import ops
dev = torch.device("cuda:0")
N, S = 16384, 512
torch.manual_seed(0)
plan = ops.Plan(log2_hashmap_size=19)
params = torch.cat([(torch.rand(3072) - 0.5) * 0.5, (torch.rand(plan.n_table_params) - 0.5) * 0.4]).to(dev)
x01 = torch.rand(N * S, 3, device=dev)
def run(tc):
    return ops.sigma_fwd(plan, params, x01=x01, tensor_cores=tc)
for tc in (True, False):
    run(tc); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run(tc)
    e1.record(); torch.cuda.synchronize()
    print("dbg", os.environ.get("LONER_B200_MLP_DBG", "0"), "tensor_cores", tc, f"{e0.elapsed_time(e1) / 5:.4f} ms")
