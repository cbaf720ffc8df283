"""Where does the ray-gradient error at the benchmark shape come from?  Runs the gradient chain of
tests/test_reference_chain_gpu.py::test_c2_shape_gradient_chain_on_reference_depths and prints error statistics of
d loss / d rays against the reference's recorded gradient (env switches select the MLP path)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import load_golden, c2_shape_random, c2_shape_params
from test_reference_chain_gpu import make_optimizer, ReplaySampler, replay_randn, DEV

g = load_golden("c2_shape_ogm_l1js")
for sc in (2000, 1):
    st = g["states"][sc]
    opt, M = make_optimizer(g, c2_shape_params(g, sc), chunk=g["chunk"])
    opt._ray_sampler = ReplaySampler(g["z"].to(DEV))
    rays = g["rays"].to(DEV).requires_grad_(True)
    with replay_randn([r[2] for r in c2_shape_random(g)]):
        loss = opt.compute_loss(None, (rays, g["depths"].to(DEV)), 0)
    loss.backward()
    a, b = rays.grad.double().cpu(), st["rays_grad"].double()
    err = (a - b).abs()
    print(f"x{sc}: loss {loss.item():.6f} ref {st['loss'].item():.6f}; |ref|max {b.abs().max():.4e}; err max {err.max():.4e} rel {err.max() / b.abs().max():.3e}")
    for c in range(6):
        print(f"   col {c}: ref max {b[:, c].abs().max():.3e} err max {err[:, c].max():.3e} rel-to-col {err[:, c].max() / b[:, c].abs().max():.3e} "
              f"median rel per-ray {(err[:, c] / b[:, c].abs().clamp_min(1e-12)).median():.3e}")
    worst = err.max(1).values.topk(5).indices
    for r in worst.tolist():
        print("   ray", r, "ours", a[r, :6].tolist(), "ref", b[r, :6].tolist())
