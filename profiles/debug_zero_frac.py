"""Fraction of exactly-zero dL/dsigma entries in the bench workload (first steps of training)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from loner_b200 import mapping as M, synth, functional as Fn
dev = torch.device("cuda:0")
torch.manual_seed(0)
kfs, scale, shift = synth.make_keyframes(4, ray_range=(1.0, 50.0), seed=0)
s = M.default_settings(n_samples_train=512, log2_hashmap_size=19, lidar=4096, sky=0, chunk=16384)
opt = M.Optimizer(s, None, M.WorldCube(scale, shift), dev)
win = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, None), k.pose, device=dev) for k in kfs]
for k in win: k.is_anchored = True
adam = M.FusedAdam([{"params": opt._model.get_sigma_parameters(), "lr": 0.01}])
orig = Fn.ops.sigma_bwd
def spy(plan, params, saved, d_sigma, **kw):
    print("  d_sigma zero fraction: %.3f   |d_sigma| mean %.3e" % ((d_sigma == 0).float().mean().item(), d_sigma.abs().mean().item()))
    return orig(plan, params, saved, d_sigma, **kw)
Fn.ops.sigma_bwd = spy
for i in range(12):
    rays, depths = opt.build_batch(win, ignore_world_cube=True)
    loss = opt.compute_loss(None, (rays, depths), i)
    print("step", i, "loss", float(loss.detach()))
    loss.backward(); adam.step(); adam.zero_grad(set_to_none=True)
    if i % 10 == 0: opt._step_occupancy_grid()
