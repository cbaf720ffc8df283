"""Profiling driver for the colour branch: N iterations of Optimizer.compute_loss_camera + backward + Adam on
8192 camera-shaped rays x 512 samples (4.2 M sample points), intensity grid T = 2^19.

    python profiles/prof_camera.py [iters] [n_rays] [n_samples]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from loner_b200 import _lib, mapping as M, synth  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n_rays = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
S = int(sys.argv[3]) if len(sys.argv) > 3 else 512
dev = torch.device("cuda:0")
torch.manual_seed(0)
kfs, scale, shift = synth.make_keyframes(4, ray_range=(1.0, 50.0), seed=0)
s = M.default_settings(n_samples_train=S, log2_hashmap_size=19, chunk=n_rays)
opt = M.Optimizer(s, None, M.WorldCube(scale, shift), dev, lidar_only=False)
with torch.no_grad():
    opt._model.nerf_model._model_sigma.params[3072:].mul_(2000.0)
    opt._model.nerf_model._pos_encoding.params.mul_(2000.0)
wc = M.WorldCube(scale.to(dev), shift.to(dev))
rays = []
for k in kfs:
    idx = torch.randint(k.distances.shape[0], (n_rays // 4,))
    r, _ = M.build_lidar_rays(k.directions[:, idx].to(dev), k.distances[idx].to(dev), k.pose.to(dev), wc,
                              torch.tensor((1.0, 50.0), device=dev), True)
    rays.append(r)
rays = torch.vstack(rays).contiguous()
target = torch.rand(rays.shape[0], 3, device=dev)
opt._model.freeze_sigma_head(True)
adam = M.FusedAdam([{"params": opt._model.get_rgb_parameters(), "lr": 0.01}])
_lib.TIMING["on"] = True
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(iters):
    if i == 1:
        e0.record()
    loss = opt.compute_loss_camera((rays, target), None, i)
    loss.backward()
    adam.step()
    adam.zero_grad(set_to_none=True)
e1.record()
torch.cuda.synchronize()
for k, (n, t) in sorted(_lib.kernel_times().items()):
    print(f"{k:24s} launches {n:3d}  avg {t / n:9.4f} ms")
if iters > 1:
    ms = e0.elapsed_time(e1) / (iters - 1)
    print(f"camera iteration: {ms:.3f} ms  {rays.shape[0] / ms / 1e3:.3f} M rays/s  loss {float(loss.detach()):.5f}")
