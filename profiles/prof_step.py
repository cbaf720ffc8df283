"""Profiling driver: N mapping iterations of one 16384-ray x 512-sample chunk (BASELINE configs[1] chunk)
through the public Optimizer API.  Used under ncu (see profiles/README.md); prints per-kernel CUDA-event times.

    python profiles/prof_step.py [iters] [log2_T] [n_rays] [n_samples]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from loner_b200 import _lib, mapping as M, synth  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 3
T = int(sys.argv[2]) if len(sys.argv) > 2 else 19
n_rays = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
S = int(sys.argv[4]) if len(sys.argv) > 4 else 512
pose_grad = len(sys.argv) > 5 and sys.argv[5] == "pose"
dev = torch.device("cuda:0")
torch.manual_seed(0)
kfs, scale, shift = synth.make_keyframes(4, ray_range=(1.0, 50.0), seed=0)
s = M.default_settings(n_samples_train=S, log2_hashmap_size=T, lidar=n_rays // 4, sky=0, chunk=n_rays)
opt = M.Optimizer(s, None, M.WorldCube(scale, shift), dev)
win = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, None), k.pose, device=dev) for k in kfs]
for k in win:
    k.is_anchored = not pose_grad
    k.get_lidar_pose().set_fixed(not pose_grad)
adam = M.FusedAdam([{"params": opt._model.get_sigma_parameters(), "lr": 0.01}])
_lib.TIMING["on"] = True
for i in range(iters):
    rays, depths = opt.build_batch(win, ignore_world_cube=True)
    loss = opt.compute_loss(None, (rays, depths), i)
    loss.backward()
    adam.step()
    adam.zero_grad(set_to_none=True)
    if i % 10 == 0:
        opt._step_occupancy_grid()
torch.cuda.synchronize()
for k, (n, t) in sorted(_lib.kernel_times().items()):
    print(f"{k:22s} launches {n:3d}  avg {t / n:9.4f} ms")
print("loss", float(loss))
