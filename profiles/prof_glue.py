"""Which PyTorch (non-library) kernels run in one bench step, by name / count / device time (torch.profiler).
    python profiles/prof_glue.py [c2|c3]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from loner_b200 import mapping as M

W = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "c2"]; bench.WORKLOAD = W
dev = torch.device("cuda:0")
torch.manual_seed(1234)
kfs, scale, shift = bench.make_scene(W["keyframes"], seed=0)
settings = M.default_settings(n_samples_train=W["n_samples"], log2_hashmap_size=W["log2_hashmap_size"], ray_range=W["ray_range"],
                              lidar=W["rays_per_keyframe"], sky=0, chunk=W["chunk"], voxel_size=W["voxel_size"])
opt = M.Optimizer(settings, None, M.WorldCube(scale, shift), dev)
window = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=dev) for k in kfs]
for i, kf in enumerate(window):
    free = bool(W["pose_grad"]) and i > 0
    kf.is_anchored = not free
    kf.get_lidar_pose().set_fixed(not free)
sigma_params = opt._model.get_sigma_parameters()
groups = [{"params": sigma_params, "lr": 0.01}]
poses = [kf.get_lidar_pose().get_pose_tensor() for kf in window if not kf.is_anchored]
if poses:
    groups.append({"params": poses, "lr": 0.001})
adam = M.FusedAdam(groups)


def step(i):
    rays, depths = opt.build_batch(window)
    loss = opt.compute_loss(None, (rays, depths), i)
    loss.backward()
    adam.step()
    adam.zero_grad(set_to_none=True)


for i in range(1, 6):
    step(i)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(6, 9):
        step(i)
    torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
rows.sort(key=lambda r: -r[2])
tot = sum(r[2] for r in rows)
print(f"3 steps: {tot / 3e3:.3f} ms of device time per step")
print(f"launches per step: {sum(r[1] for r in rows) / 3:.0f}")
for k, c, t in rows[:45]:
    print(f"{c / 3:7.1f} /step {t / 3:9.1f} us/step  {k[:110]}")
print()
print(prof.key_averages().table(sort_by="cpu_time_total", row_limit=12, max_name_column_width=60))
