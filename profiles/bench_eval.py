"""BASELINE configs[4]: eval render of depth / variance (and colour) images, 1 M rays x 2048 samples
(reference N_samples_test), 'default' and 'adjusted' (peak) render strategies, through Model.forward(testing=True).

    python profiles/bench_eval.py [n_rays] [n_samples] [log2_T]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from loner_b200 import mapping as M, synth  # noqa: E402

n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
S = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
T = int(sys.argv[3]) if len(sys.argv) > 3 else 19
dev = torch.device("cuda:0")
torch.manual_seed(0)
kfs, scale, shift = synth.make_keyframes(4, ray_range=(1.0, 50.0), seed=0)
s = M.default_settings(n_samples_train=512, n_samples_test=S, log2_hashmap_size=T, chunk=16384)
opt = M.Optimizer(s, None, M.WorldCube(scale, shift), dev, lidar_only=False)
with torch.no_grad():   # trained-like magnitudes so that rays terminate
    opt._model.nerf_model._model_sigma.params[3072:].mul_(2000.0)
    opt._model.nerf_model._pos_encoding.params.mul_(2000.0)
    opt._occupancy_grid.normal_(0.0, 0.5)
rr = (1.0, 50.0)
wc = M.WorldCube(scale.to(dev), shift.to(dev))
rays = []
per = n_rays // len(kfs)
for k in kfs:
    idx = torch.randint(k.distances.shape[0], (per,))
    r, _ = M.build_lidar_rays(k.directions[:, idx].to(dev), k.distances[idx].to(dev), k.pose.to(dev), wc,
                              torch.tensor(rr, device=dev), True)
    rays.append(r)
rays = torch.vstack(rays).contiguous()
print(f"eval render: {rays.shape[0]} rays x {S} samples, T=2^{T}")
for camera in (False, True):
    for strategy in ("default", "adjusted"):
        def run():
            with torch.no_grad():
                return opt._model(rays, opt._ray_sampler, opt._scale_factor, testing=True, camera=camera,
                                  return_variance=True, render_strategy=strategy)
        opt._model.cfg["render"]["retraw"] = False
        run(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = run(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"camera={camera!s:5} strategy={strategy:8s} {ms:9.1f} ms  {rays.shape[0] / ms / 1e3:8.3f} M rays/s  "
              f"{rays.shape[0] * S / ms / 1e6:6.2f} G samples/s  depth mean {out['depth_fine'].mean().item():.4f} "
              f"peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
