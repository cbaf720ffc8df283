"""ORACLE / TEST INFRASTRUCTURE ONLY.  Container-only script (needs /root/reference).

Golden fixture for the colour branch: runs the reference's OWN Model.forward(camera=True) and
Optimizer.compute_loss_camera (src/mapping/optimizer.py:861-894), unmodified, on torch-CPU behind the fp32
`tinycudann` shim, then loss.backward().  Records inputs (rays, intensities, the uniform / normal draws), the
sample depths, per-sample raw output [N,S,4], composited rgb, loss, and the gradients of the intensity MLP,
the intensity hash table (top entries) and the rays.

    python -m oracle.make_golden_rgb        # writes tests/golden/rgb_camera.pt
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.reference_import import import_reference  # noqa: E402
from oracle import make_golden as MG  # noqa: E402
from loner_b200 import synth  # noqa: E402


def main():
    ns = import_reference()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    seed, n_samples, log2_T, n_rays, ray_range = 3, 64, 14, 192, (1.0, 30.0)
    torch.manual_seed(seed)
    gen = torch.Generator().manual_seed(seed)
    kfs, scale, shift = synth.make_keyframes(2, ray_range=ray_range, seed=seed)
    cfg = MG.model_cfg(ns, n_samples, log2_T, 32, ray_range)
    opt = MG.bare_optimizer(ns, cfg, scale, "UNIFORM")
    params0 = MG.set_params(opt, log2_T)
    nerf = opt._model.nerf_model
    with torch.no_grad():                       # "trained-like" intensity features
        nerf._pos_encoding.params.mul_(MG.TABLE_SCALE)
    rays, _, pose, idx = MG.reference_rays(ns, kfs[1], scale, shift, ray_range, n_rays, 0, gen, True)
    rays = rays.float()
    rays.retain_grad()
    intensities = torch.rand(rays.shape[0], 3, generator=gen)
    for p in opt._model.parameters():
        p.grad = None
    with MG.capture_torch() as log:
        loss = opt.compute_loss_camera((rays, intensities), None, 0)
    loss.backward()
    res = opt._results_camera
    g_mlp = nerf._model_intensity.params.grad.detach()
    g_tab = nerf._pos_encoding.params.grad.detach()
    top = torch.topk(g_tab.abs(), 2048).indices.sort().values
    out = dict(name="rgb_camera", n_samples=n_samples, log2_T=log2_T, log2_T_int=14, ray_range=list(ray_range),
               table_scale=MG.TABLE_SCALE, param_seed=1337, scale=scale.clone(), shift=shift.clone(),
               params_checksum=params0.double().sum().item(),
               rays=rays.detach().clone(), intensities=intensities.clone(),
               u1=log["rand"][0], noise=log["randn"][0],
               z=res["samples_fine"].detach().clone(), raw=res["raw_fine"].detach().clone(),
               rgb=res["rgb_fine"].detach().clone(), depth=res["depth_fine"].detach().clone(),
               weights=res["weights_fine"].detach().clone(), loss=loss.detach().clone(),
               mlp_int_grad=g_mlp.clone(), table_int_grad_idx=top, table_int_grad_val=g_tab[top].clone(),
               table_int_grad_norm=g_tab.double().norm().item(), rays_grad=rays.grad.detach().clone(),
               sigma_grad_is_none=nerf._model_sigma.params.grad is None,
               mlp_int_checksum=nerf._model_intensity.params.detach().double().sum().item(),
               table_int_checksum=nerf._pos_encoding.params.detach().double().sum().item())
    path = os.path.join(MG.GOLDEN_DIR, "rgb_camera.pt")
    torch.save(out, path)
    print(f"rgb_camera: N={rays.shape[0]} S={n_samples} loss={float(loss):.6f} -> {path} "
          f"({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    main()
