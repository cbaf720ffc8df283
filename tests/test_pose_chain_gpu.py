"""The fused pose -> (R, origin) kernels and the window-level ray builder against the PyTorch chain they replace
(loner_b200.mapping.pose_to_rot_origin = reference pose_utils.py:354-368 + ray_utils.py:293-296 under autograd; that
chain is pinned against the reference's own Pose / build_lidar_rays in tests/test_oracle.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def test_pose_to_rot_origin_matches_torch_chain():
    from loner_b200 import mapping as M, ops
    g = torch.Generator().manual_seed(3)
    K = 37
    pose = torch.randn(K, 6, generator=g)
    pose[:, 3:] *= torch.logspace(-9, 0.4, K).unsqueeze(1)       # from far below the Taylor switch (1e-6) to large angles
    pose[0, 3:] = 0.0                                            # exactly zero rotation
    pose = pose.to(DEV)
    wc = M.WorldCube(torch.tensor(12.5, device=DEV), torch.tensor([0.3, -0.1, 0.2], device=DEV))
    p_ref = pose.clone().requires_grad_(True)
    R_ref, o_ref = M.pose_to_rot_origin(p_ref, wc)
    rot, org = ops.pose_to_rot_origin(pose.contiguous(), wc.shift.float().contiguous(), 12.5)
    assert (rot.view(K, 3, 3) - R_ref).abs().max().item() < 2e-7
    assert torch.equal(org, o_ref.detach())
    gR = torch.randn(K, 3, 3, generator=g).to(DEV)
    go = torch.randn(K, 3, generator=g).to(DEV)
    (R_ref * gR).sum().backward(retain_graph=True)
    (o_ref * go).sum().backward()
    d_pose = ops.pose_to_rot_origin_bwd(pose.contiguous(), 12.5, gR.reshape(K, 9).contiguous(), go.contiguous())
    assert rel(d_pose[:, :3], p_ref.grad[:, :3]) < 1e-6
    assert rel(d_pose[1:, 3:], p_ref.grad[1:, 3:]) < 1e-5
    assert torch.isfinite(d_pose).all()


def test_window_ray_builder_matches_per_keyframe_path():
    """Optimizer.build_batch (BuildWindowRays: one node, fused pose kernel) against KeyFrame.build_lidar_rays per keyframe
    (PyTorch pose chain + BuildLidarRays), on the same index draws: identical rays, identical pose gradients."""
    from loner_b200 import mapping as M, synth
    kfs, scale, shift = synth.make_keyframes(3, ray_range=(1.0, 50.0), seed=4, scene=synth.Scene(half_extent=(70.0, 14.0)))
    s = M.default_settings(n_samples_train=64, log2_hashmap_size=14, ray_range=(1.0, 50.0), lidar=500, sky=32, voxel_size=32)
    opt = M.Optimizer(s, None, M.WorldCube(scale, shift), DEV)
    window = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=DEV) for k in kfs]
    for i, kf in enumerate(window):
        kf.is_anchored = i == 0
        kf.get_lidar_pose().set_fixed(i == 0)
    draws = []
    orig = opt._select_rays

    def recording(kf):
        r = orig(kf)
        draws.append(r)
        return r
    opt._select_rays = recording
    rays, depths = opt.build_batch(window)
    assert rays.shape == (3 * 532, 13) and depths.shape == (3 * 532,)
    w = torch.randn(rays.shape, device=DEV)
    (rays * w).sum().backward()
    grads = [kf.get_lidar_pose().get_pose_tensor().grad for kf in window]
    assert grads[0] is None and grads[1] is not None and grads[2] is not None
    # per-keyframe reference path on the recorded draws
    ref_r, ref_d, ref_p = [], [], []
    for kf, (idx, sky_idx), k in zip(window, draws, kfs):
        p = k.pose.to(DEV).clone().requires_grad_(True)
        kf2 = M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=DEV)
        kf2._pose._pose_tensor = p
        r, d = kf2.build_lidar_rays(idx, opt._ray_range, opt._world_cube, ignore_world_cube=True, sky_indices=sky_idx)
        ref_r.append(r); ref_d.append(d); ref_p.append(p)
    ref_rays, ref_depths = torch.vstack(ref_r), torch.cat(ref_d)
    assert (rays - ref_rays).abs().max().item() < 1e-6
    assert torch.equal(depths, ref_depths)
    (ref_rays * w).sum().backward()
    for gk, p in zip(grads[1:], ref_p[1:]):
        assert rel(gk, p.grad) < 1e-4
