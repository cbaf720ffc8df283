"""Mapper / KeyFrameManager glue and the checkpoint format.  Host logic on CPU; the optimising
Mapper round trip on the GPU."""
import os

import pytest
import torch

from loner_b200 import mapper as MP, mapping as M


def _frame(t, pose6, n=16):
    d = torch.randn(3, n)
    return MP.Frame(M.LidarScanData(d / d.norm(dim=0), torch.rand(n) * 10 + 1, None), pose6.float(), time=float(t))


def _settings(tmp, kf="TEMPORAL", win="HYBRID", **kw):
    s = MP.default_mapper_settings(str(tmp), **kw)
    s["keyframe_manager"]["keyframe_selection"]["strategy"] = kf
    s["keyframe_manager"]["window_selection"]["strategy"] = win
    return s


def test_pose_algebra_round_trip():
    g = torch.Generator().manual_seed(0)
    for _ in range(50):
        aa = torch.randn(3, generator=g)
        aa = aa / aa.norm() * torch.rand(1, generator=g) * 3.1
        p = torch.cat([torch.randn(3, generator=g), aa])
        back = MP.matrix_to_pose6(MP.pose6_to_matrix(p))
        assert (back - p).abs().max() < 2e-5
    assert MP.matrix_to_axis_angle(torch.eye(3)).abs().max() == 0


def test_keyframe_selection_strategies(tmp_path):
    still, moved = torch.zeros(6), torch.tensor([0.6, 0, 0, 0, 0, 0.0])
    # TEMPORAL: first frame always, then every >= 3 s
    km = MP.KeyFrameManager(_settings(tmp_path).keyframe_manager, "cpu")
    got = [km.process_frame(_frame(t, still)) is not None for t in (0.0, 1.0, 2.9, 3.0, 5.0, 6.0)]
    assert got == [True, False, False, True, False, True] and len(km) == 3 and km.get_last_mapped_time() == 6.0
    # MOTION: translation >= 0.5 m or rotation >= 22.5 deg relative to the last keyframe
    km = MP.KeyFrameManager(_settings(tmp_path, kf="MOTION").keyframe_manager, "cpu")
    rot = torch.tensor([0, 0, 0, 0, 0, 0.45])
    got = [km.process_frame(_frame(t, p)) is not None for t, p in ((0, still), (1, still), (2, moved), (3, moved), (4, moved + rot))]
    assert got == [True, False, True, False, True]
    # HYBRID_LAZY: both criteria; a time-ok / motion-rejected frame only moves the "last mapped" clock
    km = MP.KeyFrameManager(_settings(tmp_path, kf="HYBRID_LAZY").keyframe_manager, "cpu")
    got = [km.process_frame(_frame(t, p)) is not None for t, p in ((0, still), (1, moved), (4, still), (5, moved))]
    assert got == [True, False, False, True] and len(km) == 2
    # HYBRID: time-ok / motion-rejected re-processes the last keyframe
    km = MP.KeyFrameManager(_settings(tmp_path, kf="HYBRID").keyframe_manager, "cpu")
    k0 = km.process_frame(_frame(0, still))
    assert km.process_frame(_frame(1, moved)) is None
    assert km.process_frame(_frame(3, still)) is k0 and len(km) == 1
    k1 = km.process_frame(_frame(6, moved))
    assert k1 is not k0 and len(km) == 2
    with pytest.raises(ValueError):
        MP.KeyFrameManager(_settings(tmp_path, kf="NOPE").keyframe_manager, "cpu")


def test_tracked_pose_hand_over_and_pose_state(tmp_path):
    """keyframe_manager.py:84-95: the new keyframe starts from the previous keyframe's OPTIMISED pose composed with
    the tracked relative motion."""
    km = MP.KeyFrameManager(_settings(tmp_path).keyframe_manager, "cpu")
    p0 = torch.tensor([1.0, 2.0, 0.5, 0.0, 0.0, 0.2])
    k0 = km.process_frame(_frame(0, p0))
    corr = torch.tensor([0.1, -0.05, 0.02, 0.01, -0.02, 0.03])
    k0.get_lidar_pose().get_pose_tensor().data.add_(corr)                     # "optimisation" moved keyframe 0
    p1 = torch.tensor([2.0, 2.5, 0.5, 0.0, 0.0, 0.5])
    k1 = km.process_frame(_frame(3, p1))
    T_track = torch.linalg.inv(MP.pose6_to_matrix(p0)) @ MP.pose6_to_matrix(p1)
    want = MP.pose6_to_matrix(p0 + corr) @ T_track
    assert (MP.pose6_to_matrix(k1.get_lidar_pose().get_pose_tensor().detach()) - want).abs().max() < 1e-5
    st = km.get_poses_state()
    assert len(st) == 2 and set(st[0]) == {"timestamp", "lidar_to_camera", "lidar_pose", "gt_lidar_pose", "tracked_pose"}
    assert torch.equal(st[1]["tracked_pose"], p1) and torch.allclose(st[0]["lidar_pose"], p0 + corr)
    assert float(st[1]["timestamp"]) == 3.0 and st[0]["lidar_to_camera"].shape == (6,)


def test_window_selection(tmp_path):
    torch.manual_seed(0)
    for win in ("MOST_RECENT", "RANDOM", "HYBRID"):
        s = _settings(tmp_path, win=win)
        s["keyframe_manager"]["window_selection"]["window_size"] = 4
        s["keyframe_manager"]["window_selection"]["hybrid_settings"]["num_recent_frames"] = 2
        km = MP.KeyFrameManager(s.keyframe_manager, "cpu")
        kfs = [km.process_frame(_frame(3 * i, torch.zeros(6))) for i in range(9)]
        w = km.get_active_window()
        assert len(w) == 4 and len(set(map(id, w))) == 4
        if win == "MOST_RECENT":
            assert w == kfs[-4:]
        elif win == "RANDOM":
            assert w[-1] is kfs[-1] and all(k in kfs[:-1] for k in w[:-1])
        else:
            assert w[-2:] == kfs[-2:] and all(k in kfs[:-2] for k in w[:-2])
    km = MP.KeyFrameManager(_settings(tmp_path).keyframe_manager, "cpu")
    km.process_frame(_frame(0, torch.zeros(6)))
    assert len(km.get_active_window()) == 1


@pytest.mark.gpu
def test_mapper_checkpoints_round_trip(tmp_path):
    """Mapper.process_frame -> iterate_optimizer -> ckpt_<kf>.tar / final.tar with the reference's keys; a second
    optimiser restored from final.tar renders the same depths; tcnn-named parameters load by name."""
    from loner_b200 import eval as E, synth
    dev = "cuda:0"
    kfs, scale, shift = synth.make_keyframes(3, ray_range=(1.0, 30.0), seed=2)
    s = _settings(tmp_path, n_samples_train=64, log2_hashmap_size=14, lidar=256, sky=0, voxel_size=32, num_iterations=12,
                  n_samples_test=64)
    s["keyframe_manager"]["keyframe_selection"]["temporal"]["time_diff_seconds"] = 1
    s["log_level"] = "VERBOSE"
    mp = MP.Mapper(s, None, M.WorldCube(scale, shift), dev)
    mapped = [mp.process_frame(MP.Frame(M.LidarScanData(k.directions, k.distances, None), k.pose, time=t))
              for t, k in zip((0.0, 0.5, 1.0, 2.0), (kfs[0], kfs[1], kfs[1], kfs[2]))]
    assert mapped == [True, False, True, True]
    assert sorted(os.listdir(tmp_path / "checkpoints")) == ["ckpt_0.tar", "ckpt_1.tar", "ckpt_2.tar"]
    final = mp.finish()
    ck = torch.load(final, map_location="cpu", weights_only=False)
    assert set(ck) == {"global_step", "network_state_dict", "optimizer_state_dict", "poses", "occ_model_state_dict",
                       "occ_optimizer_state_dict"}
    assert set(ck["network_state_dict"]) == {"nerf_model._model_sigma.params", "nerf_model._pos_encoding.params",
                                             "nerf_model._dir_encoding.params", "nerf_model._model_intensity.params"}
    assert set(ck["occ_model_state_dict"]) == {"occupancy_grid"} and ck["occ_model_state_dict"]["occupancy_grid"].shape == (1, 1, 32, 32, 32)
    assert set(ck["optimizer_state_dict"]) == {"state", "param_groups"} and len(ck["poses"]) == 3
    st = ck["optimizer_state_dict"]["state"][0]
    assert set(st) == {"step", "exp_avg", "exp_avg_sq"}                       # torch.optim.Adam's layout
    assert ck["global_step"] == mp._optimizer._global_step
    s["log_level"] = "DISABLED"
    # restore into a fresh optimiser and compare a render
    opt2 = M.Optimizer(s.optimizer, None, M.WorldCube(scale, shift), dev)
    MP.load_checkpoint(opt2, final)
    assert opt2._global_step == ck["global_step"]
    rays, _ = mp._keyframe_manager.get_keyframes()[0].build_lidar_rays(torch.arange(500, device=dev), (1.0, 30.0),
                                                                       mp._optimizer._world_cube, ignore_world_cube=True)
    torch.manual_seed(1)
    d1, _ = E.render_lidar_depth(mp._optimizer._model, mp._optimizer._ray_sampler, rays.detach(), scale)
    torch.manual_seed(1)
    d2, _ = E.render_lidar_depth(opt2._model, opt2._ray_sampler, rays.detach(), scale)
    assert torch.equal(d1, d2)
    with pytest.raises(Exception):
        MP.load_checkpoint(opt2, str(tmp_path / "checkpoints" / "nope.tar"))


def test_chunked_results_concatenate_on_first_read():
    """Optimizer._results_lidar keeps the reference's dict contract (optimizer.py:730-736) while concatenating the
    per-chunk pieces lazily."""
    import torch
    from loner_b200.mapping import _ChunkedResults
    chunks = [(torch.ones(2, 3), torch.zeros(2)), (2 * torch.ones(1, 3), torch.ones(1))]
    res = _ChunkedResults(chunks, {"samples_fine": 0, "depth_fine": 1}, rays=torch.arange(3))
    assert "samples_fine" in res and "rays" in res and "nope" not in res
    assert set(res.keys()) == {"rays", "samples_fine", "depth_fine"}
    z = res["samples_fine"]
    assert z.shape == (3, 3) and z[2, 0] == 2 and res["samples_fine"] is z       # cached after the first read
    assert res["depth_fine"].tolist() == [0.0, 0.0, 1.0]
    single = _ChunkedResults(chunks[:1], {"samples_fine": 0})
    assert single["samples_fine"] is chunks[0][0]                                # one chunk: no copy
    try:
        res["nope"]
        assert False
    except KeyError:
        pass
