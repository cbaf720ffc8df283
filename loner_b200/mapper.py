"""Mapper / KeyFrameManager glue and the checkpoint format.

Host-side mirror of
    KeyFrameManager      src/mapping/keyframe_manager.py:35-200  (keyframe selection TEMPORAL / MOTION / HYBRID /
                         HYBRID_LAZY, tracked-pose hand-over :84-95, window selection MOST_RECENT / RANDOM / HYBRID
                         :164-187, get_poses_state :190-194)
    Mapper.update / build_ckpt / finish   src/mapping/mapper.py:75-180  (optimise on every accepted keyframe, write
                         checkpoints/ckpt_<kf>.tar and checkpoints/final.tar)
    checkpoint loading   examples/fdt_optimize_implicit_map.py:343-359, analysis/renderer.py:160-170
without the ROS / multiprocessing plumbing around them (frame slots, signals: out of scope).  Checkpoints carry the
reference's keys -- 'global_step', 'network_state_dict' (nerf_model._model_sigma.params, ._pos_encoding.params,
._dir_encoding.params, ._model_intensity.params: the flat fp32 tcnn vectors), 'optimizer_state_dict', 'poses',
'occ_model_state_dict' (occupancy_grid), 'occ_optimizer_state_dict' -- so an `outputs/.../checkpoints/*.tar`
written by either implementation loads into the other.
"""
import os
from dataclasses import dataclass
from typing import List, Optional

import torch

from ._lib import LonerB200Error
from .mapping import KeyFrame, LidarScanData, Optimizer, axis_angle_to_matrix


# ------------------------------------------------------------------------------------------
# pose algebra on [t, axis-angle] 6-vectors (reference src/common/pose.py, pose_utils.py:321-368)
# ------------------------------------------------------------------------------------------

def pose6_to_matrix(pose6):
    m = torch.eye(4, dtype=torch.float32, device=pose6.device)
    m[:3, :3] = axis_angle_to_matrix(pose6[3:].float())
    m[:3, 3] = pose6[:3]
    return m


def matrix_to_axis_angle(R):
    """pytorch3d.transforms.matrix_to_axis_angle (matrix -> quaternion -> axis-angle, rotation angle in [0, pi])."""
    m = R.double()
    t = m[0, 0] + m[1, 1] + m[2, 2]
    cand = torch.stack([
        torch.stack([1 + t, m[2, 1] - m[1, 2], m[0, 2] - m[2, 0], m[1, 0] - m[0, 1]]),
        torch.stack([m[2, 1] - m[1, 2], 1 + m[0, 0] - m[1, 1] - m[2, 2], m[0, 1] + m[1, 0], m[0, 2] + m[2, 0]]),
        torch.stack([m[0, 2] - m[2, 0], m[0, 1] + m[1, 0], 1 - m[0, 0] + m[1, 1] - m[2, 2], m[1, 2] + m[2, 1]]),
        torch.stack([m[1, 0] - m[0, 1], m[0, 2] + m[2, 0], m[1, 2] + m[2, 1], 1 - m[0, 0] - m[1, 1] + m[2, 2]])])
    q = cand[int(torch.argmax(torch.stack([cand[i, i] for i in range(4)])))]
    q = q / q.norm()
    if q[0] < 0:
        q = -q
    n = q[1:].norm()
    half = torch.atan2(n, q[0])
    scale = torch.where(n < 1e-12, torch.tensor(2.0, dtype=torch.float64, device=R.device), 2 * half / n.clamp_min(1e-300))
    return (q[1:] * scale).float()


def matrix_to_pose6(T):
    return torch.cat([T[:3, 3].float(), matrix_to_axis_angle(T[:3, :3])])


def default_mapper_settings(log_directory, optimizer_settings=None, **optimizer_kwargs):
    """The `mapper:` block of reference cfg/defaults.yaml:37-97 (plus the debug flags / log directory Mapper reads)."""
    from .mapping import Settings, default_settings
    return Settings(dict(
        log_level="DISABLED", log_directory=log_directory,
        debug=dict(use_groundtruth_poses=False, log_times=False, pytorch_detect_anomaly=False),
        keyframe_manager=dict(
            keyframe_selection=dict(strategy="TEMPORAL", temporal=dict(time_diff_seconds=3),
                                    motion=dict(translation_threshold_m=0.5, rotation_threshold_deg=22.5)),
            window_selection=dict(strategy="HYBRID", hybrid_settings=dict(num_recent_frames=1), window_size=8)),
        optimizer=optimizer_settings if optimizer_settings is not None else default_settings(**optimizer_kwargs)))


@dataclass
class Frame:
    """The slice of reference src/common/frame.py the mapper reads: one LiDAR scan with its tracked pose (6-vector
    [t, axis-angle], start-of-scan frame), optional ground truth, image, mask and lidar-to-camera extrinsic."""
    lidar_points: LidarScanData
    lidar_pose: torch.Tensor
    time: float = 0.0
    gt_lidar_pose: Optional[torch.Tensor] = None
    image: Optional[torch.Tensor] = None
    mask: Optional[torch.Tensor] = None
    lidar_to_camera: Optional[torch.Tensor] = None

    def get_time(self):
        return self.time


class KeyFrameManager:
    """reference src/mapping/keyframe_manager.py."""

    KEYFRAME_STRATEGIES = ("TEMPORAL", "MOTION", "HYBRID", "HYBRID_LAZY")
    WINDOW_STRATEGIES = ("MOST_RECENT", "RANDOM", "HYBRID")

    def __init__(self, settings, device="cuda"):
        self._settings = settings
        self._keyframe_selection_strategy = settings.keyframe_selection.strategy
        self._window_selection_strategy = settings.window_selection.strategy
        if self._keyframe_selection_strategy not in self.KEYFRAME_STRATEGIES:
            raise ValueError(f"Can't use unknown KeyFrameSelectionStrategy {self._keyframe_selection_strategy}")
        if self._window_selection_strategy not in self.WINDOW_STRATEGIES:
            raise ValueError(f"Can't use unknown WindowSelectionStrategy {self._window_selection_strategy}")
        self._device = device
        self._last_accepted_frame_ts = None
        self._last_motion_rejected_frame_ts = None
        self._keyframes: List[KeyFrame] = []
        self._global_step = 0

    def __len__(self):
        return len(self._keyframes)

    def process_frame(self, frame: Frame):
        """keyframe_manager.py:67-123: returns the new (or, in HYBRID, the re-processed last) KeyFrame, or None."""
        strat = self._keyframe_selection_strategy
        temporal_met = True
        if strat == "TEMPORAL":
            use = self._select_frame_temporal(frame)
        else:
            motion_met = self._select_frame_motion(frame)
            temporal_met = self._select_frame_temporal(frame)
            if temporal_met and not motion_met:
                self._last_motion_rejected_frame_ts = frame.get_time()
            use = motion_met if strat == "MOTION" else (motion_met and temporal_met)
        new_kf = None
        if use:
            self._last_accepted_frame_ts = frame.get_time()
            new_kf = KeyFrame(frame.lidar_points, frame.lidar_pose, gt_pose6=frame.gt_lidar_pose, time=frame.get_time(),
                              device=self._device, image=frame.image, mask=frame.mask, lidar_to_camera=frame.lidar_to_camera)
            new_kf._tracked_pose6 = frame.lidar_pose.detach().clone().float().cpu()
            if self._keyframes:
                # hand the optimised correction of the previous keyframe over to the new one (:84-95)
                ref = self._keyframes[-1]
                T_ref_tracked = pose6_to_matrix(ref._tracked_pose6)
                T_cur_tracked = pose6_to_matrix(new_kf._tracked_pose6)
                T_track = torch.linalg.inv(T_ref_tracked) @ T_cur_tracked
                T_opt = pose6_to_matrix(ref.get_lidar_pose().get_pose_tensor().detach().cpu()) @ T_track
                new_kf.get_lidar_pose().get_pose_tensor().data.copy_(matrix_to_pose6(T_opt))
            self._keyframes.append(new_kf)
        if strat == "HYBRID":
            if temporal_met:
                self._last_accepted_frame_ts = frame.get_time()
            return self._keyframes[-1] if temporal_met else None
        return new_kf

    def get_last_mapped_time(self):
        if self._keyframe_selection_strategy in ("HYBRID_LAZY", "MOTION") and self._last_motion_rejected_frame_ts is not None:
            return max(self._last_motion_rejected_frame_ts, self._last_accepted_frame_ts)
        return self._last_accepted_frame_ts

    def _select_frame_temporal(self, frame):
        if not self._keyframes:
            return True
        return frame.get_time() - self._last_accepted_frame_ts >= self._settings.keyframe_selection.temporal.time_diff_seconds

    def _select_frame_motion(self, frame):
        if not self._keyframes:
            return True
        ref = pose6_to_matrix(self._keyframes[-1].get_lidar_pose().get_pose_tensor().detach().cpu())
        rel = torch.linalg.inv(ref) @ pose6_to_matrix(frame.lidar_pose.detach().cpu().float())
        dT = rel[:3, 3].norm()
        dR = torch.rad2deg(matrix_to_axis_angle(rel[:3, :3])).norm()
        m = self._settings.keyframe_selection.motion
        return bool(dT >= m.translation_threshold_m or dR >= m.rotation_threshold_deg)

    def get_keyframes(self, idxs=None):
        return self._keyframes if idxs is None else [self._keyframes[i] for i in idxs]

    def get_active_window(self) -> List[KeyFrame]:
        """keyframe_manager.py:164-187."""
        window_size = self._settings.window_selection.window_size
        if self._window_selection_strategy == "MOST_RECENT":
            return self._keyframes[-window_size:]
        if self._window_selection_strategy == "RANDOM":
            n_recent = 1
        else:
            n_recent = self._settings.window_selection.hybrid_settings.num_recent_frames
        n_recent = min(n_recent, len(self._keyframes), window_size)
        indices = torch.randperm(len(self._keyframes) - n_recent)[:window_size - n_recent].tolist()
        indices += list(range(-n_recent, 0))           # the recent frames come last (the reference relies on it)
        return [self._keyframes[i] for i in indices]

    def get_poses_state(self) -> List[dict]:
        """keyframe.py:130-139 per keyframe."""
        out = []
        for kf in self._keyframes:
            l2c = kf._lidar_to_camera
            if l2c is None:
                l2c6 = torch.zeros(6)
            else:
                l2c6 = (matrix_to_pose6(l2c.cpu()) if l2c.shape[-2:] == (4, 4) else l2c.detach().cpu().clone())
            out.append({"timestamp": torch.as_tensor(kf.get_time()).detach().cpu().clone(),
                        "lidar_to_camera": l2c6,
                        "lidar_pose": kf.get_lidar_pose().get_pose_tensor().detach().cpu().clone(),
                        "gt_lidar_pose": kf._gt_pose.get_pose_tensor().detach().cpu().clone(),
                        "tracked_pose": kf._tracked_pose6.clone()})
        return out


class Mapper:
    """reference src/mapping/mapper.py without the inter-process slots/signals: feed frames with `process_frame`."""

    def __init__(self, settings, calibration, world_cube, device="cuda", enable_sky_segmentation=True, lidar_only=True,
                 data_parallel=None):
        self._settings = settings
        self._world_cube = world_cube
        self._keyframe_manager = KeyFrameManager(settings.keyframe_manager, device)
        self._optimizer = Optimizer(settings.optimizer, calibration, world_cube, device,
                                    bool(settings.debug.use_groundtruth_poses), lidar_only, enable_sky_segmentation,
                                    data_parallel=data_parallel)
        self._ckpt_dir = os.path.join(settings.log_directory, "checkpoints")
        os.makedirs(self._ckpt_dir, exist_ok=True)
        self.last_ckpt = {}

    def process_frame(self, frame: Frame) -> bool:
        """mapper.py:75-137 for one frame: keyframe decision, window optimisation, checkpoint.  Returns whether the
        frame was mapped."""
        kf = self._keyframe_manager.process_frame(frame)
        if kf is None:
            return False
        if not self._settings.optimizer.enabled:
            if self._optimizer._global_step % 100 == 0:
                torch.save({'poses': self._keyframe_manager.get_poses_state()},
                           os.path.join(self._ckpt_dir, f"ckpt_{self._optimizer._keyframe_count}.tar"))
            self._optimizer._global_step += 1
            return False
        window = self._keyframe_manager.get_active_window()
        self._optimizer.iterate_optimizer(window)
        kf_idx = self._optimizer._keyframe_count - 1
        level = self._settings.log_level
        level = level[0] if isinstance(level, tuple) else level
        if (kf_idx % 10 == 0 and level == "STANDARD") or level == "VERBOSE":
            ckpt = self.build_ckpt()
        else:
            ckpt = {'global_step': self._optimizer._global_step, 'poses': self._keyframe_manager.get_poses_state()}
        torch.save(ckpt, os.path.join(self._ckpt_dir, f"ckpt_{kf_idx}.tar"))
        return True

    def build_ckpt(self) -> dict:
        """mapper.py:161-175."""
        opt = self._optimizer
        ckpt = {'global_step': opt._global_step,
                'network_state_dict': opt._model.state_dict(),
                'optimizer_state_dict': opt._optimizer.state_dict() if opt._optimizer is not None else None,
                'poses': self._keyframe_manager.get_poses_state()}
        if self._settings.optimizer.samples_selection.strategy == 'OGM':
            ckpt['occ_model_state_dict'] = opt._occupancy_grid_model.state_dict()
            # the occupancy update is a fused SGD kernel; the entry has torch.optim.SGD's (stateless) layout
            ckpt['occ_optimizer_state_dict'] = torch.optim.SGD(
                [opt._occupancy_grid], lr=opt._model_config.model.occ_model.lr).state_dict()
        return ckpt

    def finish(self) -> str:
        path = os.path.join(self._ckpt_dir, "final.tar")
        torch.save(self.build_ckpt(), path)
        return path


def load_checkpoint(optimizer: Optimizer, ckpt, strict=True):
    """examples/fdt_optimize_implicit_map.py:343-359 / analysis/renderer.py:160-170: restore the networks, the
    occupancy grid (and the sampler's copy of it) and the step counter from a checkpoint dict or path."""
    if isinstance(ckpt, (str, os.PathLike)):
        ckpt = torch.load(ckpt, map_location="cpu", weights_only=False)
    if 'network_state_dict' not in ckpt:
        raise LonerB200Error("checkpoint has no 'network_state_dict' (pose-only checkpoint: see Mapper.process_frame)")
    optimizer._model.load_state_dict(ckpt['network_state_dict'], strict=strict)
    if 'occ_model_state_dict' in ckpt and getattr(optimizer, "_occupancy_grid_model", None) is not None:
        optimizer._occupancy_grid_model.load_state_dict(ckpt['occ_model_state_dict'])
        optimizer._occupancy_grid = optimizer._occupancy_grid_model()
        optimizer._ray_sampler.update_occ_grid(optimizer._occupancy_grid.detach())
    if 'global_step' in ckpt:
        optimizer._global_step = int(ckpt['global_step'])
    return ckpt
