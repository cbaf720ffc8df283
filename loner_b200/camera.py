"""Camera rays for the colour optimisation path.

Mirrors the reference's camera-side ray code:
    get_ray_directions            src/common/ray_utils.py:62-124
    CameraRayDirections           src/common/ray_utils.py:127-224   (build_rays, fetch_chunk_rays)
    KeyFrame.build_camera_rays    src/mapping/keyframe.py:108-128   (see mapping.KeyFrame)
The per-pixel direction table is computed once on the device with torch (H*W elements, a constructor cost);
the per-iteration work -- pixel gather, rotation, normalisation, cube exit, intensity gather -- is one launch of
lb_build_camera_rays.  There is no CPU path.

Lens distortion: the reference calls kornia.geometry.calibration.undistort_points (kornia is a dependency that
is absent from /root/reference and from this image).  `undistort_points` below restates kornia's published
algorithm (OpenCV's iterative undistortPoints, 5 fixed-point iterations, up to 14 coefficients without the tilt
model); parity for distorted intrinsics is therefore unpinned, the pinhole path (dist=None) is pinned against the
reference's own code (tests/golden/camera_rays.pt).
"""
import math

import torch

from . import ops
from ._lib import LonerB200Error


def undistort_points(points, K, dist, new_K=None, num_iters=5):
    """points [..., N, 2] distorted pixel coordinates -> undistorted pixel coordinates (kornia semantics).
    K, new_K [..., 3, 3]; dist [..., 4|5|8|12|14] = (k1, k2, p1, p2[, k3[, k4, k5, k6[, s1, s2, s3, s4[, tx, ty]]]])."""
    if dist.shape[-1] not in (4, 5, 8, 12, 14):
        raise LonerB200Error(f"undistort_points: invalid number of distortion coefficients {dist.shape[-1]}")
    if dist.shape[-1] < 14:
        dist = torch.nn.functional.pad(dist, [0, 14 - dist.shape[-1]])
    if bool((dist[..., 12:] != 0).any()):
        raise LonerB200Error("undistort_points: the tilted-sensor model (tx, ty) is not built")
    if new_K is None:
        new_K = K
    cx, cy, fx, fy = K[..., 0:1, 2], K[..., 1:2, 2], K[..., 0:1, 0], K[..., 1:2, 1]
    x = (points[..., 0] - cx) / fx
    y = (points[..., 1] - cy) / fy
    x0, y0 = x, y
    d = [dist[..., i:i + 1] for i in range(12)]
    for _ in range(num_iters):
        r2 = x * x + y * y
        inv_rad = (1 + d[5] * r2 + d[6] * r2 * r2 + d[7] * r2 ** 3) / (1 + d[0] * r2 + d[1] * r2 * r2 + d[4] * r2 ** 3)
        dx = 2 * d[2] * x * y + d[3] * (r2 + 2 * x * x) + d[8] * r2 + d[9] * r2 * r2
        dy = d[2] * (r2 + 2 * y * y) + 2 * d[3] * x * y + d[10] * r2 + d[11] * r2 * r2
        x = (x0 - dx) * inv_rad
        y = (y0 - dy) * inv_rad
    nfx, nfy, ncx, ncy = new_K[..., 0:1, 0], new_K[..., 1:2, 1], new_K[..., 0:1, 2], new_K[..., 1:2, 2]
    return torch.stack([nfx * x + ncx, nfy * y + ncy], -1)


def get_ray_directions(H, W, newK, dist=None, K=None, sppd=1, with_indices=False):
    """reference ray_utils.py:62-124: per-pixel ray directions in the camera frame, (H*W, 3), pixel order
    row-major (index = row * W + col); optionally the pixel grid coordinates (H*W, 1) each."""
    xs = torch.linspace(0, W - 1. / sppd, sppd * W, device=newK.device)
    ys = torch.linspace(0, H - 1. / sppd, sppd * H, device=newK.device)
    grid_x, grid_y = torch.meshgrid([xs, ys], indexing="ij")
    grid_x = grid_x.permute(1, 0).reshape(-1, 1)
    grid_y = grid_y.permute(1, 0).reshape(-1, 1)
    if dist is not None:
        if K is None or newK is None:
            raise LonerB200Error("get_ray_directions: distortion needs K and newK")
        if isinstance(dist, (list, tuple)):
            dist = torch.tensor(dist, device=newK.device, dtype=torch.float32)
        k3 = K.reshape(-1, 3, 3)[:1].to(newK.device)
        nk3 = newK.reshape(-1, 3, 3)[:1]
        pts = torch.cat([grid_x, grid_y], dim=-1).unsqueeze(0)
        und = undistort_points(pts, k3, dist.reshape(1, -1).to(newK.device), nk3)
        new_x, new_y = und[0, :, 0:1], und[0, :, 1:2]
        newK = nk3[0]
    else:
        new_x, new_y = grid_x, grid_y
    directions = torch.cat([(new_x - newK[0, 2]) / newK[0, 0], (new_y - newK[1, 2]) / newK[1, 1],
                            torch.ones_like(grid_x)], -1)
    if with_indices:
        return directions, grid_x, grid_y
    return directions


def pose_matrix(pose):
    """4x4 transformation of a pose given as a 4x4 matrix, a [t, axis-angle] 6-vector, or a PoseHandle-like object."""
    from .mapping import axis_angle_to_matrix
    if hasattr(pose, "get_transformation_matrix"):
        return pose.get_transformation_matrix()
    if hasattr(pose, "get_pose_tensor"):
        pose = pose.get_pose_tensor()
    if pose.shape[-2:] == (4, 4):
        return pose
    m = torch.eye(4, device=pose.device, dtype=torch.float32)
    m[:3, :3] = axis_angle_to_matrix(pose[3:].float())
    m[:3, 3] = pose[:3]
    return m


class CameraRayDirections:
    """reference ray_utils.py:127-224.  `calibration.camera_intrinsic` carries k, distortion, new_k, width, height
    (reference src/common/settings.py / cfg calibration files)."""

    def __init__(self, calibration, samples_per_pixel=1, device="cuda", chunk_size=512):
        if samples_per_pixel != 1:
            raise LonerB200Error("Only 1 sample per pixel currently supported")      # ray_utils.py:140
        dev = torch.device(device)
        if dev.type != "cuda":
            raise LonerB200Error("CameraRayDirections: device must be CUDA (no CPU path)")
        ci = calibration.camera_intrinsic
        K = torch.as_tensor(ci.k, dtype=torch.float32).to(dev)
        dist = getattr(ci, "distortion", None)       # the reference always undistorts (zeros included)
        if dist is not None:
            dist = torch.as_tensor(dist, dtype=torch.float32).to(dev)
            if dist.numel() == 0:
                dist = None
        new_k = getattr(ci, "new_k", None)
        new_k = K if new_k is None else torch.as_tensor(new_k, dtype=torch.float32).to(dev)
        self.im_width, self.im_height = int(ci.width), int(ci.height)
        d, gi, gj = get_ray_directions(self.im_height, self.im_width, newK=new_k, dist=dist, K=K, sppd=1,
                                       with_indices=True)
        self.directions = d.float().contiguous()
        self.i_meshgrid = gi.float().contiguous()
        self.j_meshgrid = gj.float().contiguous()
        self._chunk_size = chunk_size
        self.num_chunks = int(math.ceil(self.directions.shape[0] / self._chunk_size))

    def __len__(self):
        return self.directions.shape[0]

    def build_rays(self, camera_indices, pose, image, world_cube, ray_range):
        """reference ray_utils.py:168-206.  camera_indices int64 [m] in [0, W*H); pose = camera-to-world (4x4,
        6-vector or pose object; used detached); image [H,W,C] tensor (or an object with `.image`) or None.
        Returns (rays [m,13] = [o, d, -d, pix_i, pix_j, near, far], intensities [m,C] | None)."""
        from .mapping import _range_host
        dev = self.directions.device
        T = pose_matrix(pose).detach().to(dev).float()
        origin = ((T[:3, 3] + torch.as_tensor(world_cube.shift).to(dev)) / world_cube.scale_host).float().contiguous()
        rot = T[:3, :3].contiguous()
        near = _range_host(ray_range)[0] / world_cube.scale_host
        img = None
        if image is not None:
            img = getattr(image, "image", image)
            img = img.to(dev).float().reshape(-1, img.shape[-1]).contiguous()
            if img.shape[0] != len(self):
                raise LonerB200Error("build_rays: image size does not match the calibration")
        idx = None if camera_indices is None else camera_indices.to(dev, torch.int64).contiguous()
        return ops.build_camera_rays(self.directions, self.i_meshgrid.view(-1), self.j_meshgrid.view(-1), idx, rot,
                                     origin, near, img)

    def fetch_chunk_rays(self, chunk_idx, pose, world_cube, ray_range):
        """reference ray_utils.py:213-219."""
        start = chunk_idx * self._chunk_size
        end = min(len(self), (chunk_idx + 1) * self._chunk_size)
        idx = torch.arange(start, end, device=self.directions.device)
        return self.build_rays(idx, pose, None, world_cube, ray_range)[0]
