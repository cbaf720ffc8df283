"""ORACLE / TEST INFRASTRUCTURE ONLY -- parity unpinned.

Stand-in for kornia.geometry.calibration (kornia is a dependency of the reference that is absent from
/root/reference and from this image; the reference calls it at src/common/ray_utils.py:110).  `undistort_points`
restates the published algorithm (OpenCV's iterative undistortPoints as kornia implements it: normalise with K,
5 fixed-point iterations of the inverse Brown-Conrady model with rational / thin-prism terms, re-project with
new_K).  The tilted-sensor terms (coefficients 13-14) are not restated.
"""
import torch


def undistort_points(points, K, dist, new_K=None, num_iters=5):
    n = dist.shape[-1]
    assert n in (4, 5, 8, 12, 14), f"Invalid number of distortion coefficients. Got {n}"
    if n < 14:
        dist = torch.nn.functional.pad(dist, [0, 14 - n])
    assert not (dist[..., 12:] != 0).any(), "tilt model not restated in the oracle stub"
    cx, cy = K[..., 0:1, 2], K[..., 1:2, 2]
    fx, fy = K[..., 0:1, 0], K[..., 1:2, 1]
    x = (points[..., 0] - cx) / fx
    y = (points[..., 1] - cy) / fy
    x0, y0 = x, y
    for _ in range(num_iters):
        r2 = x * x + y * y
        num = 1 + dist[..., 5:6] * r2 + dist[..., 6:7] * r2 * r2 + dist[..., 7:8] * r2 ** 3
        den = 1 + dist[..., 0:1] * r2 + dist[..., 1:2] * r2 * r2 + dist[..., 4:5] * r2 ** 3
        inv_rad_poly = num / den
        delta_x = 2 * dist[..., 2:3] * x * y + dist[..., 3:4] * (r2 + 2 * x * x) + dist[..., 8:9] * r2 + dist[..., 9:10] * r2 * r2
        delta_y = dist[..., 2:3] * (r2 + 2 * y * y) + 2 * dist[..., 3:4] * x * y + dist[..., 10:11] * r2 + dist[..., 11:12] * r2 * r2
        x = (x0 - delta_x) * inv_rad_poly
        y = (y0 - delta_y) * inv_rad_poly
    if new_K is None:
        new_K = K
    return torch.stack([new_K[..., 0:1, 0] * x + new_K[..., 0:1, 2], new_K[..., 1:2, 1] * y + new_K[..., 1:2, 2]], -1)
