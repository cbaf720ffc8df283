"""ORACLE / TEST INFRASTRUCTURE ONLY.  Container-only script (needs /root/reference).

Golden fixture for camera ray construction: runs the reference's OWN get_ray_directions and
CameraRayDirections.build_rays (src/common/ray_utils.py:62-124,127-206), unmodified, on torch-CPU.
  * "pinhole": get_ray_directions(dist=None) -- no third-party code involved: pins the direction table, the pixel
    order, the pose / world-cube handling, the cube exit and the intensity gather;
  * "distorted": the reference's constructor with a 5-coefficient distortion; its kornia call is served by the
    restated oracle/stubs/kornia/geometry/calibration.py (kornia itself is absent: this half is unpinned).

    python -m oracle.make_golden_camera        # writes tests/golden/camera_rays.pt
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.reference_import import import_reference  # noqa: E402


def main():
    ns = import_reference()
    RU = ns.ray_utils
    gen = torch.Generator().manual_seed(11)
    H, W = 30, 40
    K = torch.tensor([[35.0, 0.0, 19.3], [0.0, 36.5, 14.6], [0.0, 0.0, 1.0]])
    new_k = torch.tensor([[33.0, 0.0, 19.8], [0.0, 34.0, 14.9], [0.0, 0.0, 1.0]])
    dist = torch.tensor([-0.21, 0.07, 0.0011, -0.0007, -0.012])
    aa = torch.tensor([0.3, -0.5, 0.2])
    T = torch.eye(4)
    T[:3, :3] = ns.pose_utils.tensor_to_transform(torch.cat([torch.zeros(3), aa]))[:3, :3]
    T[:3, 3] = torch.tensor([1.5, -0.7, 0.4])
    scale, shift = torch.tensor(12.5), torch.tensor([0.3, 0.1, -0.2])
    ray_range = torch.tensor([0.7, 40.0])
    image = torch.rand(H, W, 3, generator=gen)
    idx = torch.randperm(H * W, generator=gen)[:257]
    out = dict(H=H, W=W, K=K, new_k=new_k, dist=dist, T=T, scale=scale, shift=shift, ray_range=ray_range, image=image,
               indices=idx)

    def run(crd):
        pose = ns.pose.Pose(T.clone())
        wc = ns.pose_utils.WorldCube(scale.clone(), shift.clone())
        rays, inten = crd.build_rays(idx, pose, ns.sensors.Image(image, 0.0), wc, ray_range)
        chunk = crd.fetch_chunk_rays(1, ns.pose.Pose(T.clone()), wc, ray_range)
        return rays, inten, chunk

    # pinhole: the reference's own functions only
    d, gi, gj = RU.get_ray_directions(H, W, newK=new_k, dist=None, K=None, sppd=1, with_indices=True)
    crd = RU.CameraRayDirections.__new__(RU.CameraRayDirections)
    crd.directions, crd.i_meshgrid, crd.j_meshgrid = d, gi, gj
    crd.im_width, crd.im_height, crd._chunk_size = W, H, 512
    crd.num_chunks = 3
    rays, inten, chunk = run(crd)
    out["pinhole"] = dict(directions=d, pix_i=gi, pix_j=gj, rays=rays, intensities=inten, chunk1=chunk)

    # distorted: the reference's constructor (kornia call served by the restated stub)
    calib = ns.settings.Settings({"camera_intrinsic": {"k": K, "distortion": dist, "new_k": new_k, "width": W, "height": H}})
    crd2 = RU.CameraRayDirections(calib, device="cpu")
    rays2, inten2, chunk2 = run(crd2)
    out["distorted"] = dict(directions=crd2.directions, rays=rays2, intensities=inten2, chunk1=chunk2)

    path = os.path.join(ROOT, "tests", "golden", "camera_rays.pt")
    torch.save(out, path)
    print("wrote", path, {k: tuple(v.shape) for k, v in out["pinhole"].items()})


if __name__ == "__main__":
    main()
