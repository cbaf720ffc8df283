"""ORACLE / TEST INFRASTRUCTURE ONLY.  Container-only script (needs /root/reference).

Golden fixtures for the eval drivers from the reference's OWN analysis code, run unmodified on
torch-CPU behind the fp32 `tinycudann` shim:

  mesher   analysis/mesher.py Mesher.get_mesh (:100-180): virtual 360-degree scans from the checkpoint poses through
           Model.forward(testing=False), validity masks, bucketize, max compositing weight per voxel.  The file is imported
           as it is; its ROS / trimesh / skimage imports are served by empty stand-ins, and the stand-in for
           skimage.measure.marching_cubes records the `volume` argument -- the array the reference hands to marching cubes
           (skimage itself is absent from this image: marching cubes is out of scope).
           `LidarScan.to(0)` inside build_lidar_scan (:49) asks for cuda:0 and is redirected to the CPU.
  renderer analysis/renderer.py render_dataset_frame (:181-215): colour / depth / peak-depth-consistency images, chunk
           loop of 512 rays through Model.forward(testing=True).  renderer.py is a script whose module level parses
           arguments and loads a checkpoint, so only the SOURCE TEXT of that function is taken from the file (ast) and
           executed in a namespace that supplies the globals it reads (model, ray_directions, world_cube, ...).

Random tensors come from oracle/portable_rng.py (nothing random is stored).

    python -m oracle.make_golden_eval          # writes tests/golden/eval_mesher.pt, tests/golden/eval_render.pt
"""
import ast
import importlib.util
import os
import sys
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.reference_import import import_reference, REFERENCE_ROOT  # noqa: E402
from oracle import make_golden as G  # noqa: E402
from oracle import portable_rng as R  # noqa: E402
from oracle import loner_oracle as O  # noqa: E402

REF_ANALYSIS = "/root/reference/analysis"
RAY_RANGE = (1.0, 30.0)
LOG2_T, LOG2_T_INT, S_TRAIN, S_TEST = 14, 14, 64, 128


def make_model(ns, scale):
    cfg = G.model_cfg(ns, S_TRAIN, LOG2_T, 32, RAY_RANGE)
    cfg["model"]["render"]["N_samples_test"] = S_TEST
    cfg["model"]["render"]["chunk"] = 512
    model = ns.model_tcnn.Model(cfg.model)
    p = O.init_sigma_params(LOG2_T, 1337)
    p[O.MLP_PARAMS:] *= G.TABLE_SCALE
    table, mlp = O.init_intensity_params(LOG2_T_INT, 3, 1337)
    with torch.no_grad():
        model.nerf_model._model_sigma.params.copy_(p)
        model.nerf_model._pos_encoding.params.copy_(table * G.TABLE_SCALE)
        model.nerf_model._model_intensity.params.copy_(mlp)
    return model, cfg


def mesher_golden(ns):
    # stand-ins for the imports of analysis/mesher.py that do not exist here
    captured = {}
    for name in ("sensor_msgs", "sensor_msgs.msg", "ros_numpy", "rosbag", "trimesh"):
        m = types.ModuleType(name)
        sys.modules[name] = m
    sys.modules["sensor_msgs.msg"].PointCloud2 = object
    sk = types.ModuleType("skimage")
    sk.__version__ = "0.19.3"
    sk.measure = types.ModuleType("skimage.measure")

    def marching_cubes(volume, level, spacing):
        captured["volume"], captured["level"], captured["spacing"] = np.array(volume), level, spacing
        raise RuntimeError("marching cubes is not available here (the volume has been recorded)")
    sk.measure.marching_cubes = marching_cubes
    sys.modules["skimage"], sys.modules["skimage.measure"] = sk, sk.measure
    spec = importlib.util.spec_from_file_location("ref_mesher", os.path.join(REF_ANALYSIS, "mesher.py"))
    mesher = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mesher)
    orig_to = ns.sensors.LidarScan.to
    mesher.LidarScan.to = lambda self, device: orig_to(self, "cpu")          # build_lidar_scan asks for device 0
    # get_grid_uniform (:86-88) calls np.linspace with 0-d torch tensors as end points; with the numpy the reference was
    # written for that returns an ndarray, numpy 2.x hands back a Tensor and :124 (torch.from_numpy) fails.  The
    # module's `np` is replaced by a proxy whose linspace converts the (float64) end points to Python floats first.
    np_proxy = types.ModuleType("numpy_proxy")
    np_proxy.__dict__.update(np.__dict__)
    np_proxy.linspace = lambda a, b, n: np.linspace(float(a), float(b), int(n))
    mesher.np = np_proxy

    scale, shift = torch.tensor(25.0), torch.tensor([0.5, -0.3, 0.2])
    model, cfg = make_model(ns, scale)
    wc = ns.pose_utils.WorldCube(scale, shift)
    poses = [torch.tensor([0.4, -0.2, 0.3, 0.02, -0.03, 0.5]), torch.tensor([1.9, 0.8, 0.1, -0.01, 0.02, -0.4])]
    ckpt = {"poses": [{"lidar_pose": p} for p in poses]}
    bound = [[-8, 8], [-6, 10], [-2, 4]]
    fov = [-10.0, 10.0]
    m = mesher.Mesher(model, ckpt, wc, torch.tensor(RAY_RANGE), resolution=0.5, marching_cubes_bound=bound,
                      lidar_vertical_fov=fov)
    sampler = ns.ray_sampling.UniformRaySampler()
    t0 = time.time()
    try:
        with R.Patch(0):
            m.get_mesh("cpu", sampler, skip_step=1, var_threshold=None)
    finally:
        mesher.LidarScan.to = orig_to
    grid = m.get_grid_uniform(0.5)
    vol = captured["volume"]
    out = dict(volume=torch.from_numpy(vol.copy()), xyz=[torch.from_numpy(a.copy()) for a in grid["xyz"]], poses=poses,
               scale=scale, shift=shift, bound=bound, fov=fov, resolution=0.5, ray_range=list(RAY_RANGE), log2_T=LOG2_T,
               n_samples_train=S_TRAIN, table_scale=G.TABLE_SCALE, param_seed=1337, chunk=512,
               nonzero=int((vol != 0).sum()), vmax=float(vol.max()))
    path = os.path.join(G.GOLDEN_DIR, "eval_mesher.pt")
    torch.save(out, path)
    print(f"eval_mesher: volume {vol.shape}, {out['nonzero']} voxels hit, max {out['vmax']:.4f}, {time.time() - t0:.1f} s -> {path} "
          f"({os.path.getsize(path) / 1e6:.2f} MB)")


def function_source(path, name):
    src = open(path).read()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            return ast.get_source_segment(src, node)
    raise RuntimeError(f"{name} not found in {path}")


def render_golden(ns):
    RU = ns.ray_utils
    H, W = 30, 40
    new_k = torch.tensor([[33.0, 0.0, 19.8], [0.0, 34.0, 14.9], [0.0, 0.0, 1.0]])
    d, gi, gj = RU.get_ray_directions(H, W, newK=new_k, dist=None, K=None, sppd=1, with_indices=True)
    crd = RU.CameraRayDirections.__new__(RU.CameraRayDirections)
    crd.directions, crd.i_meshgrid, crd.j_meshgrid = d, gi, gj
    crd.im_width, crd.im_height, crd._chunk_size = W, H, 512
    crd.num_chunks = int(np.ceil(H * W / 512))
    scale, shift = torch.tensor(25.0), torch.tensor([0.5, -0.3, 0.2])
    model, cfg = make_model(ns, scale)
    wc = ns.pose_utils.WorldCube(scale, shift)
    T = torch.eye(4)
    T[:3, :3] = ns.pose_utils.tensor_to_transform(torch.cat([torch.zeros(3), torch.tensor([0.3, -0.5, 0.2])]))[:3, :3]
    T[:3, 3] = torch.tensor([1.5, -0.7, 0.4])
    pose = ns.pose.Pose(T.clone())
    namespace = dict(torch=torch, time=time, im_size=torch.Tensor([H, W]), ray_directions=crd, world_cube=wc,
                     ray_range=torch.tensor(RAY_RANGE), model=model, ray_sampler=ns.ray_sampling.UniformRaySampler(),
                     scale_factor=scale, CHUNK_SIZE=512, _DEVICE="cpu", Pose=ns.pose.Pose)
    exec(compile(function_source(os.path.join(REF_ANALYSIS, "renderer.py"), "render_dataset_frame"), "renderer.py", "exec"), namespace)
    with R.Patch(0):
        rgb, depth, peak = namespace["render_dataset_frame"](pose)
    rays = torch.cat([crd.fetch_chunk_rays(c, ns.pose.Pose(T.clone()), wc, torch.tensor(RAY_RANGE)) for c in range(crd.num_chunks)])
    out = dict(rgb=rgb.clone(), depth=depth.clone(), peak_depth_consistency=peak.clone(), rays=rays.float().clone(), H=H, W=W,
               T=T, new_k=new_k, scale=scale, shift=shift, ray_range=list(RAY_RANGE), log2_T=LOG2_T, log2_T_int=LOG2_T_INT,
               n_samples_test=S_TEST, table_scale=G.TABLE_SCALE, param_seed=1337, chunk=512)
    path = os.path.join(G.GOLDEN_DIR, "eval_render.pt")
    torch.save(out, path)
    print(f"eval_render: rgb {tuple(rgb.shape)} depth mean {float(depth.mean()):.3f} m, peak consistency mean "
          f"{float(peak.mean()):.4f} m -> {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


def main():
    assert REFERENCE_ROOT == "/root/reference" or os.path.isdir(REF_ANALYSIS), "needs the reference's analysis/ scripts"
    ns = import_reference()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    render_golden(ns)
    mesher_golden(ns)


if __name__ == "__main__":
    main()
