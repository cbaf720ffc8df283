#!/usr/bin/env python
"""bench.py -- LONER mapping hot path on B200: rays/s (fwd+bwd+Adam) and ms per mapping iteration.

    python bench.py [--gpus N] [--steps K] [--warmup W]                 # our arm (sm_100a kernels), BASELINE configs[1]
    python bench.py --impl reference [--steps K] [--warmup W]           # CPU arm: the reference's own Python (oracle/_ref)
    torchrun ... bench.py --gpus N [--workload c2|c3|c4|c5] [--scaling weak|strong]

One "step" = one pass of the reference's hot loop body (src/mapping/optimizer.py:354-475) over one synthetic batch:
ray selection + ray construction for an 8-keyframe window (world-cube filter on, as `iterate_optimizer` runs it),
occupancy-guided sampling, hash-grid + MLP forward, compositing + JS depth loss, backward (table scatter-add, MLP),
gradient exchange (N > 1: reduce-scatter -> sharded Adam -> all-gather), Adam step, zero-grad and the occupancy
update every 10th step.  Workload at N=1 = BASELINE.json configs[1]: 8 keyframes x 4096 rays, 512 samples/ray,
16-level 2^19 hash grid.

  --scaling weak   (default) every rank processes its own 32768-ray batch (N=8 is configs[3]'s 262144 rays/iter)
  --scaling strong ONE 32768-ray batch per iteration (configs[2]), each rank takes rows [r n/N, (r+1) n/N) of every chunk;
                   random tensors are drawn for the whole chunk on every rank and sliced, so N GPUs reproduce the 1-GPU step
  --workload c5    configs[4]: eval render of 1 M rays x 2048 samples (depth / variance images), rays split over the ranks
                   like the reference's per-GPU render queues (analysis/renderer.py:433-460); no collective

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definition of every field.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1] -- the workload `metric` is quoted on (default)
    "c2": dict(name="newer_college_shaped_C2", keyframes=8, rays_per_keyframe=4096, sky_per_keyframe=0, n_samples=512,
               log2_hashmap_size=19, chunk=16384, ray_range=(1.0, 50.0), voxel_size=100, pose_grad=False),
    # configs[2]: joint keyframe pose refinement (gradients to ray origins / directions -> pose 6-vectors)
    "c3": dict(name="fusionportable_shaped_C3_pose_refinement", keyframes=8, rays_per_keyframe=4096, sky_per_keyframe=0,
               n_samples=512, log2_hashmap_size=19, chunk=16384, ray_range=(1.0, 50.0), voxel_size=100, pose_grad=True),
    # configs[3]: 2^22 hash tables, 32768 rays per GPU (262144 rays/iter at 8 GPUs); chunk and ray range as in the
    # reference's haveri_hpk configuration (cfg/haveri_hpk/02_02_04_submaps.yaml:60,71)
    "c4": dict(name="haveri_hpk_shaped_C4_T22", keyframes=8, rays_per_keyframe=4096, sky_per_keyframe=0, n_samples=512,
               log2_hashmap_size=22, chunk=32768, ray_range=(2.5, 45.0), voxel_size=100, pose_grad=False),
    # configs[4]: eval render, 1 M rays x N_samples_test = 2048 (default_model_config.yaml), sigma head (depth / variance)
    "c5": dict(name="eval_render_C5_1M_rays", keyframes=4, rays_per_keyframe=262144, sky_per_keyframe=0, n_samples=2048,
               log2_hashmap_size=19, chunk=16384, ray_range=(1.0, 50.0), voxel_size=100, pose_grad=False, eval=True),
}
WORKLOAD = WORKLOADS["c2"]
METRIC = "mapping rays/sec (fwd+bwd)"
UNIT = "rays/s"

# algorithmic bytes per sample point (encodings kept for the backward):
#   fwd: 16 levels x 8 corners x 8 B gathers + 4 B z + 128 B saved encoding + 4 B sigma
#   bwd: 16 x 8 x 8 B x 2 (read-modify-write atomics) + 128 B encoding + 4 B z + 4 B d_sigma
ALGO_BYTES = {"lb_sigma_fwd": 16 * 8 * 8 + 4 + 128 + 4, "lb_sigma_bwd": 16 * 8 * 8 * 2 + 128 + 4 + 4}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample-rays", type=int, default=512, help="rays per CPU-baseline step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS),
                    help="c2 = BASELINE configs[1] (default, the headline); c3 / c4 / c5 = configs[2] / [3] / [4]")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--no-jvp", action="store_true",
                    help="pose refinement: reverse-mode input gradient (lb_sigma_bwd d_rays) instead of lb_sigma_fwd_jvp (A/B)")
    ap.add_argument("--no-overlap", action="store_true",
                    help="optimiser step in line on the main stream (A/B against the default: side stream, overlapped with the "
                         "next step's ray build and sampler)")
    return ap.parse_args()


def load_json(name):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name)))
    except Exception:
        return {}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def config_dict(world, scaling):
    W = WORKLOAD
    n_rays = W["keyframes"] * W["rays_per_keyframe"]
    return {"workload": W["name"], **{k: W[k] for k in ("keyframes", "rays_per_keyframe", "n_samples", "log2_hashmap_size", "chunk")},
            "rays_per_step_per_gpu": n_rays if scaling == "weak" else n_rays // world, "rays_per_step": n_rays * (world if scaling == "weak" else 1),
            "parallelism": f"ray-dp{world}", "pose_refinement": bool(W["pose_grad"]),
            "world_cube_filter": "on (Optimizer.build_batch default path)", "sky_rays_per_keyframe": W["sky_per_keyframe"],
            "l2": "inputs_exceed_l2 (per-step working set ~2.4 GB of encodings + hash tables vs 126 MB L2)"}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe), polled every 20 ms."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.idx = gpu_index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                       "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
            time.sleep(0.15)          # the first sample lands before the timed region starts
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_scene(n_kf, seed=0):
    from loner_b200 import synth
    return synth.make_keyframes(n_kf, ray_range=WORKLOAD["ray_range"], seed=seed)


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's own Python (verbatim copy under oracle/_ref, or /root/reference in the build container) behind
# the fp32 CPU restatement of tiny-cuda-nn (oracle/tinycudann); falls back to the oracle port when no copy travelled
# ------------------------------------------------------------------------------------------
def cpu_reference_steps(steps, warmup, sample_rays):
    """Times `steps` mapping iterations of the reference on the host cores, each on a bounded sample (sample_rays rays of
    the workload's batch, same S / hash-grid size / occupancy grid / loss).  Returns dict."""
    import torch
    from oracle.reference_import import reference_available
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    W = WORKLOAD
    kfs, scale, shift = make_scene(W["keyframes"])
    S, T, D = W["n_samples"], W["log2_hashmap_size"], W["voxel_size"]
    per_kf = max(1, sample_rays // W["keyframes"])
    rr = torch.tensor(W["ray_range"])
    times = []
    if reference_available():
        from oracle.reference_import import import_reference
        from oracle import make_golden as G
        ns = import_reference()
        kind = "reference"
        what = ("the reference's own src/mapping/optimizer.py compute_loss / _step_occupancy_grid, src/models/*, "
                "src/common/ray_utils.py (verbatim copy, oracle/make_ref_copy.py) on torch-CPU behind oracle/tinycudann")
        cfg = G.model_cfg(ns, S, T, D, W["ray_range"], "L1_JS", 1.0, chunk=W["chunk"])
        opt = G.bare_optimizer(ns, cfg, scale, "OGM")
        wc = ns.pose_utils.WorldCube(scale, shift)
        builders, poses = [], []
        for kf in kfs:
            n = kf.distances.shape[0]
            builders.append(ns.ray_utils.LidarRayDirections(ns.sensors.LidarScan(kf.directions, kf.distances, torch.arange(n).float() * 1e-6)))
            poses.append(ns.pose.Pose(pose_tensor=kf.pose.clone(), fixed=True).get_transformation_matrix())
        adam = torch.optim.Adam([{"params": opt._model.get_sigma_parameters(), "lr": cfg.train.lrate_sigma_mlp}])
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            rays, depths = [], []
            for b, T_kf in zip(builders, poses):
                idx = torch.randint(len(b), (per_kf,))                 # RANDOM ray selection (optimizer.py:365-366)
                r, d = b.build_lidar_rays(idx, rr, wc, T_kf)
                rays.append(r); depths.append(d)
            rays, depths = torch.vstack(rays).float(), torch.cat(depths).float()
            loss = opt.compute_loss(None, (rays, depths), it)
            loss.backward()
            adam.step()
            adam.zero_grad(set_to_none=True)
            if it % cfg.model.occ_model.N_iters_acc == 0:
                opt._step_occupancy_grid()
            opt._global_step += 1
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append((dt, rays.shape[0]))
    else:
        from oracle import loner_oracle as O
        kind = "port"
        what = "oracle/loner_oracle.py: torch-CPU restatement of the reference (no reference copy on this box)"
        params = O.init_sigma_params(T).requires_grad_(True)
        m, v = torch.zeros_like(params), torch.zeros_like(params)
        occ = torch.zeros(1, 1, D, D, D)
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            rays, depths = [], []
            for kf in kfs:
                idx = torch.randint(kf.distances.shape[0], (per_kf,))
                r, d = O.build_lidar_rays(kf.directions[:, idx], kf.distances[idx], kf.pose, shift, scale, rr)
                rays.append(r); depths.append(d)
            rays, depths = torch.vstack(rays), torch.cat(depths)
            n = rays.shape[0]
            u1, u2, noise = torch.rand(n, S // 2), torch.rand(n, S // 2), torch.randn(n, S)
            out = O.mapping_forward(rays, depths, params, occ, u1, u2, noise, scale, S, log2_hashmap_size=T)
            out["loss"].backward()
            with torch.no_grad():
                O.adam_step(params, params.grad, m, v, it + 1, 0.01)
                params.grad = None
            if it % 10 == 0:
                occ = O.occ_update(occ, out["xyz"].detach(), out["z"] * scale, (depths * scale).reshape(-1, 1))
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append((dt, n))
    tot_t = sum(t for t, _ in times)
    tot_n = sum(n for _, n in times)
    return dict(value=tot_n / tot_t, unit=UNIT, cores=cores, kind=kind,
                sample=f"{per_kf * W['keyframes']} of the batch's {W['keyframes'] * W['rays_per_keyframe']} rays x {S} samples per step, "
                       f"T=2^{T}, {D}^3 occupancy grid, {steps} steps after {warmup} warm-up (fwd + bwd + Adam + occupancy update): {what}",
                ms_per_step=1e3 * tot_t / len(times), rays_per_step=tot_n // len(times))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_steps(args.steps, args.warmup, args.cpu_sample_rays)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args.gpus, args.scaling),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "sample_rays_per_step": r["rays_per_step"],
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
def count_launches(fn):
    """Kernel launches of one call of fn (CUPTI via torch.profiler): {ours, aten, nccl, names of ours}."""
    import torch
    from torch.profiler import profile, ProfilerActivity
    try:
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        ours, aten, nccl, names = 0, 0, 0, {}
        for ev in prof.events():
            if str(getattr(ev, "device_type", "")).endswith("CUDA") and ev.name and not ev.name.startswith(("Memcpy", "Memset")):
                nm = ev.name
                if "nccl" in nm.lower():
                    nccl += 1
                elif nm.startswith("void at::") or nm.startswith("at::") or "at::native" in nm or "cub::" in nm or "thrust" in nm or "curand" in nm.lower():
                    aten += 1
                else:
                    ours += 1
                    short = nm.split("<")[0].replace("void ", "").replace("(anonymous namespace)::", "")
                    names[short] = names.get(short, 0) + 1
        return {"ours": ours, "aten": aten, "nccl": nccl, "ours_by_kernel": names}
    except Exception as e:       # profiler unavailable: fall back to the C-ABI call count
        return {"error": repr(e)[:200]}


def run_eval(args, dev, world, rank, dist):
    """configs[4]: eval render (Model.forward(testing=True), sigma head: depth / variance / opacity images), rays split
    over the ranks; one step = the whole 1 M-ray render."""
    import torch
    from loner_b200 import _lib, mapping as M
    W = WORKLOAD
    torch.manual_seed(0)
    kfs, scale, shift = make_scene(W["keyframes"])
    s = M.default_settings(n_samples_train=512, n_samples_test=W["n_samples"], log2_hashmap_size=W["log2_hashmap_size"],
                           chunk=W["chunk"], ray_range=W["ray_range"])
    opt = M.Optimizer(s, None, M.WorldCube(scale, shift), dev, lidar_only=False)
    with torch.no_grad():   # trained-like magnitudes so that rays terminate
        opt._model.nerf_model._model_sigma.params[3072:].mul_(2000.0)
        opt._model.nerf_model._pos_encoding.params.mul_(2000.0)
        opt._occupancy_grid.normal_(0.0, 0.5)
    opt._model.cfg["render"]["retraw"] = False
    wc = M.WorldCube(scale.to(dev), shift.to(dev))
    rays = []
    for k in kfs:
        idx = torch.randint(k.distances.shape[0], (W["rays_per_keyframe"],))
        r, _ = M.build_lidar_rays(k.directions[:, idx].to(dev), k.distances[idx].to(dev), k.pose.to(dev), wc,
                                  torch.tensor(W["ray_range"], device=dev), True)
        rays.append(r)
    rays = torch.vstack(rays).contiguous()
    n_total = rays.shape[0]
    per = (n_total + world - 1) // world
    mine = rays[rank * per:(rank + 1) * per].contiguous()      # contiguous image tiles per rank, no collective
    host = mine.cpu().pin_memory()

    TILE = 8 * W["chunk"]          # rays per Model.forward call: an image-sized tile (131072 rays = 8 chunks)

    def render(r, camera=False, strategy="default"):
        """The whole ray set as image tiles through Model.forward(testing=True); only the images are kept (depth,
        variance, opacity, colour): the per-sample results of a tile (weights, raw: 34 KB per ray with colour) are
        dropped before the next tile, as a frame renderer does (analysis/renderer.py:181-215 keeps images only)."""
        n = r.shape[0]
        out = {"depth_fine": torch.empty(n, device=dev), "variance": torch.empty(n, device=dev),
               "opacity_fine": torch.empty(n, device=dev)}
        if camera:
            out["rgb_fine"] = torch.empty(n, 3, device=dev)
        with torch.no_grad():
            for t0 in range(0, n, TILE):
                res = opt._model(r[t0:t0 + TILE], opt._ray_sampler, opt._scale_factor, testing=True, camera=camera,
                                 return_variance=True, render_strategy=strategy)
                for k in out:
                    out[k][t0:t0 + TILE] = res[k].reshape(out[k][t0:t0 + TILE].shape)
                del res
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(1, min(args.warmup, 3))):
        render(mine)
    barrier()
    clocks = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    if rank == 0:
        clocks.start()
    _lib.reset_counters()
    _lib.TIMING["on"] = True
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = render(mine)
    e1.record()
    barrier()
    _lib.TIMING["on"] = False
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    ktimes = _lib.kernel_times()
    clk = clocks.stop() if rank == 0 else None
    # e2e: host ray buffers in, depth / variance / opacity images out (D2H) every step
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record()
    for _ in range(args.steps):
        o = render(host.to(dev, non_blocking=True))
        imgs = [o[k].cpu() for k in ("depth_fine", "variance", "opacity_fine")]
    t1.record()
    barrier()
    ms2 = torch.tensor([t0.elapsed_time(t1)], device=dev)
    extra = {}
    for camera, strategy in ((False, "adjusted"), (True, "default")):
        render(mine, camera, strategy)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        render(mine, camera, strategy)
        b.record()
        barrier()
        t = torch.tensor([a.elapsed_time(b)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        extra[("colour" if camera else "sigma") + "_" + strategy + "_ms"] = t.item()
    peak_mem = torch.cuda.max_memory_allocated() / 2 ** 30
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    if rank != 0:
        return
    launches = count_launches(lambda: render(mine[:W["chunk"]]))
    n_chunks = (mine.shape[0] + W["chunk"] - 1) // W["chunk"]
    peak, peak_src = measured_peaks()
    S = W["n_samples"]
    pts = W["chunk"] * S
    n_l, t_l = ktimes["lb_sigma_fwd"]
    l2 = load_json("r02_l2_ceiling.json").get("ncu_tex_sectors_g_per_s", {})
    sect = load_json("r02_traffic.json").get("lb_sigma_fwd_inference", {})
    roof = {"bound": "l2", "kernel": "lb_sigma_fwd (inference: no encodings written)", "unit": "GB/s",
            "avg_launch_ms": t_l / n_l, "points_per_launch": pts,
            "hbm_view": {"achieved": (ALGO_BYTES["lb_sigma_fwd"] - 128) * pts / (t_l / n_l * 1e-3) / 1e9, "peak": peak, "peak_source": peak_src,
                         "algorithmic_bytes_per_point": ALGO_BYTES["lb_sigma_fwd"] - 128},
            "kernels_ms_per_step": {k: round(v[1] / args.steps, 3) for k, v in sorted(ktimes.items())}}
    if l2 and sect.get("tex_sectors_per_point"):
        ach = sect["tex_sectors_per_point"] * pts * 32 / (t_l / n_l * 1e-3) / 1e9
        pk = l2[sect["ceiling"]] * 32.0
        roof.update(achieved=ach, peak=pk, frac=ach / pk, tex_sectors_per_point=sect["tex_sectors_per_point"], traffic=sect.get("dram_bytes"))
    else:
        roof.update(achieved=roof["hbm_view"]["achieved"], peak=peak, frac=roof["hbm_view"]["achieved"] / peak, traffic=None, bound="hbm")
    line = {"metric": "eval render rays/sec (depth + variance images)", "value": n_total * args.steps / (ms.item() * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms.item() / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": W["name"], "rays": n_total, "n_samples": S, "log2_hashmap_size": W["log2_hashmap_size"],
                       "chunk": W["chunk"], "tile_rays": TILE, "rays_per_gpu": mine.shape[0], "parallelism": f"ray-tiles{world}", "render_strategy": "default",
                       "l2": "inputs_exceed_l2"},
            "samples_per_s": n_total * S * args.steps / (ms.item() * 1e-3),
            "roofline": roof, "other_renders_ms": extra, "peak_mem_gib": peak_mem,
            "e2e": {"value": n_total * args.steps / (ms2.item() * 1e-3), "unit": UNIT, "ms_per_step": ms2.item() / args.steps,
                    "h2d_bytes_per_step": mine.shape[0] * 13 * 4, "d2h_bytes_per_step": mine.shape[0] * 3 * 4},
            "gpu_launches": (launches.get("ours", 0) * n_chunks * args.steps) if "ours" in launches else sum(_lib.LAUNCHES.values()),
            "launches_per_chunk": launches, "clocks": clk}
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from loner_b200 import _lib, mapping as M

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the CUDA path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))    # a lost collective dies fast
    _lib.load()
    W = WORKLOAD
    if W.get("eval"):
        run_eval(args, dev, world, rank, dist)
        if world > 1:
            dist.destroy_process_group()
        return
    strong = args.scaling == "strong"
    # weak: every rank owns a different set of scans and its own RNG stream; strong: identical scans and streams on every rank
    torch.manual_seed(1234 + (0 if strong else rank))
    kfs, scale, shift = make_scene(W["keyframes"], seed=0 if strong else rank)
    settings = M.default_settings(n_samples_train=W["n_samples"], log2_hashmap_size=W["log2_hashmap_size"],
                                  ray_range=W["ray_range"], lidar=W["rays_per_keyframe"], sky=W["sky_per_keyframe"],
                                  chunk=W["chunk"], voxel_size=W["voxel_size"])
    dp = M.RayDataParallel(shard_batch=strong)
    opt = M.Optimizer(settings, None, M.WorldCube(scale, shift), dev, data_parallel=dp)
    window = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=dev) for k in kfs]
    for i, kf in enumerate(window):
        # configs[1]: map optimisation with fixed poses (reference first phase, defaults.yaml:80-84);
        # pose refinement (configs[2]): every keyframe but the anchor carries a trainable pose 6-vector
        free = bool(W["pose_grad"]) and i > 0
        kf.is_anchored = not free
        kf.get_lidar_pose().set_fixed(not free)
    opt._model.freeze_sigma_head(False)
    opt._model.freeze_rgb_head(True)
    if args.no_jvp:
        opt._model.nerf_model._model_sigma.plan.jvp_supported = False
    sigma_params = opt._model.get_sigma_parameters()
    groups = [{"params": sigma_params, "lr": settings.model_config.train.lrate_sigma_mlp}]
    pose_params = [kf.get_lidar_pose().get_pose_tensor() for kf in window if not kf.is_anchored]
    if pose_params:
        groups.append({"params": pose_params, "lr": settings.model_config.train.lrate_pose})
    # the optimiser step of the field parameters (N > 1: reduce-scatter -> sharded Adam -> all-gather) runs on a side stream
    # under the ray build / sampler of step i+1 (what Optimizer.iterate_optimizer does); compute_loss waits for it before
    # the first field evaluation
    adam = M.FusedAdam(groups, data_parallel=dp, overlap_step=not args.no_overlap)
    n_acc = settings.model_config.model.occ_model.N_iters_acc
    n_rays = W["keyframes"] * W["rays_per_keyframe"]

    def step(i, host_batch=None):
        if host_batch is None:
            rays, depths = opt.build_batch(window)          # the call iterate_optimizer makes (world-cube filter on)
        else:       # e2e: this step's rays arrive in pinned host memory (the reference preps rays on the CPU)
            rays, depths = host_batch[0].to(dev, non_blocking=True), host_batch[1].to(dev, non_blocking=True)
        loss = opt.compute_loss(None, (rays, depths), i)
        loss.backward()
        adam.step()                                         # N > 1: reduce-scatter -> sharded Adam -> all-gather inside
        adam.zero_grad(set_to_none=True)
        if i % n_acc == 0:
            opt._step_occupancy_grid()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    barrier()

    # ---- timed region: device-resident inputs -------------------------------------------------
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    _lib.reset_counters()
    _lib.TIMING["on"] = True
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    dp.wait_params()                # the last step's parameter all-gather belongs to the timed region
    e1.record()
    barrier()
    _lib.TIMING["on"] = False
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = ms.item()
    abi_calls = sum(_lib.LAUNCHES.values())
    ktimes = _lib.kernel_times()
    clk = clocks.stop() if rank == 0 else None
    # collective durations: a separate short pass with every collective synchronous and bracketed by CUDA events (the
    # timed region above overlaps the parameter all-gather with the next step, which events on this stream cannot see)
    coll, coll_steps = {}, min(args.steps, 10)
    if world > 1:
        dp.events = []
        for i in range(coll_steps):
            step(args.warmup + i + 1)
        barrier()
        for name, a, b in dp.events:
            n, t = coll.get(name, (0, 0.0))
            coll[name] = (n + 1, t + a.elapsed_time(b))
        dp.events = None

    # ---- e2e: same step through the public API with HOST ray buffers --------------------------
    pool = []
    for _ in range(4):
        r, d = opt.build_batch(window)
        pool.append((r.detach().cpu().pin_memory(), d.detach().cpu().pin_memory()))
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    # D2H of EVERY step's loss into pinned memory, read by the host one step behind (after that step's event): the host
    # sees each loss inside the timed region without draining the launch queue the way a blocking .cpu() per step does
    loss_host = torch.zeros(args.steps, dtype=torch.float32).pin_memory()
    done = [torch.cuda.Event() for _ in range(args.steps)]
    losses = []
    t0.record()
    for i in range(args.steps):
        l = step(args.warmup + args.steps + i, pool[i % len(pool)])
        loss_host[i:i + 1].copy_(l.detach().reshape(1), non_blocking=True)
        done[i].record()
        if i > 0:
            done[i - 1].synchronize()
            losses.append(float(loss_host[i - 1]))
    dp.wait_params()
    done[-1].record()
    done[-1].synchronize()
    losses.append(float(loss_host[args.steps - 1]))
    t1.record()
    barrier()
    last = losses[-1]
    assert len(losses) == args.steps and all(x == x for x in losses), "a step's loss was not read back / is NaN"
    ms2 = torch.tensor([t0.elapsed_time(t1)], device=dev)
    last_t = torch.tensor([last], device=dev)
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
        dist.all_reduce(last_t)        # the loss is a sum of per-rank partial sums over GLOBAL normalisers
    e2e_ms = ms2.item()
    if world > 1:
        dist.barrier()
    # one more step under the CUPTI profiler for the real launch counts; every rank runs it (it contains collectives)
    i_prof = args.warmup + 2 * args.steps
    i_prof += 1 if i_prof % n_acc == 0 else 0                 # not a multiple of 10: no occupancy update
    if rank == 0:
        launches = count_launches(lambda: step(i_prof))
    else:
        launches = None
        step(i_prof)
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    rays_local = n_rays // world if strong else n_rays
    n_chunks = max(1, (n_rays + W["chunk"] - 1) // W["chunk"])
    per_launch_points = rays_local * W["n_samples"] / n_chunks
    dom = max((k for k in ktimes if k in ALGO_BYTES), key=lambda k: ktimes[k][1])
    # L2 request-rate roofline (profiles/README.md, round 2): both fused kernels are bound by the 32-byte sector requests
    # the SMs send to the L2-resident tables, not by DRAM.  One unit on both sides, the ncu counter
    # lts__t_sectors_srcunit_tex: achieved = sectors per point of the kernel (dense steady-state capture,
    # profiles/r02_traffic.json) x points per launch x 32 B / CUDA-event time; peak = the same counter per second for the
    # microbenchmark pattern with the kernel's request mix (profiles/l2_ceiling.cu -> r02_l2_ceiling.json: random 8-byte
    # gathers + 14 % streaming stores for the forward, red.global.add.v2.f32 + 8 % streaming loads for the backward, into
    # a 57 MB table, 'of measured').  The SURVEY 8(d) algorithmic-byte figure against the HBM copy peak stays next to it
    # as `hbm_view`.
    ceil = load_json("r02_l2_ceiling.json").get("ncu_tex_sectors_g_per_s", {})
    traffic = load_json("r02_traffic.json")

    def views(k):
        n_l, t_l = ktimes[k]
        sec = t_l / n_l * 1e-3
        hbm = ALGO_BYTES[k] * per_launch_points / sec / 1e9
        v = {"avg_launch_ms": t_l / n_l, "hbm_view": {"achieved": hbm, "peak": peak, "frac": hbm / peak, "peak_source": peak_src,
                                                       "algorithmic_bytes_per_point": ALGO_BYTES[k]}}
        tr = traffic.get(k, {})
        if ceil and tr.get("tex_sectors_per_point") and W["log2_hashmap_size"] <= 19 and not W["pose_grad"]:
            pk = ceil[tr["ceiling"]] * 32.0
            ach = tr["tex_sectors_per_point"] * per_launch_points * 32 / sec / 1e9
            v.update(bound="l2", achieved=ach, peak=pk, frac=ach / pk, tex_sectors_per_point=tr["tex_sectors_per_point"],
                     ceiling_pattern=tr["ceiling"], traffic=tr.get("dram_bytes") if abs(per_launch_points - 16384 * 512) < 1 else None)
        else:      # other table sizes / the dL/dx backward: no sector capture committed, HBM view only
            v.update(bound="hbm", achieved=hbm, peak=peak, frac=hbm / peak, traffic=None)
        return v

    roofline = {"kernel": dom, "unit": "GB/s", **views(dom), "points_per_launch": per_launch_points,
                "peak_note": "bound l2: GB/s = 32-byte sectors the SMs send to L2 per second x 32 (ncu lts__t_sectors_srcunit_tex); peak = the same "
                             "counter for the microbenchmark pattern with this kernel's request mix into a 57 MB L2-resident table "
                             "(profiles/r02_l2_ceiling.json, 'of measured'); traffic = DRAM bytes per launch; hbm_view = SURVEY 8(d) algorithmic "
                             "bytes / MEASURED_PEAKS.json copy bandwidth",
                "kernels_ms_per_step": {k: round(v[1] / args.steps, 4) for k, v in sorted(ktimes.items())}}
    for k in ALGO_BYTES:
        if k in ktimes and k != dom:
            roofline[k] = views(k)
    per_step_ours = launches.get("ours") if launches else None
    line = {"metric": METRIC, "value": n_rays * (1 if strong else world) * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(world, args.scaling),
            "roofline": roofline,
            "e2e": {"value": n_rays * (1 if strong else world) * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": n_rays * 14 * 4, "d2h_bytes_per_step": 4, "last_loss": last_t.item(),
                    "readback": "every step's loss is copied to pinned host memory and read by the host one step behind (event per step)",
                    **({"note": "host ray buffers are world-frame rays, which carry no pose gradient: this leg is the map-only step; "
                                "the pose chain (poses -> rays on the device -> dL/dpose) is in `value` only"} if W["pose_grad"] else {})},
            "gpu_launches": per_step_ours * args.steps if per_step_ours else abi_calls,
            "launches_per_step": launches, "c_abi_calls": abi_calls, "clocks": clk}
    if coll:
        line["collectives_ms_per_step"] = {k: round(v[1] / coll_steps, 4) for k, v in sorted(coll.items())}
        line["collectives_note"] = (f"separate pass of {coll_steps} steps with the optimiser step in line; in the timed region it "
                                    + ("is in line too (--no-overlap)" if args.no_overlap else
                                       "runs on a side stream under the next step's ray build and sampler"))
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_steps(10, 2, args.cpu_sample_rays)
        line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    WORKLOAD = WORKLOADS[a.workload]
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
