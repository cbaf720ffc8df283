#!/usr/bin/env python
"""bench.py -- LONER mapping hot path on B200: rays/s (fwd+bwd+Adam) and ms per mapping iteration.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (sm_100a kernels)
    python bench.py --impl reference [--steps K] [--warmup W]      # CPU arm: the oracle port of the reference

One "step" = one pass of the reference's hot loop body (src/mapping/optimizer.py:354-475) over one
synthetic batch: ray selection + ray construction for an 8-keyframe window, occupancy-guided
sampling, hash-grid + MLP forward, compositing + JS depth loss, backward (table scatter-add, MLP),
Adam step, zero-grad and the occupancy update every 10th step.  Workload at N=1 = BASELINE.json
configs[1]: 8 keyframes x 4096 rays, 512 samples/ray, 16-level 2^19 hash grid.  For N>1 every rank
processes that same batch size (weak scaling; N=8 is configs[3]'s 262144 rays/iter) and the sigma-head
gradient is all-reduced over NCCL each step.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definition of every field.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
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
                    help="c2 = BASELINE configs[1] (default, the headline); c3 / c4 = configs[2] / configs[3] variants")
    return ap.parse_args()


def ncu_traffic(kernel):
    """DRAM bytes per launch of a C-ABI call from the committed ncu capture (profiles/r01_traffic.json)."""
    p = os.path.join(ROOT, "profiles", "r01_traffic.json")
    try:
        return float(json.load(open(p)).get(kernel))
    except Exception:
        return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.idx = gpu_index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
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
# CPU arm: the oracle's restatement of the reference (oracle/loner_oracle.py), all host threads
# ------------------------------------------------------------------------------------------
def cpu_reference_steps(steps, warmup, sample_rays):
    """Times `steps` mapping iterations of the reference algorithm on the host cores on a bounded sample
    (sample_rays rays of the workload's batch, same S / hash-grid size / loss).  Returns dict."""
    import torch
    from oracle import loner_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    W = WORKLOAD
    kfs, scale, shift = make_scene(W["keyframes"])
    S, T, D = W["n_samples"], W["log2_hashmap_size"], W["voxel_size"]
    params = O.init_sigma_params(T).requires_grad_(True)
    m, v = torch.zeros_like(params), torch.zeros_like(params)
    occ = torch.zeros(1, 1, D, D, D)
    per_kf = max(1, sample_rays // W["keyframes"])
    rr = torch.tensor(W["ray_range"])
    times = []
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
    return dict(value=tot_n / tot_t, unit=UNIT, cores=cores, kind="port",
                sample=f"{per_kf * W['keyframes']} rays x {S} samples per step, T=2^{T}, {steps} steps "
                       f"(oracle/loner_oracle.py: torch-CPU restatement of the reference, fwd+bwd+Adam+occ)",
                ms_per_step=1e3 * tot_t / len(times))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 5)), min(args.warmup, 1)
    r = cpu_reference_steps(steps, warmup, args.cpu_sample_rays)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD["name"], **{k: WORKLOAD[k] for k in ("keyframes", "rays_per_keyframe", "n_samples", "log2_hashmap_size")},
                       "note": "CPU arm times a bounded sample of the batch (see cpu_baseline.sample); requested steps/warmup clamped to keep the run short"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------
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
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    W = WORKLOAD
    torch.manual_seed(1234 + rank)
    kfs, scale, shift = make_scene(W["keyframes"], seed=rank)       # every rank owns a different set of scans
    settings = M.default_settings(n_samples_train=W["n_samples"], log2_hashmap_size=W["log2_hashmap_size"],
                                  ray_range=W["ray_range"], lidar=W["rays_per_keyframe"], sky=W["sky_per_keyframe"],
                                  chunk=W["chunk"], voxel_size=W["voxel_size"])
    opt = M.Optimizer(settings, None, M.WorldCube(scale, shift), dev)
    window = [M.KeyFrame(M.LidarScanData(k.directions, k.distances, k.sky_directions), k.pose, device=dev) for k in kfs]
    for i, kf in enumerate(window):
        # configs[1]: map optimisation with fixed poses (reference first phase, defaults.yaml:80-84);
        # pose refinement (configs[2]): every keyframe but the anchor carries a trainable pose 6-vector
        free = bool(W["pose_grad"]) and i > 0
        kf.is_anchored = not free
        kf.get_lidar_pose().set_fixed(not free)
    opt._model.freeze_sigma_head(False)
    opt._model.freeze_rgb_head(True)
    sigma_params = opt._model.get_sigma_parameters()
    groups = [{"params": sigma_params, "lr": settings.model_config.train.lrate_sigma_mlp}]
    pose_params = [kf.get_lidar_pose().get_pose_tensor() for kf in window if not kf.is_anchored]
    if pose_params:
        groups.append({"params": pose_params, "lr": settings.model_config.train.lrate_pose})
    adam = M.FusedAdam(groups)
    n_acc = settings.model_config.model.occ_model.N_iters_acc
    n_rays = W["keyframes"] * W["rays_per_keyframe"]

    def step(i, host_batch=None):
        if host_batch is None:
            rays, depths = opt.build_batch(window, ignore_world_cube=True)
        else:       # e2e: this step's rays arrive in pinned host memory (the reference preps rays on the CPU)
            rays, depths = host_batch[0].to(dev, non_blocking=True), host_batch[1].to(dev, non_blocking=True)
        loss = opt.compute_loss(None, (rays, depths), i)
        loss.backward()
        opt._dp.all_reduce_grads(sigma_params + pose_params)
        adam.step()
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
    e1.record()
    barrier()
    _lib.TIMING["on"] = False
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = ms.item()
    launches = sum(_lib.LAUNCHES.values())
    ktimes = _lib.kernel_times()
    clk = clocks.stop() if rank == 0 else None

    # ---- e2e: same step through the public API with HOST ray buffers --------------------------
    pool = []
    for _ in range(4):
        r, d = opt.build_batch(window, ignore_world_cube=True)
        pool.append((r.detach().cpu().pin_memory(), d.detach().cpu().pin_memory()))
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    last = None
    t0.record()
    for i in range(args.steps):
        last = float(step(args.warmup + args.steps + i, pool[i % len(pool)]).detach().cpu())   # D2H of the loss each step
    t1.record()
    barrier()
    ms2 = torch.tensor([t0.elapsed_time(t1)], device=dev)
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_ms = ms2.item()
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    P = n_rays * W["n_samples"]
    dom = max((k for k in ktimes if k in ALGO_BYTES), key=lambda k: ktimes[k][1])
    n_l, t_l = ktimes[dom]
    per_launch_points = P / max(1, (n_rays + W["chunk"] - 1) // W["chunk"])
    achieved = ALGO_BYTES[dom] * per_launch_points / (t_l / n_l * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(dom) if abs(per_launch_points - 16384 * 512) < 1 else None,
                "traffic_note": "DRAM bytes per launch from ncu (profiles/r01_traffic.json); below the algorithmic bytes because the "
                                "57 MB of tables and their gradient stay L2-resident: the gathers / atomics are L2 traffic",
                "peak_source": peak_src, "algorithmic_bytes_per_point": ALGO_BYTES[dom],
                "points_per_launch": per_launch_points, "avg_launch_ms": t_l / n_l,
                "kernels_ms_per_step": {k: round(v[1] / args.steps, 4) for k, v in sorted(ktimes.items())}}
    for k in ALGO_BYTES:
        if k in ktimes and k != dom:
            a = ALGO_BYTES[k] * per_launch_points / (ktimes[k][1] / ktimes[k][0] * 1e-3) / 1e9
            roofline[k + "_achieved_gbs"] = a
            roofline[k + "_frac"] = a / peak
    line = {"metric": METRIC, "value": n_rays * world * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": W["name"], **{k: W[k] for k in ("keyframes", "rays_per_keyframe", "n_samples", "log2_hashmap_size", "chunk")},
                       "rays_per_step_per_gpu": n_rays, "parallelism": f"ray-dp{world}",
                       "pose_refinement": bool(W["pose_grad"]),
                       "l2": "inputs_exceed_l2 (per-step working set ~2.4 GB of encodings + hash tables vs 126 MB L2)"},
            "roofline": roofline,
            "e2e": {"value": n_rays * world * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": n_rays * 14 * 4, "d2h_bytes_per_step": 4, "last_loss": last},
            "gpu_launches": launches, "clocks": clk}
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_steps(3, 1, args.cpu_sample_rays)
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
