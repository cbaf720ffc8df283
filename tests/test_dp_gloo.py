"""Ray data-parallel host logic on CPU, world_size 2, gloo backend.

The CUDA kernels cannot run here, so each rank evaluates its ray shard with the oracle and then goes
through the SAME host-side composition the product uses: loner_b200.mapping.RayDataParallel for the
opaque-ray count / gradient all-reduce, and the loss formula of loner_b200.functional.RenderLoss
(partial sums divided by GLOBAL normalisers).  The summed result must equal the oracle's single-process
loss and gradient on the whole batch -- i.e. an N-GPU run reproduces the 1-GPU step.
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _partial_sums(O, g, params, sl):
    """loss_parts of lb_composite_loss for the ray slice `sl`, evaluated with the oracle."""
    rays, depths = g["rays"][sl], g["depths"][sl]
    cfg = dict(O.DEFAULT_LOSS_CFG, loss_selection=g["loss_selection"], min_js_score=g["min_js"])
    out = O.mapping_forward(rays, depths, params, g["occ"], g["u1"][sl], g["u2"][sl], g["noise"][sl], g["scale"],
                            g["n_samples"], log2_hashmap_size=g["log2_T"], cfg=cfg, sampler=g["sampler"])
    s = g["scale"]
    op = out["opaque"]
    gt = depths * s
    parts0 = (((out["depth"] * s) - gt)[op] ** 2).sum()
    parts1 = (out["weights"] - out["w_gt"]).abs().sum()
    parts2 = (out["opacity"][op] - 1).abs().sum()
    return parts0, parts1, parts2, op.sum().float(), cfg, out


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from conftest import load_golden, golden_params
        from loner_b200.mapping import RayDataParallel
        from oracle import loner_oracle as O
        torch.set_num_threads(2)
        g = load_golden("c1_ogm_l1js")
        dp = RayDataParallel()
        assert dp.enabled and dp.world_size == world and dp.rank == rank
        n = g["rays"].shape[0]
        per = n // world
        sl = slice(rank * per, (rank + 1) * per)
        params = golden_params(g).requires_grad_(True)
        p0, p1, p2, n_op_local, cfg, _ = _partial_sums(O, g, params, sl)
        # global normalisers exactly as Optimizer.compute_loss forms them
        n_op = dp.all_reduce_(n_op_local.clone().reshape(1))[0]
        n_total = float(per * g["n_samples"] * dp.world_size)
        loss = cfg["depthloss_lambda"] * p0 / n_op + cfg["los_lambda"] * p1 / n_total + p2 / n_op
        loss.backward()
        dp.all_reduce_grads([params])
        total = dp.all_reduce_(loss.detach().clone().reshape(1))[0]
        if rank == 0:
            ret["loss"], ret["grad"], ret["n_op"] = float(total), params.grad.clone(), float(n_op)
    finally:
        dist.destroy_process_group()


def test_two_rank_step_equals_single_process():
    sys.path.insert(0, ROOT)
    from conftest import load_golden, golden_params
    from oracle import loner_oracle as O
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    g = load_golden("c1_ogm_l1js")
    params = golden_params(g).requires_grad_(True)
    p0, p1, p2, n_op, cfg, out = _partial_sums(O, g, params, slice(0, g["rays"].shape[0]))
    out["loss"].backward()
    assert ret["n_op"] == float(n_op)
    assert abs(ret["loss"] - out["loss"].item()) < 1e-5 * abs(out["loss"].item())
    # and the reference's own recorded loss for this batch
    assert abs(ret["loss"] - g["loss"].item()) < 1e-4 * abs(g["loss"].item())
    err = (ret["grad"] - params.grad).abs().max() / params.grad.abs().max()
    assert err < 1e-4, err


def test_single_process_is_a_noop():
    sys.path.insert(0, ROOT)
    from loner_b200.mapping import RayDataParallel
    dp = RayDataParallel()
    assert not dp.enabled and dp.world_size == 1 and dp.rank == 0
    t = torch.ones(3)
    assert dp.all_reduce_(t) is t and torch.equal(t, torch.ones(3))


def _adam_worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from loner_b200.mapping import RayDataParallel
        from loner_b200.optim import FusedAdam
        dp = RayDataParallel(shard_batch=True)
        # row ranges of a sharded chunk: disjoint, ordered, cover [0, n)
        assert (dp.rows(10), dp.rows(7)) == (((0, 5), (0, 4)) if rank == 0 else ((5, 10), (4, 7)))
        lo, hi = dp.rows(1024)
        assert (lo, hi) == (rank * 512, (rank + 1) * 512)
        g0 = torch.Generator().manual_seed(5)
        big = torch.nn.Parameter(torch.randn(8192, generator=g0))          # sharded: reduce-scatter -> Adam -> all-gather
        pose = torch.nn.Parameter(torch.randn(6, generator=g0))            # small: all-reduce + full step
        # overlap_step as Optimizer passes it: on CPU tensors there is no side stream, the step runs in line and the fence
        # (wait_params) has nothing to wait for
        opt = FusedAdam([{"params": [big], "lr": 0.01}, {"params": [pose], "lr": 0.001}], data_parallel=dp, overlap_step=True)
        gr = torch.Generator().manual_seed(100 + rank)
        for _ in range(3):
            big.grad = torch.randn(8192, generator=gr)
            pose.grad = torch.randn(6, generator=gr)
            opt.step()
            opt.zero_grad(set_to_none=True)
            assert dp._param_work == []
            dp.wait_params()
        assert opt.state[big]["exp_avg"].numel() == 8192 // world          # moments exist for this rank's slice only
        if rank == 0:
            ret["big"], ret["pose"] = big.detach().clone(), pose.detach().clone()
    finally:
        dist.destroy_process_group()


def test_sharded_adam_equals_single_process_adam():
    """FusedAdam(data_parallel=...) on 2 gloo ranks (host logic + the torch update formula; the CUDA kernel is the same
    arithmetic) against torch.optim.Adam on the summed gradients."""
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_adam_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    g0 = torch.Generator().manual_seed(5)
    big = torch.nn.Parameter(torch.randn(8192, generator=g0))
    pose = torch.nn.Parameter(torch.randn(6, generator=g0))
    ref = torch.optim.Adam([{"params": [big], "lr": 0.01}, {"params": [pose], "lr": 0.001}])
    gens = [torch.Generator().manual_seed(100 + r) for r in range(world)]
    for _ in range(3):
        gb = [torch.randn(8192, generator=g) for g in gens]
        gp = [torch.randn(6, generator=g) for g in gens]
        # per rank the draw order is big then pose
        big.grad, pose.grad = gb[0] + gb[1], gp[0] + gp[1]
        ref.step()
        ref.zero_grad(set_to_none=True)
    torch.testing.assert_close(ret["big"], big.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(ret["pose"], pose.detach(), rtol=1e-5, atol=1e-6)
