"""Fused optimiser steps on the flat tcnn parameter vectors.

`FusedAdam` has torch.optim.Adam's defaults and update rule (the reference builds
`torch.optim.Adam([{'params': sigma params, 'lr': 0.01}, {'params': poses, 'lr': 0.001}])` afresh for
every phase, src/mapping/optimizer.py:255-265) but runs one lb_adam_step kernel per CUDA parameter
(p, g, m, v streamed once).  Tiny CPU/GPU tensors that are not 16-byte-aligned fp32 CUDA buffers
(the pose 6-vectors) are stepped with the same formula in torch.

Ray data parallelism (`data_parallel=` a loner_b200.mapping.RayDataParallel): the gradient exchange is
part of the step.  A flat parameter vector whose length divides by 4 * world_size is handled ZeRO-1 style:
    reduce-scatter(grad) -> Adam on this rank's 1/W slice (moments exist for the slice only) -> all-gather(params)
which moves the same bytes over NVLink as an all-reduce but cuts the optimiser's HBM traffic and moment memory by W
(T = 2^22: 2.95 GB of Adam traffic and 844 MB of moments per step and rank otherwise).  Everything else (pose 6-vectors)
is all-reduced and stepped in full on every rank.  The update of an element is the same arithmetic in either layout.
"""
import math

import torch

from . import ops


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, data_parallel=None, overlap_step=False):
        """overlap_step: the step of the large flat parameter vectors (gradient reduce-scatter, lb_adam_step,
        parameter all-gather) runs on a side stream and is still in flight when step() returns, so that it overlaps what
        the caller launches next without reading those parameters -- the next iteration's ray build, normaliser count
        and sampler (issue-bound kernels; Adam is an HBM stream, the collectives are NVLink).  The caller must run
        data_parallel.wait_params() before the parameters are read again (Optimizer.compute_loss does); `data_parallel`
        (a RayDataParallel, enabled or not) holds that fence.  Small tensors (pose 6-vectors, which the next ray build
        reads) are always stepped on the caller's stream.  Off by default: step() returns with complete parameters."""
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self._fence = data_parallel
        self._dp = data_parallel if (data_parallel is not None and data_parallel.enabled) else None
        self._overlap = bool(overlap_step) and data_parallel is not None
        self._side = {}

    @staticmethod
    def _fast(p, g):
        return p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous() \
            and p.data_ptr() % 16 == 0 and g.data_ptr() % 16 == 0 and p.numel() >= 1024

    @staticmethod
    def _update(p, g, m, v, step, lr, b1, b2, eps, grad_scale):
        """One Adam update of the flat views p, g, m, v (same length), in place on p, m, v."""
        if FusedAdam._fast(p, g):
            ops.adam_step(p, g, m, v, step, lr, b1, b2, eps, grad_scale=grad_scale)
            return
        if grad_scale != 1.0:
            g = g * grad_scale
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
        denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
        p.addcdiv_(m, denom, value=-lr / bc1)

    @torch.no_grad()
    def _step_small(self, group, params, grad_scale):
        """All small tensors of a group (pose 6-vectors) in a handful of multi-tensor launches: one flat all-reduce of
        their gradients when data-parallel, then the Adam formula through torch._foreach_* (one launch per operation for
        ALL tensors instead of ~10 launches per tensor)."""
        dp = self._dp
        b1, b2 = group["betas"]
        grads = [p.grad for p in params]
        if dp is not None:
            flat = torch.cat([g.reshape(-1) for g in grads])
            dp.all_reduce_(flat, "all_reduce_grad")
            grads = [t.view_as(g) for t, g in zip(flat.split([g.numel() for g in grads]), grads)]
        if grad_scale != 1.0:
            grads = torch._foreach_mul(grads, grad_scale)
        ms, vs, step = [], [], None
        for p in params:
            st = self.state[p]
            if not st:
                st["step"] = 0
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
            st["step"] += 1
            ms.append(st["exp_avg"]); vs.append(st["exp_avg_sq"])
        steps = {self.state[p]["step"] for p in params}
        if len(steps) != 1:                       # tensors joined the group at different times: per-tensor formula
            for p, g, m, v in zip(params, grads, ms, vs):
                self._update(p, g, m, v, self.state[p]["step"], group["lr"], b1, b2, group["eps"], 1.0)
            return
        step = steps.pop()
        torch._foreach_mul_(ms, b1)
        torch._foreach_add_(ms, grads, alpha=1 - b1)
        torch._foreach_mul_(vs, b2)
        torch._foreach_addcmul_(vs, grads, grads, value=1 - b2)
        bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
        denom = torch._foreach_sqrt(vs)
        torch._foreach_div_(denom, math.sqrt(bc2))
        torch._foreach_add_(denom, group["eps"])
        torch._foreach_addcdiv_(list(params), ms, denom, value=-group["lr"] / bc1)

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        if self._fence is not None:
            self._fence.wait_params()  # a step still in flight writes the parameters / moments this one updates
        large = []
        for group in self.param_groups:
            small = [p for p in group["params"] if p.grad is not None and p.numel() < 1024]
            if small:
                self._step_small(group, small, grad_scale)
            large += [(group, p) for p in group["params"] if p.grad is not None and p.numel() >= 1024]
        if not large:
            return None
        dev = large[0][1].device
        timing = self._dp is not None and self._dp.events is not None     # bench.py brackets the collectives: keep them in line
        if not (self._overlap and dev.type == "cuda" and not timing):
            for group, p in large:
                self._step_large(group, p, grad_scale)
            return None
        main = torch.cuda.current_stream(dev)
        side = self._side.get(dev)
        if side is None:
            side = self._side[dev] = torch.cuda.Stream(dev)
        side.wait_stream(main)                                            # the gradients are complete
        with torch.cuda.stream(side):
            for group, p in large:
                self._step_large(group, p, grad_scale)
        for _, p in large:
            p.grad.record_stream(side)         # zero_grad(set_to_none=True) may free it while the side stream still reads it
        done = torch.cuda.Event()
        done.record(side)
        self._fence.defer(done)
        return None

    def _step_large(self, group, p, grad_scale):
        dp = self._dp
        b1, b2 = group["betas"]
        st = self.state[p]
        g = p.grad
        n = p.numel()
        sharded = dp is not None and n >= 4096 and n % (4 * dp.world_size) == 0 and p.is_contiguous() and g.is_contiguous()
        if not st:
            st["step"] = 0
            if sharded:
                k = n // dp.world_size
                st["exp_avg"] = torch.zeros(k, device=p.device, dtype=p.dtype)
                st["exp_avg_sq"] = torch.zeros(k, device=p.device, dtype=p.dtype)
                st["grad_shard"] = torch.empty(k, device=p.device, dtype=p.dtype)
            else:
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
        st["step"] += 1
        args = (st["step"], group["lr"], b1, b2, group["eps"], grad_scale)
        if sharded:
            k = n // dp.world_size
            flat = p.data.view(-1)
            mine = flat[dp.rank * k:(dp.rank + 1) * k]
            dp.reduce_scatter(st["grad_shard"], g.view(-1))
            self._update(mine, st["grad_shard"], st["exp_avg"], st["exp_avg_sq"], *args)
            dp.all_gather(flat, mine)
            return
        if dp is not None:
            dp.all_reduce_(g, "all_reduce_grad")
        if self._fast(p, g):
            self._update(p.data.view(-1), g.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1), *args)
        else:
            self._update(p, g, st["exp_avg"], st["exp_avg_sq"], *args)
