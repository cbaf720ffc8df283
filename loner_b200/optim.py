"""Fused optimiser steps on the flat tcnn parameter vectors.

`FusedAdam` has torch.optim.Adam's defaults and update rule (the reference builds
`torch.optim.Adam([{'params': sigma params, 'lr': 0.01}, {'params': poses, 'lr': 0.001}])` afresh for
every phase, src/mapping/optimizer.py:255-265) but runs one lb_adam_step kernel per CUDA parameter
(p, g, m, v streamed once).  Tiny CPU/GPU tensors that are not 16-byte-aligned fp32 CUDA buffers
(the pose 6-vectors) are stepped with the same formula in torch.
"""
import math

import torch

from . import ops


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None, grad_scale=1.0):
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] += 1
                g = p.grad
                if p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and g.is_contiguous() \
                        and p.data_ptr() % 16 == 0 and g.data_ptr() % 16 == 0 and p.numel() >= 1024:
                    ops.adam_step(p.data.view(-1), g.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1),
                                  st["step"], group["lr"], b1, b2, group["eps"], grad_scale=grad_scale)
                else:
                    if grad_scale != 1.0:
                        g = g * grad_scale
                    m, v = st["exp_avg"], st["exp_avg_sq"]
                    m.mul_(b1).add_(g, alpha=1 - b1)
                    v.mul_(b2).addcmul_(g, g, value=1 - b2)
                    bc1, bc2 = 1 - b1 ** st["step"], 1 - b2 ** st["step"]
                    denom = (v.sqrt() / math.sqrt(bc2)).add_(group["eps"])
                    p.addcdiv_(m, denom, value=-group["lr"] / bc1)
        return None
