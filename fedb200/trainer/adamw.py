"""AdamW with the semantics of the (removed) ``transformers.AdamW`` the reference imports
(src/Servercase/server_IID_IMDB.py:36, used at :109): betas (0.9, 0.999), **eps 1e-6 added to
sqrt(v) before bias correction**, bias correction folded into the step size, decoupled weight
decay (0.0 by default) applied after the update.  ``torch.optim.AdamW`` differs in eps
placement and default eps/decay, so it is not a drop-in.

``FlatAdamW`` runs over the flat arena: one fused pass over ``D`` elements (the sm_100a
kernel ``adamw_flat`` on CUDA, vectorised torch ops on CPU) instead of a Python loop over 201
tensors.  The reference re-creates the optimizer on every ``train`` call, so state never
survives a round (:109) — ``reset()`` reproduces that.
"""
from __future__ import annotations

import math
from typing import Optional

import torch


def hf_adamw_reference_step(p, g, m, v, step: int, lr=5e-5, b1=0.9, b2=0.999, eps=1e-6, wd=0.0):
    """Single-tensor oracle (fp32, in place); ``step`` is 1-based."""
    m.mul_(b1).add_(g, alpha=1.0 - b1)
    v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
    denom = v.sqrt().add_(eps)
    step_size = lr * math.sqrt(1.0 - b2 ** step) / (1.0 - b1 ** step)
    p.addcdiv_(m, denom, value=-step_size)
    if wd > 0.0:
        p.add_(p, alpha=-lr * wd)


class FlatAdamW:
    def __init__(self, flat_param: torch.Tensor, flat_grad: torch.Tensor, lr=5e-5, betas=(0.9, 0.999), eps=1e-6,
                 weight_decay=0.0, bf16_shadow: Optional[torch.Tensor] = None, use_kernel: Optional[bool] = None):
        assert flat_param.shape == flat_grad.shape and flat_param.dtype == torch.float32
        self.p, self.g = flat_param, flat_grad
        self.m = torch.zeros_like(flat_param)
        self.v = torch.zeros_like(flat_param)
        self.lr, self.betas, self.eps, self.wd = lr, betas, eps, weight_decay
        self.shadow = bf16_shadow
        self.t = 0
        self.use_kernel = flat_param.is_cuda if use_kernel is None else use_kernel

    def reset(self) -> None:
        self.m.zero_()
        self.v.zero_()
        self.t = 0

    def step(self, zero_grad: bool = True, grad_scale: float = 1.0) -> None:
        self.t += 1
        b1, b2 = self.betas
        if self.use_kernel:
            from ..ops import cuda_ops
            cuda_ops.adamw_flat(self.p, self.g, self.m, self.v, self.shadow, self.t, self.lr, b1, b2, self.eps,
                                self.wd, grad_scale, zero_grad)
            return
        g = self.g if grad_scale == 1.0 else self.g * grad_scale
        hf_adamw_reference_step(self.p, g, self.m, self.v, self.t, self.lr, b1, b2, self.eps, self.wd)
        if self.shadow is not None:
            self.shadow.copy_(self.p)
        if zero_grad:
            self.g.zero_()
