"""Flat-buffer optimizers over the sm_100a kernels (single-GPU form of K7).

State layout: fp32 master weights + moments as flat 1-D tensors, bf16 gradients in,
bf16 weights out.  The multi-GPU form (gradient collective fused with this update) is
``tony_b200.parallel.fused_optimizer``; both call the same device update code.
Replaces ``torch.optim.AdamW(fused=True)`` / ``torch.optim.SGD`` in launched jobs
(reference user code: tony-examples/mnist-pytorch/mnist_distributed.py:204-206).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import check, cur_stream, ptr


class FlatAdamW:
    def __init__(self, master: torch.Tensor, lr: float = 1e-3, betas: Tuple[float, float] = (0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 0.01,
                 wd_mask: Optional[torch.Tensor] = None):
        if master.dtype != torch.float32 or master.dim() != 1 or master.numel() % 8:
            raise ValueError("master must be a flat fp32 tensor with numel % 8 == 0")
        self.master = master
        self.exp_avg = torch.zeros_like(master)
        self.exp_avg_sq = torch.zeros_like(master)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.wd_mask = wd_mask  # uint8, one byte per 256 elements; None = decay everything
        self.t = 0

    def step(self, grad_bf16: torch.Tensor, param_bf16: torch.Tensor, grad_scale: float = 1.0):
        self.t += 1
        b1, b2 = self.betas
        check(_lib.lib().tony_adamw_flat(
            ptr(self.master), ptr(self.exp_avg), ptr(self.exp_avg_sq), ptr(grad_bf16),
            ptr(param_bf16), self.master.numel(), self.lr, b1, b2, self.eps, self.weight_decay,
            1.0 - b1 ** self.t, 1.0 - b2 ** self.t, grad_scale, ptr(self.wd_mask), 0, cur_stream()),
            "tony_adamw_flat")

    def state_dict(self):
        return {"master": self.master, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq,
                "t": self.t}

    def load_state_dict(self, sd):
        self.master.copy_(sd["master"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.t = int(sd["t"])


class FlatSGD:
    def __init__(self, master: torch.Tensor, lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 0.0, nesterov: bool = False,
                 wd_mask: Optional[torch.Tensor] = None):
        if master.dtype != torch.float32 or master.dim() != 1 or master.numel() % 8:
            raise ValueError("master must be a flat fp32 tensor with numel % 8 == 0")
        self.master = master
        self.momentum_buf = torch.zeros_like(master)
        self.lr, self.momentum, self.weight_decay, self.nesterov = lr, momentum, weight_decay, nesterov
        self.wd_mask = wd_mask
        self.t = 0

    def step(self, grad_bf16: torch.Tensor, param_bf16: torch.Tensor, grad_scale: float = 1.0):
        self.t += 1
        check(_lib.lib().tony_sgd_flat(
            ptr(self.master), ptr(self.momentum_buf), ptr(grad_bf16), ptr(param_bf16),
            self.master.numel(), self.lr, self.momentum, self.weight_decay, int(self.nesterov),
            grad_scale, ptr(self.wd_mask), 0, cur_stream()), "tony_sgd_flat")

    def state_dict(self):
        return {"master": self.master, "momentum_buf": self.momentum_buf, "t": self.t}

    def load_state_dict(self, sd):
        self.master.copy_(sd["master"])
        self.momentum_buf.copy_(sd["momentum_buf"])
        self.t = int(sd["t"])
