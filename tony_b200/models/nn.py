"""nn.Module building blocks backed by the sm_100a kernels (bf16 parameters).

These are what the model zoo (BERT-large, GPT-2 medium, the wide&deep recommender, the
MNIST plumbing model) is assembled from; each forwards to ``tony_b200.ops.functional``.
They stand in for the ``nn.Linear`` / Keras ``Dense`` layers of the reference's example jobs
(tony-examples/mnist-pytorch/mnist_distributed.py:129-136,
tony-examples/horovod-on-tony/tensorflow2_mnist.py:40-52).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from ..ops import functional as F


def _param(*shape, std: Optional[float] = None, fill: Optional[float] = None, device=None):
    t = torch.empty(*shape, dtype=torch.bfloat16, device=device)
    if fill is not None:
        t.fill_(fill)
    else:
        t.normal_(0.0, std if std is not None else 0.02)
    return nn.Parameter(t)


class Linear(nn.Module):
    def __init__(self, in_features: int, out_features: int, bias: bool = True, std: float = 0.02,
                 device=None):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = _param(out_features, in_features, std=std, device=device)
        self.bias = _param(out_features, fill=0.0, device=device) if bias else None

    def forward(self, x, residual=None, act: Optional[str] = None):
        if act in (None, "none"):
            return F.linear(x, self.weight, self.bias, residual)
        y = F.linear_act(x, self.weight, self.bias, act)
        return y if residual is None else y + residual


class LayerNorm(nn.Module):
    def __init__(self, hidden: int, eps: float = 1e-5, device=None):
        super().__init__()
        self.weight = _param(hidden, fill=1.0, device=device)
        self.bias = _param(hidden, fill=0.0, device=device)
        self.eps = eps

    def forward(self, x):
        return F.layer_norm(x, self.weight, self.bias, self.eps)

    def with_residual(self, x):
        """``(layer_norm(x), x)`` as one autograd node for pre-LN blocks (see
        ``functional.layer_norm_residual``): feed the second value to the block's residual input."""
        return F.layer_norm_residual(x, self.weight, self.bias, self.eps)


class Embedding(nn.Module):
    def __init__(self, num: int, dim: int, std: float = 0.02, device=None):
        super().__init__()
        self.weight = _param(num, dim, std=std, device=device)

    def forward(self, idx):
        return F.embedding(self.weight, idx)


class MLP(nn.Module):
    """Transformer feed-forward block with both GEMM epilogues fused (see functional.mlp)."""

    def __init__(self, hidden: int, inner: int, act: str = "gelu", std: float = 0.02,
                 out_std: Optional[float] = None, device=None):
        super().__init__()
        self.fc1 = Linear(hidden, inner, std=std, device=device)
        self.fc2 = Linear(inner, hidden, std=out_std or std, device=device)
        self.act = act

    def forward(self, x, residual=None):
        return F.mlp(x, self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias, residual, self.act)


class BatchNormAct(nn.Module):
    """BatchNorm2d with the following residual add and ReLU fused in (functional.batch_norm_act).
    Training mode uses the hand-written channels-last kernels; evaluation mode (running statistics)
    and unsupported shapes use the ATen ops."""

    def __init__(self, channels: int, relu: bool = True, eps: float = 1e-5, momentum: float = 0.1,
                 zero_init: bool = False, device=None):
        super().__init__()
        self.weight = _param(channels, fill=0.0 if zero_init else 1.0, device=device)
        self.bias = _param(channels, fill=0.0, device=device)
        self.register_buffer("running_mean", torch.zeros(channels, dtype=torch.float32, device=device))
        self.register_buffer("running_var", torch.ones(channels, dtype=torch.float32, device=device))
        self.relu, self.eps, self.momentum = relu, eps, momentum

    def forward(self, x, residual=None, stats=None):
        """``stats``: fp32 ``[2, C]`` per-channel sum / sum of squares of ``x`` when the convolution that
        produced ``x`` accumulated them in its epilogue (then no statistics pass over ``x`` runs)."""
        if self.running_mean.dtype != torch.float32:     # model.to(bfloat16) also casts buffers: undo
            self.running_mean = self.running_mean.float()
            self.running_var = self.running_var.float()
        if self.training and F.batch_norm_act_supported(x):
            return F.batch_norm_act(x, self.weight, self.bias, self.running_mean, self.running_var, residual,
                                    self.eps, self.momentum, self.relu, stats)
        rm, rv = self.running_mean.to(x.dtype), self.running_var.to(x.dtype)
        y = torch.nn.functional.batch_norm(x, rm, rv, self.weight, self.bias, self.training, self.momentum, self.eps)
        if self.training:
            self.running_mean.copy_(rm)
            self.running_var.copy_(rv)
        if residual is not None:
            y = y + residual
        return torch.relu(y) if self.relu else y


class SelfAttention(nn.Module):
    """Fused QKV projection -> SDPA -> output projection with the residual in its epilogue.

    The attention core is ``functional.packed_attention``: the library flash kernel by default,
    the hand-written tcgen05 kernels of csrc/attention.cu with ``TONY_B200_ATTENTION=tcgen05``.
    """

    def __init__(self, hidden: int, heads: int, causal: bool = False, std: float = 0.02,
                 out_std: Optional[float] = None, device=None):
        super().__init__()
        self.heads, self.head_dim, self.causal = heads, hidden // heads, causal
        self.qkv = Linear(hidden, 3 * hidden, std=std, device=device)
        self.out = Linear(hidden, hidden, std=out_std or std, device=device)

    def forward(self, x, residual=None):
        B, S, H = x.shape
        if residual is x and x.is_cuda and x.dtype == torch.bfloat16:
            # post-LN block: one autograd node, residual gradient fused into the last dgrad GEMM
            return F.attention_block(x, self.qkv.weight, self.qkv.bias, self.out.weight, self.out.bias,
                                     self.heads, self.causal)
        qkv = self.qkv(x).view(B, S, 3, self.heads, self.head_dim)
        o = F.packed_attention(qkv, self.causal)
        return self.out(o, residual=residual)


def count_params(module: nn.Module) -> int:
    seen, n = set(), 0
    for p in module.parameters():
        if id(p) not in seen:
            seen.add(id(p))
            n += p.numel()
    return n


def init_scale(layers: int) -> float:
    return 0.02 / math.sqrt(2.0 * layers)
