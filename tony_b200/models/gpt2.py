"""GPT-2 (medium by default) causal LM on the sm_100a kernels.

BASELINE.json config "GPT-2 medium DDP on 8xB200 with chief-task failure and restart":
24 layers, hidden 1024, 16 heads, FFN 4096, vocab 50257 (padded to 50304), sequence 1024,
pre-LayerNorm, tanh-GELU, tied LM head.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
from torch import nn

from ..ops import functional as F
from .nn import MLP, Embedding, LayerNorm, SelfAttention, init_scale


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    hidden: int = 1024
    layers: int = 24
    heads: int = 16
    inner: int = 4096
    max_positions: int = 1024
    eps: float = 1e-5

    @property
    def vocab_padded(self) -> int:
        return (self.vocab_size + 63) // 64 * 64

    @staticmethod
    def medium(**kw):
        return GPT2Config(**kw)

    @staticmethod
    def small(**kw):
        return GPT2Config(hidden=768, layers=12, heads=12, inner=3072, **kw)

    @staticmethod
    def tiny(**kw):
        return GPT2Config(vocab_size=1000, hidden=256, layers=2, heads=4, inner=1024, max_positions=128, **kw)


class GPT2Block(nn.Module):
    def __init__(self, c: GPT2Config, device=None):
        super().__init__()
        out_std = init_scale(c.layers)
        self.ln1 = LayerNorm(c.hidden, c.eps, device=device)
        self.attn = SelfAttention(c.hidden, c.heads, causal=True, out_std=out_std, device=device)
        self.ln2 = LayerNorm(c.hidden, c.eps, device=device)
        self.mlp = MLP(c.hidden, c.inner, "gelu_tanh", out_std=out_std, device=device)

    def forward(self, x):
        # pre-LN: the residual branch leaves through the LayerNorm node, so its gradient is added inside
        # the LayerNorm backward kernel instead of by a separate elementwise add
        h, r = self.ln1.with_residual(x)
        x = self.attn(h, residual=r)
        h, r = self.ln2.with_residual(x)
        return self.mlp(h, residual=r)


class GPT2LMHead(nn.Module):
    def __init__(self, c: GPT2Config, device=None):
        super().__init__()
        self.config = c
        self.wte = Embedding(c.vocab_padded, c.hidden, device=device)
        self.wpe = Embedding(c.max_positions, c.hidden, std=0.01, device=device)
        self.blocks = nn.ModuleList(GPT2Block(c, device=device) for _ in range(c.layers))
        self.ln_f = LayerNorm(c.hidden, c.eps, device=device)
        self.shared_params = {self.wte.weight: 2}

    def forward(self, input_ids, labels):
        B, S = input_ids.shape
        if F.embedding_sum_supported(self.wte.weight) and input_ids.dtype == torch.int64:
            x = F.embedding_sum(input_ids, None, self.wte.weight, None, self.wpe.weight)
        else:
            x = self.wte(input_ids) + self.wpe(torch.arange(S, device=input_ids.device))
        for blk in self.blocks:
            x = blk(x)
        x = self.ln_f(x)
        logits = F.linear(x.view(B * S, -1), self.wte.weight)
        return F.softmax_cross_entropy(logits, labels.reshape(-1), vocab=self.config.vocab_size)


def synthetic_batch(c: GPT2Config, batch: int, seq: int, device="cpu", pin: bool = False, seed: int = 0):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, c.vocab_size, (batch, seq + 1), generator=g)
    out = {"input_ids": ids[:, :-1].contiguous(), "labels": ids[:, 1:].contiguous()}
    if pin:
        out = {k: v.pin_memory() for k, v in out.items()}
    if device != "cpu":
        out = {k: v.to(device, non_blocking=True) for k, v in out.items()}
    return out


def train_flops_per_sample(c: GPT2Config, seq: int) -> float:
    per_layer = 4 * c.hidden * c.hidden + 2 * c.hidden * c.inner
    return 6.0 * (per_layer * c.layers + c.hidden * c.vocab_padded) * seq + 3.0 * 2.0 * seq * seq * c.hidden * c.layers
