"""Wide-and-deep / MLP recommender for the ps/worker configuration.

BASELINE.json config "ps/worker job (2 ps + 6 workers) training a wide-and-deep/MLP
recommender with sharded parameter updates".  Sparse side: one embedding table per categorical
field, summed ("wide" = a 1-dim table per field, "deep" = D-dim tables).  Dense side: an MLP
tower whose big matrices are :class:`tony_b200.parallel.ps.ShardedLinear` — rows owned by the
ranks, pulled by the all-gather-fused GEMM, updated by the sharded fused optimizer.
Binary click label, logistic loss computed as a 2-class softmax cross-entropy (our CE kernel).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import torch
from torch import nn

from ..ops import functional as F
from ..parallel.ps import ShardedLinear
from .nn import Embedding, Linear


@dataclass
class WideDeepConfig:
    num_fields: int = 26
    vocab_per_field: int = 100_000
    embed_dim: int = 128
    dense_features: int = 16          # padded to a multiple of 8 by the model
    hidden: List[int] = field(default_factory=lambda: [4096, 4096, 2048, 1024])

    @staticmethod
    def tiny():
        return WideDeepConfig(num_fields=4, vocab_per_field=1000, embed_dim=64, dense_features=8,
                              hidden=[512, 256])


class WideDeep(nn.Module):
    """``sparse="bag"`` (default on CUDA) keeps the categorical tables row-sharded across the ranks
    (:class:`tony_b200.parallel.ps.ShardedEmbeddingBag`: lookups pull rows from their owners over
    NVLink, gradients are reduced into the owners' shards, each owner updates only its rows);
    ``sparse="dense"`` is the replicated nn.Embedding-style layout with dense gradients."""

    def __init__(self, c: WideDeepConfig, device=None, sharded: bool = True, sparse: str = "auto", roles=None):
        super().__init__()
        self.config = c
        if sparse == "auto":
            sparse = "bag" if (device is not None and torch.device(device).type == "cuda") else "dense"
        self.sparse = sparse
        if sparse == "bag":
            from ..parallel.ps import ShardedEmbeddingBag

            self.bag = ShardedEmbeddingBag(c.num_fields, c.vocab_per_field, c.embed_dim, 8, std=0.01, device=device,
                                           roles=roles)
        else:
            self.deep_tables = nn.ModuleList(Embedding(c.vocab_per_field, c.embed_dim, std=0.01, device=device)
                                             for _ in range(c.num_fields))
            self.wide_tables = nn.ModuleList(Embedding(c.vocab_per_field, 8, std=0.01, device=device)
                                             for _ in range(c.num_fields))
        in_dim = c.num_fields * c.embed_dim + c.dense_features
        self.in_dim = (in_dim + 7) // 8 * 8
        layers = []
        prev = self.in_dim
        for h in c.hidden:
            layers.append(ShardedLinear(prev, h, act="relu", device=device) if sharded
                          else _ActLinear(prev, h, "relu", device))
            prev = h
        self.tower = nn.ModuleList(layers)
        self.head = Linear(prev, 8, device=device)      # 2 classes padded to 8 columns

    def forward(self, sparse_ids, dense, labels):
        """sparse_ids [B, fields] int64, dense [B, dense_features] bf16, labels [B] int64."""
        B = sparse_ids.shape[0]
        if self.sparse == "bag":
            deep_cat, rest = self.bag(sparse_ids)
            x = torch.cat([deep_cat, dense], dim=1)
            wide = rest.sum(dim=1)                                                   # [B, 8], cols 0..1 used
        else:
            deep = [t(sparse_ids[:, i]) for i, t in enumerate(self.deep_tables)]
            x = torch.cat(deep + [dense], dim=1)
            wide = sum(t(sparse_ids[:, i]) for i, t in enumerate(self.wide_tables))
        if x.shape[1] < self.in_dim:
            x = torch.nn.functional.pad(x, (0, self.in_dim - x.shape[1]))
        for layer in self.tower:
            x = layer(x)
        logits = self.head(x)
        return F.softmax_cross_entropy(logits + wide, labels, vocab=2)


class _ActLinear(nn.Module):
    def __init__(self, i, o, act, device):
        super().__init__()
        self.lin = Linear(i, o, device=device)
        self.act = act

    def forward(self, x):
        return self.lin(x, act=self.act)


def synthetic_batch(c: WideDeepConfig, batch: int, device="cpu", pin: bool = False, seed: int = 0):
    g = torch.Generator().manual_seed(seed)
    out = {
        "sparse_ids": torch.randint(0, c.vocab_per_field, (batch, c.num_fields), generator=g),
        "dense": torch.randn(batch, c.dense_features, generator=g).to(torch.bfloat16),
        "labels": torch.randint(0, 2, (batch,), generator=g),
    }
    if pin:
        out = {k: v.pin_memory() for k, v in out.items()}
    if device != "cpu":
        out = {k: v.to(device, non_blocking=True) for k, v in out.items()}
    return out


class StockWideDeep(nn.Module):
    """Baseline twin on stock PyTorch modules (fp32 params, bf16 autocast, replicated weights)."""

    def __init__(self, c: WideDeepConfig):
        super().__init__()
        self.config = c
        self.deep_tables = nn.ModuleList(nn.Embedding(c.vocab_per_field, c.embed_dim) for _ in range(c.num_fields))
        self.wide_tables = nn.ModuleList(nn.Embedding(c.vocab_per_field, 2) for _ in range(c.num_fields))
        prev = c.num_fields * c.embed_dim + c.dense_features
        layers = []
        for h in c.hidden:
            layers += [nn.Linear(prev, h), nn.ReLU()]
            prev = h
        self.tower = nn.Sequential(*layers)
        self.head = nn.Linear(prev, 2)

    def forward(self, sparse_ids, dense, labels):
        deep = [t(sparse_ids[:, i]) for i, t in enumerate(self.deep_tables)]
        x = torch.cat(deep + [dense.float()], dim=1)
        logits = self.head(self.tower(x))
        wide = sum(t(sparse_ids[:, i]) for i, t in enumerate(self.wide_tables))
        return torch.nn.functional.cross_entropy((logits + wide).float(), labels)
