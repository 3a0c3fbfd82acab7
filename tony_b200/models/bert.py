"""BERT (base/large) for MLM+NSP pre-training on the sm_100a kernels.

BASELINE.json config "horovod-on-tony BERT-large pretraining bf16 with tensor fusion":
24 layers, hidden 1024, 16 heads, FFN 4096, vocab 30522 (padded to 30528 for 16-byte
rows), sequence 512, post-LayerNorm, erf-GELU, tied decoder, NSP head.  MLM logits are
computed only at the ``max_predictions`` masked positions (the phase-2 pre-training
convention); the stock arm in ``models/stock.py`` does identical math.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
from torch import nn

from ..ops import functional as F
from .nn import MLP, Embedding, LayerNorm, Linear, SelfAttention, _param


@dataclass
class BertConfig:
    vocab_size: int = 30522
    hidden: int = 1024
    layers: int = 24
    heads: int = 16
    inner: int = 4096
    max_positions: int = 512
    type_vocab: int = 2
    eps: float = 1e-12
    max_predictions: int = 80

    @property
    def vocab_padded(self) -> int:
        return (self.vocab_size + 7) // 8 * 8

    @staticmethod
    def large(**kw):
        return BertConfig(**kw)

    @staticmethod
    def base(**kw):
        return BertConfig(hidden=768, layers=12, heads=12, inner=3072, **kw)

    @staticmethod
    def tiny(**kw):
        return BertConfig(vocab_size=1000, hidden=256, layers=2, heads=4, inner=1024,
                          max_positions=128, max_predictions=8, **kw)


class BertLayer(nn.Module):
    def __init__(self, c: BertConfig, device=None):
        super().__init__()
        self.attn = SelfAttention(c.hidden, c.heads, device=device)
        self.ln1 = LayerNorm(c.hidden, c.eps, device=device)
        self.mlp = MLP(c.hidden, c.inner, "gelu", device=device)
        self.ln2 = LayerNorm(c.hidden, c.eps, device=device)

    def forward(self, x):
        x = self.ln1(self.attn(x, residual=x))
        return self.ln2(self.mlp(x, residual=x))


class BertForPreTraining(nn.Module):
    def __init__(self, c: BertConfig, device=None):
        super().__init__()
        self.config = c
        self.word = Embedding(c.vocab_padded, c.hidden, device=device)
        self.pos = Embedding(c.max_positions, c.hidden, device=device)
        self.tok_type = Embedding(c.type_vocab, c.hidden, device=device)
        self.emb_ln = LayerNorm(c.hidden, c.eps, device=device)
        self.layers = nn.ModuleList(BertLayer(c, device=device) for _ in range(c.layers))
        self.mlm_dense = Linear(c.hidden, c.hidden, device=device)
        self.mlm_ln = LayerNorm(c.hidden, c.eps, device=device)
        self.mlm_bias = _param(c.vocab_padded, fill=0.0, device=device)
        self.pooler = Linear(c.hidden, c.hidden, device=device)
        self.nsp = Linear(c.hidden, 8, device=device)  # 2 classes, padded to 8 columns
        # parameters with more than one gradient contribution per step
        self.shared_params = {self.word.weight: 2}

    def forward(self, input_ids, token_type_ids, masked_positions, masked_labels, nsp_labels):
        """Returns the summed MLM + NSP loss (scalar, fp32)."""
        B, S = input_ids.shape
        if F.embedding_sum_supported(self.word.weight, self.tok_type.weight) and input_ids.dtype == torch.int64:
            x = F.embedding_sum(input_ids, token_type_ids, self.word.weight, self.tok_type.weight,
                                self.pos.weight)
        else:
            pos_ids = torch.arange(S, device=input_ids.device)
            x = self.word(input_ids) + self.tok_type(token_type_ids)
            x = x + self.pos(pos_ids)
        x = self.emb_ln(x)
        for layer in self.layers:
            x = layer(x)
        H = x.shape[-1]
        flat = x.view(B * S, H)
        gather = (masked_positions + torch.arange(B, device=x.device).unsqueeze(1) * S).reshape(-1)
        hm = flat.index_select(0, gather)
        hm = self.mlm_ln(self.mlm_dense(hm, act="gelu"))
        logits = F.linear(hm, self.word.weight, self.mlm_bias)
        mlm_loss = F.softmax_cross_entropy(logits, masked_labels.reshape(-1),
                                           vocab=self.config.vocab_size, ignore_index=-100)
        pooled = torch.tanh(self.pooler(x[:, 0].contiguous()))
        nsp_logits = self.nsp(pooled)
        nsp_loss = F.softmax_cross_entropy(nsp_logits, nsp_labels, vocab=2)
        return mlm_loss + nsp_loss


def synthetic_batch(c: BertConfig, batch: int, seq: int, device="cpu", pin: bool = False, seed: int = 0):
    """Random token ids / masked positions of the named shape (there is no dataset access)."""
    g = torch.Generator().manual_seed(seed)
    out = {
        "input_ids": torch.randint(0, c.vocab_size, (batch, seq), generator=g),
        "token_type_ids": torch.randint(0, c.type_vocab, (batch, seq), generator=g),
        "masked_positions": torch.stack([torch.randperm(seq, generator=g)[: c.max_predictions].sort().values
                                         for _ in range(batch)]),
        "masked_labels": torch.randint(0, c.vocab_size, (batch, c.max_predictions), generator=g),
        "nsp_labels": torch.randint(0, 2, (batch,), generator=g),
    }
    if pin:
        out = {k: v.pin_memory() for k, v in out.items()}
    if device != "cpu":
        out = {k: v.to(device, non_blocking=True) for k, v in out.items()}
    return out


def train_flops_per_sample(c: BertConfig, seq: int) -> float:
    """Model FLOPs of one fwd+bwd sample (6 x params x tokens + attention), for roofline maths."""
    per_layer = 4 * c.hidden * c.hidden + 2 * c.hidden * c.inner
    dense = 6.0 * per_layer * c.layers * seq
    attn = 3.0 * 4.0 * seq * seq * c.hidden * c.layers
    head = 6.0 * (c.hidden * c.hidden + c.hidden * c.vocab_padded) * c.max_predictions
    return dense + attn + head
