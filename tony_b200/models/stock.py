"""Stock-PyTorch twins of the model zoo: the measured baseline arm.

The reference contributes no data-plane code, so "the same job launched by TonY" runs on
exactly this stack: ``nn.Linear`` (cuBLASLt), ATen LayerNorm/GELU, ``F.cross_entropy``,
``torch.autocast(bf16)`` over fp32 parameters, ``torch.optim.AdamW(fused=True)`` and
``DistributedDataParallel`` over NCCL.  Same architecture,
same synthetic inputs, same attention call as ``models/bert.py`` / ``models/gpt2.py``.
"""
from __future__ import annotations

import torch
import torch.nn.functional as TF
from torch import nn

from .bert import BertConfig
from .gpt2 import GPT2Config


class _StockAttention(nn.Module):
    def __init__(self, hidden, heads, causal):
        super().__init__()
        self.heads, self.head_dim, self.causal = heads, hidden // heads, causal
        self.qkv = nn.Linear(hidden, 3 * hidden)
        self.out = nn.Linear(hidden, hidden)

    def forward(self, x):
        B, S, H = x.shape
        qkv = self.qkv(x).view(B, S, 3, self.heads, self.head_dim)
        q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
        o = TF.scaled_dot_product_attention(q, k, v, is_causal=self.causal)
        return self.out(o.transpose(1, 2).reshape(B, S, H))


class _StockBertLayer(nn.Module):
    def __init__(self, c: BertConfig):
        super().__init__()
        self.attn = _StockAttention(c.hidden, c.heads, False)
        self.ln1 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.fc1 = nn.Linear(c.hidden, c.inner)
        self.fc2 = nn.Linear(c.inner, c.hidden)
        self.ln2 = nn.LayerNorm(c.hidden, eps=c.eps)

    def forward(self, x):
        x = self.ln1(x + self.attn(x))
        return self.ln2(x + self.fc2(TF.gelu(self.fc1(x))))


class StockBertForPreTraining(nn.Module):
    def __init__(self, c: BertConfig):
        super().__init__()
        self.config = c
        self.word = nn.Embedding(c.vocab_padded, c.hidden)
        self.pos = nn.Embedding(c.max_positions, c.hidden)
        self.tok_type = nn.Embedding(c.type_vocab, c.hidden)
        self.emb_ln = nn.LayerNorm(c.hidden, eps=c.eps)
        self.layers = nn.ModuleList(_StockBertLayer(c) for _ in range(c.layers))
        self.mlm_dense = nn.Linear(c.hidden, c.hidden)
        self.mlm_ln = nn.LayerNorm(c.hidden, eps=c.eps)
        self.mlm_bias = nn.Parameter(torch.zeros(c.vocab_padded))
        self.pooler = nn.Linear(c.hidden, c.hidden)
        self.nsp = nn.Linear(c.hidden, 2)
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                m.weight.data.normal_(0, 0.02)

    def forward(self, input_ids, token_type_ids, masked_positions, masked_labels, nsp_labels):
        B, S = input_ids.shape
        c = self.config
        x = self.word(input_ids) + self.tok_type(token_type_ids) + self.pos(torch.arange(S, device=input_ids.device))
        x = self.emb_ln(x)
        for layer in self.layers:
            x = layer(x)
        flat = x.view(B * S, -1)
        gather = (masked_positions + torch.arange(B, device=x.device).unsqueeze(1) * S).reshape(-1)
        hm = self.mlm_ln(TF.gelu(self.mlm_dense(flat.index_select(0, gather))))
        logits = TF.linear(hm, self.word.weight, self.mlm_bias)[:, : c.vocab_size]
        mlm = TF.cross_entropy(logits.float(), masked_labels.reshape(-1), ignore_index=-100)
        nsp = TF.cross_entropy(self.nsp(torch.tanh(self.pooler(x[:, 0]))).float(), nsp_labels)
        return mlm + nsp


class _StockGPT2Block(nn.Module):
    def __init__(self, c: GPT2Config):
        super().__init__()
        self.ln1 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.attn = _StockAttention(c.hidden, c.heads, True)
        self.ln2 = nn.LayerNorm(c.hidden, eps=c.eps)
        self.fc1 = nn.Linear(c.hidden, c.inner)
        self.fc2 = nn.Linear(c.inner, c.hidden)

    def forward(self, x):
        x = x + self.attn(self.ln1(x))
        return x + self.fc2(TF.gelu(self.fc1(self.ln2(x)), approximate="tanh"))


class StockGPT2LMHead(nn.Module):
    def __init__(self, c: GPT2Config):
        super().__init__()
        self.config = c
        self.wte = nn.Embedding(c.vocab_padded, c.hidden)
        self.wpe = nn.Embedding(c.max_positions, c.hidden)
        self.blocks = nn.ModuleList(_StockGPT2Block(c) for _ in range(c.layers))
        self.ln_f = nn.LayerNorm(c.hidden, eps=c.eps)
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                m.weight.data.normal_(0, 0.02)

    def forward(self, input_ids, labels):
        B, S = input_ids.shape
        x = self.wte(input_ids) + self.wpe(torch.arange(S, device=input_ids.device))
        for blk in self.blocks:
            x = blk(x)
        logits = TF.linear(self.ln_f(x).view(B * S, -1), self.wte.weight)[:, : self.config.vocab_size]
        return TF.cross_entropy(logits.float(), labels.reshape(-1))
