#!/usr/bin/env python
"""Three BERT-large GEMM shapes, twice each, for one `ncu --set full` capture (run plainly first):
FFN1 forward with bias + erf-GELU + saved pre-activation (CTA-pair kernel, TMA-store epilogue), the dGELU
data gradient, and the attention-output weight gradient (split-K + reduce)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tony_b200.ops import functional as F
from tony_b200.ops._lib import EPI_BIAS, EPI_DGELU_ERF, EPI_GELU_ERF, EPI_SAVE_PREACT

dev = torch.device("cuda")
M, H, I = 16384, 1024, 4096
x = torch.randn(M, H, device=dev).to(torch.bfloat16)
w1 = (torch.randn(I, H, device=dev) * 0.02).to(torch.bfloat16)
b1 = torch.zeros(I, device=dev, dtype=torch.bfloat16)
pre = torch.empty(M, I, device=dev, dtype=torch.bfloat16)
dy = torch.randn(M, H, device=dev).to(torch.bfloat16)
w2 = (torch.randn(H, I, device=dev) * 0.02).to(torch.bfloat16)
for _ in range(2):
    h = F.gemm(x, w1, flags=EPI_BIAS | EPI_GELU_ERF | EPI_SAVE_PREACT, bias=b1, aux_out=pre)
for _ in range(2):
    dpre = F.gemm(dy, w2, b_mn=True, flags=EPI_DGELU_ERF, aux=pre)
for _ in range(2):
    dw = F.gemm(dy, x, a_mn=True, b_mn=True)
torch.cuda.synchronize()
print("gemm_probe ok")
