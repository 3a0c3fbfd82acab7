#!/usr/bin/env python
"""Launch each memory-bound / optimizer kernel a few times at its model shape, for one `ncu --set full`
capture (run plainly first; see B200_PROFILING.md).  Shapes are the BERT-large step's: LayerNorm
16384x1024, cross-entropy 2560x30522, bias-gradient column sums 16384x{1024,4096}, flat AdamW, and the
fused reduce+optimizer kernel on a 32 Mi-element bucket at world 1 (its reduction loop degenerates to
the local shard there; the NVLink legs are timed by tools/collective_bench.py on 2-8 GPUs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tony_b200 import runtime
from tony_b200.ops import functional as F
from tony_b200.ops import optim as O

dev = torch.device("cuda")
REP = int(os.environ.get("NCU_PACK_REP", "2"))
rows, H = 16384, 1024
x = torch.randn(rows, H, device=dev).to(torch.bfloat16).requires_grad_(True)
g = torch.ones(H, device=dev, dtype=torch.bfloat16, requires_grad=True)
b = torch.zeros(H, device=dev, dtype=torch.bfloat16, requires_grad=True)
dy = torch.randn(rows, H, device=dev).to(torch.bfloat16)
for _ in range(REP):
    y = F.layer_norm(x, g, b, 1e-5)
    torch.autograd.grad(y, (x, g, b), dy)
for cols in (1024, 4096):
    m = torch.randn(rows, cols, device=dev).to(torch.bfloat16)
    for _ in range(REP):
        F.colsum(m)
logits = torch.randn(2560, 30528, device=dev).to(torch.bfloat16)
tgt = torch.randint(0, 30522, (2560,), device=dev)
for _ in range(REP):
    F.softmax_cross_entropy(logits, tgt, vocab=30522)
n = 1 << 25
st = O.FlatAdamW(torch.randn(n, device=dev), lr=1e-3)
grad = torch.randn(n, device=dev).to(torch.bfloat16)
w = torch.empty(n, device=dev, dtype=torch.bfloat16)
for _ in range(REP):
    st.step(grad, w)


class Big(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.w = torch.nn.Parameter(torch.randn(n // 1024, 1024, device=dev).to(torch.bfloat16))

    def forward(self, x):
        return F.linear(x, self.w).float().square().mean()


tr = runtime.Trainer(Big(), lr=1e-3, device=dev, bucket_mb=64.0, overlap=False)
xb = torch.randn(256, 1024, device=dev).to(torch.bfloat16)
for _ in range(REP):
    tr.step_device({"x": xb})
torch.cuda.synchronize()
print("ncu_pack ok")
