#!/usr/bin/env python
"""Device-timed micro-benchmarks of the memory-bound kernels with roofline fractions.

CUDA events on the launching stream, 3 warm-ups, inputs far larger than the 126 MB L2 or an L2
flush between iterations; achieved = algorithmic bytes / time, fraction = achieved / measured
copy bandwidth from MEASURED_PEAKS.json ("of measured").  Prints one JSON line per kernel."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tony_b200.ops import functional as F
from tony_b200.ops import optim as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    SRC = "of measured"
except Exception:
    PEAK, SRC = 6650.0, "of fallback"
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for i in range(iters + 3):
        flush.fill_(i & 1)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        if i >= 3:
            best = min(best, e0.elapsed_time(e1))
    return best


def report(name, nbytes, ms, extra=None):
    gbs = nbytes / ms / 1e6
    out = {"kernel": name, "ms": round(ms, 4), "algorithmic_MB": round(nbytes / 1e6, 1), "GBps": round(gbs, 1),
           "fraction_of_hbm_peak": round(gbs / PEAK, 3), "peak_GBps": PEAK, "peak_source": SRC}
    out.update(extra or {})
    print(json.dumps(out), flush=True)


rows, H = 16384, 1024
x = torch.randn(rows, H, device=dev).to(torch.bfloat16).requires_grad_(True)
g = torch.ones(H, device=dev, dtype=torch.bfloat16, requires_grad=True)
b = torch.zeros(H, device=dev, dtype=torch.bfloat16, requires_grad=True)
dy = torch.randn(rows, H, device=dev).to(torch.bfloat16)
report("layernorm_fwd 16384x1024", rows * H * 2 * 2, timeit(lambda: F.layer_norm(x.detach(), g.detach(), b.detach(), 1e-5)))
y = F.layer_norm(x, g, b, 1e-5)
report("layernorm_bwd 16384x1024 (+finalize)", rows * H * 2 * 3, timeit(lambda: torch.autograd.grad(y, (x, g, b), dy, retain_graph=True)))  # noqa
ref_ln = torch.nn.functional.layer_norm
xr = x.detach().clone().requires_grad_(True)
gr, br = g.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
yr = ref_ln(xr, (H,), gr, br, 1e-5)
report("ATen layer_norm bwd 16384x1024 (stock)", rows * H * 2 * 3, timeit(lambda: torch.autograd.grad(yr, (xr, gr, br), dy, retain_graph=True)))  # noqa

for cols in (1024, 4096):
    m = torch.randn(rows, cols, device=dev).to(torch.bfloat16)
    report(f"colsum 16384x{cols}", rows * cols * 2, timeit(lambda: F.colsum(m)))
    report(f"ATen sum(0) 16384x{cols} (stock)", rows * cols * 2, timeit(lambda: m.sum(0)))

for r_, v_, ld_ in ((2560, 30522, 30528), (8192, 50257, 50304)):
    logits = torch.randn(r_, ld_, device=dev).to(torch.bfloat16)
    tgt = torch.randint(0, v_, (r_,), device=dev)
    report(f"softmax_xent fwd+bwd {r_}x{v_}", r_ * ld_ * 2 * 2, timeit(lambda: F.softmax_cross_entropy(logits, tgt, vocab=v_)))
    lf = logits[:, :v_].float().requires_grad_(True)

    def stock():
        l = torch.nn.functional.cross_entropy(lf, tgt)
        torch.autograd.grad(l, lf)

    report(f"F.cross_entropy fwd+bwd fp32 {r_}x{v_} (stock)", r_ * ld_ * 2 * 2, timeit(stock, 5))

n = 1 << 27
st = O.FlatAdamW(torch.randn(n, device=dev), lr=1e-3)
grad = torch.randn(n, device=dev).to(torch.bfloat16)
w = torch.empty(n, device=dev, dtype=torch.bfloat16)
report("adamw_flat 134M params", n * (2 + 12 + 12 + 2), timeit(lambda: st.step(grad, w)))
pt = torch.randn(n, device=dev, requires_grad=True)
pt.grad = torch.randn(n, device=dev)
opt = torch.optim.AdamW([pt], lr=1e-3, fused=True)
report("torch AdamW(fused) 134M fp32 params (stock, no bf16 cast)", n * (4 + 12 + 12), timeit(lambda: opt.step()))
