#!/usr/bin/env python
"""Device-timed micro-benchmarks of the memory-bound kernels with roofline fractions.

Each case runs over several independent input sets whose total size is a multiple of the 126 MB L2, so a set is
cold again by the time it comes round.  One pass over the sets is captured in a CUDA graph and replayed; the
time per call is (CUDA events around the replays) / calls - kernel time, not the host's launch path (a 15 us
kernel behind 30 us of Python autograd measured as 45 us in the one-call-per-event form of this tool).
achieved = algorithmic bytes / time, fraction = achieved / measured copy bandwidth from MEASURED_PEAKS.json
("of measured").  One JSON line per kernel, with the SM clock read right after the timed replays."""
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
L2_BYTES = 126 << 20
BENCH_STREAM = torch.cuda.Stream()
torch.cuda.set_stream(BENCH_STREAM)

try:
    import pynvml

    pynvml.nvmlInit()
    _NVML = pynvml.nvmlDeviceGetHandleByIndex(torch.cuda.current_device())
except Exception:
    pynvml, _NVML = None, None


def sm_clock_mhz():
    if _NVML is None:
        return None
    try:
        return int(pynvml.nvmlDeviceGetClockInfo(_NVML, pynvml.NVML_CLOCK_SM))
    except Exception:
        return None


def nsets_for(bytes_per_set):
    return max(2, min(16, -(-3 * L2_BYTES // max(1, bytes_per_set))))


def time_sets(calls, reps=6):
    """calls: one zero-argument callable per input set.  -> (ms per call, SM MHz, "graph" | "eager")."""
    for f in calls:
        f()
    torch.cuda.synchronize()
    mode, graph = "graph", None
    try:
        # everything in this tool runs on BENCH_STREAM: autograd replays a node's backward on the stream its
        # forward ran on, and work sent to the legacy default stream during a capture invalidates it
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=BENCH_STREAM):
            for f in calls:
                f()
    except Exception as exc:    # noqa: BLE001 - report and fall back to timing the eager calls
        print(f"graph capture failed ({type(exc).__name__}: {str(exc)[:300]}); timing eager calls", file=sys.stderr)
        mode, graph = "eager", None
        torch.cuda.synchronize()

    def run():
        if graph is not None:
            graph.replay()
        else:
            for f in calls:
                f()

    run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    e1.synchronize()
    mhz = sm_clock_mhz()
    return e0.elapsed_time(e1) / (reps * len(calls)), mhz, mode


def report(name, nbytes, timing, extra=None):
    ms, mhz, mode = timing
    gbs = nbytes / ms / 1e6
    out = {"kernel": name, "us": round(ms * 1e3, 2), "algorithmic_MB": round(nbytes / 1e6, 1), "GBps": round(gbs, 1),
           "fraction_of_hbm_peak": round(gbs / PEAK, 3), "peak_GBps": PEAK, "peak_source": SRC, "sm_mhz": mhz,
           "timed_as": mode}
    out.update(extra or {})
    print(json.dumps(out), flush=True)


rows, H = 16384, 1024
ns = nsets_for(rows * H * 2 * 2)
xs = [torch.randn(rows, H, device=dev).to(torch.bfloat16).requires_grad_(True) for _ in range(ns)]
g = torch.ones(H, device=dev, dtype=torch.bfloat16, requires_grad=True)
b = torch.zeros(H, device=dev, dtype=torch.bfloat16, requires_grad=True)
dys = [torch.randn(rows, H, device=dev).to(torch.bfloat16) for _ in range(ns)]
report("layernorm_fwd 16384x1024", rows * H * 2 * 2,
       time_sets([(lambda x=x: F.layer_norm(x.detach(), g.detach(), b.detach(), 1e-5)) for x in xs]))
ys = [F.layer_norm(x, g, b, 1e-5) for x in xs]
report("layernorm_bwd 16384x1024 (dx + dgamma + dbeta)", rows * H * 2 * 3,
       time_sets([(lambda y=y, x=x, d=d: torch.autograd.grad(y, (x, g, b), d, retain_graph=True))
                  for y, x, d in zip(ys, xs, dys)]))
ref_ln = torch.nn.functional.layer_norm
gr, br = g.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
xrs = [x.detach().clone().requires_grad_(True) for x in xs]
yrs = [ref_ln(x, (H,), gr, br, 1e-5) for x in xrs]
report("ATen layer_norm fwd 16384x1024 (stock)", rows * H * 2 * 2,
       time_sets([(lambda x=x: ref_ln(x.detach(), (H,), gr.detach(), br.detach(), 1e-5)) for x in xrs]))
report("ATen layer_norm bwd 16384x1024 (stock)", rows * H * 2 * 3,
       time_sets([(lambda y=y, x=x, d=d: torch.autograd.grad(y, (x, gr, br), d, retain_graph=True))
                  for y, x, d in zip(yrs, xrs, dys)]))
del xs, dys, ys, xrs, yrs

for cols in (1024, 4096):
    ns = nsets_for(rows * cols * 2)
    ms_ = [torch.randn(rows, cols, device=dev).to(torch.bfloat16) for _ in range(ns)]
    report(f"colsum 16384x{cols}", rows * cols * 2, time_sets([(lambda m=m: F.colsum(m)) for m in ms_]))
    report(f"ATen sum(0) 16384x{cols} (stock)", rows * cols * 2, time_sets([(lambda m=m: m.sum(0)) for m in ms_]))
    del ms_

for r_, v_, ld_ in ((2560, 30522, 30528), (8192, 50257, 50304)):
    ns = nsets_for(r_ * ld_ * 2)
    lg = [torch.randn(r_, ld_, device=dev).to(torch.bfloat16) for _ in range(ns)]
    tgt = torch.randint(0, v_, (r_,), device=dev)
    report(f"softmax_xent fwd+bwd {r_}x{v_}", r_ * ld_ * 2 * 2,
           time_sets([(lambda l=l: F.softmax_cross_entropy(l, tgt, vocab=v_)) for l in lg]))
    lfs = [l[:, :v_].float().requires_grad_(True) for l in lg[:2]]

    def stock(lf):
        l = torch.nn.functional.cross_entropy(lf, tgt)
        torch.autograd.grad(l, lf)

    report(f"F.cross_entropy fwd+bwd fp32 {r_}x{v_} (stock)", r_ * ld_ * 2 * 2,
           time_sets([(lambda lf=lf: stock(lf)) for lf in lfs], reps=3))
    del lg, lfs

n = 1 << 27
st = O.FlatAdamW(torch.randn(n, device=dev), lr=1e-3)
grad = torch.randn(n, device=dev).to(torch.bfloat16)
w = torch.empty(n, device=dev, dtype=torch.bfloat16)
report("adamw_flat 134M params", n * (2 + 12 + 12 + 2), time_sets([lambda: st.step(grad, w)], reps=4))
pt = torch.randn(n, device=dev, requires_grad=True)
pt.grad = torch.randn(n, device=dev)
opt = torch.optim.AdamW([pt], lr=1e-3, fused=True)
report("torch AdamW(fused) 134M fp32 params (stock, no bf16 cast)", n * (4 + 12 + 12),
       time_sets([lambda: opt.step()], reps=4))
