"""Attention forward/backward: tcgen05 kernels (csrc/attention.cu) vs. the library core
(scaled_dot_product_attention) on the same packed qkv.  CUDA-event timing, L2 flushed between
iterations.  ``python tools/attn_bench.py [--batch 32 --seq 512 --heads 16] [--causal]``"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tony_b200.ops import functional as F  # noqa: E402


def timeit(fn, iters, flush, setup=None):
    for _ in range(3):
        fn(setup() if setup else None)
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        arg = setup() if setup else None
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn(arg)
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--heads", type=int, default=16)
    ap.add_argument("--causal", action="store_true")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    qkv = torch.randn(a.batch, a.seq, 3, a.heads, 64, device=dev).to(torch.bfloat16)
    fwd_flops = 4.0 * a.batch * a.heads * a.seq * a.seq * 64 * (0.5 if a.causal else 1.0)
    for mode in ("tcgen05", "library"):
        os.environ["TONY_B200_ATTENTION"] = mode
        x = qkv.clone().requires_grad_()
        out = F.packed_attention(x, a.causal)
        g = torch.randn_like(out)
        t_f = timeit(lambda _: F.packed_attention(x.detach(), a.causal), a.iters, flush)

        def fresh_forward():
            x.grad = None
            return F.packed_attention(x, a.causal)

        t_b = timeit(lambda o: o.backward(g), a.iters, flush, setup=fresh_forward)
        print(json.dumps({"impl": mode, "batch": a.batch, "seq": a.seq, "heads": a.heads, "causal": a.causal,
                          "fwd_us": round(t_f, 1), "bwd_us": round(t_b, 1),
                          "fwd_tflops": round(fwd_flops / t_f / 1e6, 1),
                          "bwd_tflops": round(2.5 * fwd_flops / t_b / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    main()
