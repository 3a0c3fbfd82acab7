#!/usr/bin/env python
"""Per-layer convolution table: this package's implicit-GEMM tcgen05 kernels against cuDNN.

    python tools/conv_bench.py [--batch 256] [--iters 5] [--stem 1] [--md out.md]

Every distinct convolution of ResNet-50 (53 layers, 24 distinct shapes incl. the stem) is timed for
forward, data gradient and weight gradient with CUDA events on the launching stream after 2 warm-up
calls; a 256 MB buffer is written between timed calls so no operand starts in L2.  The cuDNN arm is
``torch.nn.functional.conv2d`` / ``torch.ops.aten.convolution_backward`` on bf16 channels-last tensors
with ``cudnn.benchmark = True``.  Prints one JSON line per (layer, pass) and a totals line weighted by
how often the shape occurs in the network."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tony_b200.ops import functional as F

dev = torch.device("cuda")


def resnet50_layers():
    """(name, count, Cin, Cout, k, stride, Hin) for every distinct conv of ResNet-50 v1.5."""
    out = []
    cin, h = 64, 56
    for stage, (mid, n, stride) in enumerate(((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)), start=1):
        for i in range(n):
            s = stride if i == 0 else 1
            out.append((f"l{stage}.{i}.conv1", cin, mid, 1, 1, h))
            out.append((f"l{stage}.{i}.conv2", mid, mid, 3, s, h))
            out.append((f"l{stage}.{i}.conv3", mid, mid * 4, 1, 1, h // s))
            if i == 0:
                out.append((f"l{stage}.{i}.down", cin, mid * 4, 1, s, h))
            cin, h = mid * 4, h // s
    uniq = {}
    for name, ci, co, k, s, hh in out:
        key = (ci, co, k, s, hh)
        if key in uniq:
            uniq[key][1] += 1
        else:
            uniq[key] = [name, 1]
    return [(v[0], v[1]) + k for k, v in uniq.items()]


flush = None


def timeit(fn, iters):
    global flush
    if flush is None:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for i in range(iters + 2):
        flush.fill_(i & 1)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        if i >= 2:
            best = min(best, e0.elapsed_time(e1))
    return best * 1e3  # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--stem", type=int, default=1)
    ap.add_argument("--md", default="")
    ap.add_argument("--passes", default="fprop,fprop+stats,dgrad,wgrad")
    ap.add_argument("--no-lib", action="store_true", help="skip the cuDNN arm")
    a = ap.parse_args()
    want = set(a.passes.split(","))
    torch.backends.cudnn.benchmark = True
    N = a.batch
    rows = []
    tot = {"ours": 0.0, "cudnn": 0.0}
    for name, count, ci, co, k, s, h in resnet50_layers():
        pad = k // 2
        x = torch.randn(N, ci, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(co, k, k, ci, device=dev) / (k * ci ** 0.5)).to(torch.bfloat16)
        wt = w.permute(0, 3, 1, 2)          # [K, C, R, S] view, channels-last memory: what cuDNN wants
        xn = x.permute(0, 2, 3, 1)
        pointwise = k == 1 and s == 1
        if pointwise:
            M = N * h * h
            y = F.gemm(xn.reshape(M, ci), w.view(co, ci)).view(N, h, h, co)
        else:
            y = F.conv2d_nhwc(xn, w, s, pad)
        P = y.shape[1]
        yt = torch.nn.functional.conv2d(x, wt, stride=s, padding=pad)
        err = (y.permute(0, 3, 1, 2).float() - yt.float()).abs().max().item() / (yt.float().abs().max().item() + 1e-6)
        dy = torch.randn_like(y)
        dyt = dy.permute(0, 3, 1, 2)
        flops = 2.0 * N * P * P * co * ci * k * k
        stats = torch.zeros(2, co, dtype=torch.float32, device=dev)
        passes = {}
        if pointwise:
            M = N * h * h
            passes["fprop"] = (lambda: F.gemm(xn.reshape(M, ci), w.view(co, ci)),
                               lambda: torch.nn.functional.conv2d(x, wt, stride=s, padding=pad))
            passes["fprop+stats"] = (lambda: F.conv2d_nhwc(xn, w, s, pad, stats=stats), None)
            passes["dgrad"] = (lambda: F.gemm(dy.reshape(M, co), w.view(co, ci), b_mn=True),
                               lambda: torch.ops.aten.convolution_backward(dyt, x, wt, None, [s, s], [pad, pad], [1, 1],
                                                                           False, [0, 0], 1, [True, False, False]))
            passes["wgrad"] = (lambda: F.gemm(dy.reshape(M, co), xn.reshape(M, ci), a_mn=True, b_mn=True),
                               lambda: torch.ops.aten.convolution_backward(dyt, x, wt, None, [s, s], [pad, pad], [1, 1],
                                                                           False, [0, 0], 1, [False, True, False]))
        else:
            passes["fprop"] = (lambda: F.conv2d_nhwc(xn, w, s, pad),
                               lambda: torch.nn.functional.conv2d(x, wt, stride=s, padding=pad))
            passes["fprop+stats"] = (lambda: F.conv2d_nhwc(xn, w, s, pad, stats=stats), None)
            passes["dgrad"] = (lambda: F.conv2d_dgrad_nhwc(dy, w, xn.shape, s, pad),
                               lambda: torch.ops.aten.convolution_backward(dyt, x, wt, None, [s, s], [pad, pad], [1, 1],
                                                                           False, [0, 0], 1, [True, False, False]))
            passes["wgrad"] = (lambda: F.conv2d_wgrad_nhwc(dy, xn, w.shape, s, pad),
                               lambda: torch.ops.aten.convolution_backward(dyt, x, wt, None, [s, s], [pad, pad], [1, 1],
                                                                           False, [0, 0], 1, [False, True, False]))
        for pname, (ours, lib) in passes.items():
            if pname not in want:
                continue
            t_o = timeit(ours, a.iters)
            t_l = timeit(lib, a.iters) if (lib is not None and not a.no_lib) else None
            rec = {"layer": name, "count": count, "cin": ci, "cout": co, "k": k, "stride": s, "hw": h, "pass": pname,
                   "ours_us": round(t_o, 1), "cudnn_us": round(t_l, 1) if t_l else None,
                   "ours_tflops": round(flops / t_o / 1e6, 1), "ratio_cudnn_over_ours": round(t_l / t_o, 3) if t_l else None,
                   "fwd_rel_err": round(err, 5)}
            print(json.dumps(rec), flush=True)
            rows.append(rec)
            if pname != "fprop+stats":
                tot["ours"] += count * t_o
                tot["cudnn"] += count * (t_l or 0.0)
        del x, w, y, yt, dy
    if a.stem:
        x = torch.randn(N, 3, 224, 224, device=dev).to(torch.bfloat16)
        w = (torch.randn(64, 7, 7, 3, device=dev) / 12).to(torch.bfloat16).requires_grad_(True)
        xcl = torch.nn.functional.pad(x, (0, 0, 0, 0, 0, 1)).contiguous(memory_format=torch.channels_last)
        wt = torch.nn.functional.pad(w.detach().permute(0, 3, 1, 2), (0, 0, 0, 0, 0, 1)).contiguous(
            memory_format=torch.channels_last)
        for mode in ("dense", "overlap"):
            os.environ["TONY_B200_STEM_WINDOW"] = mode
            y = F.stem_conv7x7s2(x, w)
            dy = torch.randn_like(y)
            t_f = timeit(lambda: F.stem_conv7x7s2(x, w), a.iters)

            def fb():
                w.grad = None
                F.stem_conv7x7s2(x, w).backward(dy)

            t_fb = timeit(fb, a.iters)
            t_lf = timeit(lambda: torch.nn.functional.conv2d(xcl, wt, stride=2, padding=3), a.iters)
            dyt = dy
            t_lw = timeit(lambda: torch.ops.aten.convolution_backward(dyt, xcl, wt, None, [2, 2], [3, 3], [1, 1], False,
                                                                      [0, 0], 1, [False, True, False]), a.iters)
            rec = {"layer": "stem 7x7/2 3->64", "mode": mode, "ours_fwd_us": round(t_f, 1),
                   "ours_fwd_plus_wgrad_us": round(t_fb, 1), "cudnn_fwd_us_c4": round(t_lf, 1),
                   "cudnn_wgrad_us_c4": round(t_lw, 1)}
            print(json.dumps(rec), flush=True)
            rows.append(rec)
    print(json.dumps({"total_weighted_us": {k: round(v) for k, v in tot.items()},
                      "cudnn_over_ours": round(tot["cudnn"] / max(tot["ours"], 1e-9), 3), "batch": N}), flush=True)
    if a.md:
        with open(a.md, "w") as f:
            f.write("| layer | x | Cin | Cout | k | s | HW | pass | ours us | cuDNN us | ours TFLOP/s | cuDNN/ours |\n")
            f.write("|---|---|---|---|---|---|---|---|---|---|---|---|\n")
            for r in rows:
                if "pass" not in r:
                    continue
                f.write(f"| {r['layer']} | {r['count']} | {r['cin']} | {r['cout']} | {r['k']} | {r['stride']} | {r['hw']} | "
                        f"{r['pass']} | {r['ours_us']} | {r['cudnn_us']} | {r['ours_tflops']} | {r['ratio_cudnn_over_ours']} |\n")


if __name__ == "__main__":
    main()
