#!/usr/bin/env python
"""Our tcgen05 GEMM vs cuBLAS (torch.matmul) on the BERT-large step's shapes (tokens = 16384).
CUDA-event timing, 3 warm-ups, best of 10, with and without an L2 flush between iterations."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tony_b200.ops import functional as F

dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, do_flush):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for i in range(13):
        if do_flush:
            flush.fill_(i & 1)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        if i >= 3:
            best = min(best, e0.elapsed_time(e1))
    return best


T = 16384
# ResNet-50 (256 images): the pointwise convolutions' data / weight gradients run on this kernel as GEMMs over the
# [N*H*W, C] activation.  name, M, N, K, a_mn, b_mn
RESNET = []
for tag, px, cin, cout in (("l1 64->256 @56", 802816, 64, 256), ("l1 256->64 @56", 802816, 256, 64),
                           ("l2 128->512 @28", 200704, 128, 512), ("l2 512->128 @28", 200704, 512, 128),
                           ("l3 256->1024 @14", 50176, 256, 1024), ("l3 1024->256 @14", 50176, 1024, 256),
                           ("l4 512->2048 @7", 12544, 512, 2048), ("l4 2048->512 @7", 12544, 2048, 512)):
    RESNET.append((f"resnet dgrad {tag}", px, cin, cout, False, True))
    RESNET.append((f"resnet wgrad {tag}", cout, cin, px, True, True))
BERT = [("fwd qkv", T, 3072, 1024, False, False), ("fwd attn_out", T, 1024, 1024, False, False),
          ("fwd ffn1", T, 4096, 1024, False, False), ("fwd ffn2", T, 1024, 4096, False, False),
          ("dgrad ffn2", T, 4096, 1024, False, True), ("dgrad ffn1", T, 1024, 4096, False, True),
          ("dgrad qkv", T, 1024, 3072, False, True),
          ("wgrad ffn1", 4096, 1024, T, True, True), ("wgrad attn_out", 1024, 1024, T, True, True),
          ("wgrad qkv", 3072, 1024, T, True, True)]
shapes = RESNET if (len(sys.argv) > 1 and sys.argv[1] == "resnet") else BERT
for name, M, N, K, a_mn, b_mn in shapes:
    A = torch.randn(M, K, device=dev).to(torch.bfloat16)
    B = torch.randn(N, K, device=dev).to(torch.bfloat16)
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ours = lambda: F.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out)
    if a_mn and b_mn:
        ref = lambda: torch.matmul(a.t(), b, out=out)
    elif b_mn:
        ref = lambda: torch.matmul(a, b, out=out)
    else:
        ref = lambda: torch.matmul(a, b.t(), out=out)
    fl = 2.0 * M * N * K
    row = {"shape": name, "M": M, "N": N, "K": K}
    for tag, fl_ in (("warm", False), ("cold", True)):
        to, tr = timeit(ours, fl_), timeit(ref, fl_)
        row[f"ours_{tag}_tflops"] = round(fl / to / 1e9, 1)
        row[f"cublas_{tag}_tflops"] = round(fl / tr / 1e9, 1)
    print(json.dumps(row), flush=True)
