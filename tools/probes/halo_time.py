#!/usr/bin/env python
"""Time the layer-1 / layer-2 3x3 forward convolutions (halo kernel vs im2col kernel, and the halo kernel with
parts switched off through TONY_B200_CONV_DBG) - run once per setting, the switches are read at first use."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from tony_b200.ops import functional as F

dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for (ci, co, h) in ((64, 64, 56), (128, 128, 28)):
    x = torch.randn(256, h, h, ci, device=dev).to(torch.bfloat16)
    w = (torch.randn(co, 3, 3, ci, device=dev) / (3 * ci ** 0.5)).to(torch.bfloat16)
    for _ in range(3):
        F.conv2d_nhwc(x, w, 1, 1)
    ts = []
    for i in range(10):
        flush.fill_(i & 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        F.conv2d_nhwc(x, w, 1, 1)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    print(f"HALO={os.environ.get('TONY_B200_CONV_HALO', '1')} DBG={os.environ.get('TONY_B200_CONV_DBG', '0')} "
          f"{ci}->{co} {h}x{h}: median {ts[len(ts) // 2]:.1f} us, min {ts[0]:.1f} us", flush=True)
