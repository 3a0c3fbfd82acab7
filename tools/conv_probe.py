#!/usr/bin/env python
"""A handful of convolution launches for one `ncu --set full` capture (run plainly first):
layer-1 3x3 forward (N = 64 tile), layer-3 3x3 forward and weight gradient (N = 256 tiles), the stem."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tony_b200.ops import functional as F

dev = torch.device("cuda")
N = int(os.environ.get("PROBE_BATCH", "256"))
which = os.environ.get("PROBE_CASES", "l1f,l3f,l3w,stem").split(",")


def case(ci, co, k, s, h):
    x = torch.randn(N, h, h, ci, device=dev).to(torch.bfloat16)
    w = (torch.randn(co, k, k, ci, device=dev) / (k * ci ** 0.5)).to(torch.bfloat16)
    return x, w


if "l1f" in which:
    x, w = case(64, 64, 3, 1, 56)
    for _ in range(2):
        y = F.conv2d_nhwc(x, w, 1, 1)
if "l3f" in which:
    x, w = case(256, 256, 3, 1, 14)
    for _ in range(2):
        y = F.conv2d_nhwc(x, w, 1, 1)
if "l3w" in which:
    x, w = case(256, 256, 3, 1, 14)
    dy = torch.randn(N, 14, 14, 256, device=dev).to(torch.bfloat16)
    for _ in range(2):
        F.conv2d_wgrad_nhwc(dy, x, w.shape, 1, 1)
if "l3d" in which:
    x, w = case(256, 256, 3, 1, 14)
    dy = torch.randn(N, 14, 14, 256, device=dev).to(torch.bfloat16)
    for _ in range(2):
        F.conv2d_dgrad_nhwc(dy, w, x.shape, 1, 1)
if "stem" in which:
    xi = torch.randn(N, 3, 224, 224, device=dev).to(torch.bfloat16)
    ws = (torch.randn(64, 7, 7, 3, device=dev) / 12).to(torch.bfloat16)
    for _ in range(2):
        F.stem_conv7x7s2(xi, ws)
torch.cuda.synchronize()
print("conv_probe ok")
