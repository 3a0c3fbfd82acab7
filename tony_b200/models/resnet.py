"""ResNet-50 (v1.5) for the DDP configuration.

BASELINE.json config "ResNet-50 PyTorch DDP bf16, 8 workers x 1 B200 each".  Honest status
(DESIGN.md, K8): the convolutions still run on cuDNN in bf16 channels-last — no hand-written
implicit-GEMM convolution yet.  BatchNorm (+ residual add + ReLU), which is 79 % of the stock step
on B200, runs on the fused channels-last kernels of csrc/batchnorm.cu.  Stride-1 1x1 convolutions CAN
run as a tcgen05 GEMM over the NHWC activation viewed as [N*H*W, C] (no layout copy), but the
measured crossover (profiles/r1_call23_*) says cuDNN wins for every reduction depth in this
network (4494 img/s with all of them on the GEMM, 4741 with none), so by default only the
2048-channel ones do (``TONY_B200_RESNET_GEMM_MIN_CIN``).  Ours on this model: the classifier
GEMM, the fused softmax-cross-entropy kernel, and the fused reduce + sharded SGD-momentum +
broadcast kernel fed by bucketed gradient hooks (``Trainer(generic_hooks=True)``).
"""
from __future__ import annotations

import os

import torch
from torch import nn

from ..ops import functional as F
from .nn import BatchNormAct, Linear, _param


STOCK = {"on": False}   # set by stock_resnet50(): build the cuDNN-only twin for the baseline arm


class Conv1x1(nn.Module):
    """Stride-1 pointwise convolution on channels-last activations == a GEMM."""

    def __init__(self, cin: int, cout: int, device=None):
        super().__init__()
        self.weight = _param(cout, cin, std=(2.0 / cin) ** 0.5, device=device)

    def forward(self, x):                       # x: [N, C, H, W] in channels_last memory format
        n, c, h, w = x.shape
        x2 = x.permute(0, 2, 3, 1).reshape(n * h * w, c)      # a view for channels_last tensors
        y = F.linear(x2, self.weight)
        return y.view(n, h, w, -1).permute(0, 3, 1, 2)        # still channels_last


def _conv(cin, cout, k, stride=1, device=None):
    return nn.Conv2d(cin, cout, k, stride=stride, padding=k // 2, bias=False, device=device, dtype=torch.bfloat16)


def _bn(c, device=None, zero=False, relu=True):
    """Stock twin: nn.BatchNorm2d (the ReLU / add stay separate ATen kernels).  Ours: BatchNormAct,
    whose forward / backward are the fused channels-last kernels of csrc/batchnorm.cu."""
    if not STOCK["on"]:
        return BatchNormAct(c, relu=relu, zero_init=zero, device=device)
    bn = nn.BatchNorm2d(c, device=device, dtype=torch.bfloat16)
    if zero:
        nn.init.zeros_(bn.weight)
    return bn


class _MaxPool(nn.Module):
    def forward(self, x):
        return F.max_pool_3x3_s2(x)


class Bottleneck(nn.Module):
    def __init__(self, cin, mid, stride, device=None):
        super().__init__()
        cout = mid * 4
        # a pointwise conv becomes a tcgen05 GEMM only when its reduction dim is deep enough to
        # amortise the per-tile epilogue (measured crossover, see profiles/); shallower ones stay on cuDNN
        min_cin = int(os.environ.get("TONY_B200_RESNET_GEMM_MIN_CIN", "1024"))

        def C1(i, o, d):
            return _conv(i, o, 1, 1, d) if (STOCK["on"] or i < min_cin) else Conv1x1(i, o, d)

        self.conv1, self.bn1 = C1(cin, mid, device), _bn(mid, device)
        self.conv2, self.bn2 = _conv(mid, mid, 3, stride, device), _bn(mid, device)
        self.conv3, self.bn3 = C1(mid, cout, device), _bn(cout, device, zero=True)
        self.down = None
        if stride != 1 or cin != cout:
            proj = C1(cin, cout, device) if stride == 1 else _conv(cin, cout, 1, stride, device)
            self.down = nn.Sequential(proj, _bn(cout, device, relu=False))
        self.fused = not STOCK["on"]

    def forward(self, x):
        idt = x if self.down is None else self.down(x)
        if self.fused:      # ReLU and the residual add ride in the BatchNorm kernels
            y = self.bn1(self.conv1(x))
            y = self.bn2(self.conv2(y))
            return self.bn3(self.conv3(y), residual=idt)
        y = torch.relu(self.bn1(self.conv1(x)))
        y = torch.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return torch.relu(y + idt)


class ResNet50(nn.Module):
    def __init__(self, num_classes: int = 1000, device=None):
        super().__init__()
        if STOCK["on"]:
            self.stem = nn.Sequential(_conv(3, 64, 7, 2, device), _bn(64, device), nn.ReLU(inplace=True),
                                      nn.MaxPool2d(3, 2, 1))
        else:
            # the 3-channel stem convolution lands on slow non-tensor-core cuDNN kernels; padding the
            # image to 4 / 8 zero channels (the extra filter taps only ever see zeros, so their
            # gradients stay zero) lets cuDNN pick aligned kernels.  TONY_B200_RESNET_STEM_CPAD=3 disables.
            self.stem_cin = int(os.environ.get("TONY_B200_RESNET_STEM_CPAD", "3"))
            conv = _conv(self.stem_cin, 64, 7, 2, device)
            if self.stem_cin > 3:
                with torch.no_grad():
                    conv.weight[:, 3:].zero_()
            self.stem = nn.Sequential(conv, _bn(64, device), _MaxPool())
        blocks, cin = [], 64
        for mid, n, stride in ((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)):
            for i in range(n):
                blocks.append(Bottleneck(cin, mid, stride if i == 0 else 1, device))
                cin = mid * 4
        self.blocks = nn.Sequential(*blocks)
        self.stock = STOCK["on"]
        self.fc = (nn.Linear(2048, num_classes, device=device, dtype=torch.bfloat16) if self.stock
                   else Linear(2048, (num_classes + 7) // 8 * 8, device=device))
        self.num_classes = num_classes

    def forward(self, images, labels):
        if not self.stock and self.stem_cin > 3:
            images = torch.nn.functional.pad(images, (0, 0, 0, 0, 0, self.stem_cin - 3))
        x = images.contiguous(memory_format=torch.channels_last)
        x = self.blocks(self.stem(x))
        x = x.mean(dim=(2, 3))
        if self.stock:
            return torch.nn.functional.cross_entropy(self.fc(x).float(), labels)
        return F.softmax_cross_entropy(self.fc(x), labels, vocab=self.num_classes)


def stock_resnet50(num_classes: int = 1000, device=None) -> "ResNet50":
    """The baseline twin: every layer on cuDNN / ATen (bf16, channels-last), ATen cross-entropy."""
    STOCK["on"] = True
    try:
        return ResNet50(num_classes, device)
    finally:
        STOCK["on"] = False


def synthetic_batch(batch: int, device="cpu", pin: bool = False, seed: int = 0, size: int = 224, classes: int = 1000):
    g = torch.Generator().manual_seed(seed)
    out = {"images": torch.randn(batch, 3, size, size, generator=g).to(torch.bfloat16),
           "labels": torch.randint(0, classes, (batch,), generator=g)}
    if pin:
        out = {k: v.pin_memory() for k, v in out.items()}
    if device != "cpu":
        out = {k: v.to(device, non_blocking=True) for k, v in out.items()}
    return out
