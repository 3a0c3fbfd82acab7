"""ResNet-50 (v1.5) for the DDP configuration.

BASELINE.json config "ResNet-50 PyTorch DDP bf16, 8 workers x 1 B200 each".  Every layer of the own
arm runs on this package's sm_100a kernels:

* convolutions: the implicit-GEMM tcgen05 kernel of csrc/conv_tcgen05.cu (TMA im2col mode; forward,
  data gradient and weight gradient; stride-1 pointwise layers without statistics as plain GEMMs); the
  3-channel 7x7/2 stem as a 4x4 convolution over the 2x2 space-to-depth image;
* BatchNorm (+ residual add) (+ ReLU): csrc/batchnorm.cu; its batch statistics are accumulated by the
  producing convolution's epilogue, so BatchNorm forward is one elementwise pass;
* max-pool, classifier GEMM, fused softmax-cross-entropy, fused reduce + sharded SGD-momentum.

``stock_resnet50`` builds the twin on cuDNN / ATen for the baseline arm.  Filters are stored KRSC
(``[Cout, R, S, Cin]``), the layout the kernels read with no transposed copy in any of the three passes.
``TONY_B200_RESNET_CONV=cudnn`` routes the own arm's convolutions back to cuDNN (A/B measurements).
"""
from __future__ import annotations

import os

import torch
from torch import nn

from ..ops import functional as F
from .nn import BatchNormAct, Linear, _param


STOCK = {"on": False}   # set by stock_resnet50(): build the cuDNN-only twin for the baseline arm


class _StatsArena:
    """One fp32 buffer for every layer's [2, C] epilogue statistics, zeroed once per forward."""

    def __init__(self):
        self.buf = None
        self.used = 0
        self.need = 0

    def reset(self, device):
        if self.buf is None or self.buf.numel() < self.need or self.buf.device != device:
            self.buf = torch.zeros(max(self.need, 1 << 16), dtype=torch.float32, device=device)
        else:
            self.buf.zero_()
        self.used = 0

    def take(self, channels: int):
        n = 2 * channels
        if self.buf is None or self.used + n > self.buf.numel():
            self.need = max(self.need, self.used + n) * 2
            return None                  # first forward sizes the arena; this layer uses the stats kernel
        out = self.buf[self.used: self.used + n].view(2, channels)
        self.used += n
        return out


class ConvBN(nn.Module):
    """conv (KRSC filter, tcgen05 implicit GEMM) -> BatchNorm (+ residual) (+ ReLU)."""

    def __init__(self, cin, cout, k, stride=1, relu=True, zero=False, device=None, arena=None):
        super().__init__()
        self.weight = _param(cout, k, k, cin, std=(2.0 / (cin * k * k)) ** 0.5, device=device)
        self.bn = BatchNormAct(cout, relu=relu, zero_init=zero, device=device)
        self.stride, self.pad = stride, k // 2
        self._arena = [arena]            # in a list: not a submodule / buffer
        self.use_stats = os.environ.get("TONY_B200_CONV_STATS", "1") == "1"

    def forward(self, x, residual=None):
        arena = self._arena[0]
        stats = arena.take(self.weight.shape[0]) if (self.use_stats and arena is not None and self.training) else None
        y = F.conv2d(x, self.weight, self.stride, self.pad, stats)
        return self.bn(y, residual=residual, stats=stats)

    def forward_fork(self, x):
        """-> (bn(conv(x)), x): the block input also feeds the residual branch; its gradient is merged in the
        epilogue of this (pointwise) convolution's data-gradient GEMM."""
        arena = self._arena[0]
        stats = arena.take(self.weight.shape[0]) if (self.use_stats and arena is not None and self.training) else None
        y, idt = F.conv1x1_fork(x, self.weight, stats)
        return self.bn(y, stats=stats), idt


def conv_bn_pair(main: "ConvBN", short: "ConvBN", x):
    """-> (main(x), short(x)) for the first block of a stage: both pointwise convolutions read the block input and
    the shortcut's data gradient accumulates into the main path's (``F.conv1x1_pair``)."""
    arena = main._arena[0]
    take = main.use_stats and arena is not None and main.training
    s1 = arena.take(main.weight.shape[0]) if take else None
    sd = arena.take(short.weight.shape[0]) if take else None
    y1, yd = F.conv1x1_pair(x, main.weight, short.weight, short.stride, s1, sd)
    return main.bn(y1, stats=s1), short.bn(yd, stats=sd)


class StemConvBN(nn.Module):
    """7x7/2 convolution of the 3-channel image (space-to-depth form) -> BatchNorm -> ReLU."""

    def __init__(self, cout=64, device=None, arena=None):
        super().__init__()
        self.weight = _param(cout, 7, 7, 3, std=(2.0 / 147) ** 0.5, device=device)
        self.bn = BatchNormAct(cout, relu=True, device=device)
        self.register_buffer("index", F.stem_weight_index(device), persistent=False)
        self._arena = [arena]

    def forward(self, images):
        arena = self._arena[0]
        stats = arena.take(self.weight.shape[0]) if (arena is not None and self.training) else None
        y = F.stem_conv7x7s2(images, self.weight, stats, self.index)
        return self.bn(y, stats=stats)


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


class _LibConvBN(nn.Module):
    """cuDNN convolution + own BatchNorm: the round-1 arrangement, kept for A/B runs."""

    def __init__(self, cin, cout, k, stride=1, relu=True, zero=False, device=None, arena=None):
        super().__init__()
        self.conv, self.bn = _conv(cin, cout, k, stride, device), _bn(cout, device, zero=zero, relu=relu)

    def forward(self, x, residual=None):
        return self.bn(self.conv(x), residual=residual)


class Bottleneck(nn.Module):
    def __init__(self, cin, mid, stride, device=None, arena=None):
        super().__init__()
        cout = mid * 4
        self.fused = not STOCK["on"]
        if self.fused:
            CB = _LibConvBN if os.environ.get("TONY_B200_RESNET_CONV", "tcgen05") == "cudnn" else ConvBN
            self.cb1 = CB(cin, mid, 1, 1, device=device, arena=arena)
            self.cb2 = CB(mid, mid, 3, stride, device=device, arena=arena)
            self.cb3 = CB(mid, cout, 1, 1, zero=True, device=device, arena=arena)
            self.down = CB(cin, cout, 1, stride, relu=False, device=device, arena=arena) \
                if (stride != 1 or cin != cout) else None
            return
        self.conv1, self.bn1 = _conv(cin, mid, 1, 1, device), _bn(mid, device)
        self.conv2, self.bn2 = _conv(mid, mid, 3, stride, device), _bn(mid, device)
        self.conv3, self.bn3 = _conv(mid, cout, 1, 1, device), _bn(cout, device, zero=True)
        self.down = None
        if stride != 1 or cin != cout:
            self.down = nn.Sequential(_conv(cin, cout, 1, stride, device), _bn(cout, device, relu=False))

    def forward(self, x):
        if self.fused and self.down is None and isinstance(self.cb1, ConvBN) and x.requires_grad:
            y, idt = self.cb1.forward_fork(x)          # identity block: x feeds conv1 AND the residual add
            return self.cb3(self.cb2(y), residual=idt)
        if self.fused and self.down is not None and isinstance(self.cb1, ConvBN) and x.requires_grad:
            y, idt = conv_bn_pair(self.cb1, self.down, x)     # projection block: one node for both readers of x
            return self.cb3(self.cb2(y), residual=idt)
        idt = x if self.down is None else self.down(x)
        if self.fused:      # ReLU and the residual add ride in the BatchNorm kernels
            return self.cb3(self.cb2(self.cb1(x)), residual=idt)
        y = torch.relu(self.bn1(self.conv1(x)))
        y = torch.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return torch.relu(y + idt)


class ResNet50(nn.Module):
    def __init__(self, num_classes: int = 1000, device=None):
        super().__init__()
        self.stock = STOCK["on"]
        self._arena = [None if self.stock else _StatsArena()]
        arena = self._arena[0]
        self.lib_conv = os.environ.get("TONY_B200_RESNET_CONV", "tcgen05") == "cudnn"
        if self.stock:
            self.stem = nn.Sequential(_conv(3, 64, 7, 2, device), _bn(64, device), nn.ReLU(inplace=True),
                                      nn.MaxPool2d(3, 2, 1))
        elif self.lib_conv:
            self.stem = nn.Sequential(_conv(3, 64, 7, 2, device), _bn(64, device), _MaxPool())
        else:
            self.stem = nn.Sequential(StemConvBN(64, device, arena), _MaxPool())
        blocks, cin = [], 64
        for mid, n, stride in ((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)):
            for i in range(n):
                blocks.append(Bottleneck(cin, mid, stride if i == 0 else 1, device, arena))
                cin = mid * 4
        self.blocks = nn.Sequential(*blocks)
        self.fc = (nn.Linear(2048, num_classes, device=device, dtype=torch.bfloat16) if self.stock
                   else Linear(2048, (num_classes + 7) // 8 * 8, device=device))
        self.num_classes = num_classes

    def forward(self, images, labels):
        if self.stock or self.lib_conv:
            x = images.contiguous(memory_format=torch.channels_last)
        else:
            self._arena[0].reset(images.device)
            x = images          # the stem's space-to-depth kernel reads NCHW directly
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
