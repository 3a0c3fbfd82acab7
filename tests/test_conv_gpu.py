"""Implicit-GEMM convolution kernels (csrc/conv_tcgen05.cu) against fp32 ``F.conv2d`` and its autograd.

Every check bounds both the largest and the mean error: a kernel that is wrong on every small element
passes a max-norm test, and one wrong tap or one wrong image border shows up in the mean."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _close(got, want, tol=2e-2, mean_tol=4e-3, what=""):
    got, want = got.float(), want.float()
    scale = want.abs().max().item() + 1e-6
    err = (got - want).abs()
    assert err.max().item() <= tol * scale, f"{what}: max err {err.max().item():.4g} vs scale {scale:.4g}"
    assert err.mean().item() <= mean_tol * scale, f"{what}: mean err {err.mean().item():.4g} vs scale {scale:.4g}"


CASES = [
    # N, H, W, C, K, R, stride, pad
    (2, 14, 14, 64, 128, 3, 1, 1),
    (3, 7, 7, 128, 64, 3, 1, 1),        # 147 pixels: a tile spans images and ends past the last one
    (2, 16, 16, 64, 64, 3, 2, 1),
    (2, 14, 14, 128, 256, 1, 2, 0),     # strided pointwise (ResNet shortcut)
    (1, 20, 12, 192, 320, 3, 1, 1),     # non-square, several 64-channel blocks, N tail over 256
    (2, 9, 11, 64, 72, 3, 2, 1),        # odd extents with stride 2, K % 64 != 0
    (4, 8, 8, 64, 64, 1, 1, 0),         # pointwise through the conv kernel itself (statistics path)
    (3, 20, 20, 64, 64, 3, 1, 1),       # halo kernel, whole filter resident; ragged 16 x 8 tiles in both directions
    (2, 33, 9, 128, 128, 3, 1, 1),      # halo kernel, filter ring, two channel blocks, three row tiles
]


@pytest.fixture(params=["im2col", "halo"])
def conv_form(request, monkeypatch):
    """3x3 / stride-1 layers with <= 128 output channels can also run on the halo kernel (opt-in switch)."""
    monkeypatch.setenv("TONY_B200_CONV_HALO", "1" if request.param == "halo" else "0")
    return request.param


@pytest.mark.parametrize("N,H,W,C,K,R,stride,pad", CASES)
def test_conv2d_fprop_dgrad_wgrad_match_fp32(N, H, W, C, K, R, stride, pad, conv_form):
    if conv_form == "halo" and not (R == 3 and stride == 1 and pad == 1):
        pytest.skip("halo form covers 3x3 / stride 1 only")
    from tony_b200.ops import functional as F

    dev = torch.device("cuda")
    torch.manual_seed(N * 1000 + H * 10 + C)
    x = torch.randn(N, C, H, W, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(K, R, R, C, device=dev) / (R * C ** 0.5)).to(torch.bfloat16)
    xn = x.permute(0, 2, 3, 1)
    stats = torch.zeros(2, K, dtype=torch.float32, device=dev)
    y = F.conv2d_nhwc(xn, w, stride, pad, stats=stats)

    xr = x.float().requires_grad_(True)
    wr = w.float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)     # KCRS for torch
    yr = torch.nn.functional.conv2d(xr, wr, stride=stride, padding=pad)
    _close(y.permute(0, 3, 1, 2), yr, what="fprop")
    yb = y.float()
    _close(stats[0], yb.sum(dim=(0, 1, 2)), tol=2e-3, mean_tol=5e-4, what="stat sum")
    _close(stats[1], (yb * yb).sum(dim=(0, 1, 2)), tol=2e-3, mean_tol=5e-4, what="stat sqsum")

    dy = torch.randn_like(yr).to(torch.bfloat16)
    yr.backward(dy.float())
    dyn = dy.permute(0, 2, 3, 1).contiguous()
    dx = F.conv2d_dgrad_nhwc(dyn, w, xn.shape, stride, pad)
    _close(dx.permute(0, 3, 1, 2), xr.grad, what="dgrad")
    # accumulate form: dx2 = base + dgrad
    base = torch.randn_like(dx)
    dx2 = F.conv2d_dgrad_nhwc(dyn, w, xn.shape, stride, pad, out=base.clone(), accumulate=True)
    _close(dx2.permute(0, 3, 1, 2), xr.grad + base.float().permute(0, 3, 1, 2), what="dgrad accumulate")

    dw = F.conv2d_wgrad_nhwc(dyn, xn, w.shape, stride, pad)
    _close(dw.permute(0, 3, 1, 2), wr.grad, what="wgrad")
    dw2 = F.conv2d_wgrad_nhwc(dyn, xn, w.shape, stride, pad, out=w.clone(), accumulate=True)
    _close(dw2.permute(0, 3, 1, 2), wr.grad + wr.detach(), what="wgrad accumulate")


def test_conv2d_wgrad_split_k_large_pixel_count():
    """Many output pixels and one output tile row: the split-K path with > 100 splits."""
    from tony_b200.ops import functional as F

    dev = torch.device("cuda")
    torch.manual_seed(3)
    N, H, W, C, K = 16, 56, 56, 64, 64
    x = torch.randn(N, H, W, C, device=dev).to(torch.bfloat16)
    dy = (torch.randn(N, H, W, K, device=dev) * 0.05).to(torch.bfloat16)
    dw = F.conv2d_wgrad_nhwc(dy, x, (K, 3, 3, C), 1, 1)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    wr = torch.zeros(K, C, 3, 3, device=dev, requires_grad=True)
    torch.nn.functional.conv2d(xr, wr, padding=1).backward(dy.float().permute(0, 3, 1, 2))
    _close(dw.permute(0, 3, 1, 2), wr.grad, what="wgrad split-K")


def test_conv2d_autograd_function_channels_last():
    """The autograd wrapper on [N, C, H, W] channels-last tensors: 3x3, strided 3x3, pointwise (GEMM path)."""
    from tony_b200.ops import functional as F

    dev = torch.device("cuda")
    torch.manual_seed(4)
    for (C, K, R, stride) in ((64, 128, 3, 1), (128, 128, 3, 2), (128, 256, 1, 1), (64, 256, 1, 2)):
        x = torch.randn(4, C, 12, 12, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        x.requires_grad_(True)
        w = (torch.randn(K, R, R, C, device=dev) / (R * C ** 0.5)).to(torch.bfloat16).requires_grad_(True)
        y = F.conv2d(x, w, stride, R // 2)
        dy = torch.randn_like(y)
        y.backward(dy)
        xr = x.detach().float().requires_grad_(True)
        wr = w.detach().float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
        yr = torch.nn.functional.conv2d(xr, wr, stride=stride, padding=R // 2)
        yr.backward(dy.float())
        _close(y, yr, what=f"fwd {C}->{K} {R}x{R}/{stride}")
        _close(x.grad, xr.grad, what=f"dx {C}->{K} {R}x{R}/{stride}")
        _close(w.grad.permute(0, 3, 1, 2), wr.grad, what=f"dw {C}->{K} {R}x{R}/{stride}")


@pytest.mark.parametrize("stride_d,with_stats", [(1, False), (2, False), (2, True), (1, True)])
def test_conv1x1_pair_accumulates_both_data_gradients(stride_d, with_stats):
    """Projection block entry: main 1x1 and (strided) shortcut 1x1 read the same input; dx must be the SUM of
    both data gradients (the shortcut's kernel accumulates into the main path's result)."""
    from tony_b200.ops import functional as F

    dev = torch.device("cuda")
    torch.manual_seed(11)
    C, K1, Kd = 128, 64, 256
    x = torch.randn(4, C, 12, 12, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    x.requires_grad_(True)
    w1 = (torch.randn(K1, 1, 1, C, device=dev) / C ** 0.5).to(torch.bfloat16).requires_grad_(True)
    wd = (torch.randn(Kd, 1, 1, C, device=dev) / C ** 0.5).to(torch.bfloat16).requires_grad_(True)
    s1 = torch.zeros(2, K1, device=dev) if with_stats else None
    sd = torch.zeros(2, Kd, device=dev) if with_stats else None
    y1, yd = F.conv1x1_pair(x, w1, wd, stride_d, s1, sd)
    g1, gd = torch.randn_like(y1), torch.randn_like(yd)
    torch.autograd.backward([y1, yd], [g1, gd])
    xr = x.detach().float().requires_grad_(True)
    w1r = w1.detach().float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    wdr = wd.detach().float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
    r1 = torch.nn.functional.conv2d(xr, w1r)
    rd = torch.nn.functional.conv2d(xr, wdr, stride=stride_d)
    torch.autograd.backward([r1, rd], [g1.float(), gd.float()])
    _close(y1, r1, what="main fwd")
    _close(yd, rd, what="shortcut fwd")
    _close(x.grad, xr.grad, what="dx (sum of both branches)")
    _close(w1.grad.permute(0, 3, 1, 2), w1r.grad, what="dw main")
    _close(wd.grad.permute(0, 3, 1, 2), wdr.grad, what="dw shortcut")
    if with_stats:
        _close(sd[0], yd.float().sum(dim=(0, 2, 3)), tol=1e-2, what="shortcut column sums")


def test_stem_conv7x7s2_space_to_depth_matches_fp32():
    """The 3-channel 7x7/2 stem through the space-to-depth + overlapping-window tensor map."""
    from tony_b200.ops import functional as F

    dev = torch.device("cuda")
    torch.manual_seed(9)
    for (N, H, W) in ((3, 32, 32), (2, 64, 48)):
        x = torch.randn(N, 3, H, W, device=dev).to(torch.bfloat16)
        w = (torch.randn(64, 7, 7, 3, device=dev) / 12.0).to(torch.bfloat16).requires_grad_(True)
        stats = torch.zeros(2, 64, dtype=torch.float32, device=dev)
        y = F.stem_conv7x7s2(x, w, stats=stats)
        dy = torch.randn_like(y)
        y.backward(dy)
        wr = w.detach().float().permute(0, 3, 1, 2).contiguous().requires_grad_(True)
        yr = torch.nn.functional.conv2d(x.float(), wr, stride=2, padding=3)
        yr.backward(dy.float())
        _close(y, yr, what="stem fwd")
        _close(w.grad.permute(0, 3, 1, 2), wr.grad, what="stem dw")
        _close(stats[0], y.float().sum(dim=(0, 2, 3)), tol=2e-3, mean_tol=5e-4, what="stem stat sum")
