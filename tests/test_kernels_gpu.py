"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference (SURVEY.md §4 tier 4)."""
import math

import pytest
import torch

from tony_b200.ops import functional as F
from tony_b200.ops._lib import launch_count

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _rand(*shape, scale=1.0):
    return (torch.randn(*shape, device=DEV, dtype=torch.float32) * scale).to(torch.bfloat16)


def _close(out, ref, rel=2e-2, mean_rel=None):
    """Largest AND mean error, both relative to the reference's largest magnitude: a kernel that is wrong on
    every small element passes a max-norm test alone, one dropped term shows up in the mean."""
    ref = ref.float()
    scale = ref.abs().max().item()
    diff = (out.float() - ref).abs()
    err = diff.max().item()
    tol = rel * scale + 1e-3
    assert err <= tol, f"max err {err} > tol {tol}"
    mean_rel = rel / 5 if mean_rel is None else mean_rel
    mean_tol = mean_rel * scale + 2e-4
    assert diff.mean().item() <= mean_tol, f"mean err {diff.mean().item()} > tol {mean_tol}"


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (512, 1024, 512), (200, 264, 72), (1000, 136, 1032)])
def test_gemm_majors(a_mn, b_mn, M, N, K):
    from tony_b200.ops import functional as F

    A = _rand(M, K)
    B = _rand(N, K)
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    out = F.gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    _close(out, A.float() @ B.float().t())


def test_gemm_epilogues():
    from tony_b200.ops import functional as F
    from tony_b200.ops._lib import (EPI_ACCUMULATE, EPI_BIAS, EPI_DGELU_ERF, EPI_GELU_ERF,
                                    EPI_GELU_TANH, EPI_RELU, EPI_RESIDUAL, EPI_SAVE_PREACT)

    M, N, K = 384, 768, 256
    A, B, bias, aux = _rand(M, K), _rand(N, K), _rand(N), _rand(M, N)
    base = A.float() @ B.float().t()
    _close(F.gemm(A, B, flags=EPI_BIAS, bias=bias), base + bias.float())
    pre = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    out = F.gemm(A, B, flags=EPI_BIAS | EPI_GELU_ERF | EPI_SAVE_PREACT, bias=bias, aux_out=pre)
    _close(pre, base + bias.float())
    _close(out, torch.nn.functional.gelu(base + bias.float()))
    _close(F.gemm(A, B, flags=EPI_GELU_TANH), torch.nn.functional.gelu(base, approximate="tanh"))
    _close(F.gemm(A, B, flags=EPI_BIAS | EPI_RESIDUAL, bias=bias, aux=aux), base + bias.float() + aux.float())
    _close(F.gemm(A, B, flags=EPI_RELU), torch.relu(base))
    x = aux.float().requires_grad_(True)
    torch.nn.functional.gelu(x).sum().backward()
    _close(F.gemm(A, B, flags=EPI_DGELU_ERF, aux=aux), base * x.grad)
    c = _rand(M, N)
    c0 = c.clone()
    _close(F.gemm(A, B, out=c, flags=EPI_ACCUMULATE), base + c0.float())
    _close(F.gemm(A, B, alpha=0.25), base * 0.25)


def test_gemm_splitk():
    from tony_b200.ops import functional as F

    M, N, K = 1024, 1024, 8192  # wgrad-like: few tiles, long reduction
    dY, X = _rand(K, M), _rand(K, N)
    out = F.gemm(dY, X, a_mn=True, b_mn=True)
    _close(out, dY.float().t() @ X.float())


def test_linear_autograd():
    from tony_b200.ops import functional as F

    x = _rand(4, 96, 512).requires_grad_(True)
    w = _rand(768, 512, scale=0.05).requires_grad_(True)
    b = _rand(768).requires_grad_(True)
    res = _rand(4, 96, 768).requires_grad_(True)
    y = F.linear(x, w, b, res)
    g = _rand(4, 96, 768)
    y.backward(g)
    xr, wr, br, rr = (t.detach().float().requires_grad_(True) for t in (x, w, b, res))
    yr = torch.nn.functional.linear(xr, wr, br) + rr
    yr.backward(g.float())
    _close(y, yr)
    _close(x.grad, xr.grad)
    _close(w.grad, wr.grad)
    _close(b.grad, br.grad)
    _close(res.grad, rr.grad)


@pytest.mark.parametrize("act", ["gelu", "gelu_tanh", "relu"])
def test_mlp_autograd(act):
    from tony_b200.ops import functional as F

    x = _rand(512, 256).requires_grad_(True)
    w1, b1 = _rand(1024, 256, scale=0.06).requires_grad_(True), _rand(1024, scale=0.1).requires_grad_(True)
    w2, b2 = _rand(256, 1024, scale=0.03).requires_grad_(True), _rand(256, scale=0.1).requires_grad_(True)
    y = F.mlp(x, w1, b1, w2, b2, residual=x, act=act)
    g = _rand(512, 256)
    y.backward(g)
    ts = [t.detach().float().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    fn = {"gelu": torch.nn.functional.gelu, "relu": torch.relu,
          "gelu_tanh": lambda t: torch.nn.functional.gelu(t, approximate="tanh")}[act]
    # the kernel rounds the hidden activation to bf16 between the two GEMMs
    h = fn(torch.nn.functional.linear(ts[0], ts[1], ts[2]))
    yr = torch.nn.functional.linear(h, ts[3], ts[4]) + ts[0]
    yr.backward(g.float())
    _close(y, yr, rel=3e-2)
    for got, want in zip((x, w1, b1, w2, b2), ts):
        _close(got.grad, want.grad, rel=3e-2)


@pytest.mark.parametrize("H", [256, 768, 1024, 2048])
def test_layernorm(H):
    from tony_b200.ops import functional as F

    rows = 777
    x = _rand(rows, H, scale=2.0).requires_grad_(True)
    g = (_rand(H, scale=0.2).float() + 1).to(torch.bfloat16).requires_grad_(True)
    b = _rand(H, scale=0.2).requires_grad_(True)
    y = F.layer_norm(x, g, b, 1e-5)
    dy = _rand(rows, H)
    y.backward(dy)
    xr, gr, br = (t.detach().float().requires_grad_(True) for t in (x, g, b))
    yr = torch.nn.functional.layer_norm(xr, (H,), gr, br, 1e-5)
    yr.backward(dy.float())
    _close(y, yr)
    _close(x.grad, xr.grad)
    _close(g.grad, gr.grad)
    _close(b.grad, br.grad)


@pytest.mark.parametrize("V,ld", [(30522, 30528), (1000, 1000), (50257, 50304)])
def test_softmax_xent(V, ld):
    from tony_b200.ops import functional as F

    rows = 300
    logits = torch.zeros(rows, ld, device=DEV, dtype=torch.bfloat16)
    logits[:, :V] = _rand(rows, V, scale=3.0)
    tgt = torch.randint(0, V, (rows,), device=DEV)
    tgt[::7] = -100
    lr = logits[:, :V].detach().float().requires_grad_(True)
    loss_ref = torch.nn.functional.cross_entropy(lr, tgt, ignore_index=-100)
    loss_ref.backward()
    lg = logits.clone().requires_grad_(True)
    loss = F.softmax_cross_entropy(lg, tgt, vocab=V)
    loss.backward()
    assert abs(loss.item() - loss_ref.item()) <= 2e-3 * abs(loss_ref.item()) + 1e-3
    got = lg.grad
    assert got[:, V:].abs().max().item() == 0 if ld > V else True
    err = (got[:, :V].float() - lr.grad).abs().max().item()
    assert err <= 2e-2 * lr.grad.abs().max().item() + 1e-6


def test_colsum():
    from tony_b200.ops import functional as F

    x = _rand(5000, 1032)
    _close(F.colsum(x), x.float().sum(0), rel=1e-2)
    out = _rand(1032)
    o0 = out.clone()
    F.colsum(x, out=out, accumulate=True)
    _close(out, x.float().sum(0) + o0.float(), rel=1e-2)


@pytest.mark.parametrize("rows,cols", [(5000, 1032), (40000, 256), (16384, 1024), (70, 8)])
def test_colsum_two_level_fold(rows, cols):
    """Slab counts on both sides of the single-level / two-level switch, repeated so that a counter left
    non-zero by one call would show in the next."""
    x = _rand(rows, cols)
    ref = x.float().sum(0)
    for _ in range(3):
        _close(F.colsum(x), ref, rel=1e-2)


@pytest.mark.parametrize("act", ["relu", "gelu", "gelu_tanh"])
def test_act_bwd_bias_grad(act):
    rows, cols = 3000, 520
    dy, pre = _rand(rows, cols), _rand(rows, cols, scale=1.5)
    p = pre.float().requires_grad_(True)
    if act == "relu":
        y = torch.relu(p)
    else:
        y = torch.nn.functional.gelu(p, approximate="tanh" if act == "gelu_tanh" else "none")
    (ref,) = torch.autograd.grad(y, p, dy.float())
    bias = torch.zeros(cols, device=DEV, dtype=torch.bfloat16)
    dpre, db = F.act_bwd_bias_grad(dy, pre, act, bias)
    _close(dpre, ref, rel=1e-2)
    _close(db, ref.sum(0), rel=1e-2)
    dpre2, none = F.act_bwd_bias_grad(dy, pre, act, None)
    assert none is None and torch.equal(dpre2, dpre)


def test_linear_act_backward_matches_torch():
    """Fused Linear + activation against autograd on the same bf16-rounded operands."""
    x, w, b = _rand(1024, 256), _rand(512, 256, scale=0.05), _rand(512, scale=0.1)
    for act, fn in (("relu", torch.relu), ("gelu", torch.nn.functional.gelu)):
        xs = [t.clone().requires_grad_(True) for t in (x, w, b)]
        y = F.linear_act(xs[0], xs[1], xs[2], act=act)
        dy = _rand(*y.shape)
        y.backward(dy)
        rs = [t.float().clone().requires_grad_(True) for t in (x, w, b)]
        fn(rs[0] @ rs[1].t() + rs[2]).backward(dy.float())
        for got, ref in zip(xs, rs):
            _close(got.grad, ref.grad, rel=2e-2)


def test_pack3_rows_colsum_matches_separate_ops():
    """dq / dk / dv with their own row strides -> packed [rows, 3H] plus the column sums of what was written."""
    from tony_b200.ops import _lib

    rows, H = 3000, 512
    parts = [_rand(rows, H + 64)[:, :H] for _ in range(3)]          # row stride H + 64: not contiguous
    packed = torch.empty(rows, 3 * H, device=DEV, dtype=torch.bfloat16)
    db0 = _rand(3 * H)
    db = db0.clone()
    lib = _lib.lib()
    nbytes = int(lib.tony_colsum_scratch_bytes(3 * H))
    scratch = F.workspace("colsum", nbytes, DEV, counters=True)
    for acc in (0, 1):
        _lib.check(lib.tony_pack3_rows_colsum(_lib.ptr(parts[0]), _lib.ptr(parts[1]), _lib.ptr(parts[2]),
                                              parts[0].stride(0), parts[1].stride(0), parts[2].stride(0),
                                              _lib.ptr(packed), rows, H, _lib.ptr(db), _lib.ptr(scratch), nbytes, acc,
                                              _lib.cur_stream()), "tony_pack3_rows_colsum")
        want = torch.cat(parts, dim=1)
        assert torch.equal(packed, want)
        ref = want.float().sum(0) + (db0.float() if acc else 0)
        _close(db, ref, rel=1e-2)
        db = db0.clone()


def test_adamw_flat_matches_torch():
    from tony_b200.ops import optim as O

    n = 1 << 20
    p32 = torch.randn(n, device=DEV)
    ref = p32.clone().requires_grad_(True)
    opt = torch.optim.AdamW([ref], lr=1e-3, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01)
    st = O.FlatAdamW(p32.clone(), lr=1e-3, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.01)
    w = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    for _ in range(3):
        g = _rand(n)
        ref.grad = g.float()
        opt.step()
        st.step(g, w)
    assert (st.master - ref.detach()).abs().max().item() < 1e-5
    _close(w, ref.detach(), rel=1e-2)


def test_sgd_flat_matches_torch():
    from tony_b200.ops import optim as O

    n = 1 << 18
    p32 = torch.randn(n, device=DEV)
    ref = p32.clone().requires_grad_(True)
    opt = torch.optim.SGD([ref], lr=0.1, momentum=0.9, weight_decay=1e-4)
    st = O.FlatSGD(p32.clone(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    w = torch.empty(n, device=DEV, dtype=torch.bfloat16)
    for _ in range(3):
        g = _rand(n)
        ref.grad = g.float()
        opt.step()
        st.step(g, w)
    assert (st.master - ref.detach()).abs().max().item() < 1e-4


def test_gemm_from_fresh_thread():
    """Autograd runs backward on worker threads with no CUDA context bound yet; the GEMM's
    tensor-map encode (a raw driver call) must still work there."""
    import threading

    from tony_b200.ops import functional as F

    A, B = _rand(256, 128), _rand(256, 128)
    out = torch.empty(256, 256, device=DEV, dtype=torch.bfloat16)
    F.workspace("splitk", F.SPLITK_WORKSPACE_BYTES, A.device)
    torch.cuda.synchronize()
    err = []

    def run():
        try:
            F.gemm(A, B, out=out)
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            err.append(e)

    t = threading.Thread(target=run)
    t.start()
    t.join()
    assert not err, err
    _close(out, A.float() @ B.float().t())


def test_packed_attention_matches_sdpa():
    from tony_b200.ops import functional as F

    B, S, nh, hd = 2, 128, 4, 64
    qkv = _rand(B, S, 3, nh, hd).requires_grad_(True)
    ref_in = qkv.detach().clone().requires_grad_(True)
    g = _rand(B, S, nh * hd)
    for causal in (False, True):
        qkv.grad = None
        ref_in.grad = None
        out = F.packed_attention(qkv, causal)
        out.backward(g)
        q, k, v = (ref_in[:, :, i].transpose(1, 2) for i in range(3))
        ref = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=causal)
        ref = ref.transpose(1, 2).reshape(B, S, nh * hd)
        ref.backward(g)
        _close(out, ref)
        _close(qkv.grad, ref_in.grad)


@pytest.mark.parametrize("force", [128, 256, 512])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_gemm_every_kernel_variant(force, a_mn, b_mn):
    """1-CTA 128x128 / 128x256 kernels and the CTA-pair 256x256 kernel, forced through the _ex ABI."""
    from tony_b200.ops import _lib
    from tony_b200.ops._lib import EPI_BIAS, EPI_RESIDUAL

    M, N, K = 640, 776, 328
    A, B, bias, res = _rand(M, K), _rand(N, K), _rand(N), _rand(M, N)
    a = A.t().contiguous() if a_mn else A
    b = B.t().contiguous() if b_mn else B
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    rc = _lib.lib().tony_gemm_bf16_ex(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(out), M, N, K, int(a_mn), int(b_mn), a.stride(0), b.stride(0), N,
        EPI_BIAS | EPI_RESIDUAL, 1.0, _lib.ptr(bias), _lib.ptr(res), N, None, None, 0, force, 1, None, 0, 0, 0,
        _lib.cur_stream())
    _lib.check(rc, "tony_gemm_bf16_ex")
    _close(out, A.float() @ B.float().t() + bias.float() + res.float())


def test_pack3_rows():
    from tony_b200.ops import _lib

    rows, H = 1000, 1024
    a, b, c = _rand(rows, H), _rand(rows, 2 * H)[:, :H], _rand(rows, H)
    out = torch.empty(rows, 3, H, device=DEV, dtype=torch.bfloat16)
    _lib.check(_lib.lib().tony_pack3_rows(_lib.ptr(a), _lib.ptr(b), _lib.ptr(c), a.stride(0), b.stride(0),
                                          c.stride(0), _lib.ptr(out), rows, H, _lib.cur_stream()), "pack3")
    assert torch.equal(out[:, 0], a) and torch.equal(out[:, 1], b) and torch.equal(out[:, 2], c)


def test_layernorm_bwd_h4096_and_accumulate():
    from tony_b200.ops import _lib
    from tony_b200.ops import functional as F

    rows, H = 300, 4096
    x, dy = _rand(rows, H, scale=2.0), _rand(rows, H)
    g = (_rand(H, scale=0.2).float() + 1).to(torch.bfloat16)
    xr, gr = x.float().requires_grad_(True), g.float().requires_grad_(True)
    br = torch.zeros(H, device=DEV, requires_grad=True)
    torch.nn.functional.layer_norm(xr, (H,), gr, br, 1e-5).backward(dy.float())
    mean = x.float().mean(1)
    rstd = (x.float().var(1, unbiased=False) + 1e-5).rsqrt()
    dx = torch.empty_like(x)
    dg0, db0 = _rand(H), _rand(H)
    dg, db = dg0.clone(), db0.clone()
    lib = _lib.lib()
    nb = int(lib.tony_layernorm_bwd_scratch_bytes(H))
    scratch = F.workspace("ln_bwd", nb, x.device, counters=True)
    _lib.check(lib.tony_layernorm_bwd(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(g), _lib.ptr(mean), _lib.ptr(rstd),
                                      _lib.ptr(dx), _lib.ptr(dg), _lib.ptr(db), rows, H, _lib.ptr(scratch), nb, 1,
                                      _lib.cur_stream()), "ln_bwd")
    _close(dx, xr.grad)
    _close(dg, gr.grad + dg0.float())
    _close(db, br.grad + db0.float())


def _attention_reference(qkv, causal):
    q, k, v = (qkv[:, :, i].transpose(1, 2).float() for i in range(3))       # [B, nh, S, hd]
    s = q @ k.transpose(-1, -2) / math.sqrt(q.shape[-1])
    if causal:
        S = s.shape[-1]
        s = s.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device=s.device), 1), float("-inf"))
    o = torch.softmax(s, -1) @ v
    return o.transpose(1, 2).reshape(qkv.shape[0], qkv.shape[1], -1)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("B,S,nh,gain", [(2, 256, 4, 1.0), (1, 512, 2, 1.0), (1, 384, 3, 6.0)])
def test_fused_attention_matches_fp32_reference(B, S, nh, gain, causal, monkeypatch):
    """tcgen05 attention forward + backward vs. plain fp32 softmax(QK^T)V and its autograd.
    ``gain`` 6 makes the row max grow by far more than 2^8 between key blocks, which exercises
    the lazy rescale of the TMEM-resident output."""
    monkeypatch.setenv("TONY_B200_ATTENTION", "tcgen05")
    torch.manual_seed(7)
    dev = torch.device("cuda")
    qkv = (torch.randn(B, S, 3, nh, 64, device=dev) * gain).to(torch.bfloat16)
    if gain > 1:
        qkv[:, :, 2] = qkv[:, :, 2] / gain        # keep V O(1)
        # put the largest keys late in the sequence so every row's max keeps moving
        qkv[:, :, 1] = qkv[:, :, 1] * torch.linspace(0.2, 1.0, S, device=dev).view(1, S, 1, 1).to(torch.bfloat16)
    a = qkv.clone().requires_grad_()
    b = qkv.float().requires_grad_()
    out = F.packed_attention(a, causal)
    ref = _attention_reference(b, causal)
    assert out.shape == ref.shape
    err = (out.float() - ref).abs().max().item()
    assert err < 3e-2 * max(1.0, ref.abs().max().item()), err
    g = torch.randn_like(ref)
    out.backward(g.to(torch.bfloat16))
    ref.backward(g)
    for i, name in enumerate(("dq", "dk", "dv")):
        got, want = a.grad[:, :, i].float(), b.grad[:, :, i]
        scale = want.abs().max().item()
        e = (got - want).abs().max().item()
        assert e < 4e-2 * scale + 1e-3, (name, e, scale)
        # the mean error must be far smaller than the worst element (catches a wrong tile)
        assert (got - want).abs().mean().item() < 6e-3 * scale + 1e-4, name


def test_fused_attention_library_switch(monkeypatch):
    torch.manual_seed(1)
    qkv = torch.randn(1, 128, 3, 2, 64, device="cuda").to(torch.bfloat16)
    monkeypatch.setenv("TONY_B200_ATTENTION", "tcgen05")
    before = launch_count()
    ours = F.packed_attention(qkv)
    assert launch_count() == before + 1          # the own forward kernel really ran
    monkeypatch.setenv("TONY_B200_ATTENTION", "library")
    lib_out = F.packed_attention(qkv)
    assert launch_count() == before + 1
    assert (ours.float() - lib_out.float()).abs().max().item() < 2e-2


def test_residual_blocks_match_unfused_composition():
    """attention_block / mlp(residual=x) fuse the residual branch's gradient into the last dgrad
    GEMM; they must match the same block built from separate linear / attention calls."""
    torch.manual_seed(3)
    B, S, H, nh = 2, 128, 256, 4
    x = _rand(B, S, H)
    wqkv, bqkv = _rand(3 * H, H, scale=0.05), _rand(3 * H, scale=0.1)
    wo, bo = _rand(H, H, scale=0.05), _rand(H, scale=0.1)
    w1, b1 = _rand(4 * H, H, scale=0.05), _rand(4 * H, scale=0.1)
    w2, b2 = _rand(H, 4 * H, scale=0.05), _rand(H, scale=0.1)
    g = _rand(B, S, H)

    def run(fused):
        leaves = [t.clone().requires_grad_() for t in (x, wqkv, bqkv, wo, bo, w1, b1, w2, b2)]
        xi, a, ab, o, ob, f1, f1b, f2, f2b = leaves
        if fused:
            y = F.attention_block(xi, a, ab, o, ob, nh)
            z = F.mlp(y, f1, f1b, f2, f2b, residual=y)
        else:
            qkv = F.linear(xi, a, ab).view(B, S, 3, nh, H // nh)
            y = F.linear(F.packed_attention(qkv), o, ob) + xi
            z = F.mlp(y, f1, f1b, f2, f2b) + y
        z.backward(g)
        return [z] + [t.grad for t in leaves]

    for got, want in zip(run(True), run(False)):
        _close(got, want, rel=3e-2)


@pytest.mark.parametrize("with_type", [True, False])
def test_embedding_sum_forward_backward(with_type):
    """Fused word + type + position embedding vs. fp32 gathers; small vocabulary so many tokens
    collide on the same word row (exercises the bf16x2 reductions)."""
    torch.manual_seed(5)
    B, S, H, V, P = 4, 128, 512, 97, 160
    ids = torch.randint(0, V, (B, S), device=DEV)
    tt = torch.randint(0, 2, (B, S), device=DEV) if with_type else None
    word, typ, pos = _rand(V, H), (_rand(2, H) if with_type else None), _rand(P, H)
    leaves = [t.clone().requires_grad_() for t in (word, typ, pos) if t is not None]
    w, ty, ps = (leaves[0], leaves[1], leaves[2]) if with_type else (leaves[0], None, leaves[1])
    out = F.embedding_sum(ids, tt, w, ty, ps)
    ref_leaves = [t.float().requires_grad_() for t in (word, typ, pos) if t is not None]
    rw, rty, rps = (ref_leaves[0], ref_leaves[1], ref_leaves[2]) if with_type else (ref_leaves[0], None, ref_leaves[1])
    ref = rw[ids] + rps[:S].unsqueeze(0)
    if with_type:
        ref = ref + rty[tt]
    _close(out, ref, rel=1e-2)
    g = _rand(B, S, H, scale=0.1)
    out.backward(g)
    ref.backward(g.float())
    for got, want in zip(leaves, ref_leaves):
        _close(got.grad, want.grad, rel=2e-2)
    assert ps.grad[S:].abs().max().item() == 0.0


def test_sharded_embedding_bag_single_gpu():
    """K9 at world = 1: gather = row lookup into the concatenated activation, scatter = sparse
    accumulation into the gradient shard, tony_step = AdamW on the shard."""
    from tony_b200.parallel.ps import ShardedEmbeddingBag

    torch.manual_seed(2)
    Fd, V, D, R, B = 3, 50, 16, 8, 64
    bag = ShardedEmbeddingBag(Fd, V, D, R, std=0.5, device=DEV, seed=7)
    bag.tony_configure(lr=0.1)
    gen = torch.Generator(device=DEV)
    full = []
    for f in range(Fd):
        gen.manual_seed(7 + f)
        full.append((torch.randn(V, D + R, generator=gen, device=DEV) * 0.5).to(torch.bfloat16).float())
    ids = torch.randint(0, V, (B, Fd), device=DEV)
    deep, rest = bag(ids)
    ref_deep = torch.cat([full[f][ids[:, f], :D] for f in range(Fd)], dim=1)
    ref_rest = torch.stack([full[f][ids[:, f], D:] for f in range(Fd)], dim=1)
    assert torch.equal(deep.float(), ref_deep) and torch.equal(rest.float(), ref_rest)
    gd, gr = _rand(B, Fd * D, scale=0.1), _rand(B, Fd, R, scale=0.1)
    torch.autograd.backward([deep, rest], [gd, gr])
    ref_grad = [torch.zeros(V, D + R, device=DEV) for _ in range(Fd)]
    for f in range(Fd):
        ref_grad[f].index_add_(0, ids[:, f], torch.cat([gd[:, f * D:(f + 1) * D], gr[:, f]], dim=1).float())
    n = Fd * bag.rows_per_shard * bag.row_elems
    got = bag.grad[:n].view(Fd, V, D + R).float()
    for f in range(Fd):
        _close(got[f], ref_grad[f], rel=3e-2)
    before = bag.full_table(0).float().clone()
    bag.tony_step()
    after = bag.full_table(0).float()
    touched = ref_grad[0].abs().sum(1) > 0
    assert torch.equal(after[~touched], before[~touched])            # no weight decay: untouched rows stay
    step = (before - after)[touched]
    big = ref_grad[0][touched].abs() > 1e-2
    # first AdamW step moves every coordinate by ~lr in the direction of its gradient
    assert ((step[big] * torch.sign(ref_grad[0][touched][big])) > 0.05).all()
    assert bag.grad.abs().max().item() == 0.0                         # shard cleared for the next step


@pytest.mark.parametrize("C,relu,with_res", [(64, True, False), (256, True, True), (512, False, False),
                                             (2048, True, True), (128, True, True)])
def test_fused_batchnorm_matches_aten(C, relu, with_res):
    """BatchNorm (+ residual) (+ ReLU) forward / backward on a channels-last activation vs. the same
    computation in fp32 through ATen autograd (training mode, batch statistics)."""
    torch.manual_seed(4)
    N, H, W = 8, 7, 9 if C >= 512 else 14
    x = (torch.randn(N, C, H, W, device=DEV) * 1.5 + 0.3).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = _rand(N, C, H, W).contiguous(memory_format=torch.channels_last) if with_res else None
    gamma, beta = _rand(C, scale=0.5) + 1.0, _rand(C, scale=0.2)
    rm = torch.zeros(C, device=DEV)
    rv = torch.ones(C, device=DEV)
    leaves = [t.clone().requires_grad_() for t in (x, gamma, beta)] + ([res.clone().requires_grad_()] if with_res else [])
    y = F.batch_norm_act(leaves[0], leaves[1], leaves[2], rm, rv, leaves[3] if with_res else None, relu=relu)
    ref_leaves = [t.float().requires_grad_() for t in (x, gamma, beta)] + ([res.float().requires_grad_()] if with_res else [])
    rrm, rrv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    ref = torch.nn.functional.batch_norm(ref_leaves[0], rrm, rrv, ref_leaves[1], ref_leaves[2], True, 0.1, 1e-5)
    if with_res:
        ref = ref + ref_leaves[3]
    if relu:
        ref = torch.relu(ref)
    _close(y, ref, rel=2e-2)
    assert y.is_contiguous(memory_format=torch.channels_last)
    _close(rm, rrm, rel=2e-2)
    _close(rv, rrv, rel=2e-2)
    g = _rand(N, C, H, W).contiguous(memory_format=torch.channels_last)
    y.backward(g)
    ref.backward(g.float())
    for got, want in zip(leaves, ref_leaves):
        # elements whose pre-activation is within bf16 rounding of 0 may flip the ReLU mask: compare
        # in aggregate (mean error) and allow a handful of outliers in the max
        d = (got.grad.float() - want.grad).abs()
        scale = want.grad.abs().max().item() + 1e-6
        assert d.mean().item() < 4e-3 * scale, (d.mean().item(), scale)
        assert (d > 5e-2 * scale).float().mean().item() < 2e-3


@pytest.mark.parametrize("H,W", [(30, 23), (31, 16), (7, 7), (112, 112)])
def test_maxpool_3x3_s2_matches_aten(H, W):
    """Even / odd extents: the backward works on 2 x 2 pixel blocks, the last block row / column may be half empty."""
    torch.manual_seed(6)
    x = _rand(2, 64, H, W).contiguous(memory_format=torch.channels_last)
    a = x.clone().requires_grad_()
    b = x.float().requires_grad_()
    y = F.max_pool_3x3_s2(a)
    ref = torch.nn.functional.max_pool2d(b, 3, 2, 1)
    assert torch.equal(y.float(), ref)
    g = _rand(*ref.shape).contiguous(memory_format=torch.channels_last)
    y.backward(g)
    ref.backward(g.float())
    _close(a.grad, b.grad, rel=1e-2)


@pytest.mark.parametrize("M,C", [(105, 64), (1001, 8), (77, 256), (4099, 1024), (33, 2048)])
def test_batchnorm_kernels_stay_inside_their_buffers(M, C):
    """compute-sanitizer is not available on the GPU pool, so the fused BatchNorm entry points are
    called directly on guard-banded buffers with awkward row counts (odd, not foldable, smaller than
    a slab): every output must match the reference and no byte of any guard band may change."""

    from tony_b200.ops import _lib
    from tony_b200.ops._lib import check, cur_stream, ptr

    torch.manual_seed(8)
    G = 4096   # guard elements on each side

    def guarded(n, dtype, fill):
        buf = torch.full((n + 2 * G,), fill, dtype=dtype, device=DEV)
        return buf, buf[G:G + n]

    x = (torch.randn(M, C, device=DEV) * 1.3 + 0.2).to(torch.bfloat16)
    res = _rand(M, C)
    gamma, beta = _rand(C, scale=0.5) + 1.0, _rand(C, scale=0.2)
    ybuf, y = guarded(M * C, torch.bfloat16, 7.0)
    mbuf, mask = guarded(M * C // 8, torch.uint8, 0xAB)
    meanb, mean = guarded(C, torch.float32, 3.0)
    rstdb, rstd = guarded(C, torch.float32, 3.0)
    nbytes = _lib.lib().tony_bn_scratch_bytes(C)
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device=DEV)
    check(_lib.lib().tony_bn_fwd(ptr(x), ptr(res), ptr(gamma), ptr(beta), ptr(y), ptr(mask), ptr(mean), ptr(rstd),
                                 None, None, M, C, 1e-5, 0.1, 1, ptr(scratch), nbytes, cur_stream()), "tony_bn_fwd")
    xf = x.float()
    mu, var = xf.mean(0), xf.var(0, unbiased=False)
    ref = torch.relu((xf - mu) * torch.rsqrt(var + 1e-5) * gamma.float() + beta.float() + res.float())
    _close(y.view(M, C), ref, rel=2e-2)
    _close(mean, mu, rel=1e-2)
    dy = _rand(M, C)
    dxb, dx = guarded(M * C, torch.bfloat16, 7.0)
    drb, dres = guarded(M * C, torch.bfloat16, 7.0)
    dgb, dgamma = guarded(C, torch.bfloat16, 7.0)
    dbb, dbeta = guarded(C, torch.bfloat16, 7.0)
    check(_lib.lib().tony_bn_bwd(ptr(dy), ptr(x), ptr(mask), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx), ptr(dres),
                                 ptr(dgamma), ptr(dbeta), M, C, 1, ptr(scratch), nbytes, cur_stream()), "tony_bn_bwd")
    torch.cuda.synchronize()
    dym = dy.float() * (ref > 0)
    _close(dres.view(M, C), dym, rel=2e-2)
    _close(dbeta, dym.sum(0), rel=3e-2)
    for buf, fill in ((ybuf, 7.0), (dxb, 7.0), (drb, 7.0), (dgb, 7.0), (dbb, 7.0), (meanb, 3.0), (rstdb, 3.0)):
        assert (buf[:G].float() == fill).all() and (buf[-G:].float() == fill).all(), "write outside the buffer"
    assert (mbuf[:G] == 0xAB).all() and (mbuf[-G:] == 0xAB).all(), "mask written outside its buffer"
    assert torch.isfinite(dx.float()).all()


@pytest.mark.parametrize("M,N,K,acc", [(1024, 1024, 16384, False), (4096, 1024, 4096, True), (3072, 1024, 2048, False),
                                       (520, 264, 1000, True)])
def test_wgrad_gemm_carries_the_bias_gradient(M, N, K, acc):
    """dW = dY^T X with rowsum_out = colsum(dY) produced inside the GEMM (ones-tile MMA), including
    split-K shapes, ragged edges and accumulation into existing gradients."""
    torch.manual_seed(9)
    dy, x = _rand(K, M, scale=0.5), _rand(K, N, scale=0.5)        # dY [tokens, out], X [tokens, in]
    dw0, db0 = _rand(M, N), _rand(M)
    dw, db = dw0.clone(), db0.clone()
    F.gemm(dy, x, a_mn=True, b_mn=True, out=dw, flags=F.EPI_ACCUMULATE if acc else 0, rowsum_out=db,
           rowsum_accumulate=acc)
    ref_w = dy.float().t() @ x.float() + (dw0.float() if acc else 0)
    ref_b = dy.float().sum(0) + (db0.float() if acc else 0)
    _close(dw, ref_w, rel=2e-2)
    _close(db, ref_b, rel=2e-2)


def test_layer_norm_residual_fuses_the_branch_gradient():
    torch.manual_seed(12)
    x, g, b = _rand(6, 40, 1024), _rand(1024, scale=0.3) + 1.0, _rand(1024, scale=0.1)
    w = _rand(1024, 1024, scale=0.03)
    la = [t.clone().requires_grad_() for t in (x, g, b, w)]
    h, r = F.layer_norm_residual(la[0], la[1], la[2])
    ya = F.linear(h, la[3], None, residual=r)
    lb = [t.clone().requires_grad_() for t in (x, g, b, w)]
    yb = F.linear(F.layer_norm(lb[0], lb[1], lb[2]), lb[3], None) + lb[0]
    _close(ya, yb, rel=1e-2)
    go = _rand(6, 40, 1024)
    ya.backward(go)
    yb.backward(go)
    for p_, q_ in zip(la, lb):
        _close(p_.grad, q_.grad, rel=2e-2)
