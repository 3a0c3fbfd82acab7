"""Autograd-aware functional API over the sm_100a kernels.

Every op here launches hand-written kernels from ``libtony_b200_kernels.so``; none has an
eager fallback.  The ops mirror what the jobs launched by the reference spend their time in
(``nn.Linear`` / ``F.cross_entropy`` in tony-examples/mnist-pytorch/mnist_distributed.py:76-90,129-136;
dense layers in tony-examples/horovod-on-tony/tensorflow2_mnist.py), re-designed so that

* bias / GELU / residual / dGELU live in the GEMM epilogue,
* forward, dgrad and wgrad of a Linear run off the same buffers with no transpose copies,
* weight/bias/LayerNorm gradients are written *directly into the flat gradient bucket* that
  the fused all-reduce+optimizer kernel consumes (see ``tony_b200.parallel.flat``) — the
  parameter's ``_tony_slot`` attribute carries that destination.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch

from . import _lib
from ._lib import (
    EPI_ACCUMULATE,
    EPI_BIAS,
    EPI_DGELU_ERF,
    EPI_DGELU_TANH,
    EPI_DRELU,
    EPI_GELU_ERF,
    EPI_GELU_TANH,
    EPI_RELU,
    EPI_RESIDUAL,
    EPI_SAVE_PREACT,
    check,
    cur_stream,
    ptr,
)

_ACT_FWD = {None: 0, "none": 0, "gelu": EPI_GELU_ERF, "gelu_tanh": EPI_GELU_TANH, "relu": EPI_RELU}
_ACT_BWD = {"gelu": EPI_DGELU_ERF, "gelu_tanh": EPI_DGELU_TANH, "relu": EPI_DRELU}

_workspaces: dict = {}


def workspace(name: str, nbytes: int, device, counters: bool = False) -> torch.Tensor:
    """Per-device cached scratch buffer (never shrinks).  ``counters=True``: the kernels keep arrival counters in
    the first 4 KB and leave them at zero, so that head is zeroed here once, when the buffer is created."""
    key = (name, torch.device(device).index)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        if counters:
            buf[:4096].zero_()
        _workspaces[key] = buf
    return buf


SPLITK_WORKSPACE_BYTES = 128 << 20


def _check_bf16(*tensors):
    for t in tensors:
        if t is not None and (t.dtype != torch.bfloat16 or not t.is_cuda):
            raise TypeError("tony_b200 kernels take CUDA bf16 tensors")


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, flags: int = 0, bias: Optional[torch.Tensor] = None,
         aux: Optional[torch.Tensor] = None, aux_out: Optional[torch.Tensor] = None,
         alpha: float = 1.0, allow_splitk: bool = True, rowsum_out: Optional[torch.Tensor] = None,
         rowsum_accumulate: bool = False) -> torch.Tensor:
    """``out[m,n] = epi(alpha * sum_k A(m,k) B(n,k))`` on the tcgen05 tensor cores.

    ``a`` is ``[M,K]`` (``a_mn=False``) or ``[K,M]`` (``a_mn=True``); ``b`` is ``[N,K]`` or
    ``[K,N]``.  2-D, last dim contiguous, row stride a multiple of 8 elements.
    """
    _check_bf16(a, b, bias, aux, aux_out, out)
    if a.dim() != 2 or b.dim() != 2 or a.stride(1) != 1 or b.stride(1) != 1:
        raise ValueError("gemm operands must be 2-D with a contiguous last dim")
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    if K != Kb:
        raise ValueError(f"reduction dims differ: {K} vs {Kb}")
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    elif out.shape != (M, N) or out.stride(1) != 1:
        raise ValueError("bad out tensor")
    ws = workspace("splitk", SPLITK_WORKSPACE_BYTES, a.device) if allow_splitk else None
    if rowsum_out is not None:
        # rowsum_out[m] = sum_k A(m,k) rides in the GEMM (bias gradient of a wgrad GEMM)
        rc = _lib.lib().tony_gemm_bf16_rowsum(
            ptr(a), ptr(b), ptr(out), M, N, K, int(a_mn), int(b_mn), a.stride(0), b.stride(0),
            out.stride(0), flags, float(alpha), ptr(bias), ptr(aux),
            aux.stride(0) if aux is not None else 0, ptr(aux_out), ptr(ws),
            ws.numel() if ws is not None else 0, ptr(rowsum_out), int(rowsum_accumulate), cur_stream())
        check(rc, "tony_gemm_bf16_rowsum")
        return out
    rc = _lib.lib().tony_gemm_bf16(
        ptr(a), ptr(b), ptr(out), M, N, K, int(a_mn), int(b_mn), a.stride(0), b.stride(0),
        out.stride(0), flags, float(alpha), ptr(bias), ptr(aux),
        aux.stride(0) if aux is not None else 0, ptr(aux_out), ptr(ws),
        ws.numel() if ws is not None else 0, cur_stream())
    check(rc, "tony_gemm_bf16")
    return out


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """``out[c] (+)= sum_r x[r,c]`` (bias gradient)."""
    _check_bf16(x, out)
    rows, cols = x.shape
    if out is None:
        out = torch.empty(cols, dtype=torch.bfloat16, device=x.device)
        accumulate = False
    lib = _lib.lib()
    nbytes = int(lib.tony_colsum_scratch_bytes(cols))
    scratch = workspace("colsum", nbytes, x.device, counters=True)
    check(lib.tony_colsum_bf16(ptr(x), rows, cols, x.stride(0), ptr(out), ptr(scratch), nbytes,
                               int(accumulate), cur_stream()), "tony_colsum_bf16")
    return out


# ---------------------------------------------------------------------------------------
# gradient destinations: either a fresh tensor (plain autograd) or the parameter's slot in
# the flat gradient bucket (tony_b200.parallel.flat.ParamSlot)
# ---------------------------------------------------------------------------------------
class _GradDest:
    __slots__ = ("slot", "tensor", "accumulate")

    def __init__(self, param: Optional[torch.Tensor], shape=None):
        self.slot = getattr(param, "_tony_slot", None) if param is not None else None
        if self.slot is not None:
            self.tensor = self.slot.grad_view
            self.accumulate = self.slot.begin_write()
        else:
            self.tensor = None
            self.accumulate = False

    def finish(self, produced: Optional[torch.Tensor]):
        """Return what autograd should see for this parameter."""
        if self.slot is not None:
            self.slot.end_write()
            return None
        return produced


def _weight_grad(dy2: torch.Tensor, x2: torch.Tensor, weight: torch.Tensor):
    """dW[N,K] = dY^T . X  (both operands MN-major: no transposes)."""
    dest = _GradDest(weight)
    out = gemm(dy2, x2, a_mn=True, b_mn=True, out=dest.tensor,
               flags=EPI_ACCUMULATE if dest.accumulate else 0)
    return dest.finish(out)


def rowsum_gemm_supported(m: int, n: int) -> bool:
    """Shapes for which the wgrad GEMM can carry the bias gradient (CTA-pair kernel, ROWSUM
    instantiation).  Off by default: measured on BERT-large it is 1.5 % SLOWER than the separate
    column-sum kernels (736.4 vs 747.9 samples/s) — the N=16 row-sum MMA still reads the whole A tile
    from shared memory (about 32 clk next to the main MMA's 64), only the first column of tiles pays
    it, and in these single-wave GEMMs the slowest tile sets the time.  TONY_B200_GEMM_ROWSUM=1
    enables it; spreading the row-sum k-blocks over all column tiles is the open follow-up."""
    return m >= 256 and n >= 256 and os.environ.get("TONY_B200_GEMM_ROWSUM", "0") == "1" \
        and os.environ.get("TONY_B200_GEMM_2CTA", "1") != "0"


def _weight_bias_grad(dy2: torch.Tensor, x2: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]):
    """dW = dY^T X and db = colsum(dY).  When the shape allows it db is produced INSIDE the wgrad
    GEMM (dY^T is its A operand: one extra N=16 MMA per k-step against a tile of ones), so dY is
    not read a second time by a column-sum kernel."""
    if bias is None or not rowsum_gemm_supported(weight.shape[0], weight.shape[1]):
        return _weight_grad(dy2, x2, weight), _bias_grad(dy2, bias)
    wdest, bdest = _GradDest(weight), _GradDest(bias)
    db = bdest.tensor if bdest.tensor is not None else torch.empty_like(bias)
    wflags = EPI_ACCUMULATE if wdest.accumulate else 0
    try:
        out = gemm(dy2, x2, a_mn=True, b_mn=True, out=wdest.tensor, flags=wflags, rowsum_out=db,
                   rowsum_accumulate=bdest.accumulate)
    except _lib.KernelError as e:
        if "code 3" not in str(e):          # TONY_ERR_UNSUPPORTED: this launch did not take the pair kernel
            raise
        out = gemm(dy2, x2, a_mn=True, b_mn=True, out=wdest.tensor, flags=wflags)
        colsum(dy2, out=db, accumulate=bdest.accumulate)
    return wdest.finish(out), bdest.finish(db)


_ACT_BWD_KERNEL = {"relu": 1, "gelu": 2, "gelu_tanh": 3}


def act_bwd_bias_grad(dy2: torch.Tensor, pre: torch.Tensor, act: str, bias: Optional[torch.Tensor]):
    """One pass over ``dy2`` and ``pre``: returns ``(dy2 * act'(pre), bias gradient for autograd)``; the bias
    gradient goes to the parameter's flat-bucket slot when it has one (``tony_act_bwd_colsum_bf16``)."""
    _check_bf16(dy2, pre)
    if not dy2.is_contiguous():
        dy2 = dy2.contiguous()
    rows, cols = dy2.shape
    if cols % 8 or pre.stride(0) != dy2.stride(0) or pre.stride(1) != 1:
        raise ValueError("act_bwd_bias_grad: needs a multiple of 8 columns and pre laid out like dy")
    dpre = torch.empty_like(dy2)
    dest = _GradDest(bias) if bias is not None else None
    db = None
    if dest is not None:
        db = dest.tensor if dest.tensor is not None else torch.empty(cols, dtype=torch.bfloat16, device=dy2.device)
    lib = _lib.lib()
    nbytes = int(lib.tony_colsum_scratch_bytes(cols))
    scratch = workspace("colsum", nbytes, dy2.device, counters=True)
    check(lib.tony_act_bwd_colsum_bf16(ptr(dy2), ptr(pre), ptr(dpre), _ACT_BWD_KERNEL[act], rows, cols, dy2.stride(0),
                                       ptr(db), ptr(scratch), nbytes, int(dest.accumulate) if dest else 0,
                                       cur_stream()), "tony_act_bwd_colsum_bf16")
    return dpre, (dest.finish(db) if dest is not None else None)


def _bias_grad(dy2: torch.Tensor, bias: Optional[torch.Tensor]):
    if bias is None:
        return None
    dest = _GradDest(bias)
    out = colsum(dy2, out=dest.tensor, accumulate=dest.accumulate)
    return dest.finish(out)


class _LinearFn(torch.autograd.Function):
    """y = x W^T (+ b) (+ residual); activation-free so dgrad/wgrad need nothing but dy."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual):
        K = weight.shape[1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        N = weight.shape[0]
        flags = (EPI_BIAS if bias is not None else 0) | (EPI_RESIDUAL if residual is not None else 0)
        res2 = residual.reshape(-1, N) if residual is not None else None
        y = gemm(x2, weight, flags=flags, bias=bias, aux=res2)
        ctx.save_for_backward(x2, weight)
        ctx.bias = bias
        ctx.x_shape = x.shape
        ctx.has_res = residual is not None
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        N = weight.shape[0]
        dy2 = dy.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, weight, b_mn=True).view(ctx.x_shape)  # dX = dY . W
        dw, db = _weight_bias_grad(dy2, x2, weight, ctx.bias)
        return dx, dw, db, (dy if ctx.has_res else None)


def linear(x, weight, bias=None, residual=None):
    """``x @ weight.T + bias (+ residual)`` with the adds fused into the GEMM epilogue."""
    return _LinearFn.apply(x, weight, bias, residual)


class _MlpFn(torch.autograd.Function):
    """Transformer FFN: y = act(x W1^T + b1) W2^T + b2 (+ residual), fully fused.

    forward : GEMM1 epilogue = bias + activation (+ saves the pre-activation),
              GEMM2 epilogue = bias + residual.
    backward: dgrad-2 epilogue multiplies by act'(pre) so the activation gradient never
              makes its own pass over memory.
    """

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual, act, self_residual=False):
        K = w1.shape[1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        F = w1.shape[0]
        N = w2.shape[0]
        pre = torch.empty((x2.shape[0], F), dtype=torch.bfloat16, device=x.device)
        h = gemm(x2, w1, flags=EPI_BIAS | _ACT_FWD[act] | EPI_SAVE_PREACT, bias=b1, aux_out=pre)
        if self_residual:        # y = x + FFN(x): the block input is its own residual
            res2 = x2
        else:
            res2 = residual.reshape(-1, N) if residual is not None else None
        y = gemm(h, w2, flags=EPI_BIAS | (EPI_RESIDUAL if res2 is not None else 0), bias=b2, aux=res2)
        ctx.save_for_backward(x2, w1, w2, pre, h)
        ctx.b1, ctx.b2, ctx.act = b1, b2, act
        ctx.x_shape = x.shape
        ctx.has_res = residual is not None
        ctx.self_residual = self_residual
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w1, w2, pre, h = ctx.saved_tensors
        N = w2.shape[0]
        dy2 = dy.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        # Order matters with overlap=True: finishing a parameter's slot may close its bucket and launch
        # the fused reduce+optimizer on the comm stream, which REWRITES the bf16 weights.  Every GEMM
        # that reads a weight is therefore enqueued before that weight's gradient slot is finished.
        dpre = gemm(dy2, w2, b_mn=True, flags=_ACT_BWD[ctx.act], aux=pre)  # (dY.W2) * act'(pre)
        dw2, db2 = _weight_bias_grad(dy2, h, w2, ctx.b2)
        dx = None
        if ctx.needs_input_grad[0]:
            if ctx.self_residual:
                # d(x + FFN(x)) = dY + dpre . W1: the residual branch's gradient rides in the last
                # dgrad GEMM's epilogue instead of a separate elementwise add over the activation
                dx = gemm(dpre, w1, b_mn=True, flags=EPI_RESIDUAL, aux=dy2).view(ctx.x_shape)
            else:
                dx = gemm(dpre, w1, b_mn=True).view(ctx.x_shape)
        dw1, db1 = _weight_bias_grad(dpre, x2, w1, ctx.b1)
        return dx, dw1, db1, dw2, db2, (dy if ctx.has_res else None), None, None


def mlp(x, w1, b1, w2, b2, residual=None, act="gelu"):
    if residual is x and x.shape[-1] == w2.shape[0]:
        return _MlpFn.apply(x, w1, b1, w2, b2, None, act, True)
    return _MlpFn.apply(x, w1, b1, w2, b2, residual, act, False)


class _LinearActFn(torch.autograd.Function):
    """Single Linear with a fused activation (heads, recommender towers)."""

    @staticmethod
    def forward(ctx, x, weight, bias, act):
        K = weight.shape[1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        N = weight.shape[0]
        pre = torch.empty((x2.shape[0], N), dtype=torch.bfloat16, device=x.device)
        y = gemm(x2, weight, flags=(EPI_BIAS if bias is not None else 0) | _ACT_FWD[act] | EPI_SAVE_PREACT,
                 bias=bias, aux_out=pre)
        ctx.save_for_backward(x2, weight, pre)
        ctx.bias, ctx.act, ctx.x_shape = bias, act, x.shape
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, pre = ctx.saved_tensors
        N = weight.shape[0]
        dy2 = dy.reshape(-1, N)
        # the incoming gradient is not produced by a GEMM of ours, so the activation gradient cannot ride in an
        # epilogue: one own pass computes it together with the bias gradient
        dpre, db = act_bwd_bias_grad(dy2, pre, ctx.act, ctx.bias)
        dx = gemm(dpre, weight, b_mn=True).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = _weight_grad(dpre, x2, weight)
        return dx, dw, db, None


def linear_act(x, weight, bias=None, act="relu"):
    if act in (None, "none"):
        return linear(x, weight, bias)
    return _LinearActFn.apply(x, weight, bias, act)


# ---------------------------------------------------------------------------------------
class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        _check_bf16(x2, gamma, beta)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        check(_lib.lib().tony_layernorm_fwd(ptr(x2), ptr(gamma), ptr(beta), ptr(y), ptr(mean),
                                            ptr(rstd), rows, H, float(eps), cur_stream()),
              "tony_layernorm_fwd")
        ctx.save_for_backward(x2, gamma, mean, rstd)
        ctx.beta = beta
        ctx.x_shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, gamma, mean, rstd = ctx.saved_tensors
        H = x2.shape[1]
        dy2 = dy.reshape(-1, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        rows = x2.shape[0]
        lib = _lib.lib()
        dx = torch.empty_like(x2)
        dg, db = _GradDest(gamma), _GradDest(ctx.beta)
        if dg.tensor is None:
            dg_t = torch.empty(H, dtype=torch.bfloat16, device=x2.device)
            db_t = torch.empty(H, dtype=torch.bfloat16, device=x2.device)
            acc = False
        else:
            dg_t, db_t = dg.tensor, db.tensor
            acc = dg.accumulate
        nbytes = int(lib.tony_layernorm_bwd_scratch_bytes(H))
        scratch = workspace("ln_bwd", nbytes, x2.device, counters=True)
        check(lib.tony_layernorm_bwd(ptr(dy2), ptr(x2), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx),
                                     ptr(dg_t), ptr(db_t), rows, H, ptr(scratch), nbytes, int(acc),
                                     cur_stream()), "tony_layernorm_bwd")
        return dx.view(ctx.x_shape), dg.finish(dg_t), db.finish(db_t), None


def layer_norm(x, gamma, beta, eps: float = 1e-5):
    return _LayerNormFn.apply(x, gamma, beta, eps)


class _LayerNormResidualFn(torch.autograd.Function):
    """Pre-LN block entry: returns ``(layer_norm(x), x)``.  The second output is the residual branch;
    because both outputs belong to one autograd node, backward receives the normalised branch's
    gradient AND the residual branch's gradient together and adds the latter inside the LayerNorm
    backward kernel (``tony_layernorm_bwd_add``) instead of in a separate elementwise pass."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        _check_bf16(x2, gamma, beta)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        check(_lib.lib().tony_layernorm_fwd(ptr(x2), ptr(gamma), ptr(beta), ptr(y), ptr(mean),
                                            ptr(rstd), rows, H, float(eps), cur_stream()),
              "tony_layernorm_fwd")
        ctx.save_for_backward(x2, gamma, mean, rstd)
        ctx.beta = beta
        ctx.x_shape = x.shape
        return y.view(x.shape), x.view_as(x)

    @staticmethod
    def backward(ctx, dy, dres):
        x2, gamma, mean, rstd = ctx.saved_tensors
        H = x2.shape[1]
        dy2 = dy.reshape(-1, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        rows = x2.shape[0]
        lib = _lib.lib()
        dx = torch.empty_like(x2)
        dg, db = _GradDest(gamma), _GradDest(ctx.beta)
        if dg.tensor is None:
            dg_t = torch.empty(H, dtype=torch.bfloat16, device=x2.device)
            db_t = torch.empty(H, dtype=torch.bfloat16, device=x2.device)
            acc = False
        else:
            dg_t, db_t = dg.tensor, db.tensor
            acc = dg.accumulate
        nbytes = int(lib.tony_layernorm_bwd_scratch_bytes(H))
        scratch = workspace("ln_bwd", nbytes, x2.device, counters=True)
        if dres is None:
            check(lib.tony_layernorm_bwd(ptr(dy2), ptr(x2), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx),
                                         ptr(dg_t), ptr(db_t), rows, H, ptr(scratch), nbytes, int(acc),
                                         cur_stream()), "tony_layernorm_bwd")
        else:
            dr2 = dres.reshape(-1, H)
            if not dr2.is_contiguous():
                dr2 = dr2.contiguous()
            check(lib.tony_layernorm_bwd_add(ptr(dy2), ptr(x2), ptr(gamma), ptr(mean), ptr(rstd), ptr(dr2),
                                             ptr(dx), ptr(dg_t), ptr(db_t), rows, H, ptr(scratch), nbytes,
                                             int(acc), cur_stream()), "tony_layernorm_bwd_add")
        return dx.view(ctx.x_shape), dg.finish(dg_t), db.finish(db_t), None


def layer_norm_residual(x, gamma, beta, eps: float = 1e-5):
    """``(layer_norm(x), x)`` for pre-LN blocks: use the second value as the residual input."""
    return _LayerNormResidualFn.apply(x, gamma, beta, eps)


# ---------------------------------------------------------------------------------------
class _SoftmaxXentFn(torch.autograd.Function):
    """Fused softmax cross-entropy.  The logits buffer is consumed: forward overwrites it
    with the gradient, backward just hands it back — so the loss must be the root of the
    backward pass (any constant factor goes in ``loss_scale``)."""

    @staticmethod
    def forward(ctx, logits, targets, vocab, ignore_index, reduction, loss_scale):
        _check_bf16(logits)
        ld = logits.shape[-1]
        lg2 = logits.view(-1, ld)
        rows = lg2.shape[0]
        tg = targets.reshape(-1)
        if tg.dtype != torch.int64:
            tg = tg.long()
        loss_rows = torch.empty(rows, dtype=torch.float32, device=logits.device)
        if reduction == "mean":
            inv = (tg != ignore_index).sum().clamp_(min=1).to(torch.float32).reciprocal_()
        else:
            inv = None
        check(_lib.lib().tony_softmax_xent_fwd_bwd(ptr(lg2), ptr(tg), ptr(loss_rows), rows, vocab,
                                                   lg2.stride(0), ignore_index, float(loss_scale),
                                                   ptr(inv), cur_stream()),
              "tony_softmax_xent_fwd_bwd")
        ctx.save_for_backward(logits)
        loss = loss_rows.sum()
        if inv is not None:
            loss = loss * inv
        return loss * loss_scale if loss_scale != 1.0 else loss

    @staticmethod
    def backward(ctx, grad_out):
        (dlogits,) = ctx.saved_tensors
        return dlogits, None, None, None, None, None


def softmax_cross_entropy(logits, targets, *, vocab: Optional[int] = None, ignore_index: int = -100,
                          reduction: str = "mean", loss_scale: float = 1.0):
    """Cross-entropy over the last dim of ``logits`` ([..., V_padded] bf16, destroyed).

    ``vocab`` is the true class count when the last dim carries alignment padding.
    """
    if reduction not in ("mean", "sum"):
        raise ValueError("reduction must be 'mean' or 'sum'")
    return _SoftmaxXentFn.apply(logits, targets, vocab or logits.shape[-1], ignore_index, reduction,
                                loss_scale)


# ---------------------------------------------------------------------------------------
class _EmbeddingFn(torch.autograd.Function):
    """Row gather; the backward scatter-adds into the gradient bucket slot when present."""

    @staticmethod
    def forward(ctx, weight, idx):
        ctx.save_for_backward(idx)
        ctx.weight = weight
        return weight.index_select(0, idx.reshape(-1)).view(*idx.shape, weight.shape[1])

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        weight = ctx.weight
        dest = _GradDest(weight)
        dy2 = dy.reshape(-1, weight.shape[1])
        if dest.tensor is not None:
            if not dest.accumulate:
                dest.tensor.zero_()
            dest.tensor.index_add_(0, idx.reshape(-1), dy2)
            return dest.finish(None), None
        g = torch.zeros_like(weight)
        g.index_add_(0, idx.reshape(-1), dy2)
        return g, None


def embedding(weight, idx):
    return _EmbeddingFn.apply(weight, idx)


class _EmbeddingSumFn(torch.autograd.Function):
    """``word[ids] (+ type[tt]) (+ pos[0:S])`` in one kernel; backward makes all three gradients in
    one pass over dy (csrc/embedding.cu) and writes them straight into the gradient bucket slots."""

    @staticmethod
    def forward(ctx, ids, tt, word, type_w, pos_w):
        B, S = ids.shape
        H = word.shape[1]
        ids = ids.contiguous()
        tt = tt.contiguous() if tt is not None else None
        out = torch.empty((B, S, H), dtype=word.dtype, device=word.device)
        check(_lib.lib().tony_embedding_sum_fwd(ptr(word), ptr(ids), ptr(type_w) if type_w is not None else None,
                                                ptr(tt) if tt is not None else None,
                                                ptr(pos_w) if pos_w is not None else None, ptr(out), B * S, S, H,
                                                cur_stream()), "tony_embedding_sum_fwd")
        ctx.save_for_backward(ids, tt) if tt is not None else ctx.save_for_backward(ids)
        ctx.params = (word, type_w, pos_w)
        ctx.dims = (B, S, H)
        return out

    @staticmethod
    def backward(ctx, dy):
        saved = ctx.saved_tensors
        ids, tt = saved[0], (saved[1] if len(saved) > 1 else None)
        word, type_w, pos_w = ctx.params
        B, S, H = ctx.dims
        dy = dy.contiguous()
        dests, bufs, accs = [], [], []
        for prm in (word, type_w, pos_w):
            if prm is None:
                dests.append(None), bufs.append(None), accs.append(0)
                continue
            d = _GradDest(prm)
            buf = d.tensor if d.tensor is not None else torch.empty_like(prm)
            dests.append(d), bufs.append(buf), accs.append(1 if d.accumulate else 0)
        if pos_w is not None and not accs[2] and pos_w.shape[0] > S:
            bufs[2][S:].zero_()          # positions past this batch's length get no gradient
        n_types = type_w.shape[0] if type_w is not None else 0
        nbytes = _lib.lib().tony_embedding_bwd_scratch_bytes(S, n_types, H) if type_w is not None else 0
        scratch = workspace("embedding_bwd", nbytes, dy.device) if nbytes else None
        check(_lib.lib().tony_embedding_sum_bwd(
            ptr(dy), ptr(ids), ptr(tt) if tt is not None else None, ptr(bufs[0]), word.shape[0], accs[0],
            ptr(bufs[1]) if type_w is not None else None, n_types, accs[1],
            ptr(bufs[2]) if pos_w is not None else None, accs[2],
            ptr(scratch) if scratch is not None else None, nbytes, B, S, H, cur_stream()),
            "tony_embedding_sum_bwd")
        grads = [d.finish(b) if d is not None else None for d, b in zip(dests, bufs)]
        return None, None, grads[0], grads[1], grads[2]


def embedding_sum(ids, tt, word, type_w=None, pos_w=None):
    """Fused input embedding for ``ids`` ``[B, S]`` (int64): ``word[ids] + type_w[tt] + pos_w[:S]``."""
    return _EmbeddingSumFn.apply(ids, tt, word, type_w, pos_w)


def embedding_sum_supported(word, type_w=None) -> bool:
    return (word.is_cuda and word.dtype == torch.bfloat16 and word.shape[1] % 8 == 0
            and (type_w is None or type_w.shape[0] <= 4))


class _BatchNormActFn(torch.autograd.Function):
    """Training-mode BatchNorm2d (+ residual add) (+ ReLU) on a channels-last bf16 activation:
    two launches forward, two backward (csrc/batchnorm.cu)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, eps, momentum, relu, stats=None):
        N, C, H, W = x.shape
        x = x.contiguous(memory_format=torch.channels_last)
        res = residual.contiguous(memory_format=torch.channels_last) if residual is not None else None
        M = N * H * W
        y = torch.empty_like(x)
        mean = torch.empty(C, dtype=torch.float32, device=x.device)
        rstd = torch.empty(C, dtype=torch.float32, device=x.device)
        # one bit per element: backward applies the ReLU mask without re-reading y or the residual
        mask = torch.empty(M * C // 8, dtype=torch.uint8, device=x.device) if relu else None
        if stats is not None:
            # the producing convolution's epilogue already summed x and x^2 per channel: one launch
            check(_lib.lib().tony_bn_fwd_from_sums(ptr(x), ptr(res), ptr(gamma), ptr(beta), ptr(y), ptr(mask),
                                                   ptr(mean), ptr(rstd), ptr(running_mean), ptr(running_var), M, C,
                                                   eps, momentum, int(relu), ptr(stats[0]), ptr(stats[1]),
                                                   cur_stream()), "tony_bn_fwd_from_sums")
        else:
            nbytes = _lib.lib().tony_bn_scratch_bytes(C)
            scratch = workspace("batchnorm", nbytes, x.device, counters=True)
            check(_lib.lib().tony_bn_fwd(ptr(x), ptr(res), ptr(gamma), ptr(beta), ptr(y), ptr(mask), ptr(mean),
                                         ptr(rstd), ptr(running_mean), ptr(running_var), M, C, eps, momentum,
                                         int(relu), ptr(scratch), nbytes, cur_stream()), "tony_bn_fwd")
        ctx.save_for_backward(x, gamma, mean, rstd, mask)
        ctx.meta = (M, C, int(relu), residual is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, mean, rstd, mask = ctx.saved_tensors
        M, C, relu, has_res = ctx.meta
        dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if has_res else None
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        nbytes = _lib.lib().tony_bn_scratch_bytes(C)
        scratch = workspace("batchnorm", nbytes, x.device, counters=True)
        check(_lib.lib().tony_bn_bwd(ptr(dy), ptr(x), ptr(mask), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx),
                                     ptr(dres), ptr(dgamma), ptr(dbeta), M, C, relu, ptr(scratch), nbytes,
                                     cur_stream()), "tony_bn_bwd")
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None


def batch_norm_act(x, gamma, beta, running_mean=None, running_var=None, residual=None, eps: float = 1e-5,
                   momentum: float = 0.1, relu: bool = True, stats: Optional[torch.Tensor] = None):
    """``relu(batch_norm(x) + residual)`` with batch statistics, for ``x`` ``[N, C, H, W]`` bf16 in
    channels-last memory format and ``C`` a power of two in [8, 2048].  ``stats`` (fp32 ``[2, C]``): the
    per-channel sum / sum of squares of ``x`` when the producing convolution already accumulated them."""
    return _BatchNormActFn.apply(x, gamma, beta, residual, running_mean, running_var, eps, momentum, relu, stats)


def batch_norm_act_supported(x) -> bool:
    C = x.shape[1]
    return x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and 8 <= C <= 2048 and (C & (C - 1)) == 0


class _MaxPool3x3s2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        N, C, H, W = x.shape
        x = x.contiguous(memory_format=torch.channels_last)
        OH, OW = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        y = torch.empty((N, C, OH, OW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((N, OH, OW, C), dtype=torch.uint8, device=x.device)
        check(_lib.lib().tony_maxpool3x3s2_fwd(ptr(x), ptr(y), ptr(idx), N, H, W, C, cur_stream()),
              "tony_maxpool3x3s2_fwd")
        ctx.save_for_backward(idx)
        ctx.shape = (N, C, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        N, C, H, W = ctx.shape
        dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, C, H, W), dtype=dy.dtype, device=dy.device, memory_format=torch.channels_last)
        check(_lib.lib().tony_maxpool3x3s2_bwd(ptr(dy), ptr(idx), ptr(dx), N, H, W, C, cur_stream()),
              "tony_maxpool3x3s2_bwd")
        return dx


def max_pool_3x3_s2(x):
    """``max_pool2d(x, 3, stride=2, padding=1)`` for channels-last bf16 ``[N, C, H, W]`` with C % 8 == 0."""
    if x.is_cuda and x.dtype == torch.bfloat16 and x.shape[1] % 8 == 0:
        return _MaxPool3x3s2Fn.apply(x)
    return torch.nn.functional.max_pool2d(x, 3, 2, 1)


# ---------------------------------------------------------------------------------------
# convolution: implicit GEMM on tcgen05, TMA im2col mode (csrc/conv_tcgen05.cu)
# ---------------------------------------------------------------------------------------
def _conv_desc(N, H, W, C, K, R, S, stride, pad, pad_h_only=False, strides=(0, 0, 0), pad_hi=-1):
    return _lib.ConvDesc(N, H, W, C, K, R, S, stride, pad, pad_hi, int(pad_h_only), *strides)


def _conv_out(n, f, stride, pad, pad_hi=-1):
    return (n + pad + (pad if pad_hi < 0 else pad_hi) - f) // stride + 1


def conv2d_nhwc(x_nhwc: torch.Tensor, w_krsc: torch.Tensor, stride: int = 1, pad: int = 0, *,
                stats: Optional[torch.Tensor] = None, logical=None, pad_h_only: bool = False,
                pad_hi: int = -1) -> torch.Tensor:
    """Raw forward: ``x`` ``[N, H, W, C]``, ``w`` ``[K, R, S, C]`` -> ``[N, P, Q, K]`` (all bf16, contiguous).
    ``stats`` (fp32 ``[2, K]``, zeroed by the caller) receives the output's per-channel sum / sum of squares.
    ``logical=(N, H, W, C, stride_w, stride_h, stride_n)`` overrides the activation's geometry (windowed
    views whose pixel stride is smaller than C, see ``StemConv``)."""
    _check_bf16(x_nhwc, w_krsc)
    K, R, S, C = w_krsc.shape
    if logical is None:
        N, H, W, Cx = x_nhwc.shape
        strides = (0, 0, 0)
    else:
        N, H, W, Cx = logical[:4]
        strides = tuple(logical[4:])
    if Cx != C:
        raise ValueError(f"channel mismatch: activation {Cx}, filter {C}")
    d = _conv_desc(N, H, W, C, K, R, S, stride, pad, pad_h_only, strides, pad_hi)
    lib = _lib.lib()
    P = _conv_out(H, R, stride, pad, pad_hi)
    Q = _conv_out(W, S, stride, 0, 0) if pad_h_only else _conv_out(W, S, stride, pad, pad_hi)
    y = torch.empty((N, P, Q, K), dtype=torch.bfloat16, device=x_nhwc.device)
    s1 = ptr(stats[0]) if stats is not None else None
    s2 = ptr(stats[1]) if stats is not None else None
    import ctypes as _ct
    check(lib.tony_conv2d_fprop_bf16(_ct.byref(d), ptr(x_nhwc), ptr(w_krsc), ptr(y), s1, s2, cur_stream()),
          "tony_conv2d_fprop_bf16")
    return y


def conv2d_dgrad_nhwc(dy_nhwc, w_krsc, x_shape, stride, pad, out=None, accumulate=False):
    """``dx`` ``[N, H, W, C]`` from ``dy`` ``[N, P, Q, K]``; ``accumulate`` adds into ``out``."""
    import ctypes as _ct

    K, R, S, C = w_krsc.shape
    N, H, W, _ = x_shape
    d = _conv_desc(N, H, W, C, K, R, S, stride, pad)
    if out is None:
        out = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy_nhwc.device)
        accumulate = False
    check(_lib.lib().tony_conv2d_dgrad_bf16(_ct.byref(d), ptr(dy_nhwc), ptr(w_krsc), ptr(out),
                                            EPI_ACCUMULATE if accumulate else 0, cur_stream()),
          "tony_conv2d_dgrad_bf16")
    return out


def conv2d_wgrad_nhwc(dy_nhwc, x_nhwc, w_shape, stride, pad, out=None, accumulate=False, logical=None,
                      pad_h_only=False, pad_hi=-1):
    """``dw`` ``[K, R, S, C]`` from ``dy`` ``[N, P, Q, K]`` and ``x`` ``[N, H, W, C]`` (split-K over the pixels)."""
    import ctypes as _ct

    K, R, S, C = w_shape
    if logical is None:
        N, H, W, _ = x_nhwc.shape
        strides = (0, 0, 0)
    else:
        N, H, W, _ = logical[:4]
        strides = tuple(logical[4:])
    d = _conv_desc(N, H, W, C, K, R, S, stride, pad, pad_h_only, strides, pad_hi)
    lib = _lib.lib()
    if out is None:
        out = torch.empty(tuple(w_shape), dtype=torch.bfloat16, device=dy_nhwc.device)
        accumulate = False
    nbytes = int(lib.tony_conv2d_wgrad_workspace_bytes(_ct.byref(d)))
    ws = workspace("conv_wgrad", nbytes, dy_nhwc.device)
    check(lib.tony_conv2d_wgrad_bf16(_ct.byref(d), ptr(dy_nhwc), ptr(x_nhwc), ptr(out),
                                     EPI_ACCUMULATE if accumulate else 0, ptr(ws), nbytes, cur_stream()),
          "tony_conv2d_wgrad_bf16")
    return out


class _Conv2dFn(torch.autograd.Function):
    """Convolution of a channels-last ``[N, C, H, W]`` activation with a ``[K, R, S, C]`` filter.

    forward / dgrad / wgrad all run on the implicit-GEMM tcgen05 kernel; stride-1 pointwise layers
    are a plain GEMM over the activation viewed as ``[N*H*W, C]`` (CTA-pair kernel, split-K wgrad).
    The weight gradient goes straight into the parameter's slot of the flat gradient bucket."""

    @staticmethod
    def forward(ctx, x, weight, stride, pad, stats):
        x = x.contiguous(memory_format=torch.channels_last)
        xn = x.permute(0, 2, 3, 1)                      # [N, H, W, C] view, contiguous
        K, R, S, C = weight.shape
        if R == 1 and S == 1 and stride == 1 and pad == 0 and stats is None:
            N, H, W, _ = xn.shape
            yn = gemm(xn.reshape(N * H * W, C), weight.view(K, C)).view(N, H, W, K)
        else:
            yn = conv2d_nhwc(xn, weight, stride, pad, stats=stats)
        ctx.save_for_backward(xn, weight)
        ctx.geom = (stride, pad)
        return yn.permute(0, 3, 1, 2)                   # [N, K, P, Q], channels-last memory

    @staticmethod
    def backward(ctx, dy):
        xn, weight = ctx.saved_tensors
        stride, pad = ctx.geom
        K, R, S, C = weight.shape
        dyn = dy.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)
        N, H, W, _ = xn.shape
        pointwise = R == 1 and S == 1 and stride == 1 and pad == 0
        dx = None
        if ctx.needs_input_grad[0]:                      # before the weight's slot is finished (see _MlpFn)
            if pointwise:
                dxn = gemm(dyn.reshape(N * H * W, K), weight.view(K, C), b_mn=True).view(N, H, W, C)
            else:
                dxn = conv2d_dgrad_nhwc(dyn, weight, xn.shape, stride, pad)
            dx = dxn.permute(0, 3, 1, 2)
        dest = _GradDest(weight)
        if pointwise:
            out = gemm(dyn.reshape(N * H * W, K), xn.reshape(N * H * W, C), a_mn=True, b_mn=True,
                       out=dest.tensor.view(K, C) if dest.tensor is not None else None,
                       flags=EPI_ACCUMULATE if dest.accumulate else 0).view(K, R, S, C)
        else:
            out = conv2d_wgrad_nhwc(dyn, xn, weight.shape, stride, pad, out=dest.tensor, accumulate=dest.accumulate)
        return dx, dest.finish(out), None, None, None


class _PointwiseConvForkFn(torch.autograd.Function):
    """``(conv1x1(x, w), x)`` as ONE autograd node: the second output is the residual branch of a
    bottleneck block.  Backward then receives the main path's gradient and the residual branch's gradient
    together and adds the latter in the epilogue of the data-gradient GEMM (``dx = dy . W + d_residual``)
    instead of in a separate elementwise kernel over the activation."""

    @staticmethod
    def forward(ctx, x, weight, stats):
        x = x.contiguous(memory_format=torch.channels_last)
        xn = x.permute(0, 2, 3, 1)
        K, R, S, C = weight.shape
        N, H, W, _ = xn.shape
        if stats is None:
            yn = gemm(xn.reshape(N * H * W, C), weight.view(K, C)).view(N, H, W, K)
        else:
            yn = conv2d_nhwc(xn, weight, 1, 0, stats=stats)
        ctx.save_for_backward(xn, weight)
        return yn.permute(0, 3, 1, 2), x.view_as(x)

    @staticmethod
    def backward(ctx, dy, dres):
        xn, weight = ctx.saved_tensors
        K, R, S, C = weight.shape
        N, H, W, _ = xn.shape
        M = N * H * W
        dyn = dy.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1).reshape(M, K)
        dx = None
        if ctx.needs_input_grad[0]:
            if dres is not None:
                dr = dres.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1).reshape(M, C)
                dxn = gemm(dyn, weight.view(K, C), b_mn=True, flags=EPI_RESIDUAL, aux=dr)
            else:
                dxn = gemm(dyn, weight.view(K, C), b_mn=True)
            dx = dxn.view(N, H, W, C).permute(0, 3, 1, 2)
        dest = _GradDest(weight)
        out = gemm(dyn, xn.reshape(M, C), a_mn=True, b_mn=True,
                   out=dest.tensor.view(K, C) if dest.tensor is not None else None,
                   flags=EPI_ACCUMULATE if dest.accumulate else 0).view(K, R, S, C)
        return dx, dest.finish(out), None


def conv1x1_fork(x, weight, stats: Optional[torch.Tensor] = None):
    """``(conv2d(x, weight), x)`` for a stride-1 pointwise ``[K, 1, 1, C]`` filter; use the second value
    as the block's residual input (see ``_PointwiseConvForkFn``)."""
    return _PointwiseConvForkFn.apply(x, weight, stats)


class _PointwiseConvPairFn(torch.autograd.Function):
    """``(conv1x1(x, w1), conv1x1(x, wd, stride))`` as ONE autograd node: the first block of a ResNet stage feeds
    its input to the main path and to the (possibly strided) projection shortcut.  Backward computes the main
    path's data gradient and lets the shortcut's data-gradient kernel ACCUMULATE into it, so the sum of the two
    branch gradients never runs as a separate elementwise kernel over the activation."""

    @staticmethod
    def forward(ctx, x, w1, wd, stride_d, stats1, statsd):
        x = x.contiguous(memory_format=torch.channels_last)
        xn = x.permute(0, 2, 3, 1)
        N, H, W, C = xn.shape
        K1, Kd = w1.shape[0], wd.shape[0]
        if stats1 is None:
            y1 = gemm(xn.reshape(N * H * W, C), w1.view(K1, C)).view(N, H, W, K1)
        else:
            y1 = conv2d_nhwc(xn, w1, 1, 0, stats=stats1)
        if stride_d == 1 and statsd is None:
            yd = gemm(xn.reshape(N * H * W, C), wd.view(Kd, C)).view(N, H, W, Kd)
        else:
            yd = conv2d_nhwc(xn, wd, stride_d, 0, stats=statsd)
        ctx.save_for_backward(xn, w1, wd)
        ctx.stride_d = stride_d
        return y1.permute(0, 3, 1, 2), yd.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy1, dyd):
        xn, w1, wd = ctx.saved_tensors
        sd = ctx.stride_d
        N, H, W, C = xn.shape
        M = N * H * W
        K1, Kd = w1.shape[0], wd.shape[0]
        d1 = dy1.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)
        dd = dyd.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)
        dx = None
        if ctx.needs_input_grad[0]:                      # both data gradients before either weight slot closes
            dxn = gemm(d1.reshape(M, K1), w1.view(K1, C), b_mn=True).view(N, H, W, C)
            if sd == 1:
                gemm(dd.reshape(M, Kd), wd.view(Kd, C), b_mn=True, out=dxn.view(M, C), flags=EPI_ACCUMULATE)
            else:
                conv2d_dgrad_nhwc(dd, wd, xn.shape, sd, 0, out=dxn, accumulate=True)
            dx = dxn.permute(0, 3, 1, 2)
        dest = _GradDest(w1)
        g1 = gemm(d1.reshape(M, K1), xn.reshape(M, C), a_mn=True, b_mn=True,
                  out=dest.tensor.view(K1, C) if dest.tensor is not None else None,
                  flags=EPI_ACCUMULATE if dest.accumulate else 0).view(K1, 1, 1, C)
        g1 = dest.finish(g1)
        dest = _GradDest(wd)
        if sd == 1:
            gd = gemm(dd.reshape(M, Kd), xn.reshape(M, C), a_mn=True, b_mn=True,
                      out=dest.tensor.view(Kd, C) if dest.tensor is not None else None,
                      flags=EPI_ACCUMULATE if dest.accumulate else 0).view(Kd, 1, 1, C)
        else:
            gd = conv2d_wgrad_nhwc(dd, xn, wd.shape, sd, 0, out=dest.tensor, accumulate=dest.accumulate)
        return dx, g1, dest.finish(gd), None, None, None


def conv1x1_pair(x, w1, wd, stride_d: int = 1, stats1: Optional[torch.Tensor] = None,
                 statsd: Optional[torch.Tensor] = None):
    """``(conv2d(x, w1), conv2d(x, wd, stride_d))`` for two pointwise filters (see ``_PointwiseConvPairFn``)."""
    return _PointwiseConvPairFn.apply(x, w1, wd, stride_d, stats1, statsd)


def conv2d(x, weight, stride: int = 1, pad: int = 0, stats: Optional[torch.Tensor] = None):
    """``conv2d`` for a bf16 channels-last ``x`` ``[N, C, H, W]`` and a ``[K, R, S, C]`` (KRSC) filter."""
    return _Conv2dFn.apply(x, weight, stride, pad, stats)


def conv2d_supported(x, weight, stride) -> bool:
    K, R, S, C = weight.shape
    return (x.is_cuda and x.dtype == torch.bfloat16 and C % 64 == 0 and K % 8 == 0 and R * S <= 16
            and stride in (1, 2))


class _StemConvFn(torch.autograd.Function):
    """7x7 / stride 2 / pad 3 convolution of a 3-channel NCHW image batch, as a 4x4 stride-1 convolution
    on the 2x2 space-to-depth image whose 4 horizontal taps are read as one 64-element pixel (see
    ``tony_stem_space_to_depth``).  ``w2`` is the filter in that layout: ``[K, 4, 1, 64]``."""

    @staticmethod
    def forward(ctx, images, w2, stats):
        _check_bf16(images, w2)
        N, C3, H, W = images.shape
        if C3 != 3 or H % 2 or W % 2:
            raise ValueError("stem expects [N, 3, H, W] with even H, W")
        images = images.contiguous()
        H2, W2 = H // 2, W // 2
        if os.environ.get("TONY_B200_STEM_WINDOW", "dense") == "overlap":
            # 4x fewer bytes, but 128-pixel im2col loads through the overlapping map measured 16x slower
            xs = torch.empty((N, H2, W2 + 3, 16), dtype=torch.bfloat16, device=images.device)
            check(_lib.lib().tony_stem_space_to_depth(ptr(images), ptr(xs), N, H, W, cur_stream()),
                  "tony_stem_space_to_depth")
            logical = (N, H2, W2, 64, 16, (W2 + 3) * 16, H2 * (W2 + 3) * 16)
        else:
            xs = torch.empty((N, H2, W2, 64), dtype=torch.bfloat16, device=images.device)
            check(_lib.lib().tony_stem_space_to_depth_window(ptr(images), ptr(xs), N, H, W, cur_stream()),
                  "tony_stem_space_to_depth_window")
            logical = (N, H2, W2, 64, 0, 0, 0)
        yn = conv2d_nhwc(xs, w2, 1, 2, stats=stats, logical=logical, pad_h_only=True, pad_hi=1)
        ctx.save_for_backward(xs)
        ctx.logical, ctx.wshape = logical, tuple(w2.shape)
        return yn.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        (xs,) = ctx.saved_tensors
        dyn = dy.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1)
        dw2 = conv2d_wgrad_nhwc(dyn, xs, ctx.wshape, 1, 2, logical=ctx.logical, pad_h_only=True, pad_hi=1)
        return None, dw2, None


def stem_weight_index(device=None) -> torch.Tensor:
    """Gather index that lays a ``[K, 7, 7, 3]`` (KRSC) stem filter, flattened to ``[K, 147]`` with one
    zero column appended (index 147), out as the ``[K, 4*64]`` filter of the space-to-depth form:
    position ``a*64 + b*16 + (dy*2+dx)*4 + c`` holds tap ``(r, s) = (2a+dy-1, 2b+dx-1)``."""
    idx = torch.full((4, 4, 2, 2, 4), 147, dtype=torch.long)
    for a in range(4):
        for b in range(4):
            for dy in range(2):
                for dx in range(2):
                    r, s_ = 2 * a + dy - 1, 2 * b + dx - 1
                    if 0 <= r < 7 and 0 <= s_ < 7:
                        for c in range(3):
                            idx[a, b, dy, dx, c] = (r * 7 + s_) * 3 + c
    return idx.reshape(-1).to(device) if device is not None else idx.reshape(-1)


def stem_conv7x7s2(images, weight_krsc, stats: Optional[torch.Tensor] = None, index: Optional[torch.Tensor] = None):
    """``conv2d(images, W, stride=2, padding=3)`` for ``images`` ``[N, 3, H, W]`` bf16 (NCHW) and
    ``weight_krsc`` ``[K, 7, 7, 3]``; returns channels-last ``[N, K, H/2, W/2]``."""
    K = weight_krsc.shape[0]
    if index is None:
        index = stem_weight_index(weight_krsc.device)
    flat = torch.cat([weight_krsc.reshape(K, 147), weight_krsc.new_zeros(K, 1)], dim=1)
    w2 = flat.index_select(1, index).view(K, 4, 1, 64)
    return _StemConvFn.apply(images, w2, stats)


class _BroadcastAddParamFn(torch.autograd.Function):
    """x + p where p broadcasts over the leading dims (position embeddings)."""

    @staticmethod
    def forward(ctx, x, p):
        ctx.p = p
        return x + p

    @staticmethod
    def backward(ctx, dy):
        p = ctx.p
        dest = _GradDest(p)
        lead = dy.dim() - p.dim()
        g = dy.sum(dim=tuple(range(lead))) if lead > 0 else dy
        if dest.tensor is not None:
            if dest.accumulate:
                dest.tensor.add_(g)
            else:
                dest.tensor.copy_(g)
            return dy, dest.finish(None)
        return dy, g


def add_broadcast_param(x, p):
    return _BroadcastAddParamFn.apply(x, p)


# ---------------------------------------------------------------------------------------
class _PackedAttentionFn(torch.autograd.Function):
    """Attention over a packed ``qkv`` activation ``[B, S, 3, heads, head_dim]`` -> ``[B, S, H]``.

    The core is chosen by ``_attention_core_forward``: the library flash kernel behind
    ``scaled_dot_product_attention`` by default, or the tcgen05 kernels of csrc/attention.cu.  In
    both cases q/k/v are read as strided views of the packed projection output, the context comes
    back already in ``[B, S, H]`` order and backward produces ONE packed gradient (the library path
    interleaves dq/dk/dv with ``tony_pack3_rows`` instead of three zero-filled ``select_backward``
    tensors + adds)."""

    @staticmethod
    def forward(ctx, qkv, causal):
        out, saved = _attention_core_forward(qkv.detach().contiguous(), causal)
        ctx.saved_core = saved
        return out

    @staticmethod
    def backward(ctx, dout):
        saved, ctx.saved_core = ctx.saved_core, None
        return _attention_core_backward(saved, dout)[0], None


def fused_attention_supported(qkv) -> bool:
    return (qkv.is_cuda and qkv.dtype == torch.bfloat16 and qkv.dim() == 5 and qkv.shape[4] == 64
            and qkv.shape[1] % 128 == 0)


def _attention_core_forward(qkv, causal):
    """-> (context [B, S, H], saved state for ``_attention_core_backward``)."""
    if fused_attention_supported(qkv) and os.environ.get("TONY_B200_ATTENTION", "library") == "tcgen05":
        B, S, _, nh, hd = qkv.shape
        out = torch.empty((B, S, nh * hd), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((B * nh, S), dtype=torch.float32, device=qkv.device)
        scale = 1.0 / math.sqrt(hd)
        check(_lib.lib().tony_attention_fwd_bf16(ptr(qkv), ptr(out), ptr(lse), B, S, nh, hd, int(causal),
                                                 scale, cur_stream()), "tony_attention_fwd_bf16")
        return out, ("tcgen05", qkv, out, lse, (B, S, nh, hd, int(causal), scale))
    B, S, _, nh, hd = qkv.shape
    with torch.enable_grad():
        q, k, v = (qkv[:, :, i].transpose(1, 2).detach().requires_grad_() for i in range(3))
        o = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=causal)
    return o.detach().transpose(1, 2).reshape(B, S, nh * hd), ("library", q, k, v, o, (B, S, nh, hd))


_NOT_DONE = object()     # _attention_core_backward: the caller still owes the bias gradient


def _attention_core_backward(saved, dout, bias=None):
    """-> (packed gradient ``[B, S, 3, heads, head_dim]``, bias-gradient result or ``_NOT_DONE``).  With ``bias``
    (the fused QKV projection's bias) the library path packs dq / dk / dv AND sums their columns in one kernel."""
    if saved[0] == "tcgen05":
        _, qkv, out, lse, (B, S, nh, hd, causal, scale) = saved
        dout = dout.contiguous()
        delta = torch.empty_like(lse)
        dqkv = torch.empty_like(qkv)
        check(_lib.lib().tony_attention_bwd_bf16(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(delta),
                                                 ptr(dqkv), B, S, nh, hd, causal, scale, cur_stream()),
              "tony_attention_bwd_bf16")
        return dqkv, _NOT_DONE
    _, q, k, v, o, (B, S, nh, hd) = saved
    do = dout.reshape(B, S, nh, hd).transpose(1, 2)
    dq, dk, dv = torch.autograd.grad(o, (q, k, v), do)
    dqkv = torch.empty((B, S, 3, nh, hd), dtype=dout.dtype, device=dout.device)
    parts = [t.transpose(1, 2) for t in (dq, dk, dv)]          # [B, S, nh, hd] views
    H = nh * hd

    def rows_view(t):
        ok = t.stride(3) == 1 and t.stride(2) == hd and t.stride(0) == S * t.stride(1)
        return t.stride(1) if ok and t.stride(1) % 8 == 0 and t.data_ptr() % 16 == 0 else None

    strides = [rows_view(t) for t in parts]
    if dout.dtype == torch.bfloat16 and H % 8 == 0 and all(st is not None for st in strides):
        if bias is not None and H % 256 == 0:
            dest = _GradDest(bias)
            db = dest.tensor if dest.tensor is not None else torch.empty(3 * H, dtype=torch.bfloat16, device=dout.device)
            nbytes = int(_lib.lib().tony_colsum_scratch_bytes(3 * H))
            scratch = workspace("colsum", nbytes, dout.device, counters=True)
            check(_lib.lib().tony_pack3_rows_colsum(ptr(parts[0]), ptr(parts[1]), ptr(parts[2]), strides[0], strides[1],
                                                    strides[2], ptr(dqkv), B * S, H, ptr(db), ptr(scratch), nbytes,
                                                    int(dest.accumulate), cur_stream()), "tony_pack3_rows_colsum")
            return dqkv, dest.finish(db)
        check(_lib.lib().tony_pack3_rows(ptr(parts[0]), ptr(parts[1]), ptr(parts[2]), strides[0],
                                         strides[1], strides[2], ptr(dqkv), B * S, H, cur_stream()),
              "tony_pack3_rows")
    else:
        for i, t in enumerate(parts):
            dqkv[:, :, i].copy_(t)
    return dqkv, _NOT_DONE


class _AttentionBlockFn(torch.autograd.Function):
    """y = x + Wo . attention(Wqkv x + bqkv) + bo as ONE autograd node (post-LN transformer block).

    Besides the epilogue fusions of the two projections, owning the whole block lets backward
    add the residual branch's gradient in the epilogue of the LAST dgrad GEMM
    (dx = dqkv . Wqkv + dy), so autograd never runs an elementwise add over the activation to
    merge the two branches."""

    @staticmethod
    def forward(ctx, x, wqkv, bqkv, wo, bo, heads, causal):
        B, S, H = x.shape
        x2 = x.reshape(B * S, H)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        qkv = gemm(x2, wqkv, flags=EPI_BIAS if bqkv is not None else 0, bias=bqkv)
        o, saved = _attention_core_forward(qkv.view(B, S, 3, heads, H // heads), causal)
        o2 = o.reshape(B * S, H)
        y = gemm(o2, wo, flags=(EPI_BIAS if bo is not None else 0) | EPI_RESIDUAL, bias=bo, aux=x2)
        ctx.save_for_backward(x2, wqkv, wo, o2)
        ctx.saved_core = saved
        ctx.bqkv, ctx.bo = bqkv, bo
        ctx.shape = (B, S, H)
        return y.view(B, S, H)

    @staticmethod
    def backward(ctx, dy):
        x2, wqkv, wo, o2 = ctx.saved_tensors
        B, S, H = ctx.shape
        dy2 = dy.reshape(B * S, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        # dgrad GEMMs before the matching wgrad: see the ordering note in _MlpFn.backward
        do = gemm(dy2, wo, b_mn=True)
        dwo, dbo = _weight_bias_grad(dy2, o2, wo, ctx.bo)
        dqkv, dbqkv = _attention_core_backward(ctx.saved_core, do.view(B, S, H), ctx.bqkv)
        dqkv = dqkv.view(B * S, 3 * H)
        ctx.saved_core = None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dqkv, wqkv, b_mn=True, flags=EPI_RESIDUAL, aux=dy2).view(B, S, H)
        if dbqkv is _NOT_DONE:
            dwqkv, dbqkv = _weight_bias_grad(dqkv, x2, wqkv, ctx.bqkv)
        else:
            dwqkv = _weight_grad(dqkv, x2, wqkv)
        return dx, dwqkv, dbqkv, dwo, dbo, None, None


def attention_block(x, wqkv, bqkv, wo, bo, heads: int, causal: bool = False):
    """``x + out_proj(attention(qkv_proj(x)))`` for bf16 CUDA tensors (see ``_AttentionBlockFn``)."""
    return _AttentionBlockFn.apply(x, wqkv, bqkv, wo, bo, heads, causal)


def packed_attention(qkv, causal: bool = False):
    """``qkv``: ``[B, S, 3, heads, head_dim]`` -> context ``[B, S, heads*head_dim]``.

    ``TONY_B200_ATTENTION=tcgen05`` runs head_dim 64 / seq % 128 == 0 on the hand-written
    kernels of csrc/attention.cu.  The default is the library core: on B200 it is the cuDNN
    sm_100 flash kernel, which the own kernels do not beat yet (profiles/r1_attention_*.md has
    the measured comparison), and the default follows the measurement."""
    return _PackedAttentionFn.apply(qkv, causal)


# ---------------------------------------------------------------------------------------
# Optional NVTX ranges around every public op (TONY_B200_NVTX=1): a timeline tool then shows which
# framework call each kernel belongs to.  Off by default: a range is two driver calls per op.
# ---------------------------------------------------------------------------------------
def _nvtx_wrap(fn, label):
    import functools

    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        torch.cuda.nvtx.range_push(label)
        try:
            return fn(*args, **kwargs)
        finally:
            torch.cuda.nvtx.range_pop()

    return wrapped


NVTX_OPS = ("gemm", "colsum", "act_bwd_bias_grad", "linear", "mlp", "linear_act", "layer_norm", "softmax_cross_entropy",
            "embedding", "embedding_sum", "packed_attention", "attention_block", "conv2d_nhwc", "conv2d_dgrad_nhwc",
            "conv2d_wgrad_nhwc", "conv2d", "conv1x1_fork", "conv1x1_pair", "stem_conv7x7s2", "batch_norm_act",
            "max_pool_3x3_s2")
if os.environ.get("TONY_B200_NVTX", "0") == "1":
    for _name in NVTX_OPS:
        if _name in globals():
            globals()[_name] = _nvtx_wrap(globals()[_name], "tony." + _name)
