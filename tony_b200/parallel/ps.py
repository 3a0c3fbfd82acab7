"""Parameter-server (ps/worker) role on NVSwitch: sharded weights, gather fused into the GEMM.

The reference only *orchestrates* ps/worker jobs (``tony.ps.instances``; ps tasks untracked and
killed at the end — tony-core/src/main/resources/tony-default.xml:160-170,204-223); the
variable push/pull itself is TF gRPC or MXNet kvstore inside the user process
(tony-examples/mnist-tensorflow/mnist_distributed.py:213-241,
tony-examples/linearregression-mxnet/src/mxnet_dist_ex.py:35,61-68).  Here the role is a
data-plane mode:

* push = gradient reduce-scatter fused with the *sharded* optimizer update
  (``tony_fused_reduce_optim`` with ``broadcast=0``): rank r is the "server" of rows
  ``[r*N/world, (r+1)*N/world)`` of every ps-sharded weight and is the only one holding their
  fp32 master copy and moments;
* pull = ``tony_allgather_gemm_bf16``: the first GEMM that uses the weight in the next forward
  pulls the remote shards over NVLink slab by slab inside the same kernel that multiplies.

Every rank is both a worker and a shard server, so 2 ps + 6 workers of the reference's
examples map onto 8 symmetric ranks with no idle "ps" GPUs.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
from torch import nn

from ..models.nn import _param
from ..ops import _lib
from ..ops import functional as F
from ..ops._lib import EPI_BIAS, EPI_SAVE_PREACT, check, cur_stream, ptr


class _AgLinearFn(torch.autograd.Function):
    """y = act(x W^T + b) where W's row shards are pulled from their owners during the GEMM."""

    @staticmethod
    def forward(ctx, x, weight, bias, act, state):
        slot = weight._tony_slot
        store = slot.store
        comm = store.comm
        N, K = weight.shape
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        y = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
        flags = (EPI_BIAS if bias is not None else 0) | F._ACT_FWD[act]
        pre = None
        if act not in (None, "none"):
            pre = torch.empty_like(y)
            flags |= EPI_SAVE_PREACT
        world, rank = comm.world, comm.rank
        rows = N // world
        shard_ptrs = (ctypes.c_void_p * world)()
        for r in range(world):
            shard_ptrs[r] = int(comm.c.param[r]) + 2 * (slot.offset + r * rows * K)
        state.epoch += 1
        check(_lib.lib().tony_allgather_gemm_bf16(
            ptr(x2), shard_ptrs, ptr(weight), ptr(y), M, N, K, rank, world, flags, ptr(bias), ptr(pre),
            ptr(state.flags), state.epoch, state.copy_ctas, cur_stream()), "tony_allgather_gemm_bf16")
        ctx.save_for_backward(x2, weight, pre if pre is not None else y)
        ctx.bias, ctx.act, ctx.x_shape = bias, act, x.shape
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, pre = ctx.saved_tensors
        N = weight.shape[0]
        dy2 = dy.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        if ctx.act == "relu":
            dy2 = (dy2 * (pre > 0)).contiguous()
        elif ctx.act in ("gelu", "gelu_tanh"):
            p = pre.float().requires_grad_(True)
            with torch.enable_grad():
                g = torch.nn.functional.gelu(p, approximate="tanh" if ctx.act == "gelu_tanh" else "none")
            dy2 = torch.autograd.grad(g, p, dy2.float())[0].to(torch.bfloat16).contiguous()
        # the forward left a complete local copy of W in the flat parameter buffer
        dx = F.gemm(dy2, weight, b_mn=True).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = F._weight_grad(dy2, x2, weight)
        db = F._bias_grad(dy2, ctx.bias)
        return dx, dw, db, None, None


class _AgState:
    def __init__(self, n_rows: int, device, copy_ctas: int):
        self.flags = torch.zeros((n_rows + 127) // 128 + 1, dtype=torch.int32, device=device)
        self.epoch = 0
        self.copy_ctas = copy_ctas


class ShardedLinear(nn.Module):
    """Linear whose weight is owned row-wise by the ranks (register it as ``ps_sharded`` in the
    FlatParamStore, e.g. via ``Trainer(..., ps_sharded=model.ps_parameters())``)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, act: Optional[str] = None,
                 std: float = 0.02, copy_ctas: int = 24, device=None):
        super().__init__()
        self.weight = _param(out_features, in_features, std=std, device=device)
        self.bias = _param(out_features, fill=0.0, device=device) if bias else None
        self.act = act
        self._state: Optional[_AgState] = None
        self.copy_ctas = copy_ctas

    def forward(self, x):
        slot = getattr(self.weight, "_tony_slot", None)
        if slot is None or not slot.bucket.ps_sharded:
            return F.linear_act(x, self.weight, self.bias, self.act or "none")
        if self._state is None:
            self._state = _AgState(self.weight.shape[0], self.weight.device, self.copy_ctas)
        return _AgLinearFn.apply(x, self.weight, self.bias, self.act, self._state)


def ps_parameters(module: nn.Module):
    """Weights of every ShardedLinear in ``module`` (what to pass as ``ps_sharded``)."""
    return [m.weight for m in module.modules() if isinstance(m, ShardedLinear)]
