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

Two layouts.  Symmetric (default): every rank is both a worker and a shard server, so no GPU idles.
Dedicated (``PsRoles``, the reference's literal ``tony.ps.instances`` / ``tony.worker.instances``
split): the ps ranks own every shard (weights, optimizer state, embedding rows) and run only the
fused reduce + update kernels; the worker ranks run forward / backward, push gradients through the
same kernels and pull weights / rows from the servers over NVLink.
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
        roles = store.roles
        if roles is None:
            owners, my_shard = list(range(comm.world)), comm.rank
        else:
            # dedicated servers: the shards live on the ps ranks only; a worker owns none of them and
            # starts its sweep at a shard chosen by its rank so the servers are read evenly
            owners = roles.owners
            oi = roles.owner_index(comm.rank)
            my_shard = oi if oi >= 0 else comm.rank % len(owners)
        nsh = len(owners)
        rows = N // nsh
        shard_ptrs = (ctypes.c_void_p * nsh)()
        for i, r in enumerate(owners):
            shard_ptrs[i] = int(comm.c.param[r]) + 2 * (slot.offset + i * rows * K)
        state.epoch += 1
        check(_lib.lib().tony_allgather_gemm_bf16(
            ptr(x2), shard_ptrs, ptr(weight), ptr(y), M, N, K, my_shard, nsh, flags, ptr(bias), ptr(pre),
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
        db = None
        if ctx.act in ("relu", "gelu", "gelu_tanh"):
            dy2, db = F.act_bwd_bias_grad(dy2, pre, ctx.act, ctx.bias)     # `pre` is the ReLU output for relu
        elif ctx.bias is not None:
            db = F._bias_grad(dy2, ctx.bias)
        # the forward left a complete local copy of W in the flat parameter buffer
        dx = F.gemm(dy2, weight, b_mn=True).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = F._weight_grad(dy2, x2, weight)
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
                 std: float = 0.02, copy_ctas: int = 0, device=None):
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


# ---------------------------------------------------------------------------------------
# Row-sharded embedding bag (K9): ps ownership for the sparse side of the recommender
# ---------------------------------------------------------------------------------------
class SymmBarrier:
    """A device-side barrier of its own (own flag words, own epoch counter) so users of symmetric
    memory outside the gradient buckets never interleave with the fused optimizer's flag traffic."""

    def __init__(self, device, group=None):
        import torch.distributed as dist

        from .symm import SymmBuffer, VmmBuffer, symm_mode

        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.epoch = 0
        if self.world == 1:
            return
        Buf = VmmBuffer if symm_mode() == "vmm" else SymmBuffer
        self._buf = Buf(1024, torch.int32, device, group)
        self._buf.local.zero_()
        c = _lib.FusedComm()
        c.rank, c.world = self.rank, self.world
        for r in range(self.world):
            c.flags[r] = self._buf.peer_ptrs[r]
        c.counter = self._buf.ptr + 2 * _lib.MAX_RANKS * 4
        c.status = self._buf.ptr + 2 * _lib.MAX_RANKS * 4 + 64
        self.c = c
        torch.cuda.synchronize(device)
        dist.barrier(group=group)

    def __call__(self):
        if self.world == 1:
            return
        self.epoch += 1
        check(_lib.lib().tony_symm_barrier(ctypes.byref(self.c), self.epoch, cur_stream()), "tony_symm_barrier")


class _EmbBagFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, ids, bag):
        B, Fd = ids.shape
        ids = ids.contiguous()
        deep = torch.empty((B, Fd * bag.deep_dim), dtype=torch.bfloat16, device=ids.device)
        rest = torch.empty((B, Fd, bag.rest_dim), dtype=torch.bfloat16, device=ids.device)
        check(_lib.lib().tony_embbag_gather(ctypes.byref(bag.comm), ptr(ids), B, Fd, bag.rows_per_shard,
                                            bag.row_elems, bag.deep_dim, ptr(deep), ptr(rest), cur_stream()),
              "tony_embbag_gather")
        ctx.bag, ctx.ids = bag, ids
        return deep, rest

    @staticmethod
    def backward(ctx, ddeep, drest):
        bag, ids = ctx.bag, ctx.ids
        B, Fd = ids.shape
        ddeep = ddeep.contiguous()
        drest = drest.contiguous()
        check(_lib.lib().tony_embbag_scatter(ctypes.byref(bag.comm), ptr(ids), B, Fd, bag.rows_per_shard,
                                             bag.row_elems, bag.deep_dim, ptr(ddeep), ptr(drest), cur_stream()),
              "tony_embbag_scatter")
        bag._dirty = True
        return None, None, None


class ShardedEmbeddingBag(nn.Module):
    """``fields`` categorical tables of ``vocab`` rows, each row ``[deep_dim | rest_dim]`` bf16,
    row ``r`` owned by the ``r % shards``-th owner (every rank, or the ps ranks of ``roles``;
    csrc/embbag.cu).  ``forward(ids[B, fields])`` returns the
    concatenated deep activation ``[B, fields*deep_dim]`` and the per-field ``rest`` rows
    ``[B, fields, rest_dim]`` (the "wide" logit contributions).  The tables are NOT ``nn.Parameter``s:
    gradients arrive at the owners as NVLink reductions and ``tony_step`` (called by the Trainer after
    backward) runs AdamW on the local shard only."""

    def __init__(self, fields: int, vocab: int, deep_dim: int, rest_dim: int = 8, std: float = 0.01,
                 device=None, group=None, seed: int = 1234, roles=None):
        super().__init__()
        import torch.distributed as dist

        from ..ops.optim import FlatAdamW
        from .symm import SymmBuffer, VmmBuffer, symm_mode

        if deep_dim % 8 or rest_dim % 8:
            raise ValueError("deep_dim and rest_dim must be multiples of 8")
        device = torch.device(device or "cuda")
        self.fields, self.vocab, self.deep_dim, self.rest_dim = fields, vocab, deep_dim, rest_dim
        self.row_elems = deep_dim + rest_dim
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        # row r of every field lives on owners[r % len(owners)]: every rank by default, the ps ranks when
        # the job runs dedicated parameter servers (`roles`)
        self.roles = roles
        self.owners = list(roles.owners) if roles is not None else list(range(self.world))
        self.shard = self.owners.index(self.rank) if self.rank in self.owners else -1
        self.contributors = len(roles.workers) if roles is not None else self.world
        nsh = len(self.owners)
        self.rows_per_shard = (vocab + nsh - 1) // nsh
        n = fields * self.rows_per_shard * self.row_elems
        n_pad = (n + 255) // 256 * 256
        if self.world > 1:
            Buf = VmmBuffer if symm_mode() == "vmm" else SymmBuffer
            self._tbuf = Buf(n_pad * 2, torch.bfloat16, device, group)
            self._gbuf = Buf(n_pad * 2, torch.bfloat16, device, group)
            self.table, self.grad = self._tbuf.local, self._gbuf.local
            tptrs, gptrs = self._tbuf.peer_ptrs, self._gbuf.peer_ptrs
        else:
            self.table = torch.empty(n_pad, dtype=torch.bfloat16, device=device)
            self.grad = torch.empty(n_pad, dtype=torch.bfloat16, device=device)
            tptrs, gptrs = [self.table.data_ptr()], [self.grad.data_ptr()]
        self.grad.zero_()
        comm = _lib.EmbBagComm()
        comm.rank, comm.world = max(self.shard, 0), nsh          # the kernels index shards, not ranks
        for i, r in enumerate(self.owners):
            comm.table[i] = tptrs[r]
            comm.grad[i] = gptrs[r]
        self.comm = comm
        # the same full table on every rank (seeded), of which this rank keeps rows rank::world
        master = torch.zeros(n_pad, dtype=torch.float32, device=device)
        view = master[:n].view(fields, self.rows_per_shard, self.row_elems)
        gen = torch.Generator(device=device)
        for f in range(fields):
            gen.manual_seed(seed + f)
            full = torch.randn(vocab, self.row_elems, generator=gen, device=device) * std
            if self.shard >= 0:
                mine = full[self.shard::nsh]
                view[f, : mine.shape[0]].copy_(mine)
        self.table.copy_(master.to(torch.bfloat16))
        self.opt = FlatAdamW(master, lr=1e-3, weight_decay=0.0)
        self.barrier = SymmBarrier(device, group)
        self.anchor = nn.Parameter(torch.zeros(8, dtype=torch.bfloat16, device=device))
        self._dirty = False
        self.barrier()

    def tony_configure(self, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        self.opt.lr, self.opt.betas, self.opt.eps, self.opt.weight_decay = lr, betas, eps, weight_decay

    def forward(self, ids):
        return _EmbBagFn.apply(self.anchor, ids, self)

    def tony_step(self):
        """Owner-side update: wait for every rank's gradient reductions, AdamW on the local shard,
        clear the shard, and release the next step's gathers."""
        self.barrier()
        if self.shard >= 0:
            self.opt.step(self.grad, self.table, grad_scale=1.0 / self.contributors)
            self.grad.zero_()
        self.barrier()
        self._dirty = False

    def tony_state_dict(self):
        """This rank's shard: fp32 master rows + AdamW moments (the bf16 table is re-derived)."""
        return {"opt": self.opt.state_dict(), "world": self.world, "rank": self.rank}

    def tony_load_state_dict(self, sd):
        if sd["world"] != self.world or sd["rank"] != self.rank:
            raise ValueError("embedding-bag shard was saved for a different rank layout")
        self.opt.load_state_dict(sd["opt"])
        self.table.copy_(self.opt.master.to(torch.bfloat16))
        self.grad.zero_()
        self.barrier()

    def full_table(self, field: int) -> torch.Tensor:
        """Debug / tests: rows of ``field`` held by THIS rank, in global row order rank::world."""
        n = self.fields * self.rows_per_shard * self.row_elems
        return self.table[:n].view(self.fields, self.rows_per_shard, self.row_elems)[field]
