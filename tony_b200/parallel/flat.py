"""Flat parameter / gradient store and the fused (collective + optimizer) step.

All parameters of a model live in ONE bf16 buffer and all gradients in another, both in
symmetric memory, cut into buckets laid out in reverse registration order so bucket 0
is the first to complete during backward.  Gradient producers (the GEMM wgrad epilogue,
LayerNorm backward, colsum, embedding scatter) write straight into the gradient buffer
through a parameter's :class:`ParamSlot`; when the last slot of a bucket is written the
bucket's ``tony_fused_reduce_optim`` launch is queued on a side stream behind an event, so
the NVLink reduction + sharded optimizer update + weight broadcast of bucket *i* overlap
the backward GEMMs of bucket *i+1*.

This is the data-plane replacement for what launched jobs do today with stock pieces:
DDP-style per-parameter all-reduce then ``optimizer.step()``
(tony-examples/mnist-pytorch/mnist_distributed.py:113-126,83-87) or Horovod's tensor-fusion
all-reduce (tony-examples/horovod-on-tony/tensorflow2_mnist.py:73-76).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import torch

from ..ops import _lib
from ..ops._lib import OPT_ADAMW, OPT_SGD, check
from .symm import SymmComm

ALIGN = 256  # elements; also the weight-decay mask granularity


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass
class PsRoles:
    """Dedicated parameter servers: ``owners`` are the global ranks of the ``ps`` tasks (they hold the
    shards of every ps-sharded weight and run its optimizer), ``workers`` the ranks that compute
    gradients.  ``None`` everywhere means the symmetric layout (every rank is both)."""

    owners: List[int]
    workers: List[int]

    def owner_index(self, rank: int) -> int:
        return self.owners.index(rank) if rank in self.owners else -1

    def is_worker(self, rank: int) -> bool:
        return rank in self.workers


@dataclass
class Bucket:
    index: int
    begin: int
    count: int
    slots: List["ParamSlot"] = field(default_factory=list)
    pending: int = 0
    launched: bool = False
    ps_sharded: bool = False  # parameter-server bucket: weights stay sharded (no broadcast)
    state_off: int = 0        # this rank's slice of the fp32 optimizer state (set by FusedOptimizer)
    state_cnt: int = 0


class ParamSlot:
    """Where one parameter's gradient lives inside the flat gradient buffer."""

    __slots__ = ("store", "name", "offset", "numel", "shape", "grad_view", "expected", "written",
                 "bucket", "carried")

    def __init__(self, store, name, offset, numel, shape):
        self.store = store
        self.name = name
        self.offset = offset
        self.numel = numel
        self.shape = shape
        self.grad_view: Optional[torch.Tensor] = None
        self.expected = 1   # gradient contributions per step (2 for tied embeddings)
        self.written = 0
        self.carried = False  # holds the sum of earlier micro-batches (gradient accumulation, see no_sync)
        self.bucket: Optional[Bucket] = None

    def begin_write(self) -> bool:
        """True if an earlier contribution exists this step (the producer must accumulate)."""
        return self.written > 0 or self.carried

    def end_write(self) -> None:
        self.written += 1
        if self.written == self.expected and not self.store._defer:
            self.store._slot_ready(self)


class FlatParamStore:
    def __init__(self, params: Iterable[Tuple[str, torch.nn.Parameter]], device, group=None,
                 bucket_mb: float = 64.0, no_decay: Optional[Sequence[str]] = None,
                 ps_sharded: Optional[Iterable[torch.nn.Parameter]] = None, roles: Optional[PsRoles] = None):
        named = []
        seen = set()
        for name, p in params:
            if id(p) in seen or not p.requires_grad:
                continue
            seen.add(id(p))
            named.append((name, p))
        if not named:
            raise ValueError("no trainable parameters")
        import torch.distributed as dist

        self.device = torch.device(device)
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.roles = roles
        n_owners = len(roles.owners) if roles else self.world
        bucket_elems = max(ALIGN * self.world, int(bucket_mb * (1 << 20) / 2))
        bucket_align = ALIGN * self.world

        # reverse registration order ~ the order gradients become ready in backward
        self.slots: List[ParamSlot] = []
        self.buckets: List[Bucket] = []
        self._by_param: Dict[int, ParamSlot] = {}
        ps_ids = {id(p) for p in (ps_sharded or [])}
        offset = 0
        cur = Bucket(0, 0, 0)

        def close_bucket():
            nonlocal offset, cur
            if not cur.slots:
                return
            offset = cur.begin + _round_up(offset - cur.begin, bucket_align)
            cur.count = offset - cur.begin
            self.buckets.append(cur)
            cur = Bucket(len(self.buckets), offset, 0)

        for name, p in reversed(named):
            n = p.numel()
            if id(p) in ps_ids:
                # a parameter-server weight is a bucket of its own: rank r's 1/world shard of
                # the bucket is exactly rows [r*N/world, (r+1)*N/world) of the [N, K] matrix
                if p.dim() != 2 or n % bucket_align or p.shape[0] % n_owners or n % (ALIGN * n_owners):
                    raise ValueError(f"{name}: a ps-sharded weight must be 2-D with rows % shards == 0 "
                                     f"and numel % {bucket_align} == 0")
                close_bucket()
                cur.ps_sharded = True
            slot = ParamSlot(self, name, offset, n, tuple(p.shape))
            slot.bucket = cur
            cur.slots.append(slot)
            self.slots.append(slot)
            self._by_param[id(p)] = slot
            offset += _round_up(n, ALIGN)
            if cur.ps_sharded or offset - cur.begin >= bucket_elems:
                close_bucket()
        close_bucket()
        self.numel = offset

        self.comm = SymmComm(self.numel, device, group)
        self.rank = self.comm.rank
        self.param_buf = self.comm.param
        self.grad_buf = self.comm.grad

        # move parameters into the flat buffer; every rank starts from rank 0's values
        with torch.no_grad():
            for (name, p) in named:
                slot = self._by_param[id(p)]
                view = self.param_buf[slot.offset: slot.offset + slot.numel].view(slot.shape)
                view.copy_(p.data.to(torch.bfloat16))
            if self.world > 1:
                dist.broadcast(self.param_buf, src=0, group=group)
            for (name, p) in named:
                slot = self._by_param[id(p)]
                p.data = self.param_buf[slot.offset: slot.offset + slot.numel].view(slot.shape)
                slot.grad_view = self.grad_buf[slot.offset: slot.offset + slot.numel].view(slot.shape)
                p._tony_slot = slot
                p.grad = None
        self.named = named

        # weight-decay mask: one byte per ALIGN elements (0 = biases / norm scales)
        mask = torch.zeros(self.numel // ALIGN, dtype=torch.uint8)
        for (name, p) in named:
            slot = self._by_param[id(p)]
            decay = p.dim() >= 2 and not (no_decay and any(k in name for k in no_decay))
            if decay:
                mask[slot.offset // ALIGN: _round_up(slot.offset + slot.numel, ALIGN) // ALIGN] = 1
        self.wd_mask = mask.to(self.device)
        self._ready_cb = None
        self._defer = False       # inside no_sync(): gradients pile up in the flat buffer, nothing is launched
        self.reset_step()

    # -- bookkeeping -------------------------------------------------------------------
    def slot_of(self, p) -> ParamSlot:
        return self._by_param[id(p)]

    def mark_shared(self, p, uses: int) -> None:
        """Declare that ``p`` receives ``uses`` gradient contributions per step (tied weights)."""
        self._by_param[id(p)].expected = uses

    def end_micro_step(self) -> None:
        """Close a backward pass whose gradients stay local (gradient accumulation): what was written is kept
        and the next backward pass adds to it; slots it did not touch keep their earlier state."""
        self.flush_hook_grads()
        for s in self.slots:
            if s.written:
                s.carried = True
                s.written = 0

    def flush_hook_grads(self) -> None:
        """Generic (hook) path: move gradients that are still waiting for the rest of their bucket."""
        leftovers = getattr(self, "_pending_hook_grads", None)
        if not leftovers:
            return
        for idx, lst in leftovers.items():
            if lst:
                torch._foreach_copy_([s_.grad_view for s_, _ in lst], [g_ for _, g_ in lst])
                leftovers[idx] = []
                for s_, _ in lst:
                    s_.written += 1

    def reset_step(self) -> None:
        for s in self.slots:
            s.written = 0
            s.carried = False
        for b in self.buckets:
            b.pending = len(b.slots)
            b.launched = False

    def _slot_ready(self, slot: ParamSlot) -> None:
        b = slot.bucket
        b.pending -= 1
        if b.pending == 0 and self._ready_cb is not None:
            self._ready_cb(b)

    def register_autograd_hooks(self) -> None:
        """Generic path for parameters whose gradients come from stock autograd ops: collect
        ``p.grad`` as it is accumulated and move a whole bucket into the flat gradient buffer
        with ONE multi-tensor copy when its last gradient arrives."""
        pending: Dict[int, list] = {}

        for name, p in self.named:
            slot = self._by_param[id(p)]

            def hook(param, slot=slot):
                if param.grad is None:
                    return
                g = param.grad
                param.grad = None
                if slot.begin_write() or slot.expected != 1:
                    # shared / multiply-used parameter: accumulate immediately
                    if slot.begin_write():
                        slot.grad_view.add_(g.to(torch.bfloat16))
                    else:
                        slot.grad_view.copy_(g)
                    slot.end_write()
                    return
                b = slot.bucket
                lst = pending.setdefault(b.index, [])
                lst.append((slot, g))
                if len(lst) == sum(1 for s_ in b.slots if s_.expected == 1 and s_.written == 0):
                    torch._foreach_copy_([s_.grad_view for s_, _ in lst], [g_ for _, g_ in lst])
                    pending[b.index] = []
                    for s_, _ in lst:
                        s_.end_write()

            p.register_post_accumulate_grad_hook(hook)
        self._pending_hook_grads = pending

    def shard_range(self, b: Bucket) -> Tuple[int, int]:
        """(first element, length) of the part of bucket ``b`` whose optimizer state this rank holds."""
        if b.ps_sharded and self.roles is not None:
            oi = self.roles.owner_index(self.rank)
            chunk = b.count // len(self.roles.owners)
            return (b.begin + oi * chunk, chunk) if oi >= 0 else (b.begin, 0)
        chunk = b.count // self.world
        return b.begin + self.rank * chunk, chunk


class FusedOptimizer:
    """AdamW / SGD whose step IS the gradient collective (one kernel per bucket).

    ``overlap=True`` launches each bucket's kernel from inside backward as soon as the
    bucket is complete; ``step()`` then only flushes stragglers and joins the side stream.
    Optimizer state (fp32 master weights + moments) is sharded 1/world per rank.
    """

    def __init__(self, store: FlatParamStore, optimizer: str = "adamw", lr: float = 1e-3,
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, momentum: float = 0.9, overlap: bool = True,
                 comm_ctas: int = 0, broadcast: bool = True, grad_scale: float = 1.0):
        self.store = store
        self.kind = OPT_ADAMW if optimizer.lower() in ("adam", "adamw") else OPT_SGD
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.momentum = momentum
        self.overlap = overlap
        # CTAs the overlapped reduce + optimizer kernel takes from the GEMMs it runs under.  0 = by world size: 32,
        # but 64 at two ranks - each rank then owns half of the optimizer state and half of the reduction, and the
        # longer kernel costs the step more than the 32 SMs do (BERT-large, 2 GPUs: 1586 -> 1597 samples/s)
        self.comm_ctas = comm_ctas if comm_ctas > 0 else (64 if store.world == 2 else 32)
        self.broadcast = broadcast
        self.grad_scale = grad_scale
        self.t = 0
        self._step_open = False
        world = store.world
        dev = store.device
        # this rank's slices of the fp32 state, bucket after bucket (1/world of an all-reduce bucket;
        # 1/num_owners of a parameter-server bucket on its owners, nothing on pure workers)
        n_shard = 0
        for b in store.buckets:
            _, chunk = store.shard_range(b)
            b.state_off, b.state_cnt = n_shard, chunk
            n_shard += chunk
        n_alloc = max(n_shard, 8)
        self.master = torch.empty(n_alloc, dtype=torch.float32, device=dev)
        for b in store.buckets:
            lo, chunk = store.shard_range(b)
            self.master[b.state_off: b.state_off + chunk].copy_(store.param_buf[lo: lo + chunk])
        self.state1 = torch.zeros(n_alloc, dtype=torch.float32, device=dev)
        self.state2 = torch.zeros(n_alloc, dtype=torch.float32, device=dev) if self.kind == OPT_ADAMW else None
        # gradients are averaged over the ranks that produce them (dedicated servers contribute zeros)
        self.contributors = len(store.roles.workers) if store.roles else world
        self.comm_stream = torch.cuda.Stream(device=dev)
        self._join = torch.cuda.Event()
        store._ready_cb = self._bucket_ready if overlap else None

    # -- launches ----------------------------------------------------------------------
    def _open_step(self):
        if not self._step_open:
            self.t += 1
            self._step_open = True

    def _launch(self, b: Bucket, stream: torch.cuda.Stream, num_ctas: int):
        self._open_step()
        st = self.store
        if self.kind == OPT_ADAMW:
            b1, b2 = self.betas
            bc1, bc2 = 1.0 - b1 ** self.t, 1.0 - b2 ** self.t
        else:
            b1, b2, bc1, bc2 = self.momentum, 0.0, 1.0, 1.0
        comm = st.comm
        gscale = self.grad_scale / self.contributors
        if b.ps_sharded and st.roles is not None:
            # dedicated parameter servers: the bucket is cut over the owners only
            check(_lib.lib().tony_fused_reduce_optim_owners(
                ctypes.byref(comm.c), b.begin, b.count, _lib.ptr(self.master), _lib.ptr(self.state1),
                _lib.ptr(self.state2), b.state_off, self.kind, self.lr, b1, b2, self.eps, self.weight_decay,
                bc1, bc2, gscale, _lib.ptr(st.wd_mask), len(st.roles.owners), st.roles.owner_index(st.rank),
                comm.next_epoch(), num_ctas, ctypes.c_void_p(stream.cuda_stream)), "tony_fused_reduce_optim_owners")
            b.launched = True
            return
        if comm.nvls is not None:
            # in-switch reduction (multimem.ld_reduce) + multicast weight store
            check(_lib.lib().tony_fused_reduce_optim_nvls(
                ctypes.byref(comm.nvls), b.begin, b.count, _lib.ptr(self.master), _lib.ptr(self.state1),
                _lib.ptr(self.state2), b.state_off, self.kind, self.lr, b1, b2, self.eps,
                self.weight_decay, bc1, bc2, gscale, _lib.ptr(st.wd_mask),
                int(self.broadcast and not b.ps_sharded), comm.next_nvls_call(), num_ctas,
                ctypes.c_void_p(stream.cuda_stream)), "tony_fused_reduce_optim_nvls")
            b.launched = True
            return
        check(_lib.lib().tony_fused_reduce_optim(
            ctypes.byref(comm.c), b.begin, b.count, _lib.ptr(self.master), _lib.ptr(self.state1),
            _lib.ptr(self.state2), b.state_off, self.kind, self.lr, b1, b2, self.eps,
            self.weight_decay, bc1, bc2, gscale, _lib.ptr(st.wd_mask),
            int(self.broadcast and not b.ps_sharded), comm.next_epoch(), num_ctas,
            ctypes.c_void_p(stream.cuda_stream)), "tony_fused_reduce_optim")
        b.launched = True

    def _bucket_ready(self, b: Bucket):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.comm_stream.wait_event(ev)
        self._launch(b, self.comm_stream, self.comm_ctas)

    def step(self, grads_complete: bool = False):
        """Finish the step: launch whatever backward did not, then join the comm stream.
        ``grads_complete``: every gradient slot was written by device work the host did not see
        (a replayed CUDA graph), so the per-slot bookkeeping is skipped."""
        st = self.store
        cur = torch.cuda.current_stream()
        st.flush_hook_grads()
        for b in st.buckets:
            if b.launched:
                continue
            for s in b.slots:
                if s.written == 0 and not s.carried and not grads_complete:
                    s.grad_view.zero_()  # parameter unused this step
            if self.overlap:
                self._bucket_ready(b)
            else:
                self._launch(b, cur, 0 if st.world == 1 else max(self.comm_ctas, 64))
        if self.overlap:
            self._join.record(self.comm_stream)
            cur.wait_event(self._join)
        self._step_open = False
        st.reset_step()

    def zero_grad(self, set_to_none: bool = True):
        pass  # gradients are overwritten by the next step's producers (or accumulated under no_sync)

    # -- checkpoint --------------------------------------------------------------------
    def state_dict(self):
        sd = {"t": self.t, "master": self.master, "state1": self.state1, "rank": self.store.rank,
              "world": self.store.world, "kind": self.kind}
        if self.state2 is not None:
            sd["state2"] = self.state2
        return sd

    def load_state_dict(self, sd):
        if sd["world"] != self.store.world or sd["rank"] != self.store.rank:
            raise ValueError("optimizer shard belongs to a different rank/world layout")
        self.t = int(sd["t"])
        self.master.copy_(sd["master"])
        self.state1.copy_(sd["state1"])
        if self.state2 is not None:
            self.state2.copy_(sd["state2"])
