"""Horovod-style API over the fused NVLink data plane.

The reference's Horovod support is pure orchestration: it starts a rendezvous driver and
exports ``HOROVOD_*`` with ``HOROVOD_CPU_OPERATIONS=gloo`` (tony-core/src/main/java/com/linkedin/
tony/runtime/HorovodRuntime.java:312-350), i.e. on GPUs its jobs would all-reduce through host
memory.  A job launched with ``tony.application.framework=horovod`` can instead
``import tony_b200.parallel.hvd as hvd`` and keep the calls it already makes
(tony-examples/horovod-on-tony/tensorflow2_mnist.py:73-86, tensorflow2_keras_mnist.py:55-75):
``init / rank / size / local_rank / allreduce / broadcast_parameters / DistributedOptimizer``.
Tensor fusion is the flat bucket layout; the all-reduce is fused with unscale + optimizer.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from .. import runtime
from .flat import FlatParamStore, FusedOptimizer
from .symm import SymmComm

_ctx = {"rank": 0, "world": 1, "local_rank": 0, "device": None, "staging": None}
STAGING_ELEMS = 32 << 20  # 64 MB bf16 staging buffer for free-standing allreduce / broadcast


def init():
    # the launcher exports both the HOROVOD_* and the torch env:// names
    if "RANK" not in os.environ and "HOROVOD_RANK" in os.environ:
        os.environ["RANK"] = os.environ["HOROVOD_RANK"]
        os.environ["WORLD_SIZE"] = os.environ["HOROVOD_SIZE"]
    rank, world, dev = runtime.init()
    _ctx.update(rank=rank, world=world, device=dev,
                local_rank=int(os.environ.get("HOROVOD_LOCAL_RANK", os.environ.get("LOCAL_RANK", "0"))))
    return rank, world


def rank() -> int:
    return _ctx["rank"]


def size() -> int:
    return _ctx["world"]


def local_rank() -> int:
    return _ctx["local_rank"]


def _staging() -> SymmComm:
    if _ctx["staging"] is None:
        _ctx["staging"] = SymmComm(STAGING_ELEMS, _ctx["device"] or torch.device("cuda", torch.cuda.current_device()))
    return _ctx["staging"]


def _aligned(n: int, world: int) -> int:
    q = world * 8
    return (n + q - 1) // q * q


def _as_bytes(t: torch.Tensor) -> torch.Tensor:
    return t.contiguous().view(-1).view(torch.uint8)


def allreduce(tensor: torch.Tensor, average: bool = True, name: Optional[str] = None) -> torch.Tensor:
    """Sum (or mean) of ``tensor`` over all ranks through the peer-memory kernels.

    The wire format follows the tensor: bf16 tensors travel as bf16 (in-switch reduction when NVLS is
    up), fp32 / fp16 / fp64 tensors travel and accumulate as fp32 - metric averaging
    (tony-examples/horovod-on-tony/tensorflow2_keras_mnist.py:75) keeps full precision.  Integer and
    bool tensors are refused: a floating-point sum would silently corrupt them."""
    if not tensor.is_floating_point():
        raise TypeError(f"hvd.allreduce: dtype {tensor.dtype} is not supported (floating-point tensors only; "
                        "cast counters to float or use hvd.allgather / hvd.broadcast)")
    world = size()
    if world == 1:
        return tensor.clone()
    comm = _staging()
    scale = 1.0 / world if average else 1.0
    if tensor.dtype == torch.bfloat16:
        flat = tensor.reshape(-1)
        out = torch.empty_like(flat)
        for lo in range(0, flat.numel(), STAGING_ELEMS):
            n = min(STAGING_ELEMS, flat.numel() - lo)
            npad = _aligned(n, world)
            comm.grad[:n].copy_(flat[lo:lo + n])
            if npad > n:
                comm.grad[n:npad].zero_()
            comm.allreduce_(0, npad, scale=scale, num_ctas=64)
            out[lo:lo + n].copy_(comm.param[:n])
        return out.view_as(tensor)
    flat = tensor.reshape(-1).float()
    out = torch.empty_like(flat)
    grad32, param32 = comm.grad.view(torch.float32), comm.param.view(torch.float32)
    cap = STAGING_ELEMS // 2
    for lo in range(0, flat.numel(), cap):
        n = min(cap, flat.numel() - lo)
        npad = (n + world * 4 - 1) // (world * 4) * (world * 4)      # bytes % (world * 16) == 0
        grad32[:n].copy_(flat[lo:lo + n])
        if npad > n:
            grad32[n:npad].zero_()
        comm.collective_("allreduce", 0, npad * 4, fp32=True, scale=scale, num_ctas=64)
        out[lo:lo + n].copy_(param32[:n])
    return out.view_as(tensor).to(tensor.dtype)


def broadcast(tensor: torch.Tensor, root_rank: int = 0) -> torch.Tensor:
    """The root's ``tensor`` on every rank (any dtype: the payload moves as raw bytes; with NVLS the root
    publishes it through the switch with one multimem.st per 16 bytes)."""
    world = size()
    if world == 1:
        return tensor
    comm = _staging()
    src = _as_bytes(tensor)
    out = torch.empty_like(src)
    staging_bytes = comm.param.view(torch.uint8)
    cap = STAGING_ELEMS * 2
    for lo in range(0, src.numel(), cap):
        n = min(cap, src.numel() - lo)
        npad = (n + world * 16 - 1) // (world * 16) * (world * 16)
        if rank() == root_rank:
            staging_bytes[:n].copy_(src[lo:lo + n])
        comm.broadcast_(0, npad // 2, root=root_rank, num_ctas=32)
        out[lo:lo + n].copy_(staging_bytes[:n])
    return out.view(tensor.dtype).view(tensor.shape)


def allgather(tensor: torch.Tensor) -> torch.Tensor:
    """Concatenation of every rank's ``tensor`` along dim 0 (equal shapes on all ranks, any dtype)."""
    world = size()
    if world == 1:
        return tensor.clone()
    comm = _staging()
    src = _as_bytes(tensor)
    n = src.numel()
    seg = (n + 15) // 16 * 16
    if seg * world > STAGING_ELEMS * 2:
        raise ValueError("hvd.allgather: tensor larger than the staging buffer")
    gbytes, pbytes = comm.grad.view(torch.uint8), comm.param.view(torch.uint8)
    gbytes[rank() * seg: rank() * seg + n].copy_(src)
    comm.collective_("allgather", 0, seg * world, num_ctas=32)
    parts = [pbytes[r * seg: r * seg + n] for r in range(world)]
    out = torch.cat(parts).view(tensor.dtype)
    return out.view((world * tensor.shape[0],) + tuple(tensor.shape[1:])) if tensor.dim() else out


def reducescatter(tensor: torch.Tensor, average: bool = False) -> torch.Tensor:
    """This rank's 1/world slice (dim 0) of the sum of ``tensor`` over all ranks (bf16 or fp32 wire)."""
    if not tensor.is_floating_point():
        raise TypeError("hvd.reducescatter: floating-point tensors only")
    world = size()
    if world == 1:
        return tensor.clone()
    if tensor.shape[0] % world:
        raise ValueError("hvd.reducescatter: dim 0 must be divisible by the number of ranks")
    comm = _staging()
    fp32 = tensor.dtype != torch.bfloat16
    flat = tensor.reshape(-1).float() if fp32 else tensor.reshape(-1)
    esize = 4 if fp32 else 2
    n = flat.numel()
    if (n // world * esize) % 16 or n * esize > STAGING_ELEMS * 2:
        raise ValueError("hvd.reducescatter: each rank's slice must be a multiple of 16 bytes and fit the staging buffer")
    g = comm.grad.view(torch.float32) if fp32 else comm.grad
    pr = comm.param.view(torch.float32) if fp32 else comm.param
    g[:n].copy_(flat)
    comm.collective_("reduce_scatter", 0, n * esize, fp32=fp32, scale=(1.0 / world if average else 1.0), num_ctas=64)
    chunk = n // world
    out = pr[rank() * chunk:(rank() + 1) * chunk].clone()
    return out.to(tensor.dtype).view((tensor.shape[0] // world,) + tuple(tensor.shape[1:]))


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """``hvd.broadcast_parameters(model.state_dict(), root_rank=0)``: every tensor takes the root's value
    (tony-examples/horovod-on-tony/tensorflow2_mnist.py:85-86 broadcast_variables)."""
    items = params.items() if hasattr(params, "items") else params
    for _, p in items:
        if not torch.is_tensor(p):
            continue
        with torch.no_grad():
            data = p.data if hasattr(p, "data") else p
            data.copy_(broadcast(data, root_rank))


class DistributedOptimizer:
    """``hvd.DistributedOptimizer``-shaped wrapper: gradient averaging *is* the optimizer step."""

    def __init__(self, model: torch.nn.Module, optimizer: str = "adamw", lr: float = 1e-3, bucket_mb: float = 64.0,
                 generic_hooks: bool = False, **kw):
        dev = _ctx["device"] or torch.device("cuda", torch.cuda.current_device())
        self.store = FlatParamStore(model.named_parameters(), dev, bucket_mb=bucket_mb)
        for p, uses in getattr(model, "shared_params", {}).items():
            self.store.mark_shared(p, uses)
        if generic_hooks:
            self.store.register_autograd_hooks()
        self.opt = FusedOptimizer(self.store, optimizer=optimizer, lr=lr, **kw)

    def step(self):
        self.opt.step()

    def zero_grad(self, set_to_none: bool = True):
        self.opt.zero_grad(set_to_none)

    def synchronize(self):
        torch.cuda.current_stream().wait_stream(self.opt.comm_stream)

    def set_lr(self, lr: float) -> None:
        """Learning-rate schedules / warm-up callbacks (the rate is a launch argument of the fused kernel, so the
        next ``step()`` uses it)."""
        self.opt.lr = float(lr)

    @property
    def lr(self) -> float:
        return self.opt.lr

    def check(self):
        """Raise if a fused collective gave up waiting for a peer (GPU-side hang detection)."""
        self.store.comm.check_status()
