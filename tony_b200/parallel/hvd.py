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


def allreduce(tensor: torch.Tensor, average: bool = True, name: Optional[str] = None) -> torch.Tensor:
    """All-reduce any CUDA tensor through the fused kernel (bf16 on the wire, fp32 accumulation)."""
    world = size()
    if world == 1:
        return tensor.clone()
    comm = _staging()
    flat = tensor.reshape(-1)
    out = torch.empty_like(flat)
    scale = 1.0 / world if average else 1.0
    for lo in range(0, flat.numel(), STAGING_ELEMS):
        n = min(STAGING_ELEMS, flat.numel() - lo)
        npad = _aligned(n, world)
        comm.grad[:n].copy_(flat[lo:lo + n])
        if npad > n:
            comm.grad[n:npad].zero_()
        comm.allreduce_(0, npad, scale=scale, num_ctas=64)
        out[lo:lo + n].copy_(comm.param[:n])
    return out.view_as(tensor)


def broadcast(tensor: torch.Tensor, root_rank: int = 0) -> torch.Tensor:
    """Broadcast = all-reduce in which only the root contributes."""
    if size() == 1:
        return tensor
    src = tensor if rank() == root_rank else torch.zeros_like(tensor)
    return allreduce(src, average=False)


def broadcast_parameters(params, root_rank: int = 0) -> None:
    items = params.items() if hasattr(params, "items") else params
    for _, p in items:
        with torch.no_grad():
            p.copy_(broadcast(p.data, root_rank))


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
