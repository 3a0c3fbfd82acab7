"""DistributedDataParallel for arbitrary ``nn.Module``s on the fused data plane.

Stock DDP all-reduces gradient buckets with NCCL and leaves the optimizer to the user; the
reference's PyTorch example hand-rolls the same thing one parameter at a time
(tony-examples/mnist-pytorch/mnist_distributed.py:113-126).  This wrapper moves the module's
parameters into the flat symmetric buffers and registers post-accumulate-grad hooks that copy
each gradient into its bucket slot, so any model — including ones whose layers are stock ATen /
cuDNN ops, such as the ResNet-50 convolutions — gets the fused reduce + sharded optimizer +
broadcast kernel, overlapped with backward.
"""
from __future__ import annotations

import contextlib

import torch
from torch import nn

from .flat import FlatParamStore, FusedOptimizer


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device=None, bucket_mb: float = 64.0, optimizer: str = "sgd",
                 lr: float = 0.1, group=None, broadcast_buffers: bool = False, **opt_kw):
        super().__init__()
        self.module = module
        self.group = group
        self.broadcast_buffers = broadcast_buffers
        device = device or next(module.parameters()).device
        module.to(device=device, dtype=torch.bfloat16)
        self.store = FlatParamStore(module.named_parameters(), device, group=group, bucket_mb=bucket_mb,
                                    no_decay=("bias", "bn", "norm"))
        for p, uses in getattr(module, "shared_params", {}).items():
            self.store.mark_shared(p, uses)
        self.store.register_autograd_hooks()
        self.optimizer = FusedOptimizer(self.store, optimizer=optimizer, lr=lr, **opt_kw)
        # buffers (BatchNorm running stats) follow rank 0 once, like DDP's initial sync
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            for b in module.buffers():
                dist.broadcast(b, src=0, group=group)

    def forward(self, *args, **kwargs):
        if self.store._defer and any(s.written for s in self.store.slots):
            self.store.end_micro_step()       # several micro-batches inside one no_sync() block
        if self.broadcast_buffers and self.store.world > 1 and torch.is_grad_enabled():
            import torch.distributed as dist

            for b in self.module.buffers():   # torch DDP's per-forward buffer sync (BatchNorm running stats)
                dist.broadcast(b, src=0, group=self.group)
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        """Gradient accumulation, as in ``torch.nn.parallel.DistributedDataParallel.no_sync``: backward passes
        inside the block add their gradients into the flat bucket and launch nothing; the first backward pass
        after the block adds its own and runs the fused reduce + optimizer on the sum.  Scale the loss by the
        number of micro-batches yourself, as with torch's DDP."""
        prev = self.store._defer
        self.store._defer = True
        try:
            yield
        finally:
            self.store.end_micro_step()
            self.store._defer = prev

    def step(self):
        self.optimizer.step()
