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


import torch
from torch import nn

from .flat import FlatParamStore, FusedOptimizer


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device=None, bucket_mb: float = 64.0, optimizer: str = "sgd",
                 lr: float = 0.1, group=None, **opt_kw):
        super().__init__()
        self.module = module
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
        return self.module(*args, **kwargs)

    def step(self):
        self.optimizer.step()
