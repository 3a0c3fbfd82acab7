"""Horovod-shaped MNIST job written against ``tony_b200.parallel.hvd``.

Same skeleton as the reference's Horovod examples (tony-examples/horovod-on-tony/
tensorflow2_keras_mnist.py:28-36,55-56,69,75 and tensorflow2_mnist.py:73-76,85-86): ``hvd.init()``,
pin the GPU by ``hvd.local_rank()``, scale the learning rate by ``hvd.size()``, broadcast the initial
state from rank 0, wrap the optimizer in ``hvd.DistributedOptimizer`` (here the gradient averaging IS the
optimizer step: one fused NVLink kernel per gradient bucket), average the reported metric with
``hvd.allreduce`` and checkpoint on rank 0 only.  Synthetic MNIST-shaped data (no dataset access).

    python -m tony_b200.cli.cluster_submitter --conf_file examples/horovod-on-tony/tony.xml \
        --executes examples/horovod-on-tony/torch_mnist_hvd.py --python_binary_path $(which python)
"""
import argparse
import os
import sys

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")))

import torch
from torch import nn

import tony_b200.parallel.hvd as hvd
from tony_b200 import runtime
from tony_b200.models.nn import Linear


class Net(nn.Module):
    """784 -> 512 -> 256 -> 16 (10 classes padded to 16 columns) on the tcgen05 GEMM with fused ReLU."""

    def __init__(self, device):
        super().__init__()
        self.fc1 = Linear(784, 512, device=device)
        self.fc2 = Linear(512, 256, device=device)
        self.fc3 = Linear(256, 16, device=device)

    def forward(self, x):
        return self.fc3(self.fc2(self.fc1(x, act="relu"), act="relu"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--ckpt", default="")
    a = ap.parse_args()

    hvd.init()
    torch.cuda.set_device(hvd.local_rank() % max(torch.cuda.device_count(), 1))
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(1234 + hvd.rank())            # ranks start DIFFERENT on purpose ...
    model = Net(dev)
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)    # ... and take rank 0's state
    opt = hvd.DistributedOptimizer(model, optimizer="sgd", lr=0.01 * hvd.size(), momentum=0.9, weight_decay=0.0)

    from tony_b200.ops import functional as F

    g = torch.Generator().manual_seed(7)             # the same synthetic "dataset" on every rank
    protos = torch.randn(10, 784, generator=g)
    first = last = None
    for step in range(a.steps):
        gs = torch.Generator().manual_seed(1000 * step + hvd.rank())    # each rank draws its own shard
        labels = torch.randint(0, 10, (a.batch,), generator=gs)
        images = (protos[labels] + 0.5 * torch.randn(a.batch, 784, generator=gs)).to(torch.bfloat16)
        loss = F.softmax_cross_entropy(model(images.to(dev)), labels.to(dev), vocab=10)
        loss.backward()
        opt.step()                                   # fused all-reduce + SGD, overlapped with backward
        metric = hvd.allreduce(loss.detach(), average=True).item()     # fp32 metric averaging
        first = metric if first is None else first
        last = metric
        if hvd.rank() == 0 and (step % 10 == 0 or step == a.steps - 1):
            print(f"step {step} mean loss over {hvd.size()} ranks {metric:.4f}", flush=True)
        if step == 0:
            runtime.mark_first_step()
    opt.synchronize()
    opt.check()
    if hvd.rank() == 0:
        if a.ckpt:                                   # checkpoint on rank 0 only, like the reference's examples
            torch.save({k: v.cpu() for k, v in model.state_dict().items()}, a.ckpt)
        assert last < first, f"loss did not decrease: {first} -> {last}"
        print(f"done: loss {first:.4f} -> {last:.4f}", flush=True)
    if hvd.size() > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
