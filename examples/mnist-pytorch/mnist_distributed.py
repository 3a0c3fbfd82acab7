"""MNIST-style plumbing job: 2 workers, CPU, gloo — BASELINE config #1.

Same job shape as the reference's tony-examples/mnist-pytorch/mnist_distributed.py (a Linear
(784,10) net, gradients averaged across workers every step) but driven through the runtime
package: ``runtime.init()`` consumes the launcher's env contract (INIT_METHOD/RANK/WORLD or the
env:// names), gradients are averaged with ONE flattened all-reduce per step instead of the
reference's per-parameter all-reduce with a fresh process group each time (:113-126)."""
import argparse
import os
import sys

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")))

import torch
import torch.distributed as dist

from tony_b200 import runtime
from tony_b200.models.mnist import MnistNet, synthetic_mnist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.05)
    args = ap.parse_args()
    rank, world, dev = runtime.init(device="cpu")
    torch.manual_seed(0)
    model = MnistNet()
    if world > 1:
        for p in model.parameters():
            dist.broadcast(p.data, src=0)
    opt = torch.optim.SGD(model.parameters(), lr=args.lr)
    x, y = synthetic_mnist(2048, seed=rank)
    first = True
    for epoch in range(args.epochs):
        correct, total, loss_sum = 0, 0, 0.0
        for i in range(0, len(x), args.batch):
            xb, yb = x[i:i + args.batch], y[i:i + args.batch]
            out = model(xb)
            loss = torch.nn.functional.nll_loss(out, yb)
            opt.zero_grad()
            loss.backward()
            if world > 1:
                flat = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
                dist.all_reduce(flat)
                flat /= world
                off = 0
                for p in model.parameters():
                    p.grad.copy_(flat[off:off + p.numel()].view_as(p))
                    off += p.numel()
            opt.step()
            if first:
                runtime.mark_first_step()
                first = False
            correct += (out.argmax(1) == yb).sum().item()
            total += len(yb)
            loss_sum += loss.item() * len(yb)
        print(f"rank {rank} epoch {epoch} loss {loss_sum / total:.4f} acc {correct / total:.3f}", flush=True)
    assert correct / total > 0.9, "synthetic MNIST should be learnable"
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
