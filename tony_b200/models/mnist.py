"""The reference's MNIST plumbing model: ``Linear(784, 10)`` + log-softmax / NLL
(tony-examples/mnist-pytorch/mnist_distributed.py:129-136, step loop :76-90).  Plain torch so
the CPU/gloo configuration (BASELINE.json config #1) runs anywhere; on a GPU the same job can
swap in ``tony_b200.models.nn.Linear`` (784 and 10 are padded to multiples of 8)."""
from __future__ import annotations

import torch
from torch import nn


class MnistNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(784, 10)

    def forward(self, x):
        return torch.log_softmax(self.fc(x.view(x.shape[0], -1)), dim=1)


def synthetic_mnist(n: int, seed: int = 0):
    """Linearly separable fake digits: class k lights up a distinct pixel block."""
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 10, (n,), generator=g)
    x = torch.randn(n, 1, 28, 28, generator=g) * 0.1
    for k in range(10):
        x[y == k, 0, 2 * k: 2 * k + 4, 4:24] += 1.0
    return x, y
