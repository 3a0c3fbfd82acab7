"""Synchronous multi-worker training driven by TF_CONFIG: the job shape of the reference's
tony-examples/mnist-tensorflow/mnist_keras_distributed.py:58-71 (MultiWorkerMirroredStrategy: every task is a
``worker``, worker 0 is the chief, gradients are all-reduced every step).

If TensorFlow is importable the cluster resolver is exercised.  The training itself is dependency-free apart from
torch: a softmax regression on synthetic digits whose gradients are averaged with ``torch.distributed`` over the
addresses TF_CONFIG lists (gloo on CPU; on a GPU box the same loop can wrap the model in
``tony_b200.parallel.ddp.DistributedDataParallel``).  What it shows is what the launcher provides: a consistent
TF_CONFIG on every task and the gang barrier before the first step."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from tony_b200.models.mnist import synthetic_mnist  # noqa: E402

if "TF_CONFIG" not in os.environ:
    sys.exit("TF_CONFIG is not set: submit this script through the launcher with tony.application.framework=tensorflow")
cfg = json.loads(os.environ["TF_CONFIG"])
workers = cfg["cluster"]["worker"]
index = int(cfg["task"]["index"])
assert cfg["task"]["type"] == "worker", cfg["task"]
try:
    import tensorflow as tf  # noqa: F401

    print("TF cluster spec:", tf.distribute.cluster_resolver.TFConfigClusterResolver().cluster_spec().as_dict(), flush=True)
except ImportError:
    print("tensorflow not installed; all-reducing with torch.distributed over the TF_CONFIG addresses", flush=True)

# worker 0's address doubles as the rendezvous point (the launcher reserved that port for this task)
host, port = workers[0].rsplit(":", 1)
dist.init_process_group("gloo", init_method=f"tcp://{host}:{port}", rank=index, world_size=len(workers))

torch.manual_seed(0)                       # identical initial weights everywhere, as the strategy's broadcast does
w = torch.zeros(784, 10, requires_grad=True)
b = torch.zeros(10, requires_grad=True)
x, y = synthetic_mnist(2048, seed=index)   # every worker reads its own shard
first = last = None
for step in range(60):
    i = (step * 64) % 1984
    loss = torch.nn.functional.cross_entropy(x[i:i + 64].view(64, -1) @ w + b, y[i:i + 64])
    gw, gb = torch.autograd.grad(loss, (w, b))
    for g in (gw, gb):
        dist.all_reduce(g)
        g /= len(workers)
    with torch.no_grad():
        w -= 0.5 * gw
        b -= 0.5 * gb
    first = loss.item() if first is None else first
    last = loss.item()
check = w.detach().sum().reshape(1).clone()
dist.all_reduce(check)                     # replicas must have stayed identical
assert abs(check.item() / len(workers) - w.detach().sum().item()) < 1e-4, "replicas diverged"
print(f"worker {index}/{len(workers)} (chief: {index == 0}): loss {first:.3f} -> {last:.3f}", flush=True)
assert last < first
dist.destroy_process_group()
