"""chief / worker / evaluator job in the shape of ``tf.estimator.train_and_evaluate`` - the reference's
tony-examples/mnist-tensorflow/mnist_estimator_distributed.py:32-99: the chief and the workers train and the chief
writes checkpoints into a shared model directory; an ``evaluator`` task (not part of the training cluster: the
launcher drops it from everyone else's TF_CONFIG and it is untracked, so the job ends when the trainers do) scores
every checkpoint it sees.

Dependency-free apart from torch: the trainers average gradients over ``torch.distributed`` (gloo) using the
addresses in TF_CONFIG, checkpoints are ``torch.save`` files.  ``--model_dir`` must be visible to every task (any
local path on the single-node cluster)."""
import argparse
import glob
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from tony_b200.models.mnist import synthetic_mnist  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model_dir", default=os.path.join(os.environ.get("TMPDIR", "/tmp"), "tony_b200_estimator_" +
                                                    os.environ.get("JOB_ID", os.environ.get("appid", "local"))))
ap.add_argument("--steps", type=int, default=80)
ap.add_argument("--save_every", type=int, default=20)
args = ap.parse_args()
cfg = json.loads(os.environ["TF_CONFIG"])
role, index = cfg["task"]["type"], int(cfg["task"]["index"])
cluster = cfg["cluster"]
os.makedirs(args.model_dir, exist_ok=True)


def accuracy(w, b, seed):
    x, y = synthetic_mnist(1024, seed=seed)
    return ((x.view(1024, -1) @ w + b).argmax(1) == y).float().mean().item()


if role == "evaluator":
    assert "evaluator" in cluster and all(k in ("chief", "worker", "evaluator", "ps") for k in cluster), cluster
    seen, deadline = set(), time.time() + 120
    while time.time() < deadline:
        for path in sorted(glob.glob(os.path.join(args.model_dir, "ckpt-*.pt"))):
            if path not in seen:
                seen.add(path)
                state = torch.load(path)
                print(f"evaluator: {os.path.basename(path)} accuracy {accuracy(state['w'], state['b'], 999):.3f}", flush=True)
        if os.path.exists(os.path.join(args.model_dir, "DONE")):
            break
        time.sleep(0.2)
    sys.exit(0)

assert "evaluator" not in cluster, "trainers must not see the evaluator in TF_CONFIG"
trainers = list(cluster.get("chief", [])) + list(cluster.get("worker", []))
rank = 0 if role == "chief" else len(cluster.get("chief", [])) + index
host, port = trainers[0].rsplit(":", 1)
dist.init_process_group("gloo", init_method=f"tcp://{host}:{port}", rank=rank, world_size=len(trainers))
w = torch.zeros(784, 10, requires_grad=True)
b = torch.zeros(10, requires_grad=True)
x, y = synthetic_mnist(2048, seed=rank)
for step in range(1, args.steps + 1):
    i = (step * 64) % 1984
    loss = torch.nn.functional.cross_entropy(x[i:i + 64].view(64, -1) @ w + b, y[i:i + 64])
    gw, gb = torch.autograd.grad(loss, (w, b))
    for g in (gw, gb):
        dist.all_reduce(g)
        g /= len(trainers)
    with torch.no_grad():
        w -= 0.5 * gw
        b -= 0.5 * gb
    if role == "chief" and step % args.save_every == 0:
        tmp = os.path.join(args.model_dir, f".ckpt-{step:05d}.tmp")
        torch.save({"w": w.detach(), "b": b.detach(), "step": step}, tmp)
        os.replace(tmp, os.path.join(args.model_dir, f"ckpt-{step:05d}.pt"))
dist.barrier()
if role == "chief":
    acc = accuracy(w.detach(), b.detach(), 999)
    print(f"chief: final accuracy {acc:.3f}", flush=True)
    open(os.path.join(args.model_dir, "DONE"), "w").close()
    assert acc > 0.9
dist.destroy_process_group()
