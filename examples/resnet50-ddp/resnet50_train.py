"""ResNet-50 DDP bf16 (BASELINE config #2): parallel.ddp.DistributedDataParallel + fused
SGD-momentum; see tony_b200/models/resnet.py for which layers are hand-written kernels."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")))

import torch

from tony_b200 import runtime
from tony_b200.models import resnet
from tony_b200.parallel.ddp import DistributedDataParallel

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--size", type=int, default=224)
a = ap.parse_args()
rank, world, dev = runtime.init()
torch.manual_seed(0)
model = DistributedDataParallel(resnet.ResNet50(device=dev), device=dev, optimizer="sgd", lr=0.1,
                                momentum=0.9, weight_decay=1e-4)
batches = [resnet.synthetic_batch(a.batch, dev, seed=rank * 7 + i, size=a.size) for i in range(2)]
torch.cuda.synchronize()
t0 = time.time()
for step in range(a.steps):
    loss = model(**batches[step % 2])
    loss.backward()
    model.step()
    if step == 0:
        runtime.mark_first_step()
    if rank == 0 and step % 10 == 0:
        print(f"step {step} loss {loss.item():.4f}", flush=True)
torch.cuda.synchronize()
if rank == 0:
    print(f"done: {a.steps * a.batch * world / (time.time() - t0):.1f} img/s over {world} GPUs", flush=True)
if world > 1:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
