"""horovod-on-tony BERT-large pre-training (BASELINE config #3), written against the Horovod-shaped API.

Launched with ``tony.application.framework=horovod``; the script has the skeleton of the reference's Horovod jobs
(tony-examples/horovod-on-tony/tensorflow2_mnist.py:28-36,73-76,85-86): ``hvd.init()``, pin the GPU by
``hvd.local_rank()``, ``hvd.broadcast_parameters`` from rank 0, ``hvd.DistributedOptimizer`` around the model,
``hvd.allreduce`` of the reported loss.  "Tensor fusion" is the flat gradient bucket: the wgrad / LayerNorm /
bias-gradient kernels write straight into it and one NVLink kernel per 64 MB bucket does all-reduce + AdamW +
weight broadcast, launched from inside backward."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")))

import torch

import tony_b200.parallel.hvd as hvd
from tony_b200 import runtime
from tony_b200.models import bert

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--seq", type=int, default=512)
ap.add_argument("--tiny", action="store_true")
a = ap.parse_args()

hvd.init()
torch.cuda.set_device(hvd.local_rank() % max(torch.cuda.device_count(), 1))
dev = torch.device("cuda", torch.cuda.current_device())
cfg = bert.BertConfig.tiny() if a.tiny else bert.BertConfig.large()
seq = min(a.seq, cfg.max_positions)
torch.manual_seed(1234 + hvd.rank())                     # ranks start different; rank 0's weights win below
model = bert.BertForPreTraining(cfg, device=dev)
hvd.broadcast_parameters(model.state_dict(), root_rank=0)
opt = hvd.DistributedOptimizer(model, optimizer="adamw", lr=1e-4)
batches = [{k: v.to(dev) for k, v in bert.synthetic_batch(cfg, a.batch, seq, seed=100 + hvd.rank() * 17 + i).items()}
           for i in range(4)]
t0 = time.time()
for step in range(a.steps):
    loss = model(**batches[step % 4])
    loss.backward()
    opt.step()                                           # fused all-reduce + AdamW, overlapped with backward
    if step == 0:
        runtime.mark_first_step()
    if step % 10 == 0 or step == a.steps - 1:
        mean = hvd.allreduce(loss.detach(), average=True).item()      # fp32 metric averaging
        if hvd.rank() == 0:
            print(f"step {step} mean loss over {hvd.size()} ranks {mean:.4f}", flush=True)
opt.synchronize()
torch.cuda.synchronize()
opt.check()
dt = time.time() - t0
grad_bytes = 2 * opt.store.numel
world = hvd.size()
runtime.report_metrics({"STEP_TIME_MS": 1000.0 * dt / a.steps, "SAMPLES_PER_SEC": a.steps * a.batch / dt,
                        "NVLINK_BYTES_PER_STEP": 2.0 * grad_bytes * (world - 1) / max(world, 1)})
if hvd.rank() == 0:
    print(f"done: {a.steps * a.batch * world / dt:.1f} samples/s over {world} GPUs", flush=True)
if world > 1:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
