"""ps/worker wide-and-deep recommender (BASELINE config #4), in either layout:

* dedicated servers — ``tony.ps.instances=2`` + ``tony.worker.instances=6``, one GPU each (``tony.xml``):
  the ``ps`` tasks own every shard (tower matrices, fp32 master + Adam moments, embedding rows) and run
  only their side of the fused reduce + update kernels; the ``worker`` tasks run forward / backward, push
  gradients through the same kernels and pull weights and rows from the servers over NVLink.  The ``ps``
  tasks are untracked: when the workers finish, the ApplicationMaster tears them down (the lifecycle of
  tony-core/src/main/resources/tony-default.xml:160-170,204-223 and the ``server.join()`` of
  tony-examples/mnist-tensorflow/mnist_distributed.py:206-241);
* symmetric — ``tony-symmetric.xml`` (8 workers, no ps): every rank is both a shard server and a worker.
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")))

import torch

from tony_b200 import runtime
from tony_b200.models import widedeep
from tony_b200.parallel.ps import ps_parameters

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--batch", type=int, default=8192)
ap.add_argument("--tiny", action="store_true")
ap.add_argument("--ps_ranks", default="", help="outside the launcher (torchrun): comma list of ranks that act "
                                               "as dedicated parameter servers, e.g. 6,7")
a = ap.parse_args()
if a.ps_ranks and "TONY_B200_ALL_RANK" not in os.environ:
    os.environ["TONY_B200_ALL_RANK"] = os.environ.get("RANK", "0")
    os.environ["TONY_B200_ALL_WORLD"] = os.environ.get("WORLD_SIZE", "1")
    os.environ["TONY_B200_PS_RANKS"] = a.ps_ranks
    if int(os.environ["TONY_B200_ALL_RANK"]) in [int(x) for x in a.ps_ranks.split(",")]:
        os.environ["JOB_NAME"] = "ps-torchrun"       # not "ps": no server.join() wait without an AM to end it

rank, world, dev = runtime.init(include_ps=True)
roles = runtime.ps_roles()                      # None in the symmetric layout
job = os.environ.get("JOB_NAME", "worker")
torch.manual_seed(1234)
cfg = widedeep.WideDeepConfig.tiny() if a.tiny else widedeep.WideDeepConfig()
model = widedeep.WideDeep(cfg, device=dev, roles=roles)
trainer = runtime.Trainer(model, lr=1e-3, weight_decay=0.0, device=dev, ps_sharded=ps_parameters(model), roles=roles)
n_workers = len(roles.workers) if roles else world
batches = [widedeep.synthetic_batch(cfg, a.batch, pin=True, seed=100 + rank * 17 + i) for i in range(4)]
torch.cuda.synchronize()
t0 = time.time()
first = last = None
for step in range(a.steps):
    if trainer.is_worker:
        loss = trainer.step(batches[step % 4])
        first = loss if first is None else first
        last = loss
        if rank == 0 and (step % 10 == 0 or step == a.steps - 1):
            print(f"step {step} loss {loss:.4f}", flush=True)
    else:
        trainer.serve_step()                    # a parameter server: reduce + update + barriers only
torch.cuda.synchronize()
trainer.check()
dt = time.time() - t0
if world > 1:
    torch.distributed.barrier()
if rank == 0:
    layout = f"{len(roles.owners)} ps + {n_workers} workers" if roles else f"{world} symmetric ranks"
    print(f"done ({layout}): {a.steps * a.batch * n_workers / dt:.1f} samples/s, loss {first:.4f} -> {last:.4f}",
          flush=True)
if world > 1:
    torch.distributed.destroy_process_group()
if roles is not None and job == "ps":
    # like a TensorFlow ps task's server.join(): stay up until the ApplicationMaster tears the job down
    print(f"ps task {os.environ.get('TASK_INDEX')}: serving done, waiting for teardown", flush=True)
    while True:
        time.sleep(1)
