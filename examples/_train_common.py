"""Shared driver for the GPU examples: build a model, train N steps through runtime.Trainer,
checkpoint / resume when the launcher retried the session, print throughput."""
import os
import sys
import time

sys.path.insert(0, os.environ.get("TONY_REPO", os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")))

import torch

from tony_b200 import runtime
from tony_b200.runtime import checkpoint


def train(build_model, make_batch, *, steps, lr, optimizer="adamw", ckpt_dir=None, ckpt_every=0,
          trainer_kw=None, samples_per_step=1, crash_at=None):
    rank, world, dev = runtime.init()
    torch.manual_seed(1234)
    model = build_model(dev)
    kw = trainer_kw(model) if callable(trainer_kw) else (trainer_kw or {})
    trainer = runtime.Trainer(model, lr=lr, optimizer=optimizer, device=dev, **kw)
    start = 0
    if ckpt_dir and checkpoint.should_resume():
        resumed = checkpoint.load_latest(ckpt_dir, trainer)
        if resumed is not None:
            start = resumed
            print(f"rank {rank}: resumed from step {resumed} (NUM_AM_RETRIES={os.environ.get('NUM_AM_RETRIES')})", flush=True)
    batches = [make_batch(100 + rank * 17 + i) for i in range(4)]
    t0 = time.time()
    for step in range(start, steps):
        loss = trainer.step(batches[step % 4])
        if rank == 0 and (step % 10 == 0 or step == steps - 1):
            print(f"step {step} loss {loss:.4f}", flush=True)
        if ckpt_dir and ckpt_every and (step + 1) % ckpt_every == 0:
            checkpoint.save(ckpt_dir, trainer)
        if crash_at is not None and step + 1 == crash_at and not checkpoint.should_resume() \
                and os.environ.get("IS_CHIEF") == "true":
            print("chief: simulating a crash", flush=True)
            os._exit(17)
    trainer.check()
    dt = time.time() - t0
    n_steps = max(1, steps - start)
    grad_bytes = 2 * trainer.store.numel
    runtime.report_metrics({"STEP_TIME_MS": 1000.0 * dt / n_steps,
                            "SAMPLES_PER_SEC": n_steps * samples_per_step / dt,
                            # bytes this rank moves over NVLink per step: it pulls (world-1)/world of its
                            # gradient shard set and publishes its weight shard to the other ranks
                            "NVLINK_BYTES_PER_STEP": 2.0 * grad_bytes * (world - 1) / max(world, 1)})
    if rank == 0:
        print(f"done: {(steps - start) * samples_per_step * world / dt:.1f} samples/s over {world} GPUs", flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
