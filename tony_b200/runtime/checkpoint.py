"""Sharded checkpoint / auto-resume for whole-session restarts.

The reference restarts a failed attempt by re-launching every task with ``SESSION_ID`` bumped
and ``NUM_AM_RETRIES`` exported (tony-core/src/main/java/com/linkedin/tony/ApplicationMaster.java
:407-421,612-628); resuming is left to the user script (tf.train.Checkpoint / ModelCheckpoint in
tony-examples/horovod-on-tony/tensorflow2_mnist.py:62-63,100-101).  Here it is part of the
runtime: every rank writes its optimizer shard (fp32 master + moments), rank 0 writes the bf16
parameters and step counter, files appear atomically (tmp + rename), and ``load_latest`` picks
the newest step for which every shard of the current world size exists.
"""
from __future__ import annotations

import glob
import json
import os
import re
from typing import Optional

import torch


def _atomic_save(obj, path: str) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(obj, tmp)
    os.replace(tmp, path)


def save(directory: str, trainer, step: Optional[int] = None) -> str:
    os.makedirs(directory, exist_ok=True)
    step = trainer.steps if step is None else step
    st = trainer.store
    base = os.path.join(directory, f"step{step:09d}")
    opt_state = trainer.opt.state_dict()
    # modules with state outside the flat store (row-sharded embedding bags): each rank saves its shard
    extra = [m.tony_state_dict() for m in getattr(trainer, "_step_modules", []) if hasattr(m, "tony_state_dict")]
    if extra:
        opt_state = dict(opt_state, _step_modules=extra)
    _atomic_save(opt_state, f"{base}.opt.rank{st.rank:03d}of{st.world:03d}.pt")
    if st.rank == 0:
        # module buffers (BatchNorm running statistics) are model state outside the flat store
        buffers = {k: v.detach().clone() for k, v in trainer.model.named_buffers()}
        _atomic_save({"param_buf": st.param_buf, "numel": st.numel, "step": step, "buffers": buffers},
                     f"{base}.params.pt")
        tmp = f"{base}.meta.json.tmp"
        with open(tmp, "w") as f:
            json.dump({"step": step, "world": st.world, "numel": st.numel}, f)
        os.replace(tmp, f"{base}.meta.json")
    return base


def latest_complete(directory: str, world: int) -> Optional[int]:
    steps = []
    for meta in glob.glob(os.path.join(directory, "step*.meta.json")):
        m = re.search(r"step(\d+)\.meta\.json$", meta)
        if not m:
            continue
        step = int(m.group(1))
        base = os.path.join(directory, f"step{step:09d}")
        ok = os.path.exists(f"{base}.params.pt") and all(
            os.path.exists(f"{base}.opt.rank{r:03d}of{world:03d}.pt") for r in range(world))
        if ok:
            steps.append(step)
    return max(steps) if steps else None


def load_latest(directory: str, trainer) -> Optional[int]:
    """Restore parameters + this rank's optimizer shard; returns the step, or None if none."""
    st = trainer.store
    step = latest_complete(directory, st.world) if os.path.isdir(directory) else None
    if step is None:
        return None
    base = os.path.join(directory, f"step{step:09d}")
    params = torch.load(f"{base}.params.pt", map_location=st.device)
    if params["numel"] != st.numel:
        raise ValueError("checkpoint was written for a different model layout")
    st.param_buf.copy_(params["param_buf"])
    own = dict(trainer.model.named_buffers())
    with torch.no_grad():
        for k, v in params.get("buffers", {}).items():
            if k in own and own[k].shape == v.shape:
                own[k].copy_(v.to(own[k].dtype))
    opt_state = torch.load(f"{base}.opt.rank{st.rank:03d}of{st.world:03d}.pt", map_location=st.device)
    extra = opt_state.pop("_step_modules", [])
    trainer.opt.load_state_dict(opt_state)
    # rank 0's bf16 copy is exact only where weights are broadcast; a parameter-server bucket keeps
    # each rank's own shard current and nothing else, so re-derive this rank's shard of every bucket
    # from its fp32 master (bit-identical to what the optimizer kernel last stored)
    with torch.no_grad():
        for b in st.buckets:
            lo, chunk = st.shard_range(b)
            st.param_buf[lo: lo + chunk].copy_(trainer.opt.master[b.state_off: b.state_off + chunk])
    mods = [m for m in getattr(trainer, "_step_modules", []) if hasattr(m, "tony_load_state_dict")]
    if len(extra) != len(mods):
        raise ValueError("checkpoint was written for a different set of sharded modules")
    for m, sd in zip(mods, extra):
        m.tony_load_state_dict(sd)
    trainer.steps = step
    torch.cuda.synchronize()
    return step


def should_resume() -> bool:
    """True inside a retried session (the launcher exports NUM_AM_RETRIES after a reset)."""
    return int(os.environ.get("NUM_AM_RETRIES", "0") or 0) > 0
