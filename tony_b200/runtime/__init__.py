"""The PyTorch runtime package handed to launched jobs.

``init()`` turns the per-task environment contract the launcher exports into an
initialised ``torch.distributed`` process group + CUDA device.  It accepts both the
reference's names (``INIT_METHOD`` / ``RANK`` / ``WORLD``:
tony-core/src/main/java/com/linkedin/tony/runtime/PyTorchRuntime.java:45-56,
Constants.java:56-60) and torch's ``env://`` names (``MASTER_ADDR``/``MASTER_PORT``/
``WORLD_SIZE``/``LOCAL_RANK``), which torchrun and our launcher also set.

``Trainer`` is the user-facing training step: it owns the flat parameter store and the
fused collective+optimizer, copies each batch from pinned host memory, runs
forward/backward on the sm_100a kernels and returns the loss as a Python float.
"""
from __future__ import annotations

import os
import time
from typing import Callable, Dict, Optional

_T0 = time.time()
_state = {"initialised": False, "first_step_marked": False}


def env_rank_world():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", os.environ.get("WORLD", "1")))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def ps_roles():
    """``PsRoles`` of a job launched with dedicated parameter servers (``tony.ps.instances > 0`` and
    ``runtime.init(include_ps=True)``), else ``None``.  The launcher exports the all-task numbering as
    ``TONY_B200_ALL_RANK`` / ``TONY_B200_ALL_WORLD`` / ``TONY_B200_PS_RANKS``."""
    ps = os.environ.get("TONY_B200_PS_RANKS", "")
    if not ps or not _state.get("include_ps"):
        return None
    from ..parallel.flat import PsRoles

    world = int(os.environ["TONY_B200_ALL_WORLD"])
    owners = sorted(int(x) for x in ps.split(",") if x != "")
    return PsRoles(owners=owners, workers=[r for r in range(world) if r not in owners])


def init(backend: Optional[str] = None, device: Optional[str] = None, include_ps: bool = False):
    """Initialise torch.distributed from the TonY env contract; returns (rank, world, device).
    ``include_ps``: number the ``ps`` tasks into the same process group (dedicated parameter servers on
    their own GPUs; see ``ps_roles``) instead of the tracked-tasks-only ranks of the reference."""
    import torch
    import torch.distributed as dist

    if include_ps and "TONY_B200_ALL_RANK" in os.environ:
        os.environ["RANK"] = os.environ["TONY_B200_ALL_RANK"]
        os.environ["WORLD_SIZE"] = os.environ["TONY_B200_ALL_WORLD"]
        _state["include_ps"] = True
    rank, world, local = env_rank_world()
    use_cuda = torch.cuda.is_available() and device != "cpu"
    if use_cuda:
        visible = torch.cuda.device_count()
        torch.cuda.set_device(local % max(visible, 1))
        dev = torch.device("cuda", local % max(visible, 1))
    else:
        dev = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        backend = backend or ("nccl" if use_cuda else "gloo")
        init_method = os.environ.get("INIT_METHOD")
        if "MASTER_ADDR" in os.environ and "MASTER_PORT" in os.environ:
            init_method = "env://"
            os.environ.setdefault("WORLD_SIZE", str(world))
        kwargs = {"device_id": dev} if (use_cuda and backend == "nccl") else {}
        dist.init_process_group(backend=backend, init_method=init_method, rank=rank,
                                world_size=world, **kwargs)
    _state["initialised"] = True
    return rank, world, dev


def mark_first_step(extra: Optional[Dict[str, float]] = None) -> None:
    """Tell the launcher the first training step finished (submit-to-first-step latency).

    Writes ``<TONY_B200_MARK_DIR>/first_step.<JOB_NAME>.<TASK_INDEX>`` with the wall time;
    the TaskExecutor forwards it as a metric.  A no-op outside a launched task.
    """
    if _state["first_step_marked"]:
        return
    _state["first_step_marked"] = True
    d = os.environ.get("TONY_B200_MARK_DIR")
    if not d:
        return
    name = f"first_step.{os.environ.get('JOB_NAME', 'task')}.{os.environ.get('TASK_INDEX', '0')}"
    try:
        os.makedirs(d, exist_ok=True)
        tmp = os.path.join(d, name + ".tmp")
        with open(tmp, "w") as f:
            f.write(f"{time.time():.6f}\n")
            for k, v in (extra or {}).items():
                f.write(f"{k}={v}\n")
        os.replace(tmp, os.path.join(d, name))
    except OSError:
        pass


def report_metrics(metrics: Dict[str, float]) -> None:
    """Hand training metrics to the launcher: the TaskExecutor's monitor forwards them over the
    updateMetrics RPC and the ApplicationMaster writes them into this task's ``TaskFinished`` history
    event (e.g. ``STEP_TIME_MS``, ``SAMPLES_PER_SEC``, ``NVLINK_BYTES_PER_STEP``).  A no-op outside a
    launched task; later calls overwrite earlier values of the same name."""
    d = os.environ.get("TONY_B200_MARK_DIR")
    if not d:
        return
    name = f"metrics.{os.environ.get('JOB_NAME', 'task')}.{os.environ.get('TASK_INDEX', '0')}"
    try:
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, name)
        cur: Dict[str, float] = {}
        try:
            with open(path) as f:
                for line in f:
                    k, _, v = line.strip().partition("=")
                    if k and v:
                        cur[k] = float(v)
        except (OSError, ValueError):
            pass
        cur.update({k: float(v) for k, v in metrics.items()})
        tmp = path + ".tmp"
        with open(tmp, "w") as f:
            for k, v in cur.items():
                f.write(f"{k}={v}\n")
        os.replace(tmp, path)
    except OSError:
        pass


class LossFuture:
    """Loss of one :meth:`Trainer.step_async` call (see there)."""

    def __init__(self, trainer, host_buf, event, step_no):
        self._trainer, self._host, self._event, self._step = trainer, host_buf, event, step_no

    def done(self) -> bool:
        return self._event.query()

    def result(self) -> float:
        self._event.synchronize()
        value = float(self._host[0])
        t = self._trainer
        if t.store.world > 1:
            t.store.comm.check_status()
        mark_first_step()
        return value


class Trainer:
    """model + fused optimizer + host<->device batch plumbing, one call per step."""

    def __init__(self, model, *, lr: float = 1e-4, optimizer: str = "adamw", weight_decay: float = 0.01,
                 betas=(0.9, 0.999), eps: float = 1e-8, momentum: float = 0.9, bucket_mb: float = 64.0,
                 overlap: bool = True, comm_ctas: int = 0, device=None, group=None,
                 broadcast: bool = True, generic_hooks: bool = False, ps_sharded=None,
                 cuda_graph: Optional[bool] = None, roles=None):
        import torch

        from ..parallel.flat import FlatParamStore, FusedOptimizer

        self.torch = torch
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.model = model.to(self.device)
        # dedicated parameter servers (PsRoles): ps ranks own the shards and run no forward / backward;
        # every rank launches the bucket kernels in bucket order at step() so the flag epochs line up
        self.roles = roles
        self.is_worker = roles is None or roles.is_worker(self._rank())
        if roles is not None:
            overlap = False
        self.store = FlatParamStore(model.named_parameters(), self.device, group=group,
                                    bucket_mb=bucket_mb, no_decay=("bias", "ln", "LayerNorm", "norm", "bn"),
                                    ps_sharded=ps_sharded, roles=roles)
        for p, uses in getattr(model, "shared_params", {}).items():
            self.store.mark_shared(p, uses)
        if generic_hooks:
            self.store.register_autograd_hooks()
        self.opt = FusedOptimizer(self.store, optimizer=optimizer, lr=lr, betas=betas, eps=eps,
                                  weight_decay=weight_decay, momentum=momentum, overlap=overlap,
                                  comm_ctas=comm_ctas, broadcast=broadcast)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.steps = 0
        # modules that own state outside the flat parameter store (row-sharded embedding bags)
        self._step_modules = [m for m in self.model.modules() if hasattr(m, "tony_step")]
        for m in self._step_modules:
            m.tony_configure(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        # CUDA-graph mode: forward + backward (about 600 launches for BERT-large) are captured once
        # and replayed; the fused reduce+optimizer launches stay outside the graph because their
        # arguments change every step (bias corrections, flag epochs).  Needs a model whose
        # gradients are all produced by this package's ops (no autograd hooks) and static shapes.
        if cuda_graph is None:
            cuda_graph = os.environ.get("TONY_B200_CUDA_GRAPH", "0") == "1"
        self.cuda_graph = bool(cuda_graph) and not generic_hooks and not ps_sharded and not self._step_modules
        self._graph = None
        self._graph_kernels = 0
        self._static_in: Optional[Dict[str, "object"]] = None
        self._static_loss = None
        self._eager_before_capture = 2

    @staticmethod
    def _rank() -> int:
        import torch.distributed as dist

        return dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0

    def serve_step(self):
        """One step of a dedicated parameter server: no forward / backward, only its side of every
        bucket's fused reduce + update (its own gradients are zero) and of the sharded tables."""
        for s in self.store.slots:
            s.written = 0
        self.opt.step()
        for m in self._step_modules:
            m.tony_step()
        self.steps += 1
        return self.torch.zeros((), device=self.device)

    def to_device(self, batch_cpu: Dict[str, "object"]):
        """Async H2D of a (pinned) host batch into FRESH device tensors (copy stream; the compute stream
        waits for the copy's event).  ``step`` / ``step_async`` use the resident ring of
        :meth:`_stage_inputs` instead."""
        torch = self.torch
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.copy_stream):
            out = {k: v.to(self.device, non_blocking=True) for k, v in batch_cpu.items()}
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        cur.wait_event(ev)
        for v in out.values():
            v.record_stream(cur)
        return out

    def _stage_inputs(self, batch_cpu: Dict[str, "object"]):
        """Async H2D of a (pinned) host batch into one of three resident device input sets, on the copy
        stream: the transfer of step i+1's inputs runs under step i's kernels (the host is a step ahead
        of the GPU).  The compute stream waits for the copy's event; the copy waits for the step that
        last consumed the set.  No per-step device allocation (a fresh 77 MB image tensor per step made
        the caching allocator's cross-stream bookkeeping the end-to-end bottleneck for ResNet-50)."""
        torch = self.torch
        cur = torch.cuda.current_stream(self.device)
        if not hasattr(self, "_in_sets"):
            self._in_sets, self._in_done, self._in_next = [None, None, None], [None, None, None], 0
        slot = self._in_next
        self._in_next = (slot + 1) % 3
        bufs = self._in_sets[slot]
        if bufs is None or any(k not in bufs or bufs[k].shape != v.shape or bufs[k].dtype != v.dtype
                               for k, v in batch_cpu.items()):
            # a block handed out by the caching allocator may still be in use by kernels already queued
            # on the compute stream: allocate from the copy stream's pool and order after that work
            self.copy_stream.wait_stream(cur)
            with torch.cuda.stream(self.copy_stream):
                bufs = {k: torch.empty(v.shape, dtype=v.dtype, device=self.device) for k, v in batch_cpu.items()}
            self._in_sets[slot] = bufs
            self._in_done[slot] = None
        if self._in_done[slot] is not None:
            self.copy_stream.wait_event(self._in_done[slot])
        with torch.cuda.stream(self.copy_stream):
            for k, v in batch_cpu.items():
                bufs[k].copy_(v, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        cur.wait_event(ev)
        self._in_live = (slot, bufs)
        return bufs

    def _mark_inputs_consumed(self, batch_dev) -> None:
        live = getattr(self, "_in_live", None)
        if live is not None and live[1] is batch_dev:
            ev = self.torch.cuda.Event()
            ev.record(self.torch.cuda.current_stream(self.device))
            self._in_done[live[0]] = ev

    def _capture(self, batch_dev) -> None:
        torch = self.torch
        from ..ops import _lib

        self._static_in = {k: v.clone() for k, v in batch_dev.items()}
        self.store._ready_cb = None          # no optimizer launches from inside the captured backward
        self.store.reset_step()
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        before = _lib.launch_count()
        with torch.cuda.graph(graph):
            loss = self.model(**self._static_in)
            loss.backward()
        self._graph_kernels = _lib.launch_count() - before
        for s in self.store.slots:           # parameters the model never touches keep a zero gradient
            if s.written == 0:
                s.grad_view.zero_()
        self._graph, self._static_loss = graph, loss.detach()

    def _graph_step(self, batch_dev):
        from ..ops import _lib

        if self._graph is None:
            self._capture(batch_dev)
        for k, v in batch_dev.items():
            if v is not self._static_in[k]:
                self._static_in[k].copy_(v, non_blocking=True)
        self._graph.replay()
        _lib.lib().tony_launch_count_add(self._graph_kernels)
        self.opt.step(grads_complete=True)
        self._mark_inputs_consumed(batch_dev)
        self.steps += 1
        return self._static_loss

    def step_device(self, batch_dev) -> "object":
        """One optimisation step on device-resident inputs; returns the loss tensor (no sync).
        In CUDA-graph mode the returned tensor is overwritten by the next step."""
        if not self.is_worker:
            return self.serve_step()
        if self.cuda_graph and self.steps >= self._eager_before_capture:
            return self._graph_step(batch_dev)
        nvtx = self.torch.cuda.nvtx
        nvtx.range_push("tony.forward")
        loss = self.model(**batch_dev)
        nvtx.range_pop()
        nvtx.range_push("tony.backward+fused_reduce_optim")
        loss.backward()
        nvtx.range_pop()
        nvtx.range_push("tony.optimizer_join")
        self.opt.step()
        for m in self._step_modules:
            m.tony_step()
        nvtx.range_pop()
        self._mark_inputs_consumed(batch_dev)
        self.steps += 1
        # detached: a caller holding the loss must not keep this step's autograd graph (and the
        # AccumulateGrad nodes bound to this stream) alive into a later CUDA-graph capture
        return loss.detach()

    def step(self, batch_cpu) -> float:
        """The end-to-end user call: pinned host batch in, Python float loss out."""
        loss = self.step_device(self._stage_inputs(batch_cpu))
        value = float(loss.item())  # D2H read of the step's result; also the host sync point
        if self.store.world > 1:
            # GPU-side hang detection: a fused collective that gave up waiting for a dead peer
            # set the status word; fail the task so the launcher's failure policy takes over
            self.store.comm.check_status()
        mark_first_step()       # no-op after this process's first completed step (also after a resume)
        return value

    def step_async(self, batch_cpu) -> "LossFuture":
        """Pipelined form of :meth:`step`: enqueue the H2D copy, the whole step and a 4-byte D2H copy
        of the loss into pinned memory, and return at once.  ``future.result()`` blocks until THAT
        step has finished.  Reading step i's loss after step i+1 has been enqueued keeps the GPU
        busy across the host's per-step Python work (about 1.5 ms per BERT-large step otherwise)."""
        torch = self.torch
        if self._graph is not None:
            # H2D straight into the captured graph's input tensors
            for k, v in batch_cpu.items():
                self._static_in[k].copy_(v, non_blocking=True)
            loss = self._graph_step(self._static_in)
        else:
            loss = self.step_device(self._stage_inputs(batch_cpu))
        if not hasattr(self, "_loss_ring"):
            self._loss_ring = [torch.empty(1, dtype=torch.float32, pin_memory=True) for _ in range(4)]
        host = self._loss_ring[self.steps % len(self._loss_ring)]
        src = loss.detach().reshape(1)
        host.copy_(src if src.dtype == torch.float32 else src.float(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        return LossFuture(self, host, ev, self.steps)

    def check(self):
        self.store.comm.check_status()
