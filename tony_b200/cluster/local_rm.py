"""Local resource manager + container launcher for a single 8xB200 box.

Plays the parts YARN's RM and NM play for the reference (AMRMClientAsync/NMClientAsync use
in tony-core/src/main/java/com/linkedin/tony/ApplicationMaster.java:443-527,1096-1115 and the
in-process MiniYARNCluster of tony-mini/.../MiniCluster.java:43-60), re-designed for one
NVSwitch node:

* inventory = GPUs (one task per GPU it asks for), memory, vcores;
* GPUs are leased across concurrently running applications with ``flock`` on
  ``<cluster-dir>/gpu<N>.lock`` — no daemon needed;
* a container is a process group (``start_new_session``), its exit is observed by a reaper
  thread the instant it happens (no 1 s RM heartbeat), a container stopped by the AM
  reports the YARN constant KILLED_BY_APPMASTER (-105);
* each container gets ``CUDA_VISIBLE_DEVICES`` for its GPUs and is CPU-pinned near them.
"""
from __future__ import annotations

import fcntl
import os
import shutil
import subprocess
import threading
from collections import deque
from dataclasses import dataclass, field
from typing import Callable, Deque, Dict, List, Optional, Tuple

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration, parse_memory_string
from ..utils.utils import JobContainerRequest, kill_process_group
from . import topology


@dataclass
class Container:
    container_id: str
    priority: int
    memory_mb: int
    vcores: int
    gpu_ids: List[int] = field(default_factory=list)
    host: str = "127.0.0.1"
    proc: Optional[subprocess.Popen] = None
    work_dir: Optional[str] = None
    killed_by_am: bool = False
    exit_status: Optional[int] = None
    _locks: List[object] = field(default_factory=list)


class NodeInventory:
    def __init__(self, conf: TonyConfiguration, cluster_dir: str):
        n = conf.get_int(K.B200_NUM_GPUS, -1)
        self.gpus: List[int] = list(range(n)) if n >= 0 else topology.discover_gpu_indices()
        mem = conf.get(K.B200_CLUSTER_MEMORY, "") or ""
        self.memory_mb = parse_memory_string(mem) if mem.strip() else topology.physical_memory_mb()
        vc = conf.get_int(K.B200_CLUSTER_VCORES, -1)
        self.vcores = vc if vc > 0 else max(1, (os.cpu_count() or 1)) * 16  # oversubscribable
        self.cluster_dir = cluster_dir
        os.makedirs(cluster_dir, exist_ok=True)


class LocalResourceManager:
    def __init__(self, conf: TonyConfiguration, app_id: str, cluster_dir: str,
                 on_allocated: Callable[[Container], None],
                 on_completed: Callable[[Container, int, str], None]):
        self.conf = conf
        self.app_id = app_id
        self.inv = NodeInventory(conf, cluster_dir)
        self.on_allocated = on_allocated
        self.on_completed = on_completed
        self._lock = threading.Condition()
        self._asks: Deque[Tuple[JobContainerRequest, int]] = deque()  # (request, attempt)
        self._free_mem = self.inv.memory_mb
        self._free_vcores = self.inv.vcores
        self._containers: Dict[str, Container] = {}
        self._seq = 0
        self._stop = False
        self.attempt = 1
        self.pin_numa = conf.get_bool(K.B200_PIN_NUMA, True)
        self._alloc_thread = threading.Thread(target=self._allocate_loop, name="rm-alloc", daemon=True)
        self._alloc_thread.start()

    # -- asks ---------------------------------------------------------------------------
    def validate(self, req: JobContainerRequest) -> Optional[str]:
        if req.gpus > len(self.inv.gpus):
            return (f"job '{req.job_name}' asks for {req.gpus} GPUs per task but the node has "
                    f"{len(self.inv.gpus)}")
        if req.memory_mb > self.inv.memory_mb:
            return f"job '{req.job_name}' asks for {req.memory_mb} MB but the node has {self.inv.memory_mb} MB"
        return None

    def add_container_request(self, req: JobContainerRequest) -> None:
        with self._lock:
            for _ in range(req.num_instances):
                self._asks.append((req, self.attempt))
            self._lock.notify_all()

    def _try_lease_gpus(self, n: int) -> Optional[Tuple[List[int], List[object]]]:
        if n == 0:
            return [], []
        got, locks = [], []
        for g in self.inv.gpus:
            path = os.path.join(self.inv.cluster_dir, f"gpu{g}.lock")
            f = open(path, "a+")
            try:
                fcntl.flock(f, fcntl.LOCK_EX | fcntl.LOCK_NB)
            except OSError:
                f.close()
                continue
            got.append(g)
            locks.append(f)
            if len(got) == n:
                return got, locks
        for f in locks:
            fcntl.flock(f, fcntl.LOCK_UN)
            f.close()
        return None

    def _allocate_loop(self) -> None:
        while True:
            with self._lock:
                while not self._stop and not self._asks:
                    self._lock.wait()
                if self._stop:
                    return
                req, attempt = self._asks[0]
                if attempt != self.attempt:  # stale ask from a previous session
                    self._asks.popleft()
                    continue
                ok = req.memory_mb <= self._free_mem and req.vcores <= self._free_vcores
                lease = self._try_lease_gpus(req.gpus) if ok else None
                if not ok or lease is None:
                    self._lock.wait(timeout=0.2)  # resources free up on container exit
                    continue
                self._asks.popleft()
                self._free_mem -= req.memory_mb
                self._free_vcores -= req.vcores
                self._seq += 1
                ts, seq = self.app_id.split("_")[1:3]
                cid = f"container_{ts}_{seq}_{self.attempt:02d}_{self._seq:06d}"
                c = Container(cid, req.priority, req.memory_mb, req.vcores, lease[0])
                c._locks = lease[1]
                self._containers[cid] = c
            self.on_allocated(c)

    # -- containers ---------------------------------------------------------------------
    def start_container(self, c: Container, command: List[str], env: Dict[str, str], work_dir: str) -> None:
        os.makedirs(work_dir, exist_ok=True)
        c.work_dir = work_dir
        full_env = dict(os.environ)
        full_env.update(env)
        if c.gpu_ids:
            full_env["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in c.gpu_ids)
            full_env["TONY_B200_GPU_IDS"] = full_env["CUDA_VISIBLE_DEVICES"]
        cmd = list(command)
        if self.pin_numa and c.gpu_ids and shutil.which("taskset"):
            cpus = topology.cpus_near_gpus(c.gpu_ids)
            if cpus:
                cmd = ["taskset", "-c", cpus] + cmd
        out = open(os.path.join(work_dir, "stdout"), "ab")
        err = open(os.path.join(work_dir, "stderr"), "ab")
        c.proc = subprocess.Popen(cmd, env=full_env, cwd=work_dir, stdout=out, stderr=err,
                                  start_new_session=True)
        out.close()
        err.close()
        threading.Thread(target=self._reap, args=(c,), name=f"reap-{c.container_id}", daemon=True).start()

    def _reap(self, c: Container) -> None:
        code = c.proc.wait()
        # children the task left behind must not outlive the container
        try:
            os.killpg(c.proc.pid, 9)
        except (ProcessLookupError, PermissionError):
            pass
        self._release(c)
        if c.killed_by_am:
            status, diag = K.KILLED_BY_APPMASTER, "Container killed by the ApplicationMaster."
        elif code < 0:
            status, diag = 128 - code, f"Container killed by signal {-code}."
        else:
            status, diag = code, ("" if code == 0 else f"Container exited with status {code}.")
        c.exit_status = status
        self.on_completed(c, status, diag)

    def _release(self, c: Container) -> None:
        with self._lock:
            for f in c._locks:
                try:
                    fcntl.flock(f, fcntl.LOCK_UN)
                    f.close()
                except OSError:
                    pass
            c._locks = []
            self._free_mem += c.memory_mb
            self._free_vcores += c.vcores
            self._lock.notify_all()

    def release_unlaunched(self, c: Container) -> None:
        """Give back a container that was allocated but never started."""
        self._containers.pop(c.container_id, None)
        self._release(c)

    def stop_container(self, container_id: str) -> None:
        c = self._containers.get(container_id)
        if c is None or c.proc is None or c.proc.poll() is not None:
            return
        c.killed_by_am = True
        kill_process_group(c.proc)

    def stop_all(self, new_attempt: bool = False) -> None:
        with self._lock:
            self._asks.clear()
            if new_attempt:
                self.attempt += 1
            live = [c for c in self._containers.values() if c.proc is not None and c.proc.poll() is None]
        threads = []
        for c in live:
            c.killed_by_am = True
            t = threading.Thread(target=kill_process_group, args=(c.proc,), daemon=True)
            t.start()
            threads.append(t)
        for t in threads:
            t.join(timeout=10)

    def shutdown(self) -> None:
        self.stop_all()
        with self._lock:
            self._stop = True
            self._lock.notify_all()

    def container(self, cid: str) -> Optional[Container]:
        return self._containers.get(cid)
