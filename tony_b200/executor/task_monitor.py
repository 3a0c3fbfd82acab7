"""Per-task resource monitor: process-tree RSS + GPU gauges, pushed to the AM's metrics RPC.

Same eight gauges as the reference (tony-core/src/main/java/com/linkedin/tony/TaskMonitor.java
:34-46,101-170): MAX/AVG of MEMORY_BYTES, GPU_UTILIZATION, GPU_FB_MEMORY_USAGE,
GPU_MAIN_MEMORY_USAGE, running average ``avg += (x - avg) / n`` (:173-177), GPU sampling
disabled after 10 consecutive nvidia-smi errors (:160-169).  Extended with the first-step
marker so the launcher can report submit-to-first-step latency, and with the training metrics the
user process reports through ``tony_b200.runtime.report_metrics`` (STEP_TIME_MS, SAMPLES_PER_SEC,
NVLINK_BYTES_PER_STEP, ...): they travel the same updateMetrics RPC and end up in
``TaskFinished.metrics`` (tony-core/src/main/avro/TaskFinished.avsc; emission at
tony-core/src/main/java/com/linkedin/tony/ApplicationMaster.java:1246-1258).
"""
from __future__ import annotations

import os
import threading
from typing import Callable, Dict, List, Optional

from ..cluster.topology import GpuDeviceInformation, GpuDiscoverer

MAX_MEMORY_BYTES = "MAX_MEMORY_BYTES"
AVG_MEMORY_BYTES = "AVG_MEMORY_BYTES"
MAX_GPU_UTILIZATION = "MAX_GPU_UTILIZATION"
AVG_GPU_UTILIZATION = "AVG_GPU_UTILIZATION"
MAX_GPU_FB_MEMORY_USAGE = "MAX_GPU_FB_MEMORY_USAGE"
AVG_GPU_FB_MEMORY_USAGE = "AVG_GPU_FB_MEMORY_USAGE"
MAX_GPU_MAIN_MEMORY_USAGE = "MAX_GPU_MAIN_MEMORY_USAGE"
AVG_GPU_MAIN_MEMORY_USAGE = "AVG_GPU_MAIN_MEMORY_USAGE"
METRIC_NAMES = [MAX_MEMORY_BYTES, AVG_MEMORY_BYTES, MAX_GPU_UTILIZATION, AVG_GPU_UTILIZATION,
                MAX_GPU_FB_MEMORY_USAGE, AVG_GPU_FB_MEMORY_USAGE, MAX_GPU_MAIN_MEMORY_USAGE,
                AVG_GPU_MAIN_MEMORY_USAGE]


def process_tree_rss_bytes(root_pid: int) -> int:
    """RSS of ``root_pid`` and all its descendants, from /proc (no psutil dependency)."""
    children: Dict[int, List[int]] = {}
    rss: Dict[int, int] = {}
    page = os.sysconf("SC_PAGE_SIZE")
    for name in os.listdir("/proc"):
        if not name.isdigit():
            continue
        pid = int(name)
        try:
            with open(f"/proc/{pid}/stat") as f:
                stat = f.read()
            rest = stat[stat.rindex(")") + 2:].split()
            ppid = int(rest[1])
            rss[pid] = int(rest[21]) * page
            children.setdefault(ppid, []).append(pid)
        except (OSError, ValueError, IndexError):
            continue
    total, stack, seen = 0, [root_pid], set()
    while stack:
        p = stack.pop()
        if p in seen:
            continue
        seen.add(p)
        total += rss.get(p, 0)
        stack.extend(children.get(p, []))
    return total


class TaskMonitor:
    def __init__(self, push: Callable[[Dict[str, float]], None], interval_ms: int = 5000,
                 gpu_enabled: bool = True, gpu_exec_path: Optional[str] = None,
                 visible_gpus: Optional[List[int]] = None, pid: Optional[int] = None,
                 user_metrics_path: Optional[str] = None):
        self.push = push
        self.user_metrics_path = user_metrics_path
        self.interval_s = max(0.05, interval_ms / 1000.0)
        self.pid = pid or os.getpid()
        self.visible_gpus = visible_gpus
        self.discoverer = GpuDiscoverer(gpu_exec_path) if gpu_enabled else None
        self.metrics: Dict[str, float] = {n: 0.0 for n in METRIC_NAMES}
        self.num_refreshes = 0
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name="task-monitor", daemon=True)

    def start(self) -> None:
        self._thread.start()

    def stop(self) -> None:
        self._stop.set()

    def _run(self) -> None:
        while not self._stop.wait(self.interval_s):
            try:
                self.refresh()
                self.push(dict(self.metrics))
            except Exception:  # noqa: BLE001 - monitoring must never kill the task
                pass

    def read_user_metrics(self) -> None:
        """Merge ``name=value`` lines the user process wrote (atomically) into the pushed gauges."""
        if not self.user_metrics_path:
            return
        try:
            with open(self.user_metrics_path) as f:
                for line in f:
                    k, _, v = line.strip().partition("=")
                    if k and v:
                        self.metrics[k] = float(v)
        except (OSError, ValueError):
            pass

    def flush(self) -> None:
        """Final push at task end, so a job shorter than one interval still reports its metrics."""
        try:
            self.read_user_metrics()
            self.push(dict(self.metrics))
        except Exception:  # noqa: BLE001
            pass

    def set_max(self, name: str, value: float) -> None:
        if value > self.metrics[name]:
            self.metrics[name] = value

    def set_avg(self, name: str, value: float) -> None:
        self.metrics[name] += (value - self.metrics[name]) / self.num_refreshes

    def refresh(self, gpu_info: Optional[GpuDeviceInformation] = None, rss: Optional[int] = None) -> None:
        self.num_refreshes += 1
        self.read_user_metrics()
        mem = process_tree_rss_bytes(self.pid) if rss is None else rss
        self.set_max(MAX_MEMORY_BYTES, mem)
        self.set_avg(AVG_MEMORY_BYTES, mem)
        if gpu_info is None and self.discoverer is not None and self.discoverer.usable:
            gpu_info = self.discoverer.sample()
        if gpu_info is None or not gpu_info.gpus:
            return
        gpus = gpu_info.gpus
        if self.visible_gpus is not None:
            picked = [g for i, g in enumerate(gpus) if (g.minor_number if g.minor_number >= 0 else i) in self.visible_gpus]
            gpus = picked or gpus
        util = sum(g.gpu_util_pct for g in gpus) / len(gpus)
        fb = sum(g.fb_usage_pct for g in gpus) / len(gpus)
        main = sum(g.bar1_usage_pct for g in gpus) / len(gpus)
        self.set_max(MAX_GPU_UTILIZATION, util)
        self.set_avg(AVG_GPU_UTILIZATION, util)
        self.set_max(MAX_GPU_FB_MEMORY_USAGE, fb)
        self.set_avg(AVG_GPU_FB_MEMORY_USAGE, fb)
        self.set_max(MAX_GPU_MAIN_MEMORY_USAGE, main)
        self.set_avg(AVG_GPU_MAIN_MEMORY_USAGE, main)
