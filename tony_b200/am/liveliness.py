"""Heartbeat liveliness monitor.

A task is registered on its first ``registerWorkerSpec``, pinged by every
``taskExecutorHeartbeat`` and unregistered by ``registerExecutionResult``; if no ping
arrives within ``heartbeat-interval * max(3, max-missed-heartbeats)`` the expiry callback
fires.  Same arithmetic as the reference's AbstractLivelinessMonitor wiring
(tony-core/src/main/java/com/linkedin/tony/ApplicationMaster.java:202-222; expiry handler
:1225-1232).  One daemon thread sleeps exactly until the earliest deadline instead of
polling every 3 heartbeats.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, Dict


class LivelinessMonitor:
    def __init__(self, hb_interval_ms: int, max_missed: int, on_expire: Callable[[str], None]):
        self.expire_s = hb_interval_ms * max(3, max_missed) / 1000.0
        self.on_expire = on_expire
        self._last: Dict[str, float] = {}
        self._cv = threading.Condition()
        self._stop = False
        self._thread = threading.Thread(target=self._run, name="hb-monitor", daemon=True)

    def start(self) -> None:
        self._thread.start()

    def stop(self) -> None:
        with self._cv:
            self._stop = True
            self._cv.notify_all()

    def register(self, task_id: str) -> None:
        with self._cv:
            self._last[task_id] = time.monotonic()
            self._cv.notify_all()

    def unregister(self, task_id: str) -> None:
        with self._cv:
            self._last.pop(task_id, None)

    def received_ping(self, task_id: str) -> None:
        with self._cv:
            if task_id in self._last:
                self._last[task_id] = time.monotonic()

    def reset(self) -> None:
        with self._cv:
            self._last.clear()

    def _run(self) -> None:
        while True:
            expired = []
            with self._cv:
                if self._stop:
                    return
                now = time.monotonic()
                nxt = None
                for tid, t in list(self._last.items()):
                    deadline = t + self.expire_s
                    if deadline <= now:
                        expired.append(tid)
                        del self._last[tid]
                    elif nxt is None or deadline < nxt:
                        nxt = deadline
                if not expired:
                    self._cv.wait(timeout=None if nxt is None else max(0.01, nxt - now))
            for tid in expired:
                try:
                    self.on_expire(tid)
                except Exception:  # noqa: BLE001
                    pass
