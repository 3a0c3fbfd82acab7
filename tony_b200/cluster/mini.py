"""MiniCluster: a private, disposable local cluster for tests and LocalSubmitter.

Reference: tony-mini/src/main/java/com/linkedin/minitony/cluster/MiniCluster.java:39-73
(in-process MiniYARNCluster + MiniDFSCluster).  The local resource manager needs no daemons,
so a mini cluster is just an isolated work directory (application dirs, GPU lease files,
history tree) plus factory helpers that point clients at it.
"""
from __future__ import annotations

import os
import shutil
import tempfile
from typing import Optional

from ..client.tony_client import TonyClient
from ..conf import keys as K
from ..conf.configuration import TonyConfiguration


class MiniCluster:
    def __init__(self, num_node_managers: int = 1, gpus: Optional[int] = None, keep: bool = False):
        self.num_node_managers = num_node_managers  # kept for API parity; one box == one node
        self.gpus = gpus
        self.keep = keep
        self.work_dir: Optional[str] = None

    def start(self) -> "MiniCluster":
        self.work_dir = tempfile.mkdtemp(prefix="minitony_")
        os.makedirs(os.path.join(self.work_dir, "history", "intermediate"), exist_ok=True)
        os.makedirs(os.path.join(self.work_dir, "history", "finished"), exist_ok=True)
        return self

    def stop(self) -> None:
        if self.work_dir and not self.keep:
            shutil.rmtree(self.work_dir, ignore_errors=True)

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()

    @property
    def history_dir(self) -> str:
        return os.path.join(self.work_dir, "history")

    def base_conf(self) -> TonyConfiguration:
        c = TonyConfiguration()
        c.set(K.B200_WORK_DIR, self.work_dir, "minicluster")
        c.set(K.HISTORY_LOCATION, self.history_dir, "minicluster")
        c.set(K.HISTORY_INTERMEDIATE, os.path.join(self.history_dir, "intermediate"), "minicluster")
        c.set(K.HISTORY_FINISHED, os.path.join(self.history_dir, "finished"), "minicluster")
        if self.gpus is not None:
            c.set(K.B200_NUM_GPUS, str(self.gpus), "minicluster")
        return c

    def new_client(self, callback_handler=None) -> TonyClient:
        return TonyClient(self.base_conf(), callback_handler)
