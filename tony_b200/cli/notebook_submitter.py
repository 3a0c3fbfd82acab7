"""``python -m tony_b200.cli.notebook_submitter`` — run a long-lived single-node service
(Jupyter etc.) inside the cluster and proxy a local port to it.

Reference: tony-cli/src/main/java/com/linkedin/tony/cli/NotebookSubmitter.java:71-144:
single-node job executed by the AM itself, 24 h timeout, wait for the ``notebook`` TaskInfo
and start a local ProxyServer towards its URL.
"""
from __future__ import annotations

import sys
import threading
from typing import List, Optional, Set

from ..am.session import TaskInfo
from ..client.tony_client import TaskUpdateListener, TonyClient, UsageError, usage
from ..conf import keys as K
from ..proxy.proxy_server import ProxyServer
from .cluster_submitter import TonySubmitter

NOTEBOOK_TIMEOUT_MS = 24 * 60 * 60 * 1000


class NotebookSubmitter(TonySubmitter, TaskUpdateListener):
    def __init__(self, client: Optional[TonyClient] = None, local_port: int = 0):
        self.client = client or TonyClient()
        self.local_port = local_port
        self.proxy: Optional[ProxyServer] = None
        self.ready = threading.Event()

    def on_task_infos_updated(self, task_infos: Set[TaskInfo]) -> None:
        if self.proxy is not None:
            return
        for t in task_infos:
            if t.name == K.NOTEBOOK_JOB_NAME and t.url:
                hostport = t.url.split("://", 1)[-1].rstrip("/")
                host, _, port = hostport.rpartition(":")
                self.proxy = ProxyServer(host, int(port), self.local_port).start()
                print(f"Notebook is reachable at http://127.0.0.1:{self.proxy.local_port} (-> {hostport})")
                self.ready.set()

    def submit(self, argv: List[str]) -> int:
        argv = list(argv) + ["--conf", f"{K.APPLICATION_TIMEOUT}={NOTEBOOK_TIMEOUT_MS}"]
        try:
            if not self.client.init(argv):
                return 0
        except (UsageError, RuntimeError, ValueError) as exc:
            sys.stderr.write(f"{exc}\n{usage()}\n")
            return -1
        self.client.add_listener(self)
        try:
            return self.client.start()
        finally:
            if self.proxy:
                self.proxy.stop()


def main(argv=None) -> int:
    return NotebookSubmitter().submit(sys.argv[1:] if argv is None else argv)


if __name__ == "__main__":
    sys.exit(main() & 0xFF)
