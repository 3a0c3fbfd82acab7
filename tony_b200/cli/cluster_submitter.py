"""``python -m tony_b200.cli.cluster_submitter <TonyClient flags>`` — submit to the cluster.

Reference: tony-cli/src/main/java/com/linkedin/tony/cli/ClusterSubmitter.java:49-93 (uploads
its fat jar as a resource, installs a shutdown hook that force-kills the application, then
``client.start()``).  There is no jar to ship here — tasks import this package from the same
checkout — so what remains is the kill-on-interrupt guarantee and the exit code contract.
"""
from __future__ import annotations

import signal
import sys
from typing import List, Optional

from ..client.tony_client import TonyClient, UsageError, usage


class TonySubmitter:
    def submit(self, argv: List[str]) -> int:  # pragma: no cover - interface
        raise NotImplementedError


class ClusterSubmitter(TonySubmitter):
    def __init__(self, client: Optional[TonyClient] = None):
        self.client = client or TonyClient()

    def submit(self, argv: List[str]) -> int:
        try:
            if not self.client.init(argv):
                return 0
        except (UsageError, RuntimeError, ValueError, OSError) as exc:
            sys.stderr.write(f"{exc}\n{usage()}\n")
            return -1

        def _kill(signum, frame):
            self.client.force_kill_application()
            sys.exit(-1)

        old = {s: signal.signal(s, _kill) for s in (signal.SIGINT, signal.SIGTERM)}
        try:
            return self.client.start()
        finally:
            for s, h in old.items():
                signal.signal(s, h)


def main(argv=None) -> int:
    return ClusterSubmitter().submit(sys.argv[1:] if argv is None else argv)


if __name__ == "__main__":
    sys.exit(main() & 0xFF)
