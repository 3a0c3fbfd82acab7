"""``python -m tony_b200.cli.local_submitter`` — run a job on a throw-away mini cluster.

Reference: tony-cli/src/main/java/com/linkedin/tony/cli/LocalSubmitter.java:39-70 starts a
2-NodeManager MiniCluster, dumps its configuration and submits to it.  Equivalent here: a
:class:`tony_b200.cluster.mini.MiniCluster` with its own private work directory (history and
GPU leases included) that is torn down afterwards.
"""
from __future__ import annotations

import sys
from typing import List

from ..cluster.mini import MiniCluster
from .cluster_submitter import ClusterSubmitter, TonySubmitter


class LocalSubmitter(TonySubmitter):
    def submit(self, argv: List[str]) -> int:
        with MiniCluster(num_node_managers=2) as cluster:
            return ClusterSubmitter(cluster.new_client()).submit(argv)


def main(argv=None) -> int:
    return LocalSubmitter().submit(sys.argv[1:] if argv is None else argv)


if __name__ == "__main__":
    sys.exit(main() & 0xFF)
