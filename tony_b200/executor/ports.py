"""Port reservation for task RPC / TensorBoard ports.

``EphemeralPort`` holds a bound listening socket on an OS-picked port until released
(reference: tony-core/src/main/java/com/linkedin/tony/EphemeralPort.java:43-56).
``ReusablePort`` binds with ``SO_REUSEPORT`` so the user process can bind the same port while
the reservation is still held — no window in which another process can steal it
(reference: ReusablePort.java:151-237 + resources/reserve_reusable_port.py:52-89, which need a
helper Python process and a marker file because the JVM cannot set SO_REUSEPORT; here the
executor is Python, so it is one setsockopt).  Selected by the user env ``TF_GRPC_REUSE_PORT``
/ ``TB_SERVER_REUSE_PORT`` (TaskExecutor.java:123-153).
"""
from __future__ import annotations

import socket
from typing import Optional


class ServerPort:
    def __init__(self, sock: socket.socket):
        self._sock: Optional[socket.socket] = sock
        self.port: int = sock.getsockname()[1]

    def close(self) -> None:
        if self._sock is not None:
            try:
                self._sock.close()
            finally:
                self._sock = None

    @property
    def held(self) -> bool:
        return self._sock is not None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class EphemeralPort(ServerPort):
    @staticmethod
    def create() -> "EphemeralPort":
        s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        s.bind(("", 0))
        s.listen(1)
        return EphemeralPort(s)


class ReusablePort(ServerPort):
    @staticmethod
    def create(port: int = 0) -> "ReusablePort":
        if not hasattr(socket, "SO_REUSEPORT"):
            raise OSError("SO_REUSEPORT is not available on this platform")
        s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEPORT, 1)
        s.bind(("", port))
        # deliberately NOT listening: connections must go to the user's socket, not ours
        return ReusablePort(s)


def is_port_available(port: int) -> bool:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        try:
            s.bind(("", port))
            return True
        except OSError:
            return False


def reserve(reuse: bool) -> ServerPort:
    return ReusablePort.create() if reuse else EphemeralPort.create()


def main(argv=None) -> int:
    """``python -m tony_b200.executor.ports -p PORT [-d SECONDS]``: hold PORT with SO_REUSEPORT and
    create ``<PORT>___PORT___`` in the working directory once it is held — the contract of the
    reference's reserve_reusable_port.py (resources/reserve_reusable_port.py:52-89) for callers
    that want the reservation to live in a separate process."""
    import argparse
    import os
    import time

    ap = argparse.ArgumentParser()
    ap.add_argument("-p", "--port", type=int, required=True)
    ap.add_argument("-d", "--duration", type=float, default=3600.0)
    a = ap.parse_args(argv)
    port = ReusablePort.create(a.port)
    marker = f"{port.port}___PORT___"
    open(marker, "w").close()
    try:
        time.sleep(a.duration)
    except KeyboardInterrupt:
        pass
    finally:
        port.close()
        try:
            os.unlink(marker)
        except OSError:
            pass
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
