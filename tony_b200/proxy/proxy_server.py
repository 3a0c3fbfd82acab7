"""TCP relay used by the notebook submitter to reach a service running inside the cluster.

Reference: tony-proxy/src/main/java/com/linkedin/tonyproxy/ProxyServer.java:27-90
(thread-per-connection relay).  Here one selector loop moves bytes for all connections.
"""
from __future__ import annotations

import selectors
import socket
import threading
from typing import Dict, Optional


class ProxyServer:
    def __init__(self, remote_host: str, remote_port: int, local_port: int = 0):
        self.remote = (remote_host, int(remote_port))
        self._lsock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self._lsock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self._lsock.bind(("127.0.0.1", local_port))
        self._lsock.listen(64)
        self.local_port = self._lsock.getsockname()[1]
        self._sel = selectors.DefaultSelector()
        self._peer: Dict[socket.socket, socket.socket] = {}
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "ProxyServer":
        self._lsock.setblocking(False)
        self._sel.register(self._lsock, selectors.EVENT_READ)
        self._thread = threading.Thread(target=self._loop, name="proxy", daemon=True)
        self._thread.start()
        return self

    def _close_pair(self, s: socket.socket) -> None:
        for x in (s, self._peer.pop(s, None)):
            if x is None:
                continue
            self._peer.pop(x, None)
            try:
                self._sel.unregister(x)
            except (KeyError, ValueError):
                pass
            x.close()

    def _loop(self) -> None:
        while not self._stop.is_set():
            for key, _ in self._sel.select(timeout=0.2):
                s = key.fileobj
                if s is self._lsock:
                    try:
                        c, _ = self._lsock.accept()
                        r = socket.create_connection(self.remote, timeout=5)
                    except OSError:
                        continue
                    r.settimeout(None)
                    self._peer[c], self._peer[r] = r, c
                    self._sel.register(c, selectors.EVENT_READ)
                    self._sel.register(r, selectors.EVENT_READ)
                    continue
                try:
                    data = s.recv(65536)
                    if data:
                        self._peer[s].sendall(data)
                    else:
                        self._close_pair(s)
                except (OSError, KeyError):
                    self._close_pair(s)

    def stop(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)
        for s in list(self._peer):
            self._close_pair(s)
        self._lsock.close()
