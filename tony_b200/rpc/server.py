"""Threaded RPC server + ACL, and the retrying client.

Server side mirrors ApplicationRpcServer (an OS-picked port, handler object supplies the
methods: rpc/ApplicationRpcServer.java:36-43) and the service ACL of TonyPolicyProvider.java
:15-22 (only the declared protocol methods are callable).  Client side mirrors
ApplicationRpcClient's retry policy — 10 attempts, 2 s apart
(rpc/impl/ApplicationRpcClient.java:48-77) — but keeps one persistent connection.
"""
from __future__ import annotations

import socket
import socketserver
import threading
import time
from typing import Any, Callable, Dict, Iterable, Optional

from .wire import AuthError, RpcError, recv_frame, send_frame

APPLICATION_RPC_METHODS = (
    "getTaskInfos", "getClusterSpec", "registerWorkerSpec", "registerTensorBoardUrl",
    "registerExecutionResult", "finishApplication", "taskExecutorHeartbeat", "registerCallbackInfo",
    # additions of this implementation
    "getApplicationStatus", "markFirstStep",
)
METRICS_RPC_METHODS = ("updateMetrics",)


class RpcServer:
    def __init__(self, handler: Any, methods: Iterable[str], secret: Optional[bytes] = None,
                 host: str = "127.0.0.1", port: int = 0):
        self.handler = handler
        self.methods = set(methods)
        self.secret = secret
        outer = self

        class _Conn(socketserver.BaseRequestHandler):
            def handle(self):
                sock = self.request
                sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                while True:
                    try:
                        req = recv_frame(sock, outer.secret)
                    except AuthError:
                        try:
                            send_frame(sock, {"ok": False, "e": "authentication failed"}, None)
                        except OSError:
                            pass
                        return
                    except (RpcError, OSError, ValueError):
                        return
                    if req is None:
                        return
                    resp = outer._dispatch(req)
                    try:
                        send_frame(sock, resp, outer.secret)
                    except OSError:
                        return

        class _Server(socketserver.ThreadingTCPServer):
            allow_reuse_address = True
            daemon_threads = True

        self._server = _Server((host, port), _Conn)
        self.host, self.port = self._server.server_address[:2]
        self._thread = threading.Thread(target=self._server.serve_forever, kwargs={"poll_interval": 0.05},
                                        name=f"rpc-{self.port}", daemon=True)

    def _dispatch(self, req: Dict[str, Any]) -> Dict[str, Any]:
        m = req.get("m")
        if m not in self.methods:
            return {"ok": False, "e": f"method {m!r} is not part of this protocol"}
        fn: Optional[Callable] = getattr(self.handler, m, None)
        if fn is None:
            return {"ok": False, "e": f"method {m!r} not implemented"}
        try:
            return {"ok": True, "r": fn(**(req.get("a") or {}))}
        except Exception as exc:  # noqa: BLE001 - surfaced to the caller
            return {"ok": False, "e": f"{type(exc).__name__}: {exc}"}

    def start(self) -> "RpcServer":
        self._thread.start()
        return self

    def stop(self) -> None:
        self._server.shutdown()
        self._server.server_close()


class RpcClient:
    """Blocking client with one persistent connection and reconnect-and-retry."""

    def __init__(self, host: str, port: int, secret: Optional[bytes] = None, retries: int = 10,
                 retry_interval_s: float = 2.0, timeout_s: Optional[float] = None):
        self.host, self.port, self.secret = host, int(port), secret
        self.retries, self.retry_interval_s, self.timeout_s = retries, retry_interval_s, timeout_s
        self._sock: Optional[socket.socket] = None
        self._lock = threading.Lock()

    def _connect(self) -> socket.socket:
        s = socket.create_connection((self.host, self.port), timeout=10)
        s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        s.settimeout(self.timeout_s)
        return s

    def call(self, method: str, **args) -> Any:
        last: Optional[Exception] = None
        for attempt in range(max(1, self.retries)):
            with self._lock:
                try:
                    if self._sock is None:
                        self._sock = self._connect()
                    send_frame(self._sock, {"m": method, "a": args}, self.secret)
                    resp = recv_frame(self._sock, self.secret)
                    if resp is None:
                        raise RpcError("server closed the connection")
                    if not resp.get("ok"):
                        raise RpcError(resp.get("e", "remote error"))
                    return resp.get("r")
                except (OSError, AuthError) as exc:
                    last = exc
                    self._close_locked()
                except RpcError as exc:
                    if "closed" in str(exc):
                        last = exc
                        self._close_locked()
                    else:
                        raise
            if attempt + 1 < self.retries:
                time.sleep(self.retry_interval_s)
        raise RpcError(f"{method}: no response from {self.host}:{self.port} after {self.retries} tries: {last}")

    def _close_locked(self) -> None:
        if self._sock is not None:
            try:
                self._sock.close()
            except OSError:
                pass
            self._sock = None

    def close(self) -> None:
        with self._lock:
            self._close_locked()
