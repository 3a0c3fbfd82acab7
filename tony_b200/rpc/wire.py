"""RPC transport: length-prefixed, HMAC-authenticated JSON frames over TCP.

The reference speaks Hadoop IPC + protobuf 2.5 between client, ApplicationMaster and
TaskExecutors (tony-core/src/main/proto/tensorflow_cluster_service_protos.proto:11-20,
rpc/ApplicationRpcServer.java:36-43,123-143) and authenticates with ClientToAM tokens
(ApplicationMaster.java:484-504).  The method names and fields are kept (``rpc/api.py``);
the transport is rebuilt lean: one frame = 4-byte big-endian length, 32-byte
HMAC-SHA256(secret, payload), UTF-8 JSON payload.  With security disabled the MAC is zeros.
"""
from __future__ import annotations

import hashlib
import hmac
import json
import socket
import struct
from typing import Any, Optional

MAC_LEN = 32
MAX_FRAME = 64 << 20


class RpcError(Exception):
    pass


class AuthError(RpcError):
    pass


def _mac(secret: Optional[bytes], payload: bytes) -> bytes:
    if not secret:
        return b"\0" * MAC_LEN
    return hmac.new(secret, payload, hashlib.sha256).digest()


def send_frame(sock: socket.socket, obj: Any, secret: Optional[bytes]) -> None:
    payload = json.dumps(obj, separators=(",", ":")).encode("utf-8")
    sock.sendall(struct.pack(">I", len(payload)) + _mac(secret, payload) + payload)


def _recv_exact(sock: socket.socket, n: int) -> Optional[bytes]:
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(n - len(buf))
        if not chunk:
            return None
        buf.extend(chunk)
    return bytes(buf)


def recv_frame(sock: socket.socket, secret: Optional[bytes]) -> Optional[Any]:
    """Return the decoded object, or None on a clean EOF before a frame starts."""
    head = _recv_exact(sock, 4)
    if head is None:
        return None
    (n,) = struct.unpack(">I", head)
    if n > MAX_FRAME:
        raise RpcError(f"frame of {n} bytes exceeds limit")
    mac = _recv_exact(sock, MAC_LEN)
    payload = _recv_exact(sock, n)
    if mac is None or payload is None:
        raise RpcError("connection closed mid-frame")
    if secret and not hmac.compare_digest(mac, _mac(secret, payload)):
        raise AuthError("bad message authentication code")
    return json.loads(payload.decode("utf-8"))
