"""Arbitrary job types: a ``head`` and N ``worker`` tasks that find each other through the
cluster spec, the pattern of the reference's tony-examples/ray-on-tony/discovery.py:27-37 (which
starts ``ray start --head`` / ``ray start --address``).  Without ray installed the head runs a
tiny TCP registry and each worker checks in, which exercises the same contract: per-type
commands, TF_CONFIG-style discovery of the head's address, head tracked as chief-like role."""
import json
import os
import socket
import sys
import time

spec = json.loads(os.environ["CLUSTER_SPEC"])
role, index = os.environ["JOB_NAME"], int(os.environ["TASK_INDEX"])
head_host, head_port = spec["head"][0].rsplit(":", 1)
n_workers = len(spec.get("worker", []))
if role == "head":
    srv = socket.socket()
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("0.0.0.0", int(head_port)))
    srv.listen(16)
    seen = set()
    while len(seen) < n_workers:
        c, _ = srv.accept()
        seen.add(c.recv(64).decode())
        c.sendall(b"ok")
        c.close()
    print("head saw", sorted(seen), flush=True)
else:
    for _ in range(200):
        try:
            c = socket.create_connection((head_host, int(head_port)), timeout=2)
            break
        except OSError:
            time.sleep(0.05)
    c.sendall(f"worker:{index}".encode())
    assert c.recv(8) == b"ok"
sys.exit(0)
