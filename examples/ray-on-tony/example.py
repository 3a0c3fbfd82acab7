"""A program for the head + workers cluster that ``discovery.py`` brings up - the counterpart of the reference's
tony-examples/ray-on-tony/example.py:13-60 (wait until every node has joined, then run tasks on all of them and
count which host answered).  With ray installed it is that program; without it the same check runs over the
registry the head task of ``discovery.py`` keeps: every worker registers, the head waits for all of them and
then hands each a small task (a checksum of a byte range) whose results it verifies."""
import hashlib
import json
import os
import socket
import sys
import time

spec = json.loads(os.environ["CLUSTER_SPEC"])
role, index = os.environ["JOB_NAME"], int(os.environ["TASK_INDEX"])
head_host, head_port = spec["head"][0].rsplit(":", 1)
n_workers = len(spec.get("worker", []))
try:
    import ray  # noqa: F401

    HAVE_RAY = True
except ImportError:
    HAVE_RAY = False


def task(lo, hi):
    return hashlib.sha1(bytes(range(lo % 256, 256)) * (hi - lo)).hexdigest()


if HAVE_RAY and role == "head":
    import ray

    ray.init(address="auto")
    while len(ray.nodes()) < n_workers + 1:
        time.sleep(0.5)
    remote = ray.remote(lambda i: (socket.gethostname(), task(i, i + 3)))
    out = ray.get([remote.remote(i) for i in range(4 * (n_workers + 1))])
    assert all(h == task(i, i + 3) for i, (_, h) in enumerate(out))
    print("ray tasks ran on", sorted({h for h, _ in out}), flush=True)
elif role == "head":
    srv = socket.socket()
    srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
    srv.bind(("0.0.0.0", int(head_port)))
    srv.listen(16)
    conns = []
    while len(conns) < n_workers:                      # wait_for_nodes
        c, _ = srv.accept()
        conns.append((c, c.recv(64).decode()))
    for i, (c, name) in enumerate(conns):              # one task per node, result checked by the head
        c.sendall(json.dumps({"lo": i, "hi": i + 3}).encode())
        assert c.recv(64).decode() == task(i, i + 3), name
        c.close()
    print("head ran tasks on", sorted(n for _, n in conns), flush=True)
elif not HAVE_RAY:
    for _ in range(400):
        try:
            c = socket.create_connection((head_host, int(head_port)), timeout=2)
            break
        except OSError:
            time.sleep(0.05)
    c.sendall(f"worker:{index}@{socket.gethostname()}".encode())
    job = json.loads(c.recv(256).decode())
    c.sendall(task(job["lo"], job["hi"]).encode())
sys.exit(0)
