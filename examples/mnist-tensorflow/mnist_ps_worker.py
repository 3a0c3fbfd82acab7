"""ps/worker asynchronous training driven by TF_CONFIG — the job shape of the reference's
tony-examples/mnist-tensorflow/mnist_distributed.py:206-241 (replica_device_setter +
MonitoredTrainingSession).  If TensorFlow is importable the cluster resolver is exercised;
either way the training itself is a dependency-free softmax regression against a tiny TCP
parameter server started by the ``ps`` tasks, so the example runs on the bare B200 image and
shows exactly what the launcher provides: TF_CONFIG / CLUSTER_SPEC, untracked ps tasks that the
AM stops when the workers are done."""
import json
import os
import pickle
import socket
import socketserver
import struct
import sys
import threading

import numpy as np

cfg = json.loads(os.environ["TF_CONFIG"])
cluster, task = cfg["cluster"], cfg["task"]
role, index = task["type"], int(task["index"])
try:
    import tensorflow as tf  # noqa: F401

    resolver = tf.distribute.cluster_resolver.TFConfigClusterResolver()
    print("TF cluster spec:", resolver.cluster_spec().as_dict(), flush=True)
except ImportError:
    print("tensorflow not installed; using the built-in numpy parameter server", flush=True)


def send(sock, obj):
    b = pickle.dumps(obj)
    sock.sendall(struct.pack(">I", len(b)) + b)


def recv(sock):
    n = struct.unpack(">I", sock.recv(4, socket.MSG_WAITALL))[0]
    return pickle.loads(sock.recv(n, socket.MSG_WAITALL))


if role == "ps":
    host, port = cluster["ps"][index].rsplit(":", 1)
    W = np.zeros((784, 10), np.float32)
    lock = threading.Lock()

    class H(socketserver.BaseRequestHandler):
        def handle(self):
            global W
            while True:
                try:
                    op, payload = recv(self.request)
                except Exception:
                    return
                with lock:
                    if op == "push":
                        W -= 0.5 * payload
                    send(self.request, W)

    srv = socketserver.ThreadingTCPServer(("0.0.0.0", int(port)), H)
    srv.daemon_threads = True
    print(f"ps:{index} serving on {port}", flush=True)
    srv.serve_forever()  # untracked: the ApplicationMaster stops us when the workers finish
else:
    rng = np.random.default_rng(index)
    y = rng.integers(0, 10, 4096)
    x = rng.normal(0, 0.1, (4096, 784)).astype(np.float32)
    for k in range(10):
        x[y == k, 70 * k:70 * k + 40] += 1.0
    host, port = cluster["ps"][0].rsplit(":", 1)
    for attempt in range(100):
        try:
            s = socket.create_connection((host, int(port)), timeout=5)
            break
        except OSError:
            import time

            time.sleep(0.1)
    s.settimeout(None)
    send(s, ("pull", None))
    W = recv(s)
    for step in range(60):
        idx = rng.integers(0, len(x), 128)
        logits = x[idx] @ W
        p = np.exp(logits - logits.max(1, keepdims=True))
        p /= p.sum(1, keepdims=True)
        p[np.arange(128), y[idx]] -= 1
        send(s, ("push", x[idx].T @ p / 128))
        W = recv(s)
    acc = ((x @ W).argmax(1) == y).mean()
    print(f"worker:{index} accuracy {acc:.3f}", flush=True)
    sys.exit(0 if acc > 0.9 else 1)
