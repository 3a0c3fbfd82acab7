"""MXNet-style kvstore job: scheduler / server / worker roles wired by the DMLC_* variables the
mxnet runtime exports (reference job: tony-examples/linearregression-mxnet/src/mxnet_dist_ex.py:35,
61-68 with kvstore 'dist_async').  MXNet is not on the B200 image, so each role only validates its
contract and the workers fit y = 2x + 1 locally; with MXNet installed the same variables drive
``mx.kv.create('dist_async')`` unchanged."""
import os
import sys

import numpy as np

role = os.environ["DMLC_ROLE"]
need = ["DMLC_PS_ROOT_URI", "DMLC_PS_ROOT_PORT", "DMLC_NUM_SERVER", "DMLC_NUM_WORKER", "DMLC_LOCAL"]
missing = [k for k in need if k not in os.environ]
assert not missing, missing
print(role, {k: os.environ[k] for k in need}, flush=True)
try:
    import mxnet as mx  # noqa: F401

    kv = mx.kv.create("dist_async")
    print("kvstore rank", kv.rank, "of", kv.num_workers, flush=True)
except ImportError:
    pass
if role == "worker":
    rng = np.random.default_rng(int(os.environ["TASK_INDEX"]))
    x = rng.normal(size=(256, 1))
    y = 2 * x + 1
    w, b = 0.0, 0.0
    for _ in range(200):
        err = w * x + b - y
        w -= 0.1 * float((err * x).mean())
        b -= 0.1 * float(err.mean())
    assert abs(w - 2) < 1e-2 and abs(b - 1) < 1e-2
sys.exit(0)
