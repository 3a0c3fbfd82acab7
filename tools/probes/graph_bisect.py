"""Which part of the BERT step breaks CUDA-graph capture?  Runs small variants and prints ok/error."""
import os
import sys
import traceback

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tony_b200 import runtime  # noqa: E402
from tony_b200.models import bert  # noqa: E402


def trial(name, **kw):
    env = kw.pop("env", {})
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        cfg = bert.BertConfig(vocab_size=kw.pop("vocab", 30522), hidden=kw.pop("hidden", 1024),
                              layers=kw.pop("layers", 2), heads=kw.pop("heads", 16), inner=kw.pop("inner", 4096),
                              max_positions=512, max_predictions=kw.pop("preds", 80))
        B, S = kw.pop("B", 32), kw.pop("S", 512)
        model = bert.BertForPreTraining(cfg, device="cuda")
        tr = runtime.Trainer(model, lr=1e-4, cuda_graph=True, bucket_mb=kw.pop("bucket_mb", 64.0))
        batches = [bert.synthetic_batch(cfg, B, S, device="cuda", seed=s) for s in range(2)]
        for i in range(5):
            loss = tr.step_device(batches[i % 2])
        torch.cuda.synchronize()
        print(f"{name}: ok loss={float(loss.item()):.3f} graph_kernels={tr._graph_kernels}", flush=True)
    except Exception as e:  # noqa: BLE001
        msg = str(e).splitlines()[0][:150]
        print(f"{name}: FAILED {type(e).__name__}: {msg}", flush=True)
        tb = traceback.format_exc().splitlines()
        print("   ", [l.strip() for l in tb if "tony_b200" in l][-3:], flush=True)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        torch.cuda.synchronize()


if __name__ == "__main__":
    which = sys.argv[1]
    if which == "a":
        trial("2 layers, large dims")
    elif which == "b":
        trial("2 layers, large dims, no PDL", env={"TONY_B200_PDL": "0"})
    elif which == "c":
        trial("2 layers, large dims, own attention", env={"TONY_B200_ATTENTION": "tcgen05"})
    elif which == "d":
        trial("2 layers, small batch", B=4, S=128)
    elif which == "e":
        trial("2 layers, small vocab", vocab=1000)
    elif which == "f":
        trial("2 layers, tiny buckets", bucket_mb=8.0, hidden=256, inner=1024, heads=4, vocab=1000, B=4, S=128)
