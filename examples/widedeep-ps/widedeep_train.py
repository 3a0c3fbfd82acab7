"""ps/worker wide-and-deep recommender (BASELINE config #4).  The reference's 2 ps + 6 workers
become 8 symmetric ranks: each owns a row shard of every tower matrix (its "ps" role: sharded
fp32 master weights + Adam moments, updated by the reduce-scatter-fused optimizer) and trains
on its own mini-batch (its "worker" role), pulling the other shards with the all-gather-fused
GEMM at the start of each forward."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from _train_common import train  # noqa: E402

from tony_b200.models import widedeep  # noqa: E402
from tony_b200.parallel.ps import ps_parameters  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--batch", type=int, default=8192)
ap.add_argument("--tiny", action="store_true")
a = ap.parse_args()
cfg = widedeep.WideDeepConfig.tiny() if a.tiny else widedeep.WideDeepConfig()
train(lambda dev: widedeep.WideDeep(cfg, device=dev),
      lambda seed: widedeep.synthetic_batch(cfg, a.batch, pin=True, seed=seed),
      steps=a.steps, lr=1e-3, trainer_kw=lambda model: {"ps_sharded": ps_parameters(model)},
      samples_per_step=a.batch)
