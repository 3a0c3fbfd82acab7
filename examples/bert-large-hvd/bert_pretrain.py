"""horovod-on-tony BERT-large pre-training (BASELINE config #3): launched with
tony.application.framework=horovod, gradients fused into flat buckets ("tensor fusion") and
all-reduced + optimised by one NVLink kernel per bucket, overlapped with backward."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from _train_common import train  # noqa: E402

from tony_b200.models import bert  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--seq", type=int, default=512)
ap.add_argument("--tiny", action="store_true")
a = ap.parse_args()
cfg = bert.BertConfig.tiny() if a.tiny else bert.BertConfig.large()
seq = min(a.seq, cfg.max_positions)
train(lambda dev: bert.BertForPreTraining(cfg, device=dev),
      lambda seed: bert.synthetic_batch(cfg, a.batch, seq, pin=True, seed=seed),
      steps=a.steps, lr=1e-4, samples_per_step=a.batch)
