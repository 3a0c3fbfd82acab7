"""GPT-2 medium DDP with chief failure + restart (BASELINE config #5): run with
tony.am.retry-count=1; --crash_at N makes the chief die after N steps of the first session, the
launcher restarts the whole session and the job resumes from the last sharded checkpoint."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from _train_common import train  # noqa: E402

from tony_b200.models import gpt2  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--seq", type=int, default=1024)
ap.add_argument("--ckpt_dir", default=os.path.join(os.environ.get("TMPDIR", "/tmp"), "tony_gpt2_ckpt"))
ap.add_argument("--ckpt_every", type=int, default=10)
ap.add_argument("--crash_at", type=int, default=None)
ap.add_argument("--tiny", action="store_true")
a = ap.parse_args()
cfg = gpt2.GPT2Config.tiny() if a.tiny else gpt2.GPT2Config.medium()
seq = min(a.seq, cfg.max_positions)
train(lambda dev: gpt2.GPT2LMHead(cfg, device=dev),
      lambda seed: gpt2.synthetic_batch(cfg, a.batch, seq, pin=True, seed=seed),
      steps=a.steps, lr=1e-4, ckpt_dir=a.ckpt_dir, ckpt_every=a.ckpt_every, crash_at=a.crash_at,
      samples_per_step=a.batch)
