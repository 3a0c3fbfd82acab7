"""cProfile of the host side of ResNet-50 / BERT-large training steps (how much Python sits between two launches)."""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from tony_b200 import runtime

which = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
dev = torch.device("cuda")
torch.manual_seed(0)
if which == "resnet50":
    from tony_b200.models import resnet

    model = resnet.ResNet50(device=dev)
    batch = {k: v.to(dev) for k, v in resnet.synthetic_batch(256, "cpu", False, 0).items()}
else:
    from tony_b200.models import bert

    cfg = bert.BertConfig.large()
    model = bert.BertForPreTraining(cfg, device=dev)
    batch = {k: v.to(dev) for k, v in bert.synthetic_batch(cfg, 32, 512, seed=0).items()}
tr = runtime.Trainer(model, lr=1e-4, device=dev)
for _ in range(4):
    tr.step_device(batch)
torch.cuda.synchronize()
t = time.perf_counter()
for _ in range(6):
    tr.step_device(batch)
host = (time.perf_counter() - t) / 6 * 1e3
torch.cuda.synchronize()
print(f"{which}: host enqueue {host:.2f} ms/step")
pr = cProfile.Profile()
pr.enable()
for _ in range(4):
    tr.step_device(batch)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
