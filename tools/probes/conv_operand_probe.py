import os, sys, json
sys.path.insert(0, '/root/repo')
import torch
from tony_b200.ops import functional as F
dev = torch.device('cuda')
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, iters=5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for i in range(iters + 2):
        flush.fill_(i & 1); e0.record(); fn(); e1.record(); e1.synchronize()
        if i >= 2: best = min(best, e0.elapsed_time(e1))
    return best * 1e3
N = 256
for (ci, co, k, h) in ((64, 64, 3, 56), (128, 128, 3, 28), (256, 256, 3, 14), (512, 512, 3, 7)):
    x = torch.randn(N, h, h, ci, device=dev).to(torch.bfloat16)
    w = (torch.randn(co, k, k, ci, device=dev) / (k * ci ** 0.5)).to(torch.bfloat16)
    dy = torch.randn(N, h, h, co, device=dev).to(torch.bfloat16)
    print(json.dumps({"dbg": os.environ.get("TONY_B200_CONV_DBG", "0"), "cin": ci, "cout": co, "hw": h,
                      "fprop_us": round(timeit(lambda: F.conv2d_nhwc(x, w, 1, 1)), 1),
                      "wgrad_us": round(timeit(lambda: F.conv2d_wgrad_nhwc(dy, x, w.shape, 1, 1)), 1)}), flush=True)
