"""Force each tiling / split count of the GEMM on a few small-tile-count shapes (ResNet-50's 7x7 and 14x14 pointwise
gradients) and print the time of each, next to cuBLAS: input for the cost model in csrc/host_math.h."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from tony_b200.ops import _lib
from tony_b200.ops._lib import cur_stream, ptr

dev = torch.device("cuda")
ws = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
L = _lib.lib()


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 10 * 1e3


# name, M, N, K, a_mn, b_mn
shapes = [("dgrad 512->2048 @7", 12544, 512, 2048, 0, 1), ("dgrad 2048->512 @7", 12544, 2048, 512, 0, 1),
          ("wgrad 512->2048 @7", 2048, 512, 12544, 1, 1), ("wgrad 2048->512 @7", 512, 2048, 12544, 1, 1),
          ("dgrad 1024->256 @14", 50176, 1024, 256, 0, 1), ("wgrad 256->1024 @14", 1024, 256, 50176, 1, 1)]
for name, M, N, K, a_mn, b_mn in shapes:
    A = torch.randn((K, M) if a_mn else (M, K), device=dev).to(torch.bfloat16)
    B = torch.randn((K, N) if b_mn else (N, K), device=dev).to(torch.bfloat16)
    C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ref = (lambda: torch.matmul(A.t(), B, out=C)) if a_mn else (lambda: torch.matmul(A, B, out=C))
    line = [f"{name}: cuBLAS {timed(ref):.1f} us"]
    for bn in (0, 512, 256, 128):
        for sp in ((0, 1) if not a_mn else (0, 2, 4, 8, 16)):
            def run(bn=bn, sp=sp):
                rc = L.tony_gemm_bf16_ex(ptr(A), ptr(B), ptr(C), M, N, K, a_mn, b_mn, A.stride(0), B.stride(0), N, 0, 1.0,
                                         None, None, 0, None, ptr(ws), ws.numel(), bn, sp, None, 0, 0, 0, cur_stream())
                if rc != 0:
                    raise RuntimeError(f"rc {rc}")
            try:
                line.append(f"bn{bn}/s{sp} {timed(run):.1f}")
            except RuntimeError as e:
                line.append(f"bn{bn}/s{sp} {e}")
    print(" | ".join(line), flush=True)
