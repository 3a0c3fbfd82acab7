"""A/B two builds of the kernel library on the BERT GEMM shapes (TONY_AB_LIB = path of the other .so)."""
import ctypes, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tony_b200.ops import _lib
from tony_b200.ops._lib import ptr, cur_stream
libs = {"new": _lib.lib()}
old = ctypes.CDLL(os.environ["TONY_AB_LIB"])
old.tony_gemm_bf16.argtypes = _lib.lib().tony_gemm_bf16.argtypes
libs["old"] = old
shapes = [(16384, 1024, 4096), (16384, 4096, 1024), (16384, 3072, 1024), (4096, 1024, 16384)]
ws = torch.empty(128 << 20, dtype=torch.uint8, device="cuda")
for M, N, K in shapes:
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    b = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for name, L in libs.items():
        def run():
            L.tony_gemm_bf16(ptr(a), ptr(b), ptr(c), M, N, K, 0, 0, K, K, N, 0, 1.0, None, None, 0, None, ptr(ws), ws.numel(), cur_stream())  # noqa
        for _ in range(5): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(M, N, K, name, round(2 * M * N * K / ms / 1e9, 1), "TFLOP/s")
