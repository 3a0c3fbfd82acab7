#!/usr/bin/env python
"""Roofline numbers for the fused compute + collective kernels (launch with torchrun, one rank per GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29671 tools/collective_bench.py [--sizes-mb 16,64,256,672]

Every figure is timed on the device with CUDA events after 3 warm-up calls, bracketed by a barrier +
synchronize, and is the MAX over ranks.  One JSON line per measurement (rank 0):

* K1  fused all-reduce + AdamW + weight broadcast (``tony_fused_reduce_optim[_nvls]``), per bucket size;
* K2  reduce-scatter + sharded AdamW, weights stay sharded (``broadcast = 0``);
* K0b plain all-reduce through the same kernels, and NCCL's all-reduce for reference;
* K3  all-gather fused into a GEMM (``tony_allgather_gemm_bf16``) against the same GEMM on local weights;
* overlap: K1 on the side stream while GEMMs run on the main stream - how much of K1 is hidden.

Roofline of a bucket of S bytes on N ranks (B200_PROFILING.md): the slower of the NVLink term and the HBM
term.  NVLink, P2P form: every rank pulls (N-1)/N * S of gradients and pushes (N-1)/N * S of weights,
per direction (N-1)/N * S; NVLS form: the switch reduces and replicates, S/N in and S/N out per rank (the
multimem traffic the GPU's own links carry).  HBM: the optimizer touches 28 bytes per owned element
(bf16 grad + fp32 master/m/v read+write + bf16 weight) = 14 * S / N.  Link bandwidth = the measured
770 GB/s peer copy per direction (900 nominal), HBM = MEASURED_PEAKS.json."""
import argparse
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from tony_b200 import runtime
from tony_b200.ops import _lib
from tony_b200.ops import functional as F
from tony_b200.parallel.symm import SymmComm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINK_GBS, LINK_NOMINAL = 770.0, 900.0
try:
    PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:  # noqa: BLE001
    PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", default="16,64,256,672")
    ap.add_argument("--ctas", type=int, default=32)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    rank, world, dev = runtime.init()
    lib = _lib.lib()
    sizes = [int(float(x) * (1 << 20)) // 2 for x in a.sizes_mb.split(",")]        # elements (bf16)
    align = world * 256
    sizes = [(n + align - 1) // align * align for n in sizes]
    nmax = max(sizes)
    comm = SymmComm(nmax, dev)
    comm.grad.normal_(0, 0.01)
    comm.param.normal_(0, 0.02)
    chunk_max = nmax // world
    master = comm.param[:chunk_max].float().clone()
    m, v = torch.zeros_like(master), torch.zeros_like(master)
    side = torch.cuda.Stream(device=dev)
    step = [0]

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, iters=a.iters, stream=None):
        for _ in range(3):
            fn()
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(iters):
            fn()
        e1.record(stream)
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) * 1e3          # us

    def fused(n, optimizer, broadcast, ctas, stream=None):
        step[0] += 1
        s = ctypes.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
        t = step[0]
        if comm.nvls is not None:
            _lib.check(lib.tony_fused_reduce_optim_nvls(
                ctypes.byref(comm.nvls), 0, n, _lib.ptr(master), _lib.ptr(m), _lib.ptr(v), 0, optimizer, 1e-4, 0.9,
                0.999, 1e-8, 0.01, 1 - 0.9 ** min(t, 50), 1 - 0.999 ** min(t, 5000), 1.0 / world, None, broadcast,
                comm.next_nvls_call(), ctas, s), "nvls")
        else:
            if optimizer < 0:
                _lib.check(lib.tony_symm_allreduce_bf16(ctypes.byref(comm.c), 0, n, 1.0 / world, comm.next_epoch(),
                                                        ctas, s), "allreduce")
                return
            _lib.check(lib.tony_fused_reduce_optim(
                ctypes.byref(comm.c), 0, n, _lib.ptr(master), _lib.ptr(m), _lib.ptr(v), 0, optimizer, 1e-4, 0.9,
                0.999, 1e-8, 0.01, 1 - 0.9 ** min(t, 50), 1 - 0.999 ** min(t, 5000), 1.0 / world, None, broadcast,
                comm.next_epoch(), ctas, s), "fused")

    def emit(**kw):
        if rank == 0:
            print(json.dumps(kw), flush=True)

    form = "nvls" if comm.nvls is not None else "p2p"
    for n in sizes:
        S = 2.0 * n
        # NVLink bytes per direction per GPU.  Reduce phase: every GPU's copy of the other (world-1)/world of the bucket
        # leaves it (the switch pulls it for multimem.ld_reduce, or the owners read it with P2P loads) and its own
        # reduced shard comes in; weight push: the updated shard goes out once through the switch (multimem.st) or
        # world-1 times with P2P stores, and the other (world-1)/world of the bucket comes in.
        if world > 1:
            red_out, red_in = S * (world - 1) / world, (S / world if form == "nvls" else S * (world - 1) / world)
            bc_out, bc_in = (S / world if form == "nvls" else S * (world - 1) / world), S * (world - 1) / world
            link, link_nobc = max(red_out + bc_out, red_in + bc_in), max(red_out, red_in)
        else:
            link = link_nobc = 0.0
        hbm = 14.0 * S / world
        for name, opt, bc in (("K1 all-reduce + AdamW + broadcast", _lib.OPT_ADAMW, 1),
                              ("K2 reduce-scatter + sharded AdamW", _lib.OPT_ADAMW, 0),
                              ("K0b all-reduce", -1, 1)):
            for ctas in sorted({a.ctas, 148 if world > 1 else a.ctas}):
                us = timed(lambda: fused(n, opt, bc, ctas))
                link_b = link_nobc if name.startswith("K2") else link      # K2 has no weight push
                hbm_b = hbm if opt >= 0 else (S / world) * 2
                t_link = link_b / (LINK_GBS * 1e3)                        # us
                t_hbm = hbm_b / (PEAKS["hbm_gbs"] * 1e3)
                roof = max(t_link, t_hbm)
                emit(kernel=name, form=form, world=world, bucket_mb=round(S / 2 ** 20, 1), ctas=ctas, us=round(us, 1),
                     nvlink_bytes_per_direction_mb=round(link_b / 2 ** 20, 1),
                     nvlink_gbs_per_direction=round(link_b / us / 1e3, 1) if us else None,
                     fraction_of_770=round(link_b / us / 1e3 / LINK_GBS, 3) if us else None,
                     fraction_of_900=round(link_b / us / 1e3 / LINK_NOMINAL, 3) if us else None,
                     allreduce_busbw_gbs=round(S * 2 * (world - 1) / world / us / 1e3, 1) if world > 1 else None,
                     hbm_bytes_mb=round(hbm_b / 2 ** 20, 1), roofline_us=round(roof, 1),
                     bound="nvlink" if t_link >= t_hbm else "hbm", fraction_of_roofline=round(roof / us, 3))
        if world > 1:
            buf = comm.grad[:n]
            us = timed(lambda: dist.all_reduce(buf))
            emit(kernel="NCCL all_reduce (reference)", world=world, bucket_mb=round(S / 2 ** 20, 1), us=round(us, 1),
                 allreduce_busbw_gbs=round(S * 2 * (world - 1) / world / us / 1e3, 1))

    # ---- overlap: K1 on the side stream under a train of GEMMs on the main stream ----------------
    n = sizes[min(1, len(sizes) - 1)]
    A = torch.randn(16384, 1024, device=dev).to(torch.bfloat16)
    B = torch.randn(4096, 1024, device=dev).to(torch.bfloat16)
    reps = 24

    def gemms():
        for _ in range(reps):
            F.gemm(A, B)

    t_g = timed(gemms, iters=3)
    t_k = timed(lambda: fused(n, _lib.OPT_ADAMW, 1, a.ctas), iters=5)

    def both():
        ev = torch.cuda.Event()
        ev.record()
        side.wait_event(ev)
        fused(n, _lib.OPT_ADAMW, 1, a.ctas, stream=side)
        gemms()
        torch.cuda.current_stream().wait_stream(side)

    t_b = timed(both, iters=3)
    emit(kernel="overlap: K1 under 24 GEMMs 16384x4096x1024", form=form, world=world,
         bucket_mb=round(2.0 * n / 2 ** 20, 1), ctas=a.ctas, gemms_alone_us=round(t_g, 1), k1_alone_us=round(t_k, 1),
         together_us=round(t_b, 1), hidden_fraction_of_k1=round(max(0.0, min(1.0, (t_g + t_k - t_b) / t_k)), 3),
         gemm_slowdown=round(t_b / t_g, 3))

    # ---- K3: all-gather fused into the GEMM vs the same GEMM on local weights ------------------
    from tony_b200.parallel.ps import ShardedLinear

    M, N, K = 16384, 4096, 3344
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    for copy_ctas in (4, 8, 24):
        lin = ShardedLinear(K, N, act="relu", device=dev, copy_ctas=copy_ctas)
        tr = runtime.Trainer(lin, lr=1e-3, device=dev, ps_sharded=[lin.weight])
        with torch.no_grad():
            t_ag = timed(lambda: lin(x), iters=5)
            t_plain = timed(lambda: F.linear_act(x, lin.weight, lin.bias, "relu"), iters=5)
        wbytes = 2.0 * N * K * (world - 1) / world
        t_link = wbytes / (LINK_GBS * 1e3)
        t_flop = 2.0 * M * N * K / (PEAKS["bf16_tflops"] * 1e6)
        roof = max(t_link, t_flop)
        emit(kernel="K3 all-gather + GEMM 16384x4096x3344 (+bias+ReLU)", world=world, copy_ctas=copy_ctas,
             us=round(t_ag, 1), same_gemm_local_weights_us=round(t_plain, 1),
             weight_bytes_over_nvlink_mb=round(wbytes / 2 ** 20, 1), nvlink_term_us=round(t_link, 1),
             compute_term_us=round(t_flop, 1), roofline_us=round(roof, 1), fraction_of_roofline=round(roof / t_ag, 3))
        del tr, lin
        sync()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
