"""Multi-rank worker (launched by torchrun) validating the fused collective kernels against
torch.distributed/NCCL + a plain fp32 AdamW reference.  Prints 'MP_OK' on rank 0 on success."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from tony_b200 import runtime
from tony_b200.ops import _lib
from tony_b200.parallel.symm import SymmComm


def main():
    rank, world, dev = runtime.init()
    torch.manual_seed(1234 + rank)
    n = world * 8 * 1000 * 37          # odd-ish size, multiple of world*8
    comm = SymmComm(n, dev)

    # ---- 1. symmetric all-reduce vs NCCL -------------------------------------------------
    g = (torch.randn(n, device=dev) * 0.5).to(torch.bfloat16)
    comm.grad.copy_(g)
    ref = g.float().clone()
    dist.all_reduce(ref)
    torch.cuda.synchronize()
    dist.barrier()
    comm.allreduce_(0, n, scale=1.0 / world, num_ctas=16)
    torch.cuda.synchronize()
    comm.check_status()
    err = (comm.param.float() - ref / world).abs().max().item()
    assert err <= 2e-2 * (ref / world).abs().max().item() + 1e-3, f"allreduce err {err}"

    # partial range with offset
    comm.param.zero_()
    dist.barrier()
    lo, cnt = world * 8 * 100, world * 8 * 500
    comm.allreduce_(lo, cnt, scale=1.0, num_ctas=4)
    torch.cuda.synchronize()
    assert (comm.param[:lo].abs().max().item() == 0) and (comm.param[lo + cnt:].abs().max().item() == 0)
    err = (comm.param[lo:lo + cnt].float() - ref[lo:lo + cnt]).abs().max().item()
    assert err <= 2e-2 * ref.abs().max().item() + 1e-3, f"ranged allreduce err {err}"

    # ---- 2. device barrier, many epochs ------------------------------------------------------
    for _ in range(50):
        comm.barrier()
    torch.cuda.synchronize()
    comm.check_status()

    # ---- 3. fused reduce + sharded AdamW + broadcast vs reference ---------------------------
    import ctypes

    torch.manual_seed(7)                       # same initial params on every rank
    p0 = torch.randn(n, device=dev) * 0.1
    comm.param.copy_(p0.to(torch.bfloat16))
    chunk = n // world
    master = comm.param[rank * chunk:(rank + 1) * chunk].float().clone()
    m = torch.zeros_like(master)
    v = torch.zeros_like(master)
    ref_p = comm.param.float().clone().requires_grad_(True)
    opt = torch.optim.AdamW([ref_p], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01)
    lib = _lib.lib()
    for step in range(1, 4):
        torch.manual_seed(100 * step + rank)
        g = (torch.randn(n, device=dev) * 0.3).to(torch.bfloat16)
        comm.grad.copy_(g)
        gsum = g.float().clone()
        dist.all_reduce(gsum)
        ref_p.grad = gsum / world
        opt.step()
        torch.cuda.synchronize()
        dist.barrier()
        _lib.check(lib.tony_fused_reduce_optim(
            ctypes.byref(comm.c), 0, n, _lib.ptr(master), _lib.ptr(m), _lib.ptr(v), 0, _lib.OPT_ADAMW,
            1e-2, 0.9, 0.95, 1e-8, 0.01, 1 - 0.9 ** step, 1 - 0.95 ** step, 1.0 / world, None, 1,
            comm.next_epoch(), 16, _lib.cur_stream()), "fused")
        torch.cuda.synchronize()
        comm.check_status()
        err = (comm.param.float() - ref_p.detach()).abs().max().item()
        tol = 1.5e-2 * ref_p.detach().abs().max().item() + 2e-3
        assert err <= tol, f"step {step}: fused param err {err} > {tol}"
    # every rank holds identical parameters
    mine = comm.param.float().sum()
    lo_, hi_ = mine.clone(), mine.clone()
    dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
    assert lo_.item() == hi_.item(), "ranks diverged"

    # ---- 4. end-to-end: Trainer with overlap on a tiny BERT, loss identical across ranks -------
    from tony_b200.models import bert

    torch.manual_seed(0)
    cfg = bert.BertConfig.tiny()
    model = bert.BertForPreTraining(cfg, device=dev)
    tr = runtime.Trainer(model, lr=1e-3, device=dev, bucket_mb=0.5, overlap=True, comm_ctas=8)
    batch = bert.synthetic_batch(cfg, 4, 128, pin=True, seed=rank)
    losses = [tr.step(batch) for _ in range(5)]
    tr.check()
    assert losses[-1] < losses[0], losses
    chk = tr.store.param_buf.float().sum()
    a, b = chk.clone(), chk.clone()
    dist.all_reduce(a, op=dist.ReduceOp.MIN)
    dist.all_reduce(b, op=dist.ReduceOp.MAX)
    assert a.item() == b.item(), "trainer params diverged across ranks"

    # ---- 5. parameter-server path: sharded buckets + all-gather fused into the GEMM ---------
    from tony_b200.models import widedeep
    from tony_b200.parallel.ps import ShardedLinear, ps_parameters

    torch.manual_seed(3)
    lin = ShardedLinear(512, 1024, act="relu", device=dev, copy_ctas=8)
    lin.weight.data.normal_(0, 0.05)
    trl = runtime.Trainer(lin, lr=1e-3, device=dev, ps_sharded=[lin.weight])
    full = trl.store.param_buf.clone()                       # identical on all ranks after broadcast
    w_ref = lin.weight.detach().float().clone()
    b_ref = lin.bias.detach().float().clone()
    # wipe every row this rank does not own: the fused gather must bring them back
    rows = 1024 // world
    keep = lin.weight.data[rank * rows:(rank + 1) * rows].clone()
    lin.weight.data.zero_()
    lin.weight.data[rank * rows:(rank + 1) * rows].copy_(keep)
    torch.cuda.synchronize()
    dist.barrier()
    x = torch.randn(300, 512, device=dev).to(torch.bfloat16)
    for _ in range(3):
        y = lin(x)
    torch.cuda.synchronize()
    ref = torch.relu(x.float() @ w_ref.t() + b_ref)
    err = (y.float() - ref).abs().max().item()
    assert err <= 2e-2 * ref.abs().max().item() + 1e-3, f"allgather-gemm err {err}"
    assert torch.equal(lin.weight.detach().float(), w_ref), "gathered weight differs from the original"
    dist.barrier()

    torch.manual_seed(0)
    wcfg = widedeep.WideDeepConfig.tiny()
    wmodel = widedeep.WideDeep(wcfg, device=dev)
    wtr = runtime.Trainer(wmodel, lr=3e-3, device=dev, bucket_mb=0.25, ps_sharded=ps_parameters(wmodel), comm_ctas=8)
    wb = widedeep.synthetic_batch(wcfg, 512, pin=True, seed=rank)
    wl = [wtr.step(wb) for _ in range(8)]
    wtr.check()
    assert wl[-1] < wl[0], wl
    # after a final gather every rank must hold the same tower weights
    with torch.no_grad():
        wmodel(**wtr.to_device(wb))
    torch.cuda.synchronize()
    for p_ in ps_parameters(wmodel):
        sm = p_.detach().float().sum()
        a2, b2 = sm.clone(), sm.clone()
        dist.all_reduce(a2, op=dist.ReduceOp.MIN)
        dist.all_reduce(b2, op=dist.ReduceOp.MAX)
        assert a2.item() == b2.item(), "ps-sharded weights diverged after gather"

    # ---- 5b. row-sharded embedding bag (K9): lookups and gradient reductions over peer memory ----
    from tony_b200.parallel.ps import ShardedEmbeddingBag

    Fd, V, D, R, B = 3, 101, 16, 8, 96           # vocab not divisible by the world size on purpose
    bag = ShardedEmbeddingBag(Fd, V, D, R, std=0.5, device=dev, seed=21)
    bag.tony_configure(lr=0.05)
    gen = torch.Generator(device=dev)
    full = []
    for f in range(Fd):
        gen.manual_seed(21 + f)
        full.append((torch.randn(V, D + R, generator=gen, device=dev) * 0.5).to(torch.bfloat16).float())
    torch.manual_seed(100 + rank)
    ids = torch.randint(0, V, (B, Fd), device=dev)
    deep, rest = bag(ids)
    ref_deep = torch.cat([full[f][ids[:, f], :D] for f in range(Fd)], dim=1)
    ref_rest = torch.stack([full[f][ids[:, f], D:] for f in range(Fd)], dim=1)
    assert torch.equal(deep.float(), ref_deep) and torch.equal(rest.float(), ref_rest), "bag gather mismatch"
    gd = (torch.randn(B, Fd * D, device=dev) * 0.1).to(torch.bfloat16)
    gr = (torch.randn(B, Fd, R, device=dev) * 0.1).to(torch.bfloat16)
    torch.autograd.backward([deep, rest], [gd, gr])
    ref_grad = torch.zeros(Fd, V, D + R, device=dev)
    for f in range(Fd):
        ref_grad[f].index_add_(0, ids[:, f], torch.cat([gd[:, f * D:(f + 1) * D], gr[:, f]], dim=1).float())
    dist.all_reduce(ref_grad)                      # what every owner should have received in total
    bag.barrier()                                  # every rank's NVLink reductions have landed
    torch.cuda.synchronize()
    n_el = Fd * bag.rows_per_shard * bag.row_elems
    got = bag.grad[:n_el].view(Fd, bag.rows_per_shard, D + R).float()
    mine = ref_grad[:, rank::world]
    err = (got[:, : mine.shape[1]] - mine).abs().max().item()
    assert err <= 3e-2 * mine.abs().max().item() + 1e-3, f"bag scatter err {err}"
    before = bag.full_table(0).float().clone()
    bag.tony_step()
    deep2, _ = bag(ids)                            # pulls freshly updated rows from the owners
    torch.cuda.synchronize()
    assert (bag.full_table(0).float() - before).abs().max().item() > 1e-3, "owner shard did not move"
    assert torch.isfinite(deep2.float()).all()
    dist.barrier()

    # ---- 5b. dedicated parameter servers: only the LAST rank owns the bucket --------------------
    torch.manual_seed(11)
    n2 = 8 * 256 * 12
    pinit = torch.randn(n2, device=dev) * 0.1
    comm.param[:n2].copy_(pinit.to(torch.bfloat16))
    owner = world - 1
    is_owner = rank == owner
    m2 = comm.param[:n2].float().clone() if is_owner else torch.zeros(8, device=dev)
    s1, s2 = torch.zeros_like(m2), torch.zeros_like(m2)
    refp = comm.param[:n2].float().clone().requires_grad_(True)
    ropt = torch.optim.AdamW([refp], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.0)
    for step in range(1, 3):
        torch.manual_seed(300 * step + rank)
        g2 = (torch.randn(n2, device=dev) * 0.3).to(torch.bfloat16)
        if is_owner:
            g2.zero_()                       # a server computes no gradient
        comm.grad[:n2].copy_(g2)
        gs = g2.float().clone()
        dist.all_reduce(gs)
        refp.grad = gs / (world - 1)         # averaged over the workers
        ropt.step()
        torch.cuda.synchronize()
        dist.barrier()
        _lib.check(lib.tony_fused_reduce_optim_owners(
            ctypes.byref(comm.c), 0, n2, _lib.ptr(m2), _lib.ptr(s1), _lib.ptr(s2), 0, _lib.OPT_ADAMW, 1e-2, 0.9,
            0.95, 1e-8, 0.0, 1 - 0.9 ** step, 1 - 0.95 ** step, 1.0 / (world - 1), None, 1, 0 if is_owner else -1,
            comm.next_epoch(), 16, _lib.cur_stream()), "fused owners")
        torch.cuda.synchronize()
        comm.check_status()
    if is_owner:
        err = (comm.param[:n2].float() - refp.detach()).abs().max().item()
        assert err <= 1e-2 * refp.detach().abs().max().item() + 2e-3, f"owner-sharded update err {err}"
    dist.barrier()

    # ---- 6. hvd shim ---------------------------------------------------------------------------
    import tony_b200.parallel.hvd as hvd

    hvd.init()
    t = torch.full((1000,), float(rank + 1), device=dev)
    out = hvd.allreduce(t, average=False)
    assert abs(out[0].item() - world * (world + 1) / 2) < 1e-2
    bt = hvd.broadcast(torch.full((17,), float(rank + 5), device=dev), root_rank=0)
    assert abs(bt[3].item() - 5.0) < 1e-2

    # fp32 wire: values a bf16 wire would round (1 + 2^-12) must survive exactly
    fine = torch.full((4099,), 1.0 + 2.0 ** -12, device=dev)
    assert torch.equal(hvd.allreduce(fine, average=True), fine), "fp32 all-reduce lost precision"
    v = torch.randn(3, 1001, device=dev)
    ref32 = v.clone()
    dist.all_reduce(ref32)
    got32 = hvd.allreduce(v, average=False)
    assert (got32 - ref32).abs().max().item() <= 1e-5 * ref32.abs().max().item() + 1e-6
    # bf16 wire against NCCL
    vb = (torch.randn(5000, device=dev) * 0.5).to(torch.bfloat16)
    refb = vb.float().clone()
    dist.all_reduce(refb)
    gotb = hvd.allreduce(vb, average=False).float()
    assert (gotb - refb).abs().max().item() <= 2e-2 * refb.abs().max().item() + 1e-3
    assert (gotb - refb).abs().mean().item() <= 4e-3 * refb.abs().max().item() + 1e-4
    try:
        hvd.allreduce(torch.ones(4, dtype=torch.int64, device=dev))
        raise AssertionError("integer all-reduce must be refused")
    except TypeError:
        pass
    # broadcast moves raw bytes: integers, odd sizes, a non-zero root
    root = world - 1
    ib = torch.arange(1003, device=dev, dtype=torch.int64) * (rank + 1)
    gb = hvd.broadcast(ib, root_rank=root)
    assert torch.equal(gb, torch.arange(1003, device=dev, dtype=torch.int64) * (root + 1))
    # all-gather / reduce-scatter against NCCL
    part = torch.randn(7, 33, device=dev) + rank
    gathered = hvd.allgather(part)
    ref_parts = [torch.empty_like(part) for _ in range(world)]
    dist.all_gather(ref_parts, part)
    assert torch.equal(gathered, torch.cat(ref_parts))
    full = torch.randn(world * 16, 8, device=dev)
    rs = hvd.reducescatter(full)
    ref_full = full.clone()
    dist.all_reduce(ref_full)
    assert (rs - ref_full[rank * 16:(rank + 1) * 16]).abs().max().item() <= 1e-5 * ref_full.abs().max().item() + 1e-6

    # broadcast_parameters + DistributedOptimizer: ranks start from different weights, the root's win,
    # then identical fused steps keep every rank's parameters bit-identical and match an fp32 AdamW
    # fed the NCCL-averaged gradients
    torch.manual_seed(50 + rank)
    net = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 32)).to(dev).to(torch.bfloat16)
    hvd.broadcast_parameters(net.state_dict(), root_rank=0)
    w0 = net[0].weight.detach().float().clone()
    chk = w0.clone()
    dist.broadcast(chk, src=0)
    assert torch.equal(w0, chk), "broadcast_parameters did not deliver rank 0's weights"
    ref_net = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.ReLU(), torch.nn.Linear(128, 32)).to(dev).float()
    ref_net.load_state_dict({k: v.float() for k, v in net.state_dict().items()})
    w2_start = ref_net[2].weight.detach().clone()
    # SGD: the update is linear in the gradient, so bf16-vs-fp32 gradient noise stays proportionally small
    # (Adam's sign-like first steps turn a rounding difference of a tiny gradient into a full +-lr move)
    ref_opt = torch.optim.SGD(ref_net.parameters(), lr=0.05, momentum=0.9, weight_decay=0.0)
    hopt = hvd.DistributedOptimizer(net, optimizer="sgd", lr=0.05, momentum=0.9, weight_decay=0.0, bucket_mb=0.01,
                                    generic_hooks=True)
    for step in range(3):
        torch.manual_seed(900 + 10 * step + rank)
        xb = torch.randn(32, 64, device=dev)
        net(xb.to(torch.bfloat16)).float().square().mean().backward()
        hopt.step()
        hopt.synchronize()
        ref_net(xb).square().mean().backward()
        for q in ref_net.parameters():
            dist.all_reduce(q.grad)
            q.grad /= world
        ref_opt.step()
        ref_opt.zero_grad()
    torch.cuda.synchronize()
    hopt.check()
    mine = net[2].weight.detach().float()
    same = mine.clone()
    dist.broadcast(same, src=0)
    assert torch.equal(mine, same), "ranks diverged under DistributedOptimizer"
    refw = ref_net[2].weight.detach()
    moved = (refw - w2_start).abs().max().item()
    assert moved > 1e-4, "reference did not move: the comparison would be vacuous"
    err = (mine - refw).abs()
    assert err.max().item() <= 0.1 * moved + 1.5e-2 * refw.abs().max().item(), f"DistributedOptimizer max err {err.max().item()}"
    assert err.mean().item() <= 0.03 * moved + 4e-3 * refw.abs().max().item(), f"DistributedOptimizer mean err {err.mean().item()}"  # noqa

    dist.barrier()
    if rank == 0:
        from tony_b200.parallel.symm import symm_mode

        print(f"symm_mode={symm_mode()} nvls={'on' if comm.nvls is not None else 'off'}", flush=True)
        print("MP_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
