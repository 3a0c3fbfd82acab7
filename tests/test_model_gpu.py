"""End-to-end model checks on a GPU: our BERT/GPT-2 vs the stock-PyTorch twin, and the
fused optimizer step through the public Trainer API."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _copy_bert(mine, ref):
    with torch.no_grad():
        def lin(a, b):
            b.weight.copy_(a.weight.float()); b.bias.copy_(a.bias.float())
        ref.word.weight.copy_(mine.word.weight.float())
        ref.pos.weight.copy_(mine.pos.weight.float())
        ref.tok_type.weight.copy_(mine.tok_type.weight.float())
        lin(mine.emb_ln, ref.emb_ln)
        for a, b in zip(mine.layers, ref.layers):
            lin(a.attn.qkv, b.attn.qkv); lin(a.attn.out, b.attn.out); lin(a.ln1, b.ln1)
            lin(a.mlp.fc1, b.fc1); lin(a.mlp.fc2, b.fc2); lin(a.ln2, b.ln2)
        lin(mine.mlm_dense, ref.mlm_dense); lin(mine.mlm_ln, ref.mlm_ln)
        ref.mlm_bias.copy_(mine.mlm_bias.float())
        lin(mine.pooler, ref.pooler)
        ref.nsp.weight.copy_(mine.nsp.weight[:2].float()); ref.nsp.bias.copy_(mine.nsp.bias[:2].float())


def test_bert_matches_stock_forward_backward():
    from tony_b200.models import bert, stock

    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = bert.BertConfig.tiny()
    mine = bert.BertForPreTraining(cfg, device=dev)
    ref = stock.StockBertForPreTraining(cfg).to(dev)
    _copy_bert(mine, ref)
    batch = bert.synthetic_batch(cfg, 4, 128, device=dev)
    loss = mine(**batch)
    loss.backward()
    loss_ref = ref(**batch)
    loss_ref.backward()
    assert abs(loss.item() - loss_ref.item()) < 3e-2 * abs(loss_ref.item())
    for got, want in [(mine.layers[0].mlp.fc1.weight.grad, ref.layers[0].fc1.weight.grad),
                      (mine.layers[1].attn.qkv.weight.grad, ref.layers[1].attn.qkv.weight.grad),
                      (mine.layers[0].ln1.weight.grad, ref.layers[0].ln1.weight.grad),
                      (mine.word.weight.grad, ref.word.weight.grad)]:
        err = (got.float() - want).abs().max().item()
        assert err <= 6e-2 * want.abs().max().item() + 1e-4, err


def test_trainer_fused_step_decreases_loss():
    from tony_b200 import runtime
    from tony_b200.models import bert
    from tony_b200.ops import _lib

    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = bert.BertConfig.tiny()
    tr = runtime.Trainer(bert.BertForPreTraining(cfg, device=dev), lr=1e-3, device=dev, bucket_mb=0.5)
    assert len(tr.store.buckets) > 1
    batch = bert.synthetic_batch(cfg, 4, 128, pin=True)
    n0 = _lib.launch_count()
    losses = [tr.step(batch) for _ in range(6)]
    tr.check()
    assert _lib.launch_count() > n0
    assert losses[-1] < losses[0] - 0.1, losses


def test_trainer_matches_torch_adamw_one_step():
    """Fused bucket step == torch AdamW on the same bf16 gradients (single GPU)."""
    from tony_b200 import runtime
    from tony_b200.models import gpt2

    dev = torch.device("cuda")
    torch.manual_seed(1)
    cfg = gpt2.GPT2Config.tiny()
    model = gpt2.GPT2LMHead(cfg, device=dev)
    tr = runtime.Trainer(model, lr=1e-2, weight_decay=0.0, device=dev, bucket_mb=0.25, overlap=False)
    p = model.blocks[0].mlp.fc1.weight
    before = p.detach().float().clone()
    batch = gpt2.synthetic_batch(cfg, 2, 64, device=dev)
    loss = model(**batch)
    loss.backward()
    g = tr.store.slot_of(p).grad_view.float().clone()
    tr.opt.step()
    torch.cuda.synchronize()
    ref = before.clone().requires_grad_(True)
    opt = torch.optim.AdamW([ref], lr=1e-2, weight_decay=0.0)
    ref.grad = g
    opt.step()
    err = (p.detach().float() - ref.detach()).abs().max().item()
    assert err <= 1e-2 * ref.detach().abs().max().item() + 1e-3, err


def test_widedeep_ps_mode_single_gpu_trains():
    """PS path at world=1: per-weight sharded buckets + all-gather-fused GEMM (copy CTAs idle)."""
    from tony_b200 import runtime
    from tony_b200.models import widedeep
    from tony_b200.parallel.ps import ps_parameters

    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = widedeep.WideDeepConfig.tiny()
    model = widedeep.WideDeep(cfg, device=dev)
    tr = runtime.Trainer(model, lr=3e-3, device=dev, bucket_mb=0.25, ps_sharded=ps_parameters(model))
    assert sum(b.ps_sharded for b in tr.store.buckets) == len(cfg.hidden)
    batch = widedeep.synthetic_batch(cfg, 512, pin=True)
    losses = [tr.step(batch) for _ in range(8)]
    tr.check()
    assert losses[-1] < losses[0] - 0.05, losses


def test_allgather_gemm_matches_plain_gemm_single_gpu():
    from tony_b200.parallel.ps import ShardedLinear
    from tony_b200 import runtime

    dev = torch.device("cuda")
    torch.manual_seed(0)
    lin = ShardedLinear(512, 1024, act="relu", device=dev)
    lin.weight.data.normal_(0, 0.05)
    tr = runtime.Trainer(lin, lr=1e-3, device=dev, ps_sharded=[lin.weight])
    x = (torch.randn(300, 512, device=dev)).to(torch.bfloat16)
    for _ in range(3):   # repeated calls exercise the per-slab epoch flags
        y = lin(x)
    ref = torch.relu(x.float() @ lin.weight.float().t() + lin.bias.float())
    err = (y.float() - ref).abs().max().item()
    assert err <= 2e-2 * ref.abs().max().item() + 1e-3, err


def test_resnet50_conv1x1_and_generic_ddp_hooks():
    from tony_b200.models import resnet
    from tony_b200.parallel.ddp import DistributedDataParallel

    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = DistributedDataParallel(resnet.ResNet50(num_classes=16, device=dev), device=dev, optimizer="sgd",
                                    lr=0.05, momentum=0.9, weight_decay=1e-4, bucket_mb=8.0)
    batch = resnet.synthetic_batch(8, dev, size=64, classes=16)
    losses = []
    for _ in range(6):
        loss = model(**batch)
        loss.backward()
        model.step()
        losses.append(loss.item())
    model.store.comm.check_status()
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses


@pytest.mark.parametrize("fused_layers", [False, True])
def test_ddp_no_sync_accumulates_like_one_big_batch(fused_layers):
    """Two micro-batches under no_sync() + one synchronising backward must move the parameters like a single
    backward over the three batches at once; covers the hook path (stock nn.Linear) and the producers that
    write straight into the flat bucket (models.nn.Linear + LayerNorm), and a step after it starts clean."""
    import copy

    from tony_b200.models import nn as tnn
    from tony_b200.parallel.ddp import DistributedDataParallel

    dev = torch.device("cuda")
    torch.manual_seed(5)
    if fused_layers:
        net = torch.nn.Sequential(tnn.Linear(64, 256, device=dev), tnn.LayerNorm(256, device=dev),
                                  tnn.Linear(256, 32, device=dev))
    else:
        net = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.Tanh(), torch.nn.Linear(128, 32)).to(dev)
    a = DistributedDataParallel(copy.deepcopy(net), device=dev, optimizer="sgd", lr=0.05, momentum=0.0,
                                weight_decay=0.0, bucket_mb=0.01)
    b = DistributedDataParallel(copy.deepcopy(net), device=dev, optimizer="sgd", lr=0.05, momentum=0.0,
                                weight_decay=0.0, bucket_mb=0.01)
    xs = [torch.randn(256, 64, device=dev).to(torch.bfloat16) for _ in range(3)]
    start = a.store.param_buf.clone()
    for _round in range(2):
        with a.no_sync():
            for x in xs[:2]:
                (a(x).float().pow(2).mean() / 3).backward()
        (a(xs[2]).float().pow(2).mean() / 3).backward()
        a.step()
        b(torch.cat(xs)).float().pow(2).mean().backward()      # one backward over the same 768 rows
        b.step()
        torch.cuda.synchronize()
        moved = (b.store.param_buf.float() - start.float()).abs().max().item()
        diff = (a.store.param_buf.float() - b.store.param_buf.float()).abs().max().item()
        assert moved > 0 and diff <= 0.05 * moved + 2e-3, (diff, moved)
    a.store.comm.check_status()


def test_checkpoint_roundtrip(tmp_path):
    from tony_b200 import runtime
    from tony_b200.models import gpt2
    from tony_b200.runtime import checkpoint

    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = gpt2.GPT2Config.tiny()
    tr = runtime.Trainer(gpt2.GPT2LMHead(cfg, device=dev), lr=1e-3, device=dev, bucket_mb=0.5)
    batch = gpt2.synthetic_batch(cfg, 2, 64, pin=True)
    for _ in range(3):
        tr.step(batch)
    checkpoint.save(str(tmp_path), tr)
    want_params = tr.store.param_buf.clone()
    want_m = tr.opt.state1.clone()
    next_loss = tr.step(batch)
    tr.store.param_buf.zero_()
    tr.opt.state1.zero_()
    assert checkpoint.load_latest(str(tmp_path), tr) == 3
    assert torch.equal(tr.store.param_buf, want_params) and torch.equal(tr.opt.state1, want_m)
    assert abs(tr.step(batch) - next_loss) < 1e-3


def test_cuda_graph_mode_matches_eager():
    """Trainer(cuda_graph=True) replays forward+backward as one graph; parameters after a few
    steps must match the eager trainer fed the same batches."""
    import copy

    from tony_b200 import runtime
    from tony_b200.models import bert

    torch.manual_seed(11)
    cfg = bert.BertConfig(vocab_size=1000, hidden=256, layers=2, heads=4, inner=1024, max_positions=128,
                          max_predictions=8)
    base = bert.BertForPreTraining(cfg, device="cuda")
    twin = copy.deepcopy(base)
    eager = runtime.Trainer(base, lr=1e-3, cuda_graph=False)
    graph = runtime.Trainer(twin, lr=1e-3, cuda_graph=True)
    batches = [bert.synthetic_batch(cfg, 4, 128, device="cuda", seed=s) for s in range(3)]
    losses_e, losses_g = [], []
    for i in range(6):
        losses_e.append(float(eager.step_device(batches[i % 3]).item()))
        losses_g.append(float(graph.step_device(batches[i % 3]).item()))
    assert graph._graph is not None and graph._graph_kernels > 20
    for a, b in zip(losses_e, losses_g):
        assert abs(a - b) < 2e-2 * max(1.0, abs(a)), (losses_e, losses_g)
    pe, pg = eager.store.param_buf.float(), graph.store.param_buf.float()
    assert (pe - pg).abs().max().item() < 2e-2 * pe.abs().max().item()


def test_checkpoint_roundtrip_includes_sharded_embedding_bag(tmp_path):
    """save / load_latest must carry the row-sharded tables (state outside the flat store)."""
    from tony_b200 import runtime
    from tony_b200.models import widedeep
    from tony_b200.parallel.ps import ps_parameters
    from tony_b200.runtime import checkpoint

    dev = torch.device("cuda")
    cfg = widedeep.WideDeepConfig.tiny()

    def make():
        torch.manual_seed(0)
        m = widedeep.WideDeep(cfg, device=dev)
        return m, runtime.Trainer(m, lr=3e-3, device=dev, bucket_mb=0.25, ps_sharded=ps_parameters(m))

    m1, t1 = make()
    batch = widedeep.synthetic_batch(cfg, 256, pin=True)
    for _ in range(3):
        t1.step(batch)
    checkpoint.save(str(tmp_path), t1)
    nxt = t1.step(batch)
    m2, t2 = make()
    assert checkpoint.load_latest(str(tmp_path), t2) == 3
    assert abs(t2.step(batch) - nxt) < 2e-2 * max(1.0, abs(nxt))


@pytest.mark.parametrize("with_attention", [False, True])
def test_overlap_launch_does_not_race_with_dgrad(with_attention, monkeypatch):
    """overlap=True launches a bucket's fused reduce+optimizer from inside backward.  That kernel
    rewrites the bf16 weights, so every dgrad GEMM reading a weight must already be enqueued when
    the weight's gradient slot is finished.  Tiny buckets make every bucket close on a single fc1 /
    fc2 / qkv / out weight; the parameters must then be bit-identical to the overlap=False run."""
    import copy

    from torch import nn

    from tony_b200 import runtime
    from tony_b200.models import bert

    if with_attention:
        monkeypatch.setenv("TONY_B200_ATTENTION", "tcgen05")   # own kernels: no atomics, deterministic
    dev = torch.device("cuda")
    cfg = bert.BertConfig(vocab_size=1000, hidden=256, layers=3, heads=4, inner=1024, max_positions=128)

    class Stack(nn.Module):
        def __init__(self):
            super().__init__()
            self.layers = nn.ModuleList(bert.BertLayer(cfg, device=dev) for _ in range(cfg.layers))

        def forward(self, x):
            for layer in self.layers:
                x = layer(x) if with_attention else layer.ln2(layer.mlp(x, residual=x))
            return x.float().square().mean()

    torch.manual_seed(5)
    base = Stack()
    twin = copy.deepcopy(base)
    a = runtime.Trainer(base, lr=1e-2, device=dev, bucket_mb=0.01, overlap=True)
    b = runtime.Trainer(twin, lr=1e-2, device=dev, bucket_mb=0.01, overlap=False)
    assert len(a.store.buckets) >= 4 * cfg.layers
    xs = [torch.randn(4, 128, cfg.hidden, device=dev).to(torch.bfloat16) for _ in range(2)]
    for i in range(5):
        a.step_device({"x": xs[i % 2]})
        b.step_device({"x": xs[i % 2]})
    torch.cuda.synchronize()
    a.check()
    assert torch.equal(a.store.param_buf, b.store.param_buf)
