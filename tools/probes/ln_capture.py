import os, sys, torch
sys.path.insert(0, '/root/repo')
from tony_b200.ops import functional as F
dev = torch.device('cuda')
x = torch.randn(4096, 1024, device=dev).to(torch.bfloat16).requires_grad_(True)
g = torch.ones(1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
b = torch.zeros(1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
dy = torch.randn_like(x)
y = F.layer_norm(x, g, b, 1e-5)
def f(): return torch.autograd.grad(y, (x, g, b), dy, retain_graph=True)
f(); torch.cuda.synchronize()
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    f()
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
gr = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(gr):
        out = f()
    gr.replay(); torch.cuda.synchronize()
    print(os.environ.get('TONY_B200_LN_BWD', 'fused'), os.environ.get('TONY_B200_PDL', '1'), 'capture OK')
except Exception as e:
    print(os.environ.get('TONY_B200_LN_BWD', 'fused'), os.environ.get('TONY_B200_PDL', '1'), 'capture FAILED', str(e)[:120])
