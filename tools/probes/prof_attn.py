import torch, sys, os
sys.path.insert(0, '/root/repo')
import os
os.environ['TONY_B200_NO_BUILD']='1'
from pathlib import Path
from tony_b200.ops import _lib as _l
_l.LIB_PATH = Path('/root/repo/tony_b200/lib/libtony_prof.so')
qkv = torch.randn(32, 512, 3, 16, 64, device='cuda').to(torch.bfloat16)
from tony_b200.ops import _lib
from tony_b200.ops._lib import ptr, check, cur_stream
out = torch.empty(32, 512, 1024, device='cuda', dtype=torch.bfloat16)
lse = torch.empty(32*16, 512, device='cuda', dtype=torch.float32)
for _ in range(3):
    check(_lib.lib().tony_attention_fwd_bf16(ptr(qkv), ptr(out), ptr(lse), 32, 512, 16, 64, 0, 0.125, cur_stream()), 'x')
torch.cuda.synchronize()
t = lse.view(torch.int64).reshape(-1)[:16*8].view(16, 8).cpu()
names = ['epilogue+lse(0)', 'wait s_full', 'tmem ld + free', 'max+vote', 'exp loop', 'wait pv_done', 'st P + arrive', 'loop ovh']
print('per warp totals (clk):')
for w in range(8):
    print(w, [int(x) for x in t[w]])
tot = t[:8].float().mean(0)
print('mean per block (55.4 blocks/CTA approx 7 items*8):')
for nme, v in zip(names, tot): print(f'{nme:20s} {v.item()/56:8.1f}')
print('sum', tot.sum().item())
