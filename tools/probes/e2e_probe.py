"""ResNet-50 step time with device-resident inputs vs the pipelined end-to-end call, the staging alone and a
raw 77 MB host-to-device copy (was the end-to-end path itself slow, or its first calls?)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from tony_b200 import runtime
from tony_b200.models import resnet
dev = torch.device('cuda')
torch.manual_seed(0)
model = resnet.ResNet50(device=dev)
tr = runtime.Trainer(model, lr=0.05, device=dev)
host = [resnet.synthetic_batch(256, 'cpu', True, s) for s in range(4)]
devb = {k: v.to(dev) for k, v in host[0].items()}
def timed(fn, n=8):
    torch.cuda.synchronize(); t=time.perf_counter(); fn(n); torch.cuda.synchronize(); return (time.perf_counter()-t)/n*1e3
def dev_loop(n):
    for i in range(n): tr.step_device(devb)
def async_loop(n):
    p=[]
    for i in range(n):
        p.append(tr.step_async(host[i%4]))
        if len(p)>2: p.pop(0).result()
    while p: p.pop(0).result()
def copy_only(n):
    for i in range(n): tr._stage_inputs(host[i%4])
for _ in range(2): dev_loop(3); async_loop(3)
print('device-resident step ms', round(timed(dev_loop),2))
print('step_async ms', round(timed(async_loop),2))
print('stage_inputs only ms', round(timed(copy_only),2))
x = host[0]['images']; d = torch.empty_like(x, device=dev)
def raw(n):
    for i in range(n): d.copy_(x, non_blocking=True)
print('raw H2D 77MB ms', round(timed(raw),2), 'pinned', x.is_pinned())
t=time.perf_counter()
for i in range(8): tr.step_device(devb)
print('host enqueue ms/step', round((time.perf_counter()-t)/8*1e3,2)); torch.cuda.synchronize()
