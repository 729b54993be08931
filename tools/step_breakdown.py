# -*- coding: utf-8 -*-
"""Per-kernel durations of ONE warm fwd+bwd step of the flagship layer on one GPU (torch.profiler / CUPTI activity
records: kernels run back to back as in production, unlike an ncu replay).  For shares, not bench values.

    python tools/step_breakdown.py [T] [dtype: bf16|mixed]
"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product.module import DistributedDotProductAttn  # noqa: E402

t = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
mode = sys.argv[2] if len(sys.argv) > 2 else 'bf16'
dev = torch.device('cuda', 0)
torch.manual_seed(0)
if mode == 'mixed':
    model = DistributedDotProductAttn(1024, num_heads=16, offset=1024, compute_dtype=torch.bfloat16).to(dev)
    dt = torch.float32
else:
    model = DistributedDotProductAttn(1024, num_heads=16, offset=1024).to(dev, torch.bfloat16)
    dt = torch.bfloat16
x = torch.rand(1, t, 1024, device=dev, dtype=dt)
y = torch.rand(1, t, 1024, device=dev, dtype=dt)
mask = torch.zeros(1, t, t, device=dev).bool()


def step():
    for p in model.parameters():
        p.grad = None
    out = model(x, x, x, mask)
    loss = torch.nn.functional.mse_loss(out.float(), y.float())
    loss.backward()


for _ in range(5):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    step()
e1.record()
torch.cuda.synchronize()
print('plain: %.3f ms/step' % (e0.elapsed_time(e1) / 10))
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
tot = 0.0
t_first = evs[0].time_range.start
t_last = max(e.time_range.end for e in evs)
prev_end = None
for e in evs:
    dur = e.time_range.end - e.time_range.start
    gap = (e.time_range.start - prev_end) if prev_end is not None else 0.0
    prev_end = e.time_range.end
    tot += dur
    print('%8.1f us  gap %6.1f  %s' % (dur, gap, e.name[:110]))
print('sum of kernels %.1f us, first start to last end %.1f us, %d kernels' % (tot, t_last - t_first, len(evs)))
