# -*- coding: utf-8 -*-
"""Short driver for ncu: a few fwd+bwd steps of the flagship layer on one GPU (T = T/N = 8192, 16 heads,
d_model 1024, bf16).  Used as
    python tools/profile_step.py 3 > plain.log && \
    ncu --set full --clock-control none --import-source on -k regex:ddp -s 40 -c 20 -o gpurun_out/prof \
        python tools/profile_step.py 3
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product.module import DistributedDotProductAttn  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
t = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
dev = torch.device('cuda', 0)
torch.manual_seed(0)
model = DistributedDotProductAttn(1024, num_heads=16, offset=1024).to(dev, torch.bfloat16)
x = torch.rand(1, t, 1024, device=dev, dtype=torch.bfloat16)
mask = torch.zeros(1, t, t, device=dev).bool()
for _ in range(steps):
    for p in model.parameters():
        p.grad = None
    out = model(x, x, x, mask)
    out.float().pow(2).mean().backward()
torch.cuda.synchronize()
print('ok', float(out.float().abs().mean()))
