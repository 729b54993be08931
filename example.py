# -*- coding: utf-8 -*-
"""Usage example (same script as the reference's example.py:1-33, launched with torchrun instead of horovodrun):

    python example.py                                   # one process
    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 example.py
"""
import torch
import torch.nn as nn

from distributed_dot_product.utils.comm import get_rank, get_world_size, get_local_rank, shutdown
from distributed_dot_product.module import DistributedDotProductAttn

torch.manual_seed(111)

device = torch.device('cpu')

if torch.cuda.is_available():
    torch.cuda.set_device(get_local_rank())
    device = torch.device('cuda')

module = DistributedDotProductAttn(768, num_heads=2, offset=64)
module.to(device)

criterion = nn.MSELoss()

length = 4096
world_size = get_world_size()
x = torch.rand(1, length // world_size, 768, device=device)
y = torch.rand(1, length // world_size, 768, device=device)
mask = torch.zeros(1, length // world_size, length, device=device).bool()

out = module(x, x, x, mask)
loss = criterion(out, y)
loss.backward()
if get_rank() == 0:
    print('loss %.6f, out %s, grad norm %.4f' % (loss.item(), tuple(out.shape),
                                                  module.keys.weight.grad.norm().item()))
shutdown()
