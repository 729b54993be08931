# -*- coding: utf-8 -*-
"""Prints which symmetric-heap back end the runtime picked on this box (one line per rank) and times the three
fused products with multicast pushes against unicast pushes.  Run under torchrun:
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/heap_report.py"""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product.utils import comm
from distributed_dot_product.parallel import runtime
from distributed_dot_product.multiplication.functions import distributed_matmul_nt

comm.init()
rank, world = comm.get_rank(), comm.get_world_size()
dev = torch.device('cuda', torch.cuda.current_device())
torch.manual_seed(3)
t, d = 8192, 768
a = torch.randint(-2, 3, (1, t // world, d), device=dev).to(torch.bfloat16)
out = distributed_matmul_nt(a, a, offset=256)
hp = runtime.heap()
import torch.distributed as dist
parts = [torch.empty_like(a) for _ in range(world)]
dist.all_gather(parts, a)
ref = a.float() @ torch.cat(parts, 1).float().transpose(-1, -2)
ok = bool((out.float() - ref).abs().max() <= 1e-2 * ref.abs().max())  # bf16 output: rounding only
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for _ in range(5):
    distributed_matmul_nt(a, a, offset=256)
torch.cuda.synchronize(); dist.barrier()
ev[0].record()
for _ in range(20):
    distributed_matmul_nt(a, a, offset=256)
ev[1].record(); torch.cuda.synchronize()
print('rank %d/%d: heap backend=%s multicast=%s note=%r correct=%s nt(T=%d,d=%d) %.3f ms'
      % (rank, world, hp.backend, bool(hp.mc_base), hp.heap_note, ok, t, d, ev[0].elapsed_time(ev[1]) / 20), flush=True)
runtime.check_status()
comm.shutdown()
