# -*- coding: utf-8 -*-
"""Per-kernel CUDA-event timing of the flagship layer's slab kernels at T/N = 8192 rows per GPU, any N
(torchrun).  Prints ms (max over ranks), slab GB moved and achieved TB/s per kernel - the multi-GPU counterpart
of the single-GPU ncu launch list (ncu must not wrap multi-rank commands)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product import _native  # noqa: E402
from distributed_dot_product.utils import comm  # noqa: E402
import distributed_dot_product.ops.gemm as G  # noqa: E402
from distributed_dot_product.ops.softmax import softmax_forward  # noqa: E402

comm.init()
world, rank = comm.get_world_size(), comm.get_rank()
dev = torch.device('cuda', torch.cuda.current_device())
H, TN, D = 16, 8192, 64
T = TN * world
torch.manual_seed(rank)
k = torch.randn(1, TN, H * D, device=dev).bfloat16()
q = torch.randn(1, TN, H * D, device=dev).bfloat16()
k4 = k.view(1, TN, H, D).transpose(1, 2)
q4 = q.view(1, TN, H, D).transpose(1, 2)
slab = torch.empty(1, H, TN, T, device=dev, dtype=torch.bfloat16)
slab2 = torch.empty_like(slab)
out = torch.empty(1, TN, H * D, device=dev, dtype=torch.bfloat16)
out4 = out.view(1, TN, H, D).transpose(1, 2)
delta = torch.randn(1, H, TN, device=dev)
mask = torch.zeros(1, TN, T, device=dev).bool()
gb = slab.numel() * 2 / 1e9


def timed(name, fn, passes, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({'kernel': name, 'n_gpus': world, 'T': T, 'ms': round(ms.item(), 4),
                          'slab_GB': round(gb * passes, 2), 'TBps': round(gb * passes / ms.item(), 2)}), flush=True)


timed('nt  K.gather(Q)^T -> S', lambda: G.gemm(_native.OP_NT, k4, q4, offset=1024, out=slab), 1)
timed('softmax in place', lambda: softmax_forward(slab, mask, 0.125, inplace=True), 2)
timed('all P.gather(V) -> O', lambda: G.gemm(_native.OP_NN, slab, q4, offset=1024, out=out4), 1)
timed('nt + fused softmax bwd -> dS', lambda: G.gemm(_native.OP_NT, k4, q4, offset=1024, out=slab2,
                                                   softmax_bwd=(slab, delta, 0.125)), 2)
timed('tn P^T.dO (reduce-scatter)', lambda: G.gemm(_native.OP_TN, slab, k4, out=out4), 1)
timed('all dS.gather(Q)', lambda: G.gemm(_native.OP_NN, slab2, q4, offset=1024, out=out4), 1)
comm.synchronize()
comm.shutdown()
