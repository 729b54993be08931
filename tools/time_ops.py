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

# ---- slab-free attention kernels on the same operands (no slab traffic: report algorithmic TFLOP/s instead)
import distributed_dot_product.ops.attention as A  # noqa: E402


def timed_flops(name, fn, flops, iters=5):
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
                          'TFLOPs': round(flops / ms.item() * 1e-9, 1)}), flush=True)


del slab, slab2
v4 = q4
g4 = k4
pair = 2.0 * H * TN * T * D  # one 128x128xd product over the whole block of the score matrix
stat_ld = (TN + 63) // 64 * 64
neg_lse = torch.zeros(1, H, stat_ld, device=dev)
nds = torch.zeros(1, H, stat_ld, device=dev)
dk = torch.empty_like(k)
dq = torch.empty_like(k)
dv = torch.empty_like(k)
h4 = lambda t: t.view(1, TN, H, D).transpose(1, 2)  # noqa: E731
timed_flops('attn_fwd (slab-free)', lambda: A.flash_attention_forward(k4, q4, v4, mask, 0.125, 1024, out4), 2 * pair)
lse = A.flash_attention_forward(k4, q4, v4, mask, 0.125, 1024, out4, want_lse=True)
neg_lse[..., :TN] = -lse
timed_flops('attn_bwd rows + cols (slab-free)',
            lambda: A.flash_attention_backward(k4, q4, v4, g4, neg_lse, nds, mask, 0.125, 1024, h4(dk), h4(dq), h4(dv)),
            7 * pair)
comm.synchronize()
comm.shutdown()
