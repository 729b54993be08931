# -*- coding: utf-8 -*-
"""Per-kernel CUDA-event timings of the slab-free attention kernels at the flagship per-GPU shape (16 heads x 64,
T/N = 8192), any world size, max over ranks; every kernel is timed in isolation behind a barrier (no launch skew).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/time_flash_ops.py"""
import json
import math
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product.utils import comm
from distributed_dot_product.parallel import runtime
import distributed_dot_product.ops.attention as att

comm.init()
rank, world = comm.get_rank(), comm.get_world_size()
dev = torch.device('cuda', torch.cuda.current_device())
torch.manual_seed(5 + rank)
h, d, tn = 16, 64, 8192
t = tn * world
k, q, v, g = (torch.randn(1, tn, h * d, device=dev).to(torch.bfloat16) for _ in range(4))
mask = torch.zeros(1, tn, t, device=dev).bool()
scale = 1.0 / math.sqrt(d)
H = lambda x: att._heads(x, h)
bits, _, _ = att._pack_mask(mask, 1, h, tn, t, dev)
out = torch.empty_like(k)
dk, dq, dv = torch.empty_like(k), torch.empty_like(k), torch.empty_like(k)
lse = att.flash_attention_forward(H(k), H(q), H(v), mask, scale, 1024, H(out), want_lse=True, bits=bits)
stat_ld = tn
neg_lse = (-lse).contiguous()
nds = torch.zeros_like(neg_lse)
import ctypes
from distributed_dot_product import _native
import distributed_dot_product.ops.gemm as _g
lib = _native.load()


def fwd():
    att.flash_attention_forward(H(k), H(q), H(v), mask, scale, 1024, H(out), want_lse=True, bits=bits)


def bwd():
    att.flash_attention_backward(H(k), H(q), H(v), H(g), neg_lse, nds, mask, scale, 1024, H(dk), H(dq), H(dv), bits=bits)


def bwd_rows():
    att.flash_attention_backward(H(k), H(q), H(v), H(g), neg_lse, nds, mask, scale, 1024, H(dk), H(dq), H(dv), bits=bits,
                                 which=1)


def bwd_cols():
    att.flash_attention_backward(H(k), H(q), H(v), H(g), neg_lse, nds, mask, scale, 1024, H(dk), H(dq), H(dv), bits=bits,
                                 which=2)


def fwd_bwd(reuse):
    """forward, then both backward kernels reading the forward's gather in place (``reuse``) or gathering again"""
    def run():
        fwd()
        tag = None
        if reuse and world > 1:
            hp = runtime.heap()
            tag = (hp.generation, hp.seq)
        att.flash_attention_backward(H(k), H(q), H(v), H(g), neg_lse, nds, mask, scale, 1024, H(dk), H(dq), H(dv),
                                     bits=bits, gather_tag=tag)
    return run


def timed(fn, iters=5):
    for _ in range(2):
        fn()
    ms = 0.0
    for _ in range(iters):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ms += e0.elapsed_time(e1)
    x = torch.tensor([ms / iters], device=dev)
    if world > 1:
        dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return x.item()


res = {'world': world, 'heap': runtime.heap().backend if world > 1 else 'none',
       'multicast': bool(runtime.heap().mc_base) if world > 1 else False,
       'fwd_ms': timed(fwd), 'bwd_rows_plus_cols_ms': timed(bwd), 'bwd_rows_ms': timed(bwd_rows),
       'bwd_cols_ms': timed(bwd_cols), 'fwd_bwd_regather_ms': timed(fwd_bwd(False)),
       'fwd_bwd_gather_once_ms': timed(fwd_bwd(True))}
# the two backward kernels separately (dbg-free: call the C entry points through the same helper with one disabled)
if rank == 0:
    print(json.dumps(res), flush=True)
runtime.check_status()
comm.shutdown()
