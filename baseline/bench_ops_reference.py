# -*- coding: utf-8 -*-
"""Reference arm of the operator benchmarks: the UNMODIFIED reference ops (baseline/_ref) on
torch.distributed NCCL through the horovod/mpi4py stand-ins (baseline/shim), timed like
benchmarks/benchmark.py (CUDA events, warm-up, max over ranks).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node N baseline/bench_ops_reference.py \
        --mode nt|all|tn|attn --length T --dim d [--heads H] [--offset 1000] [--backward]
"""
import argparse
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path[:] = [p for p in sys.path if os.path.abspath(p or '.') != ROOT]
sys.path.insert(0, os.path.join(HERE, '_ref'))
sys.path.insert(0, os.path.join(HERE, 'shim'))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import horovod.torch as hvd  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--mode', default='nt')
    ap.add_argument('--length', type=int, default=8192)
    ap.add_argument('--dim', type=int, default=768)
    ap.add_argument('--heads', type=int, default=1)
    ap.add_argument('--offset', type=int, default=1000)
    ap.add_argument('--backward', action='store_true')
    ap.add_argument('--iters', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    args = ap.parse_args()
    hvd.init()
    world, rank = hvd.size(), hvd.rank()
    dev = torch.device('cuda', torch.cuda.current_device())
    torch.backends.cuda.matmul.allow_tf32 = True
    from distributed_dot_product.multiplication.ops import (
        FullMultiplication, LeftTransposeMultiplication, RightTransposeMultiplication)
    from distributed_dot_product.module import DistributedDotProductAttn
    torch.set_grad_enabled(args.backward)
    tn = args.length // world
    t = tn * world
    d = args.dim
    lead = (1,) if args.heads == 1 else (1, args.heads)

    def mk(*s):
        return torch.rand(*s, device=dev).requires_grad_(args.backward)

    try:
        if args.mode == 'nt':
            a, b = mk(*lead, tn, d), mk(*lead, tn, d)
            op = lambda: RightTransposeMultiplication.apply(a, b, args.offset)  # noqa: E731
        elif args.mode in ('all', 'tn'):
            a, b = mk(*lead, tn, t), mk(*lead, tn, d)
            cls = FullMultiplication if args.mode == 'all' else LeftTransposeMultiplication
            op = lambda: cls.apply(a, b, args.offset)  # noqa: E731
        else:
            model = DistributedDotProductAttn(d, num_heads=args.heads, offset=args.offset).to(dev)
            a = mk(1, tn, d)
            mask = torch.zeros(1, tn, t, device=dev).bool()
            op = lambda: model(a, a, a, mask)  # noqa: E731

        def run():
            out = op()
            if args.backward:
                out.backward(torch.ones_like(out))

        for _ in range(args.warmup):
            run()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            run()
        e1.record()
        e1.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / args.iters], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        rec = {'impl': 'reference', 'mode': args.mode, 'world_size': world, 'T': t, 'd': d, 'heads': args.heads,
               'backward': args.backward, 'dtype': 'fp32 (tf32 matmul)', 'ms': ms.item()}
    except torch.cuda.OutOfMemoryError:
        rec = {'impl': 'reference', 'mode': args.mode, 'world_size': world, 'T': t, 'd': d, 'unavailable': 'CUDA OOM'}
    if rank == 0:
        print(json.dumps(rec))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
