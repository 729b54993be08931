# -*- coding: utf-8 -*-
"""Operator benchmark CLI: same flags and JSON record keys as the reference's benchmark.py
(reference: benchmark.py:29-39 flags, :241-250 record), fixed where the reference's methodology is unusable
(SURVEY.md Q13): CUDA-event timing with synchronisation, warm-up, any world size (torchrun), max over ranks,
optional backward, optional dtype, and an ``attn`` mode for the whole layer.

    python benchmarks/benchmark.py --mode nt --offset 1000 --scale 1 --file nt_benchmark.json
    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 benchmarks/benchmark.py --mode all --backward

The record keeps the reference's eight keys (dense numbers measured on rank 0 with torch.matmul) and adds
``distributed_time_max``, ``world_size``, ``backward``, ``dtype``, ``T``, ``d``, ``iters``, ``tflops_per_gpu``.
"""
import argparse
import json
import os
import os.path as osp
import sys

import torch

sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))

from distributed_dot_product.utils.comm import (  # noqa: E402
    get_local_rank, get_rank, get_world_size, is_main_process, shutdown, synchronize)
from distributed_dot_product.multiplication.ops import (  # noqa: E402
    FullMultiplication, LeftTransposeMultiplication, RightTransposeMultiplication)
from distributed_dot_product.module import DistributedDotProductAttn  # noqa: E402

parser = argparse.ArgumentParser(description='Benchmark suite for distributed operations')
parser.add_argument('--mode', type=str, default='nt', help='Benchmark mode: nt | all | tn | attn')
parser.add_argument('--offset', type=int, default=1000, help='Offset used to balance time/memory')
parser.add_argument('--scale', type=int, default=1, help='Scale factor used for reduction')
parser.add_argument('--file', type=str, default='nt_benchmark.json', help='JSON file used to append results')
parser.add_argument('--length', type=int, default=75000, help='Total sequence length T before --scale')
parser.add_argument('--dim', type=int, default=768, help='Feature dimension d')
parser.add_argument('--heads', type=int, default=1)
parser.add_argument('--dtype', type=str, default='fp32', choices=['fp32', 'bf16', 'fp16'])
parser.add_argument('--backward', action='store_true', help='time forward + backward')
parser.add_argument('--iters', type=int, default=10)
parser.add_argument('--warmup', type=int, default=3)
parser.add_argument('--no-dense', action='store_true', help='skip the dense single-GPU torch.matmul run')


def timed(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
        synchronize()
        torch.cuda.reset_peak_memory_stats()
        start_mem = torch.cuda.max_memory_allocated()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        e1.synchronize()
        return e0.elapsed_time(e1) / 1e3 / iters, torch.cuda.max_memory_allocated() - start_mem
    import time
    synchronize()
    t0 = time.time()
    for _ in range(iters):
        fn()
    return (time.time() - t0) / iters, 0


def main():
    args = parser.parse_args()
    values = []
    if osp.exists(args.file):
        values = json.load(open(args.file, 'r'))
    torch.manual_seed(111)
    device = torch.device('cpu')
    if torch.cuda.is_available():
        torch.cuda.set_device(get_local_rank())
        device = torch.device('cuda')
    dtype = {'fp32': torch.float32, 'bf16': torch.bfloat16, 'fp16': torch.float16}[args.dtype]
    world = get_world_size()
    t_total = args.length // args.scale
    tn = t_total // world
    t_total = tn * world
    d = args.dim
    grad = args.backward
    torch.set_grad_enabled(grad)

    def mk(*shape):
        return torch.rand(*shape, device=device).to(dtype).requires_grad_(grad)

    lead = (1,) if args.heads == 1 else (1, args.heads)
    flops = 2.0 * tn * t_total * d * args.heads
    if args.mode == 'nt':
        left, right = mk(*lead, tn, d), mk(*lead, tn, d)
        op = lambda: RightTransposeMultiplication.apply(left, right, args.offset)  # noqa: E731
        dense = (lambda a, b: a.matmul(b.transpose(-1, -2)), (*lead, t_total, d), (*lead, t_total, d))
    elif args.mode in ('all', 'tn'):
        left, right = mk(*lead, tn, t_total), mk(*lead, tn, d)
        cls = FullMultiplication if args.mode == 'all' else LeftTransposeMultiplication
        op = lambda: cls.apply(left, right, args.offset)  # noqa: E731
        if args.mode == 'all':
            dense = (lambda a, b: a.matmul(b), (*lead, t_total, t_total), (*lead, t_total, d))
        else:
            dense = (lambda a, b: a.transpose(-1, -2).matmul(b), (*lead, t_total, t_total), (*lead, t_total, d))
    elif args.mode == 'attn':
        model = DistributedDotProductAttn(d, num_heads=args.heads, offset=args.offset).to(device, dtype)
        x = mk(1, tn, d)
        mask = torch.zeros(1, tn, t_total, device=device).bool()
        op = lambda: model(x, x, x, mask)  # noqa: E731
        dense = None
        flops = 2 * 2.0 * tn * t_total * d + 4 * 2.0 * tn * d * d
    else:
        raise SystemExit('unknown mode ' + args.mode)

    def run():
        out = op()
        if grad:
            out.backward(torch.ones_like(out))
            left_grads = [left, right] if args.mode != 'attn' else [x] + list(model.parameters())
            for t in left_grads:
                t.grad = None

    input_memory = torch.cuda.memory_allocated() if device.type == 'cuda' else 0
    dist_time, dist_peak = timed(run, args.iters, args.warmup)
    out = op()
    output_memory = out.numel() * out.element_size()
    del out

    rec = {'input_memory': None, 'total_time': None, 'peak_memory': None, 'output_memory': None}
    if dense is not None and not args.no_dense and is_main_process() and device.type == 'cuda':
        fn, sa, sb = dense
        need = (torch.Size(sa).numel() + torch.Size(sb).numel()) * (4 if dtype == torch.float32 else 2)
        free = torch.cuda.mem_get_info()[0]
        if need * 3 < free:
            torch.set_grad_enabled(False)
            a, b = torch.rand(*sa, device=device).to(dtype), torch.rand(*sb, device=device).to(dtype)
            rec['input_memory'] = torch.cuda.memory_allocated()
            dt, peak = 0.0, 0
            try:
                for _ in range(2):
                    fn(a, b)
                torch.cuda.synchronize()
                torch.cuda.reset_peak_memory_stats()
                m0 = torch.cuda.max_memory_allocated()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                r = fn(a, b)
                e1.record()
                e1.synchronize()
                dt, peak = e0.elapsed_time(e1) / 1e3, torch.cuda.max_memory_allocated() - m0
                rec['output_memory'] = r.numel() * r.element_size()
            except torch.cuda.OutOfMemoryError:
                dt, peak = None, None
            rec['total_time'], rec['peak_memory'] = dt, peak
            del a, b
    # gather per-rank numbers (reference: benchmark.py:104-113 gathers and averages)
    stats = torch.tensor([float(input_memory), dist_time, float(dist_peak), float(output_memory)],
                         dtype=torch.float64, device=device)
    if world > 1:
        import torch.distributed as dist
        all_stats = [torch.empty_like(stats) for _ in range(world)]
        dist.all_gather(all_stats, stats)
        all_stats = torch.stack(all_stats)
    else:
        all_stats = stats.unsqueeze(0)
    if is_main_process():
        mean = all_stats.mean(0).tolist()
        tmax = all_stats[:, 1].max().item()
        rec.update({'distributed_input_memory': mean[0], 'distributed_time': mean[1],
                    'distributed_peak_memory': mean[2], 'distributed_output_memory': mean[3],
                    'distributed_time_max': tmax, 'world_size': world, 'mode': args.mode, 'backward': grad,
                    'dtype': args.dtype, 'T': t_total, 'd': d, 'heads': args.heads, 'offset': args.offset,
                    'iters': args.iters,
                    'tflops_per_gpu': flops * (3 if grad else 1) / tmax / 1e12})
        values.append(rec)
        json.dump(values, open(args.file, 'w'))
        print(json.dumps(rec))
    synchronize()
    shutdown()


if __name__ == '__main__':
    main()
