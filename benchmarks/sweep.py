# -*- coding: utf-8 -*-
"""BASELINE.json configs 2, 3 and 5 in ONE process per rank (a torchrun launch per point costs more GPU time
than the points themselves): both arms, any world size, CUDA-event timing after warm-up, max over ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 benchmarks/sweep.py \
        --impl ours|reference [--out file.jsonl] [--reduced]

    config 2  distributed_matmul_nt fwd+bwd, d = 768, T in {1000, 4096, 8192, 16384, 25000}   (reference: benchmark.py:70-117)
    config 3  distributed_matmul_all / _tn fwd+bwd, T = 8192, d in {2, 16, 64, 256, 768}      (reference: benchmark.py:120-227)
    config 5  DistributedDotProductAttn, 1 head, d = 512, T = 50176 (non-local block), fwd+bwd

``--impl reference`` runs the UNMODIFIED reference from baseline/_ref through the horovod / mpi4py stand-ins of
baseline/shim, in its native fp32 (TF32 matmul); ``--impl ours`` runs fp32 storage / TF32 math for configs 2 and 3
(same precision as the reference arm) and bf16 for config 5.  T is truncated to a multiple of the world size.
"""
import argparse
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

ap = argparse.ArgumentParser()
ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
ap.add_argument('--out', default='')
ap.add_argument('--reduced', action='store_true', help='end points of every sweep only')
ap.add_argument('--iters', type=int, default=10)
ap.add_argument('--warmup', type=int, default=3)
ap.add_argument('--offset', type=int, default=1024)
args = ap.parse_args()

if args.impl == 'reference':
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or '.') not in (ROOT, HERE)]
    sys.path.insert(0, os.path.join(ROOT, 'baseline', '_ref'))
    sys.path.insert(0, os.path.join(ROOT, 'baseline', 'shim'))
else:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

if args.impl == 'reference':
    import horovod.torch as hvd
    hvd.init()
    world, rank = hvd.size(), hvd.rank()
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
else:
    from distributed_dot_product.utils import comm
    comm.init()
    world, rank = comm.get_world_size(), comm.get_rank()
from distributed_dot_product.multiplication.ops import (  # noqa: E402
    FullMultiplication, LeftTransposeMultiplication, RightTransposeMultiplication)
from distributed_dot_product.module import DistributedDotProductAttn  # noqa: E402

dev = torch.device('cuda', torch.cuda.current_device())
torch.manual_seed(111 + rank)


def timed(run):
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
    return ms.item()


def point(mode, length, d, heads=1, dtype=torch.float32):
    tn = length // world
    t = tn * world
    rec = {'impl': args.impl, 'mode': mode, 'world_size': world, 'T': t, 'd': d, 'heads': heads, 'backward': True,
           'dtype': {torch.float32: 'fp32 (tf32 matmul)', torch.bfloat16: 'bf16'}[dtype], 'offset': args.offset}
    try:
        def mk(*s):
            return torch.rand(*s, device=dev).to(dtype).requires_grad_(True)

        if mode == 'nt':
            a, b = mk(1, tn, d), mk(1, tn, d)
            op = lambda: RightTransposeMultiplication.apply(a, b, args.offset)  # noqa: E731
            flops = 3 * 2.0 * tn * t * d
        elif mode in ('all', 'tn'):
            a, b = mk(1, tn, t), mk(1, tn, d)
            cls = FullMultiplication if mode == 'all' else LeftTransposeMultiplication
            op = lambda: cls.apply(a, b, args.offset)  # noqa: E731
            flops = 3 * 2.0 * tn * t * d
        else:
            model = DistributedDotProductAttn(d, num_heads=heads, offset=args.offset).to(dev, dtype)
            a = mk(1, tn, d)
            b = None
            mask = torch.zeros(1, tn, t, device=dev).bool()
            op = lambda: model(a, a, a, mask)  # noqa: E731
            flops = 3 * (2 * 2.0 * tn * t * d + 4 * 2.0 * tn * d * d)

        def run():
            out = op()
            out.backward(torch.ones_like(out))
            a.grad = None
            if b is not None:
                b.grad = None

        ms = timed(run)
        rec.update({'ms': ms, 'tflops_per_gpu': flops / (ms * 1e-3) / 1e12})
    except torch.cuda.OutOfMemoryError:
        rec['unavailable'] = 'CUDA out of memory'
    except RuntimeError as e:  # keep the sweep going; the record says what happened
        rec['unavailable'] = str(e).splitlines()[0][:200]
    torch.cuda.empty_cache()
    if rank == 0:
        line = json.dumps(rec)
        print(line, flush=True)
        if args.out:
            with open(args.out, 'a') as f:
                f.write(line + '\n')


ts = [1000, 25000] if args.reduced else [1000, 4096, 8192, 16384, 25000]
ds = [2, 768] if args.reduced else [2, 16, 64, 256, 768]
for t_ in ts:
    point('nt', t_, 768)
for d_ in ds:
    for m_ in ('all', 'tn'):
        point(m_, 8192, d_)
point('attn', 50176, 512, heads=1, dtype=torch.float32 if args.impl == 'reference' else torch.bfloat16)

if world > 1:
    dist.barrier()
if args.impl == 'ours':
    comm.shutdown()
elif world > 1:
    dist.destroy_process_group()
