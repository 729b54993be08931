# -*- coding: utf-8 -*-
"""A few training steps of one sequence-parallel attention layer (the scenario of the reference's example.py:1-33,
extended to what a user actually runs: several steps, a causal mask, gradient all-reduce, an optimiser, and a check
against the dense single-device twin).

    python example.py                                                       # one process (CPU or one GPU)
    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 example.py [--length 65536 --dtype bf16]

Every rank holds T / N consecutive time steps; the layer's weights are replicated.
"""
import argparse

import torch

from distributed_dot_product.module import DistributedDotProductAttn
from distributed_dot_product.parallel import allreduce_gradients, broadcast_parameters
from distributed_dot_product.utils import comm


def parse():
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument('--length', type=int, default=4096, help='total sequence length T (split over the ranks)')
    ap.add_argument('--dim', type=int, default=768)
    ap.add_argument('--heads', type=int, default=2)
    ap.add_argument('--offset', type=int, default=64, help='rows per all-gather chunk (results do not depend on it)')
    ap.add_argument('--steps', type=int, default=2)
    ap.add_argument('--dtype', choices=['fp32', 'bf16'], default='fp32')
    ap.add_argument('--causal', action='store_true', help='causal mask instead of the all-visible one')
    return ap.parse_args()


def my_rows(total, world, rank):
    per_rank = total // world
    return per_rank, slice(rank * per_rank, (rank + 1) * per_rank)


def build_mask(rows, total, causal, device):
    """bool [1, T/N, T], True = this (key row, query column) pair is masked out"""
    if not causal:
        return torch.zeros(1, rows.stop - rows.start, total, dtype=torch.bool, device=device)
    row_idx = torch.arange(rows.start, rows.stop, device=device).view(1, -1, 1)
    return torch.arange(total, device=device).view(1, 1, -1) > row_idx


def main():
    args = parse()
    rank, world = comm.get_rank(), comm.get_world_size()
    on_gpu = torch.cuda.is_available()
    if on_gpu:
        torch.cuda.set_device(comm.get_local_rank())
    device = torch.device('cuda' if on_gpu else 'cpu')
    dtype = torch.bfloat16 if (args.dtype == 'bf16' and on_gpu) else torch.float32

    torch.manual_seed(111)                       # same initial weights everywhere ...
    layer = DistributedDotProductAttn(args.dim, num_heads=args.heads, offset=args.offset).to(device, dtype)
    broadcast_parameters(layer)                  # ... and made sure of it
    optimiser = torch.optim.SGD(layer.parameters(), lr=0.05)

    per_rank, rows = my_rows(args.length, world, rank)
    mask = build_mask(rows, args.length, args.causal, device)
    torch.manual_seed(1000 + rank)               # every rank has its own slice of the sequence
    inputs = torch.rand(1, per_rank, args.dim, device=device, dtype=dtype)
    target = torch.rand(1, per_rank, args.dim, device=device, dtype=dtype)

    for it in range(args.steps):
        optimiser.zero_grad(set_to_none=True)
        out = layer(inputs, inputs, inputs, mask)                 # [1, T/N, dim]: this rank's rows of the result
        loss = torch.nn.functional.mse_loss(out.float(), target.float())
        loss.backward()
        allreduce_gradients(layer)               # weight gradients are rank-partial sums, as in the reference
        optimiser.step()
        if comm.is_main_process():
            grad_norm = layer.keys.weight.grad.float().norm().item()
            print('step %d: loss %.6f, out %s, |dW_keys| %.4f' % (it, loss.item(), tuple(out.shape), grad_norm))
    comm.shutdown()


if __name__ == '__main__':
    main()
