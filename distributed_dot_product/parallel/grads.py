# -*- coding: utf-8 -*-
"""Helpers for the replicated weights of a sequence-parallel layer.

The layer's ``param.grad`` on each rank is the partial sum over that rank's ``T/N`` rows
(reference behaviour: tests/test_gradient.py:116-121 all-reduces them by hand, and
tests/test_gradient.py:48 broadcasts the initial weights).  These two helpers are that plumbing.
"""
import torch
import torch.distributed as dist

from distributed_dot_product.parallel import runtime


def broadcast_parameters(module_or_state, root_rank=0):
    """Makes every rank hold rank ``root_rank``'s parameters and buffers."""
    if runtime.get_world_size() == 1:
        return
    state = module_or_state.state_dict() if hasattr(module_or_state, 'state_dict') else module_or_state
    for _, t in sorted(state.items()):
        if torch.is_tensor(t):
            dist.broadcast(t.data, src=root_rank)


def allreduce_gradients(module, average=False):
    """Sums (or averages) ``param.grad`` over the ranks, one flat bucket per dtype."""
    world = runtime.get_world_size()
    if world == 1:
        return
    buckets = {}
    for p in module.parameters():
        if p.grad is not None:
            buckets.setdefault((p.grad.dtype, p.grad.device), []).append(p.grad)
    for grads in buckets.values():
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat)
        if average:
            flat.div_(world)
        off = 0
        for g in grads:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n
