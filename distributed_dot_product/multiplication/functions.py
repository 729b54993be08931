# -*- coding: utf-8 -*-
"""Functional distributed products over a time axis split as ``T/N`` rows per rank.

Same names, argument order and result shapes as the reference
(reference: distributed_dot_product/multiplication/functions.py:44-212):

* ``distributed_matmul_nt(left[*,Tn,D], right[*,Tn,D], offset) -> [*,Tn,T]``   (``A B^T``)
* ``distributed_matmul_all(left[*,Tn,T], right[*,Tn,D], offset) -> [*,Tn,D]``  (``A B``)
* ``distributed_matmul_tn(left[*,Tn,T], right[*,Tn,D]) -> [*,Tn,D]``           (``A^T B``)

CUDA tensors run ONE fused sm_100a kernel per call (``ddp_gemm``: TMA-fed tcgen05 tiles with the
all-gather / reduce-scatter done inside the kernel over NVLink peer memory); CPU tensors run a
``torch.distributed`` (gloo) implementation used for plumbing tests.  Differences from the
reference, all deliberate: results keep the input dtype (fp32 accumulate;
fp32 inputs use TF32 tensor-core math), ``offset`` is honoured by every caller, no host barrier,
one reduce-scatter instead of N all-reduces for ``tn``, no full-slab layout copies.
"""
import functools
import os
import time

import torch
import torch.distributed as dist
from torch import Tensor

from distributed_dot_product import _native
from distributed_dot_product.utils.comm import get_rank, get_world_size

DEBUG = bool(int(os.environ.get('DISTRIBUTED_DOT_DEBUG', 0)))


def measure(f):
    """Debug decorator (reference: functions.py:24-41): with ``DISTRIBUTED_DOT_DEBUG=1`` prints the
    operand sizes, the peak-memory delta and the elapsed time of the call - timed with CUDA events
    on the launching stream for CUDA tensors, host clock otherwise."""
    @functools.wraps(f)
    def wrapper(*args, **kwargs):
        if not DEBUG:
            return f(*args, **kwargs)
        cuda = torch.is_tensor(args[0]) and args[0].is_cuda
        print(f'{f.__name__} - Left: {args[0].size()}, Right: {args[1].size()}')
        if cuda:
            start_memory = torch.cuda.max_memory_allocated()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
        start_time = time.time()
        result = f(*args, **kwargs)
        if cuda:
            e1.record()
            e1.synchronize()
            total_memory = torch.cuda.max_memory_allocated() - start_memory
            print(f'{f.__name__} GPU memory: {total_memory}')
            print(f'{f.__name__} elapsed time: {e0.elapsed_time(e1) / 1e3}')
        else:
            print(f'{f.__name__} elapsed time: {time.time() - start_time}')
        return result
    return wrapper


def _gpu(t):
    return t.is_cuda


def _gemm(op, left, right, offset, out=None, alpha=1.0):
    import distributed_dot_product.ops.gemm as _g
    return _g.gemm(op, left, right, offset=offset, out=out, alpha=alpha)


# --------------------------------------------------------------------------- CPU / gloo plumbing
def _all_gather(t):
    world = get_world_size()
    if world == 1:
        return [t]
    t = t.contiguous()
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return out


def _cpu_nt(left, right, offset):
    world = get_world_size()
    rows = right.size(-2)
    result = left.new_empty(tuple(left.shape[:-2]) + (left.size(-2), rows * world))
    offset = max(int(offset), 1)
    for row in range(0, rows, offset):
        end = min(row + offset, rows)
        chunks = _all_gather(right[..., row:end, :])
        for r, chunk in enumerate(chunks):
            result[..., r * rows + row:r * rows + end] = torch.matmul(left, chunk.transpose(-1, -2))
    return result


def _cpu_all(left, right, offset):
    world = get_world_size()
    rows = right.size(-2)
    cols = right.size(-1)
    if left.size(-1) != rows * world:
        raise ValueError('left.size(-1) must equal world_size * right.size(-2)')
    result = left.new_zeros(tuple(left.shape[:-1]) + (cols,))
    offset = max(int(offset), 1)
    for col in range(0, cols, offset):
        end = min(col + offset, cols)
        chunks = _all_gather(right[..., col:end])
        for r, chunk in enumerate(chunks):
            result[..., col:end] += torch.matmul(left[..., r * rows:(r + 1) * rows], chunk)
    return result


def _cpu_tn(left, right):
    world, rank = get_world_size(), get_rank()
    if left.size(-1) % world:
        raise ValueError('left.size(-1) must be divisible by the world size')
    split = left.size(-1) // world
    mine = None
    for r in range(world):
        part = torch.matmul(left[..., r * split:(r + 1) * split].transpose(-1, -2), right).contiguous()
        if world > 1:
            dist.reduce(part, dst=r)
        if r == rank:
            mine = part
    return mine


# --------------------------------------------------------------------------- public functions
@measure
def distributed_matmul_nt(left: Tensor, right: Tensor, offset=32, out: Tensor = None) -> Tensor:
    """``A B^T`` for row-split sequences (reference: functions.py:44-99).

    ``left``: ``[*, T/N, D]``, ``right``: ``[*, T/N, D]`` -> ``[*, T/N, T]`` whose column
    ``r * T/N + j`` is ``left . right_r[j]`` (global time order).  ``offset`` is the number of rows of
    ``right`` published per communication step (rounded up to the 64-row staging atom on GPU).
    """
    if _gpu(left):
        return _gemm(_native.OP_NT, left, right, offset, out=out)
    res = _cpu_nt(left, right, offset)
    if out is not None:
        out.copy_(res)
        return out
    return res


@measure
def distributed_matmul_tn(left: Tensor, right: Tensor, out: Tensor = None) -> Tensor:
    """``A^T B`` (reference: functions.py:102-148): ``left``: ``[*, T/N, T]``, ``right``:
    ``[*, T/N, D]`` -> ``[*, T/N, D]``, the rank's row block of ``A^T B``.  On GPU the partial tiles
    are stored straight into the owning rank's buffer by the GEMM epilogue (one reduce-scatter)."""
    if _gpu(left):
        return _gemm(_native.OP_TN, left, right, 32, out=out)
    res = _cpu_tn(left, right)
    if out is not None:
        out.copy_(res)
        return out
    return res


@measure
def distributed_matmul_all(left: Tensor, right: Tensor, offset=32, out: Tensor = None) -> Tensor:
    """``A B`` (reference: functions.py:160-212): ``left``: ``[*, T/N, T]``, ``right``:
    ``[*, T/N, D]`` -> ``[*, T/N, D]``.  The sum over ranks happens in the TMEM accumulator on GPU
    (the rank boundary is just a K-chunk boundary); ``offset`` sets the staging granularity."""
    if _gpu(left):
        return _gemm(_native.OP_NN, left, right, offset, out=out)
    res = _cpu_all(left, right, offset)
    if out is not None:
        out.copy_(res)
        return out
    return res


def distributed_matmul_block(left: Tensor, right: Tensor, transpose: bool = False) -> Tensor:
    """Local product summed over ranks.  The reference's version is dead and broken
    (functions.py:151-157: ``.tranpose`` typo, Average instead of Sum); this is the working helper."""
    block = torch.matmul(left, right)
    if transpose:
        block = block.transpose(-1, -2).contiguous()
    if get_world_size() > 1:
        dist.all_reduce(block)
    return block
