# -*- coding: utf-8 -*-
"""Operator parity on exact integers, N real processes over gloo.

Same idea as the reference's only operator test (reference: tests/test_multiplication.py:27-144):
``arange`` tensors make every product exact, so a layout or ordering bug cannot hide behind a
tolerance.  Six modes (nt / tn / all x 3-D and 4-D with two heads), several world sizes, offsets that
do and do not divide the chunked dimension.
"""
import pytest
import torch

from dist_utils import run_distributed

LENGTH = 4
DIM = 6


def _global_pair(mode, world, heads):
    t = LENGTH * world
    lead = (1,) if heads is None else (1, heads)
    if mode == 'nt':
        left = torch.arange(int(torch.tensor(lead + (t, DIM)).prod())).view(*lead, t, DIM)
        right = left.flip(-1) - 3
    else:
        left = torch.arange(int(torch.tensor(lead + (t, t)).prod())).view(*lead, t, t) - 11
        right = torch.arange(int(torch.tensor(lead + (t, DIM)).prod())).view(*lead, t, DIM) - 5
    return left, right


def _dense(mode, left, right):
    if mode == 'nt':
        return left.matmul(right.transpose(-1, -2))
    if mode == 'tn':
        return left.transpose(-1, -2).matmul(right)
    return left.matmul(right)


def _worker(rank, world, mode, heads, offset, dtype):
    from distributed_dot_product.multiplication.functions import (
        distributed_matmul_all, distributed_matmul_nt, distributed_matmul_tn)
    left, right = _global_pair(mode, world, heads)
    left, right = left.to(dtype), right.to(dtype)
    gt = _dense(mode, left, right)
    rows = slice(rank * LENGTH, (rank + 1) * LENGTH)
    l_loc, r_loc = left[..., rows, :].contiguous(), right[..., rows, :].contiguous()
    if mode == 'nt':
        out = distributed_matmul_nt(l_loc, r_loc, offset=offset)
    elif mode == 'tn':
        out = distributed_matmul_tn(l_loc, r_loc)
    else:
        out = distributed_matmul_all(l_loc, r_loc, offset=offset)
    assert out.dtype == dtype
    assert out.shape == gt[..., rows, :].shape
    assert (out == gt[..., rows, :]).all(), (mode, rank, world)


def _all_modes(rank, world, offsets, dtype):
    for mode in ('nt', 'tn', 'all'):
        for heads in (None, 2):
            for offset in offsets:
                _worker(rank, world, mode, heads, offset, dtype)


@pytest.mark.parametrize('world', [2, 3])
def test_integer_parity(world):
    """nt / tn / all x 3-D / 4-D (six modes), exact integer equality."""
    run_distributed(_all_modes, world, (2,), torch.int64)


def test_ragged_offsets():
    run_distributed(_all_modes, 2, (1, 3, 5, 64), torch.float64)
