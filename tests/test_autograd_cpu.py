# -*- coding: utf-8 -*-
"""Gradients of the three autograd ops against dense fp64 autograd, N processes over gloo.

Covers what the reference never tested: ``LeftTransposeMultiplication`` - whose
reference ``left`` gradient is the transpose of the true one (reference: ops.py:69) - and random cotangents.
"""
import pytest
import torch

from dist_utils import run_distributed

TN_ROWS = 5
DIM = 7


def _worker(rank, world, which):
    from distributed_dot_product.multiplication.functions import distributed_matmul_nt
    from distributed_dot_product.multiplication.ops import (
        FullMultiplication, LeftTransposeMultiplication, RightTransposeMultiplication)
    torch.manual_seed(1234)  # same global tensors on every rank
    t = TN_ROWS * world
    rows = slice(rank * TN_ROWS, (rank + 1) * TN_ROWS)
    if which == 'nt':
        gl = torch.randn(2, t, DIM, dtype=torch.float64, requires_grad=True)
        gr = torch.randn(2, t, DIM, dtype=torch.float64, requires_grad=True)
        cot = torch.randn(2, t, t, dtype=torch.float64)
        dense = gl.matmul(gr.transpose(-1, -2))
        op = RightTransposeMultiplication
    else:
        gl = torch.randn(2, t, t, dtype=torch.float64, requires_grad=True)
        gr = torch.randn(2, t, DIM, dtype=torch.float64, requires_grad=True)
        cot = torch.randn(2, t, DIM, dtype=torch.float64)
        dense = gl.matmul(gr) if which == 'all' else gl.transpose(-1, -2).matmul(gr)
        op = FullMultiplication if which == 'all' else LeftTransposeMultiplication
    (dense * cot).sum().backward()
    left = gl.detach()[:, rows].clone().requires_grad_(True)
    right = gr.detach()[:, rows].clone().requires_grad_(True)
    out = op.apply(left, right, 3)
    assert torch.allclose(out, dense.detach()[:, rows], atol=1e-10)
    (out * cot[:, rows]).sum().backward()
    assert torch.allclose(left.grad, gl.grad[:, rows], atol=1e-10), which
    assert torch.allclose(right.grad, gr.grad[:, rows], atol=1e-10), which
    if which == 'tn':
        # pin the documented deviation: the reference formula nt(dOut, right) is the TRANSPOSE of dA
        ref_formula = distributed_matmul_nt(cot[:, rows].contiguous(), right.detach(), offset=3)
        full = [torch.empty_like(ref_formula) for _ in range(world)]
        import torch.distributed as dist
        if world > 1:
            dist.all_gather(full, ref_formula)
        else:
            full = [ref_formula]
        ref_global = torch.cat(full, dim=-2)
        assert torch.allclose(ref_global.transpose(-1, -2), gl.grad, atol=1e-10)
        assert not torch.allclose(ref_global, gl.grad, atol=1e-3)


def _all_ops(rank, world):
    for which in ('nt', 'all', 'tn'):
        _worker(rank, world, which)


@pytest.mark.parametrize('world', [1, 2, 3])
def test_op_gradients(world):
    if world == 1:
        _all_ops(0, 1)
    else:
        run_distributed(_all_ops, world)
