# -*- coding: utf-8 -*-
"""Autograd wrappers of the three distributed products (reference: multiplication/ops.py:19-71).

Every backward pass is built from the other two products, exactly as in the reference; on GPU each of
them is one fused sm_100a kernel (all-gather->GEMM or GEMM->reduce-scatter).  Deviations, documented in
SURVEY.md section 4.5: ``offset`` reaches every product (the reference drops it in three calls), and
``LeftTransposeMultiplication`` returns the mathematically correct ``left`` gradient (the reference
returns its transpose, ops.py:69).
"""
from typing import Any

import torch.autograd as autograd
from torch import Tensor

from distributed_dot_product.multiplication.functions import (
    distributed_matmul_all, distributed_matmul_nt, distributed_matmul_tn)


class RightTransposeMultiplication(autograd.Function):
    """``left @ gather(right)^T`` -> ``[*, T/N, T]`` (reference: ops.py:19-37)."""

    @staticmethod
    def forward(ctx: Any, left: Tensor, right: Tensor, offset: int = 32) -> Tensor:
        ctx.save_for_backward(left, right)
        ctx.offset = offset
        return distributed_matmul_nt(left, right, offset=offset)

    @staticmethod
    def backward(ctx: Any, output_grad: Tensor):
        left, right = ctx.saved_tensors
        offset = ctx.offset
        grad_left = grad_right = None
        if ctx.needs_input_grad[1]:
            grad_right = distributed_matmul_tn(output_grad, left)
        if ctx.needs_input_grad[0]:
            grad_left = distributed_matmul_all(output_grad, right, offset=offset)
        return grad_left, grad_right, None


class FullMultiplication(autograd.Function):
    """``left @ gather(right)`` -> ``[*, T/N, D]`` (reference: ops.py:40-54)."""

    @staticmethod
    def forward(ctx: Any, left: Tensor, right: Tensor, offset: int = 32) -> Tensor:
        ctx.save_for_backward(left, right)
        ctx.offset = offset
        return distributed_matmul_all(left, right, offset=offset)

    @staticmethod
    def backward(ctx: Any, output_grad: Tensor):
        left, right = ctx.saved_tensors
        offset = ctx.offset
        left_grad = right_grad = None
        if ctx.needs_input_grad[0]:
            left_grad = distributed_matmul_nt(output_grad, right, offset=offset)
        if ctx.needs_input_grad[1]:
            right_grad = distributed_matmul_tn(left, output_grad)
        return left_grad, right_grad, None


class LeftTransposeMultiplication(autograd.Function):
    """``gather(left)^T @ gather(right)`` row block -> ``[*, T/N, D]`` (reference: ops.py:57-71).

    With ``A`` the global ``[T, T]`` left matrix (row-split) and ``Out = A^T B``:
    ``dA = B dOut^T`` -> row block ``nt(right, dOut)``; ``dB = A dOut`` -> row block ``all(left, dOut)``.
    The reference computes ``nt(dOut, right)`` for ``dA``, which is the transpose.
    """

    @staticmethod
    def forward(ctx: Any, left: Tensor, right: Tensor, offset: int = 32) -> Tensor:
        ctx.save_for_backward(left, right)
        ctx.offset = offset
        return distributed_matmul_tn(left, right)

    @staticmethod
    def backward(ctx: Any, output_grad: Tensor):
        left, right = ctx.saved_tensors
        offset = ctx.offset
        left_grad = right_grad = None
        if ctx.needs_input_grad[0]:
            left_grad = distributed_matmul_nt(right, output_grad, offset=offset)
        if ctx.needs_input_grad[1]:
            right_grad = distributed_matmul_all(left, output_grad, offset=offset)
        return left_grad, right_grad, None
