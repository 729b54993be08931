# -*- coding: utf-8 -*-
"""Projection layers on the framework's own tcgen05 GEMM (reference: module.py:43-45,75 use
``nn.Linear`` -> cuBLAS).

``y = x W^T + b`` is the NT kernel with the bias fused in the epilogue; ``dx = dy W`` is the NN kernel;
``dW = dy^T x`` is the TN kernel (both operands MN-major straight from their row-major storage, no
transpose pass); ``db`` is a column-sum kernel.  Weight gradients stay rank-partial, as in the
reference (tests/test_gradient.py:116-121).
"""
import ctypes

import torch
import torch.nn.functional as F

from distributed_dot_product import _native
import distributed_dot_product.ops.gemm as _g


def colsum(x2):
    """Column sum of a 2-D CUDA tensor (row-split two-stage reduction through an fp32 scratch buffer)."""
    lib = _native.load()
    out = torch.empty(x2.size(-1), dtype=x2.dtype, device=x2.device)
    v = _g.as_view(x2.unsqueeze(0).unsqueeze(0))
    nbytes = lib.ddp_colsum_scratch_bytes(x2.size(0), x2.size(1))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x2.device) if nbytes else None
    _native.check(lib.ddp_colsum(ctypes.byref(v), ctypes.c_void_p(out.data_ptr()),
                                 ctypes.c_void_p(scratch.data_ptr()) if scratch is not None else None,
                                 nbytes, _g._stream()))
    _native.count_launch(2 if nbytes else 1)
    return out


def _weight_grad(g2, x2):
    """``dW = dy^T x`` on the TN kernel.  A tall-skinny problem (many rows, few output tiles) would leave most
    SMs idle, so the row (= K) dimension is split into S batches that run as one batched TN launch and the S
    partial products are summed afterwards (split-K)."""
    rows, n_out, n_in = g2.size(0), g2.size(1), x2.size(1)
    tiles = ((n_out + 127) // 128) * ((n_in + 255) // 256)
    split = 1
    if g2.is_contiguous() and x2.is_contiguous():
        while split < 8 and tiles * split * 2 <= 148 and rows % (split * 2 * 64) == 0 and rows // (split * 2) >= 1024:
            split *= 2
    if split == 1:
        return _g.gemm(_native.OP_TN, g2, x2, distributed=False)
    part = _g.gemm(_native.OP_TN, g2.view(split, rows // split, n_out), x2.view(split, rows // split, n_in),
                   distributed=False)
    # fp32 sum of the S partial products in one native launch
    n = n_out * n_in
    if (n * part.element_size()) % 16:
        return colsum(part.view(split, n)).view(n_out, n_in)
    out = torch.empty(n_out, n_in, dtype=part.dtype, device=part.device)
    _native.check(_native.load().ddp_sum_partials(ctypes.c_void_p(part.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                                                  _g._DT[part.dtype], split, n, _g._stream()))
    _native.count_launch()
    return out


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        x2 = x.reshape(-1, x.size(-1))
        y = _g.gemm(_native.OP_NT, x2, weight, bias=bias, distributed=False)
        return y.reshape(tuple(x.shape[:-1]) + (weight.size(0),))

    @staticmethod
    def backward(ctx, gy):
        x, weight = ctx.saved_tensors
        g2 = gy.reshape(-1, gy.size(-1))
        x2 = x.reshape(-1, x.size(-1))
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = _g.gemm(_native.OP_NN, g2, weight, distributed=False).reshape(x.shape)
        if ctx.needs_input_grad[1]:
            gw = _weight_grad(g2, x2)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            g2c = g2 if g2.stride(-1) == 1 else g2.contiguous()
            gb = colsum(g2c)
        return gx, gw, gb


def linear(x, weight, bias=None):
    """``F.linear`` semantics; CUDA tensors of a tensor-core dtype run on the sm_100a GEMM kernels."""
    if x.is_cuda and x.dtype in _g._DT and weight.dtype == x.dtype:
        return _Linear.apply(x, weight, bias)
    return F.linear(x, weight, bias)
