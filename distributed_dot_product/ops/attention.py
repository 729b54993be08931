# -*- coding: utf-8 -*-
"""The attention core of ``DistributedDotProductAttn`` as ONE autograd node on the native kernels.

Forward (reference: module.py:47-74): head split -> ``K Q^T`` (all-gather->GEMM) -> scale, mask, softmax
(in place) -> ``attn V`` (all-gather->GEMM) -> head merge.  Backward (reference: ops.py:29-54 + autograd of
module.py:65-67): ``dP = dO V^T`` (gather), ``dV = P^T dO`` (reduce-scatter), ``dS`` (in place over ``dP``),
``dQ = dS^T K`` (reduce-scatter), ``dK = dS Q`` (gather).

All tensors cross this node in the merged ``[B, T/N, H*d]`` layout of the projections; the per-head
``[B, H, T/N, d]`` operands are strided *views* that go straight into the TMA tensor maps, and results are
written in the merged layout by the kernels, so no head split/merge copies exist in either direction.
Memory: the probability slab ``P`` is kept for backward (as in the reference, which saves ``attn``);
the gradient slab reuses one buffer for ``dP`` and ``dS``.
"""
import torch

from distributed_dot_product import _native
import distributed_dot_product.ops.gemm as _g
from distributed_dot_product.ops.softmax import softmax_backward, softmax_forward


def _heads(x, h):
    """[B, Tn, H*d] -> strided view [B, H, Tn, d]."""
    b, tn, hd = x.shape
    return x.view(b, tn, h, hd // h).transpose(1, 2)


class _AttentionCore(torch.autograd.Function):
    @staticmethod
    def forward(ctx, keys, queries, values, mask, num_heads, scale, offset):
        k4, q4, v4 = _heads(keys, num_heads), _heads(queries, num_heads), _heads(values, num_heads)
        scores = _g.gemm(_native.OP_NT, k4, q4, offset=offset)          # [B, H, Tn, T]
        probs = softmax_forward(scores, mask, scale, inplace=True)
        out = torch.empty_like(values)
        _g.gemm(_native.OP_NN, probs, v4, offset=offset, out=_heads(out, num_heads))
        ctx.save_for_backward(keys, queries, values, probs)
        ctx.num_heads, ctx.scale, ctx.offset = num_heads, scale, offset
        return out

    @staticmethod
    def backward(ctx, grad_out):
        keys, queries, values, probs = ctx.saved_tensors
        h, scale, offset = ctx.num_heads, ctx.scale, ctx.offset
        if not grad_out.is_contiguous():
            grad_out = grad_out.contiguous()
        g4 = _heads(grad_out, h)
        k4, q4, v4 = _heads(keys, h), _heads(queries, h), _heads(values, h)
        d_probs = _g.gemm(_native.OP_NT, g4, v4, offset=offset)          # dP  [B, H, Tn, T]
        d_values = torch.empty_like(values)
        _g.gemm(_native.OP_TN, probs, g4, out=_heads(d_values, h))       # dV = P^T dO
        d_scores = softmax_backward(probs, d_probs, scale, inplace=True)  # dS overwrites dP
        d_queries = torch.empty_like(queries)
        _g.gemm(_native.OP_TN, d_scores, k4, out=_heads(d_queries, h))   # dQ = dS^T K
        d_keys = torch.empty_like(keys)
        _g.gemm(_native.OP_NN, d_scores, q4, offset=offset, out=_heads(d_keys, h))  # dK = dS Q
        return d_keys, d_queries, d_values, None, None, None, None


def attention_core(keys, queries, values, mask, num_heads, scale, offset=32):
    """``softmax(scale * K Q^T masked) V`` with K/Q/V given as ``[B, T/N, H*d]`` projections."""
    return _AttentionCore.apply(keys, queries, values, mask, num_heads, scale, offset)
