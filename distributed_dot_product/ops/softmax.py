# -*- coding: utf-8 -*-
"""Fused scale + boolean mask + row softmax (reference: module.py:65-67 and its autograd).

``masked_softmax(scores, mask, scale)`` equals
``softmax((scores * scale).masked_fill(mask, -inf), dim=-1)``.  On CUDA it is one sm_100a kernel
per direction (``ddp_softmax_fwd`` / ``ddp_softmax_bwd``): the forward makes a single pass over the
``[B, H, T/N, T]`` slab with the row held in registers, reads the mask once per ``(B, T/N, T)`` row for
all heads, and can run in place; the backward is ``dS = scale * P * (dP - rowsum(dP * P))``.
"""
import ctypes

import torch

from distributed_dot_product import _native
from distributed_dot_product.ops.gemm import _DT, _aligned, _stream, as_view


def _canon4(t):
    if t.dim() == 2:
        return t.unsqueeze(0).unsqueeze(0)
    if t.dim() == 3:
        return t.unsqueeze(1)
    if t.dim() == 4:
        return t
    return t.reshape((-1,) + tuple(t.shape[-3:]))


def _prep_mask(mask, x4):
    """Returns (uint8 tensor [b1, rows, cols] view, ld, sb1) or (None, 0, 0)."""
    if mask is None:
        return None, 0, 0
    if mask.dtype != torch.bool:
        raise RuntimeError('masked_softmax only supports boolean masks (True = masked out), as masked_fill')
    b1, b0, rows, cols = x4.shape
    m = mask
    if m.dim() == 4:
        # [B, H, Tn, T] expanded over heads (module.py:48-50) -> drop the broadcast head dim
        if m.size(1) == 1 or m.stride(1) == 0:
            m = m[:, 0]
        else:
            raise RuntimeError('per-head masks are not supported; pass a [B, T/N, T] mask')
    if m.dim() == 2:
        m = m.unsqueeze(0)
    if m.dim() != 3:
        raise RuntimeError('mask must be [B, T/N, T]')
    m = m.expand(b1, rows, cols)
    if m.stride(-1) != 1:
        m = m.contiguous()
    mu8 = m.view(torch.uint8)
    return mu8, mu8.stride(1), (mu8.stride(0) if b1 > 1 else 0)


def softmax_forward(scores, mask, scale, inplace=False):
    lib = _native.load()
    if scores.dtype not in _DT:
        raise RuntimeError('masked_softmax: unsupported dtype %s on CUDA' % scores.dtype)
    x4 = _canon4(scores)
    if x4.stride(-1) != 1:
        x4 = x4.contiguous()
        inplace = False if x4.data_ptr() != scores.data_ptr() else inplace
    y4 = x4 if inplace else torch.empty_like(x4)
    mu8, m_ld, m_sb1 = _prep_mask(mask, x4)
    vx, vy = as_view(x4), as_view(y4)
    _native.check(lib.ddp_softmax_fwd(ctypes.byref(vx), ctypes.byref(vy),
                                      ctypes.c_void_p(mu8.data_ptr()) if mu8 is not None else None,
                                      m_ld, m_sb1, float(scale), _stream()))
    _native.count_launch()
    if inplace and x4.data_ptr() == scores.data_ptr():
        return scores
    return y4.reshape(scores.shape)


def softmax_backward(probs, grad, scale, inplace=False, delta=None):
    lib = _native.load()
    y4 = _canon4(probs)
    g4 = _canon4(grad)
    if g4.stride(-1) != 1:
        g4 = g4.contiguous()
        inplace = False
    if y4.stride(-1) != 1:
        y4 = y4.contiguous()
    d4 = g4 if inplace else torch.empty_like(g4)
    vy, vg, vd = as_view(y4), as_view(g4), as_view(d4)
    dptr = None
    if delta is not None:
        delta = delta.to(torch.float32).contiguous()
        dptr = ctypes.c_void_p(delta.data_ptr())
    _native.check(lib.ddp_softmax_bwd(ctypes.byref(vy), ctypes.byref(vg), ctypes.byref(vd), dptr, float(scale),
                                      _stream()))
    _native.count_launch()
    return d4.reshape(grad.shape)


class _MaskedSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scores, mask, scale, inplace):
        if scores.is_cuda:
            out = softmax_forward(scores, mask, scale, inplace=inplace)
            if inplace and out is scores:
                ctx.mark_dirty(scores)
        else:
            s = scores * scale
            if mask is not None:
                s = s.masked_fill(mask, -float('inf'))
            out = torch.softmax(s, dim=-1)
        ctx.save_for_backward(out)
        ctx.scale = scale
        ctx.inplace = inplace
        return out

    @staticmethod
    def backward(ctx, grad):
        (out,) = ctx.saved_tensors
        if grad.is_cuda:
            # The incoming gradient slab is consumed by this node only (it is produced by the
            # FullMultiplication backward), so the in-place variant may overwrite it.
            gin = softmax_backward(out, grad, ctx.scale, inplace=ctx.inplace and grad.is_contiguous())
        else:
            dot = (grad * out).sum(dim=-1, keepdim=True)
            gin = ctx.scale * out * (grad - dot)
        return gin, None, None, None


def masked_softmax(scores, mask=None, scale=1.0):
    """Out-of-place ``softmax((scores * scale).masked_fill(mask, -inf), -1)``."""
    return _MaskedSoftmax.apply(scores, mask, float(scale), False)


def masked_softmax_(scores, mask=None, scale=1.0):
    """In-place variant: overwrites ``scores`` with the probabilities (saves one slab of memory and
    its backward overwrites the incoming gradient slab)."""
    return _MaskedSoftmax.apply(scores, mask, float(scale), True)
