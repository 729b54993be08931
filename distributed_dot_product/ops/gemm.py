# -*- coding: utf-8 -*-
"""Python front end of the fused collective GEMM kernels (``ddp_gemm`` in include/ddp_b200.h).

Turns torch tensors into strided ``ddp_tensor`` views without copying whenever TMA can address them
directly (unit inner stride, 16-byte aligned base/strides, at most two leading batch dims with
arbitrary strides - so head-split ``[B, H, Tn, d]`` *views* of a ``[B, Tn, H*d]`` projection go
straight into the tensor map), and falls back to one padded staging copy otherwise.
"""
import ctypes

import torch

from distributed_dot_product import _native
from distributed_dot_product.parallel import runtime

import os as _os
_DBG = int(_os.environ.get('DDP_GEMM_DBG', '0'))  # kernel debug bits (see ddp_gemm); 0 in production
_DT = {torch.float16: _native.F16, torch.bfloat16: _native.BF16, torch.float32: _native.F32}


def compute_dtype(dtype):
    """dtype the kernels run in for tensors of ``dtype`` (fp32 -> TF32 tensor-core math)."""
    return dtype if dtype in _DT else torch.float32


def _aligned(t):
    es = t.element_size()
    if t.data_ptr() % 16:
        return False
    if t.stride(-1) != 1 and t.size(-1) != 1:
        return False
    for d in range(t.dim() - 1):
        if t.size(d) != 1 and (t.stride(d) * es) % 16:
            return False
    return True


def _padded_empty(shape, dtype, device):
    """Row-major tensor of ``shape`` whose row stride is a multiple of 16 bytes (a view if padded)."""
    es = torch.empty((), dtype=dtype).element_size()
    q = 16 // es
    cols = shape[-1]
    cols_p = (cols + q - 1) // q * q
    if cols_p == cols:
        return torch.empty(shape, dtype=dtype, device=device)
    buf = torch.empty(tuple(shape[:-1]) + (cols_p,), dtype=dtype, device=device)
    return buf[..., :cols]


def _canon(t, pad_cols_zero=False):
    """Returns a 4-D tensor [b1, b0, rows, cols] that TMA can address (copying only if needed)."""
    if t.dim() < 2:
        raise ValueError('operands must have at least 2 dims')
    if t.dim() == 2:
        t = t.unsqueeze(0).unsqueeze(0)
    elif t.dim() == 3:
        t = t.unsqueeze(1)
    elif t.dim() > 4:
        t = t.reshape((-1,) + tuple(t.shape[-3:]))
    es = t.element_size()
    q = 16 // es
    need_copy = not _aligned(t)
    if pad_cols_zero and t.size(-1) % q:
        need_copy = True
    if need_copy:
        cols = t.size(-1)
        cols_p = (cols + q - 1) // q * q
        buf = torch.zeros(tuple(t.shape[:-1]) + (cols_p,), dtype=t.dtype, device=t.device)
        buf[..., :cols].copy_(t)
        t = buf if pad_cols_zero else buf[..., :cols]
    return t


def as_view(t):
    """ddp_tensor struct for a canonical 4-D tensor."""
    v = _native.Tensor()
    v.ptr = t.data_ptr()
    v.b1, v.b0, v.rows, v.cols = t.shape
    v.sb1, v.sb0, v.ld = t.stride(0), t.stride(1), t.stride(2)
    if v.b1 == 1:
        v.sb1 = 0
    if v.b0 == 1:
        v.sb0 = 0
    if v.rows == 1:
        v.ld = (v.cols + 15) // 16 * 16
    v.dtype = _DT[t.dtype]
    return v


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def chunk_rows_from_offset(offset, rows, batch=1):
    """Staging granularity of the in-kernel all-gather: ``offset`` rows per published chunk (reference
    semantics of ``offset``: functions.py:64-68), rounded up to 64 rows - but never so coarse that the push
    runs on a handful of warps: one warp pushes one (batch, chunk) item, so small problems get finer chunks
    (down to 16 rows) until there are about two items per SM."""
    offset = max(int(offset), 1)
    c = (offset + 63) // 64 * 64
    c = max(64, min(c, (rows + 63) // 64 * 64))
    want_items = 2 * 148
    if batch * ((rows + c - 1) // c) < want_items:
        fine = (rows * batch) // want_items // 16 * 16
        c = min(c, max(16, fine))
    return c


def _unsqueeze4(t):
    if t.dim() == 2:
        return t.unsqueeze(0).unsqueeze(0)
    if t.dim() == 3:
        return t.unsqueeze(1)
    return t


def _gemm_padded_segments(op, left, right, offset, out, alpha, world, seg, q, in_dtype):
    """Slow path for ``T/N`` not a multiple of 16 bytes: zero-pad every rank segment of ``left`` (and the
    rows of ``right`` for ``all``) to the next aligned width, run the kernel, slice the result."""
    seg_p = (seg + q - 1) // q * q
    lp = left.new_zeros(tuple(left.shape[:-1]) + (world, seg_p))
    lp[..., :seg].copy_(left.reshape(tuple(left.shape[:-1]) + (world, seg)))
    lp = lp.reshape(tuple(left.shape[:-1]) + (world * seg_p,))
    if op == _native.OP_NN:
        rp = right.new_zeros(tuple(right.shape[:-2]) + (seg_p, right.size(-1)))
        rp[..., :seg, :].copy_(right)
        res = gemm(op, lp, rp, offset=offset, alpha=alpha)
    else:
        res = gemm(op, lp, right, offset=offset, alpha=alpha)[..., :seg, :]
    if res.dtype != in_dtype:
        res = res.to(in_dtype)
    if out is not None:
        out.copy_(res)
        return out
    return res


def gemm(op, left, right, offset=32, out=None, alpha=1.0, bias=None, distributed=True, softmax_bwd=None,
         epilogue_builder=None):
    """Runs one fused product.  ``left``/``right`` follow the reference shapes
    (functions.py:44-212).  ``out`` may be a preallocated, possibly strided result tensor (e.g. the
    ``[B, H, Tn, d]`` view of a ``[B, Tn, H*d]`` buffer): the kernel then writes the final layout
    directly.  ``softmax_bwd=(P, delta, scale)`` (nt only) fuses the softmax backward into the epilogue:
    the result is ``scale * P * (left @ gather(right)^T - delta[..., None])``."""
    lib = _native.load()
    world = runtime.get_world_size() if distributed else 1
    in_dtype = left.dtype
    cdt = compute_dtype(in_dtype)
    if left.dtype != cdt:
        left = left.to(cdt)
    if right.dtype != cdt:
        right = right.to(cdt)
    lead = tuple(left.shape[:-2])
    if world > 1 and op != _native.OP_NT:
        # A rank segment of the slab must start on a 16-byte boundary for TMA: repack odd T/N once.
        q = 16 // left.element_size()
        seg = left.size(-1) // world
        if seg % q:
            return _gemm_padded_segments(op, left, right, offset, out, alpha, world, seg, q, in_dtype)
    gathered = op != _native.OP_TN and world > 1
    a = _canon(left, pad_cols_zero=gathered and op == _native.OP_NT)
    b = _canon(right, pad_cols_zero=gathered)
    rows_l, rows_r = a.size(-2), b.size(-2)
    if op == _native.OP_NT:
        out_rows, out_cols = rows_l, world * rows_r
        kernel_cols = out_cols
    elif op == _native.OP_NN:
        out_rows, out_cols = rows_l, right.size(-1)
        kernel_cols = b.size(-1)
    else:
        if a.size(-1) % world:
            raise ValueError('left.size(-1) must be divisible by the world size')
        out_rows, out_cols = a.size(-1) // world, right.size(-1)
        kernel_cols = b.size(-1)
    c4 = None
    if out is not None and out.dtype == cdt and kernel_cols == out_cols and out.dim() <= 4:
        cand = _unsqueeze4(out)
        if _aligned(cand) and tuple(cand.shape) == (a.size(0), a.size(1), out_rows, out_cols):
            c4 = cand
    direct = c4 is not None
    if not direct:
        c4 = _padded_empty((a.size(0), a.size(1), out_rows, kernel_cols), cdt, left.device)
    va, vb, vc = as_view(a), as_view(b), as_view(c4)
    comm_ptr = None
    if world > 1:
        chunk = chunk_rows_from_offset(offset, rows_r, a.size(0) * a.size(1))
        ws = ctypes.c_size_t()
        nf = ctypes.c_size_t()
        _native.check(lib.ddp_gemm_workspace(op, ctypes.byref(va), ctypes.byref(vb), world, chunk,
                                             ctypes.byref(ws), ctypes.byref(nf)))
        hp = runtime.heap()
        hp.ensure(ws.value, nf.value)
        comm_ptr = ctypes.byref(hp.next_comm(chunk))
    bias_ptr = None
    if bias is not None:
        if bias.dtype != cdt or not bias.is_contiguous():
            bias = bias.to(cdt).contiguous()
        bias_ptr = ctypes.c_void_p(bias.data_ptr())
    epi_ptr = None
    if softmax_bwd is not None:
        probs, delta, scale = softmax_bwd
        p4 = _unsqueeze4(probs)
        if p4.dim() != 4 or not _aligned(p4) or tuple(p4.shape) != tuple(c4.shape) or p4.dtype != cdt:
            raise ValueError('softmax_bwd: P must be a TMA-addressable tensor with the shape of the result')
        vp = as_view(p4)
        epi = _native.Epilogue()
        epi.mode, epi.scale = _native.EPI_SOFTMAX_BWD, float(scale)
        epi.P = ctypes.pointer(vp)
        epi.delta = delta.data_ptr()
        epi_ptr = ctypes.byref(epi)
    elif epilogue_builder is not None:
        epi = epilogue_builder(va, vb, vc, world)   # returns a filled _native.Epilogue
        epi_ptr = ctypes.byref(epi)
    _native.check(lib.ddp_gemm_ex(op, ctypes.byref(va), ctypes.byref(vb), ctypes.byref(vc), float(alpha), bias_ptr,
                                  comm_ptr, epi_ptr, _DBG, _stream()))
    _native.count_launch()
    if direct:
        return out
    res = c4 if kernel_cols == out_cols else c4[..., :out_cols]
    if len(lead) == 0:
        res = res[0, 0]
    elif len(lead) == 1:
        res = res[:, 0]
    elif len(lead) > 2:
        res = res.reshape(lead + (out_rows, out_cols))
    if out is not None:
        out.copy_(res)
        return out
    if res.dtype != in_dtype:
        res = res.to(in_dtype)
    return res
