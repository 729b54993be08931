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
``dP`` never exists in memory: the softmax backward ``dS = scale * P * (dP - delta)`` runs in the epilogue
of the ``dO V^T`` kernel (``delta = rowsum(dO * O)``, the flash-attention identity), so the backward makes
five passes over a slab instead of nine.
"""
import torch

from distributed_dot_product import _native
import distributed_dot_product.ops.gemm as _g
from distributed_dot_product.ops.softmax import softmax_backward, softmax_forward


#: fuse the softmax backward into the epilogue of the dP product (set False to run the two-kernel path)
FUSE_SOFTMAX_BWD = True
#: fuse scale + mask + softmax into the K Q^T product (two passes over the tiles, no score slab).  Off by
#: default: on B200 the two exp passes are SFU-bound (measured 0.59 + 1.02 ms vs 0.41 + 0.93 ms for the
#: nt kernel + one-pass softmax kernel at T/N = T = 8192, 16 heads), so the unfused path is faster; the fused
#: path saves one slab of peak memory traffic and is kept for experiments (DDP_FUSE_SOFTMAX_FWD=1).
FUSE_SOFTMAX_FWD = bool(int(__import__('os').environ.get('DDP_FUSE_SOFTMAX_FWD', '0')))


def fused_softmax_nt(k4, q4, mask, scale, offset):
    """``softmax(scale * k4 @ gather(q4)^T masked)`` without ever materialising the scores.

    Launch 1 (``DDP_EPI_SOFTMAX_STATS``) computes every score tile on the tensor cores and keeps only the
    online-softmax (max, sum) per row and column tile; a tiny combine kernel folds them into per-row
    (max, 1/sum); launch 2 (``DDP_EPI_SOFTMAX_APPLY``) recomputes the tiles and writes the probabilities
    once.  The slab is written exactly once and never read: 1 slab pass instead of 3.
    """
    import ctypes
    from distributed_dot_product.ops.softmax import _prep_mask
    from distributed_dot_product.parallel import runtime
    lib = _native.load()
    world = runtime.get_world_size()
    b, h, tn, _ = k4.shape
    total = world * q4.size(-2)
    probs = _g._padded_empty((b, h, tn, total), k4.dtype, k4.device)
    bits, ldw, sb1w = None, 0, 0
    if mask is not None:
        mu8, m_ld, m_sb1 = _prep_mask(mask, probs)
        ldw = (total + 31) // 32
        bits = torch.empty(b, tn, ldw, dtype=torch.int32, device=k4.device)
        _native.check(lib.ddp_pack_mask(ctypes.c_void_p(mu8.data_ptr()), b, tn, total, m_ld, m_sb1,
                                        ctypes.c_void_p(bits.data_ptr()), ldw, _g._stream()))
        _native.count_launch()
        sb1w = tn * ldw
    n_rows = b * h * tn
    keep = {}

    def base(mode):
        epi = _native.Epilogue()
        epi.mode, epi.scale = mode, float(scale)
        epi.mask_bits = bits.data_ptr() if bits is not None else None
        epi.mask_ld_words, epi.mask_sb1_words = ldw, sb1w
        return epi

    def build_stats(va, vb, vc, world_):
        nt = lib.ddp_gemm_stats_tiles(ctypes.byref(va), ctypes.byref(vb), world_)
        keep['nt'] = nt
        keep['stats'] = torch.empty(n_rows * nt * 2, dtype=torch.float32, device=k4.device)
        epi = base(_native.EPI_SOFTMAX_STATS)
        epi.stats = keep['stats'].data_ptr()
        return epi

    _g.gemm(_native.OP_NT, k4, q4, offset=offset, out=probs, epilogue_builder=build_stats)
    rowstat = torch.empty(n_rows * 2, dtype=torch.float32, device=k4.device)
    _native.check(lib.ddp_softmax_combine(ctypes.c_void_p(keep['stats'].data_ptr()), n_rows, keep['nt'],
                                          ctypes.c_void_p(rowstat.data_ptr()), _g._stream()))
    _native.count_launch()

    def build_apply(va, vb, vc, world_):
        epi = base(_native.EPI_SOFTMAX_APPLY)
        epi.rowstat = rowstat.data_ptr()
        return epi

    _g.gemm(_native.OP_NT, k4, q4, offset=offset, out=probs, epilogue_builder=build_apply)
    return probs


#: run the no-grad forward through the slab-free attention kernel (``ddp_attn_fwd``) when the shapes allow it
FLASH_FWD = bool(int(__import__('os').environ.get('DDP_FLASH_FWD', '1')))


def flash_eligible(k4, q4, v4, mask, world):
    """True if ``ddp_attn_fwd`` covers these ``[B, H, rows, d]`` operands (16-bit, head dims <= 64, zero-copy
    views); everything else takes the three-kernel path."""
    if k4.dtype not in (torch.bfloat16, torch.float16) or q4.dtype != k4.dtype or v4.dtype != k4.dtype:
        return False
    d, dv = k4.size(-1), v4.size(-1)
    if d != q4.size(-1) or d % 8 or dv % 8 or not 8 <= d <= 64 or not 8 <= dv <= 64:
        return False
    if q4.size(-2) != v4.size(-2) or k4.dim() != 4:
        return False
    if mask is not None and (mask.dtype != torch.bool or (world > 1 and q4.size(-2) % 32)):
        return False
    return all(_g._aligned(t) for t in (k4, q4, v4))


def flash_attention_forward(k4, q4, v4, mask, scale, offset, out4, want_lse=False):
    """``out4 = softmax(scale * k4 @ gather(q4)^T masked) @ gather(v4)`` in ONE kernel (reference:
    module.py:61-69): scores stay in TMEM, probabilities go through shared memory back into the tensor
    core, the ``[T/N, T]`` slab is never written; the all-gather of ``q4`` / ``v4`` is fused in.
    Returns the fp32 row log-sum-exp (log2 domain) if ``want_lse``."""
    import ctypes
    from distributed_dot_product.ops.softmax import _prep_mask
    from distributed_dot_product.parallel import runtime
    lib = _native.load()
    world = runtime.get_world_size()
    b, h, tn, _ = k4.shape
    total = world * q4.size(-2)
    bits, ldw, sb1w = None, 0, 0
    if mask is not None:
        mu8, m_ld, m_sb1 = _prep_mask(mask, torch.empty((b, h, tn, total), device='meta'))
        ldw = (total + 31) // 32
        bits = torch.empty(b, tn, ldw, dtype=torch.int32, device=k4.device)
        _native.check(lib.ddp_pack_mask(ctypes.c_void_p(mu8.data_ptr()), b, tn, total, m_ld, m_sb1,
                                        ctypes.c_void_p(bits.data_ptr()), ldw, _g._stream()))
        _native.count_launch()
        sb1w = tn * ldw
    lse = torch.empty(b, h, tn, dtype=torch.float32, device=k4.device) if want_lse else None
    va, vb, vv, vo = _g.as_view(k4), _g.as_view(q4), _g.as_view(v4), _g.as_view(out4)
    comm_ptr = None
    if world > 1:
        chunk = _g.chunk_rows_from_offset(offset, q4.size(-2), 2 * b * h)
        ws, nf = ctypes.c_size_t(), ctypes.c_size_t()
        _native.check(lib.ddp_attn_fwd_workspace(ctypes.byref(vb), ctypes.byref(vv), world, chunk,
                                                 ctypes.byref(ws), ctypes.byref(nf)))
        hp = runtime.heap()
        hp.ensure(ws.value, nf.value)
        comm_ptr = ctypes.byref(hp.next_comm(chunk))
    _native.check(lib.ddp_attn_fwd(ctypes.byref(va), ctypes.byref(vb), ctypes.byref(vv), ctypes.byref(vo),
                                   ctypes.c_void_p(lse.data_ptr()) if lse is not None else None,
                                   ctypes.c_void_p(bits.data_ptr()) if bits is not None else None, ldw, sb1w,
                                   float(scale), comm_ptr, 0, _g._stream()))
    _native.count_launch()
    return lse


def _rowdot(a, b):
    """fp32 ``(a * b).sum(-1)`` of two 2-D CUDA tensors (``ddp_rowdot``)."""
    import ctypes
    lib = _native.load()
    out = torch.empty(a.size(0), dtype=torch.float32, device=a.device)
    va, vb = _g.as_view(a.unsqueeze(0).unsqueeze(0)), _g.as_view(b.unsqueeze(0).unsqueeze(0))
    _native.check(lib.ddp_rowdot(ctypes.byref(va), ctypes.byref(vb), ctypes.c_void_p(out.data_ptr()), _g._stream()))
    _native.count_launch()
    return out


def _heads(x, h):
    """[B, Tn, H*d] -> strided view [B, H, Tn, d]."""
    b, tn, hd = x.shape
    return x.view(b, tn, h, hd // h).transpose(1, 2)


class _AttentionCore(torch.autograd.Function):
    @staticmethod
    def forward(ctx, keys, queries, values, mask, num_heads, scale, offset):
        k4, q4, v4 = _heads(keys, num_heads), _heads(queries, num_heads), _heads(values, num_heads)
        from distributed_dot_product.parallel import runtime
        world = runtime.get_world_size()
        tn, d = k4.size(-2), k4.size(-1)
        es = keys.element_size()
        fuse_fwd = (FUSE_SOFTMAX_FWD and scale > 0 and q4.size(-2) == tn and (d * es) % 16 == 0 and
                    (world == 1 or tn % 32 == 0))
        if fuse_fwd:
            probs = fused_softmax_nt(k4, q4, mask, scale, offset)       # P, written once
        else:
            scores = _g.gemm(_native.OP_NT, k4, q4, offset=offset)      # [B, H, Tn, T]
            probs = softmax_forward(scores, mask, scale, inplace=True)
        out = torch.empty_like(values)
        _g.gemm(_native.OP_NN, probs, v4, offset=offset, out=_heads(out, num_heads))
        ctx.save_for_backward(keys, queries, values, probs, out)
        ctx.num_heads, ctx.scale, ctx.offset = num_heads, scale, offset
        return out

    @staticmethod
    def backward(ctx, grad_out):
        keys, queries, values, probs, out = ctx.saved_tensors
        h, scale, offset = ctx.num_heads, ctx.scale, ctx.offset
        if not grad_out.is_contiguous():
            grad_out = grad_out.contiguous()
        g4 = _heads(grad_out, h)
        k4, q4, v4 = _heads(keys, h), _heads(queries, h), _heads(values, h)
        b, tn, hd = grad_out.shape
        fused = FUSE_SOFTMAX_BWD and (probs.stride(-2) * probs.element_size()) % 16 == 0
        if fused:
            # delta_i = sum_j P_ij dP_ij = sum_d dO_id O_id: a [B*Tn*H, d] row dot instead of a slab pass
            delta = _rowdot(grad_out.view(b * tn * h, hd // h), out.view(b * tn * h, hd // h))
            delta = delta.view(b, tn, h).transpose(1, 2).contiguous()          # [B, H, Tn]
            d_scores = _g.gemm(_native.OP_NT, g4, v4, offset=offset,           # dS straight from the dP tiles
                               softmax_bwd=(probs, delta, scale))
        else:
            d_probs = _g.gemm(_native.OP_NT, g4, v4, offset=offset)            # dP  [B, H, Tn, T]
            d_scores = softmax_backward(probs, d_probs, scale, inplace=True)   # dS overwrites dP
        d_values = torch.empty_like(values)
        _g.gemm(_native.OP_TN, probs, g4, out=_heads(d_values, h))             # dV = P^T dO
        d_queries = torch.empty_like(queries)
        _g.gemm(_native.OP_TN, d_scores, k4, out=_heads(d_queries, h))         # dQ = dS^T K
        d_keys = torch.empty_like(keys)
        _g.gemm(_native.OP_NN, d_scores, q4, offset=offset, out=_heads(d_keys, h))  # dK = dS Q
        return d_keys, d_queries, d_values, None, None, None, None


def attention_core(keys, queries, values, mask, num_heads, scale, offset=32):
    """``softmax(scale * K Q^T masked) V`` with K/Q/V given as ``[B, T/N, H*d]`` projections.

    When no gradient is required (inference) and the shapes allow it, the whole thing is ONE slab-free
    kernel (``flash_attention_forward``); the training path keeps the probability slab for its backward."""
    needs_grad = torch.is_grad_enabled() and any(t.requires_grad for t in (keys, queries, values))
    if FLASH_FWD and not needs_grad and scale > 0:
        from distributed_dot_product.parallel import runtime
        k4, q4, v4 = _heads(keys, num_heads), _heads(queries, num_heads), _heads(values, num_heads)
        if flash_eligible(k4, q4, v4, mask, runtime.get_world_size()):
            out = torch.empty(keys.shape[:-1] + (values.size(-1),), dtype=values.dtype, device=values.device)
            flash_attention_forward(k4, q4, v4, mask, scale, offset, _heads(out, num_heads))
            return out
    return _AttentionCore.apply(keys, queries, values, mask, num_heads, scale, offset)
