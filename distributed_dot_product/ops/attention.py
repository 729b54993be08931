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
import weakref

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


def flash_attention_forward(k4, q4, v4, mask, scale, offset, out4, want_lse=False, bits=None):
    """``out4 = softmax(scale * k4 @ gather(q4)^T masked) @ gather(v4)`` in ONE kernel (reference:
    module.py:61-69): scores stay in TMEM, probabilities go through shared memory back into the tensor
    core, the ``[T/N, T]`` slab is never written; the all-gather of ``q4`` / ``v4`` is fused in.
    Returns the fp32 row log-sum-exp (log2 domain) if ``want_lse``."""
    import ctypes
    from distributed_dot_product.parallel import runtime
    lib = _native.load()
    world = runtime.get_world_size()
    b, h, tn, _ = k4.shape
    total = world * q4.size(-2)
    ldw = sb1w = 0
    if mask is not None:
        if bits is None:
            bits, ldw, sb1w = _pack_mask(mask, b, h, tn, total, k4.device)
        else:
            ldw, sb1w = (total + 31) // 32, bits.numel() // b
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


#: packed masks of the most recent calls, keyed by mask object identity (see _pack_mask)
_MASK_CACHE = {}
#: ``DDP_MASK_CACHE=0`` re-packs the mask on every call.  The cache trusts the tensor's version counter: writes that
#: bypass it (``mask.data[...] = x``, a custom kernel, DLPack / numpy aliases) are not seen - disable the cache then.
MASK_CACHE = bool(int(__import__('os').environ.get('DDP_MASK_CACHE', '1')))


def _pack_mask(mask, b, h, tn, total, device, transposed=False):
    """Bit-packs a bool ``[B, T/N, T]`` mask (True = masked out) for the attention kernels
    (``ddp_attn_pack_mask``): one word per 32 columns of a row for the row-side kernels, or - ``transposed`` - per
    32 rows of a column for the column-side kernel, words of 32 neighbouring rows (columns) interleaved so that a
    warp's loads coalesce.  Returns ``(bits, ld_words, words_per_batch_entry)``."""
    import ctypes
    from distributed_dot_product.ops.softmax import _prep_mask
    lib = _native.load()
    t = 1 if transposed else 0
    ldw = ((tn if transposed else total) + 31) // 32
    # A mask is usually the same tensor step after step (causal / padding masks): the packed words are cached per
    # (storage, version counter, geometry) so that steady-state steps do not re-read the [T/N, T] byte mask.
    # Identity of the tensor OBJECT (weak reference) + its version counter: a new mask that merely lands on the
    # address of a freed one never matches.
    # (the stream is part of the key: words packed on one stream are not handed to a kernel on another one)
    key = (id(mask), t, b, tn, total, torch.cuda.current_stream().cuda_stream)
    hit = _MASK_CACHE.get(key) if MASK_CACHE else None
    if hit is not None:
        ref, version, cached = hit
        if ref() is mask and version == mask._version:
            return cached, ldw, cached.numel() // b
        del _MASK_CACHE[key]
    mu8, m_ld, m_sb1 = _prep_mask(mask, torch.empty((b, h, tn, total), device='meta'))
    words = lib.ddp_attn_mask_words(b, tn, total, t)
    bits = torch.empty(words, dtype=torch.int32, device=device)
    _native.check(lib.ddp_attn_pack_mask(ctypes.c_void_p(mu8.data_ptr()), b, tn, total, m_ld, m_sb1,
                                         ctypes.c_void_p(bits.data_ptr()), t, _g._stream()))
    _native.count_launch()
    if MASK_CACHE:
        if len(_MASK_CACHE) >= 4:
            _MASK_CACHE.pop(next(iter(_MASK_CACHE)))
        # the entry (and its device memory) goes away with the mask tensor, not only on the next lookup
        _MASK_CACHE[key] = (weakref.ref(mask, lambda _r, k=key: _MASK_CACHE.pop(k, None)), mask._version, bits)
    return bits, ldw, words // b


#: the backward kernels read the forward's all-gather of q / v in place when nothing else has used the symmetric
#: heap since (``DDP_GATHER_REUSE=0``: every kernel gathers for itself)
GATHER_REUSE = bool(int(__import__('os').environ.get('DDP_GATHER_REUSE', '1')))
#: how many backward passes skipped their two all-gathers (tests)
gather_reuse_count = 0


def gather_reusable(gather_tag, generation, seq, which):
    """May the two backward kernels read the all-gather of the forward call tagged ``gather_tag`` in place?

    Only when (a) reuse is enabled, (b) BOTH kernels run (the row-side call at ``seq + 1`` reads the other parity's
    workspace; it is the column-side call at ``seq + 2``, which owns that parity and does not push either, that keeps
    every pushing call into it at ``>= seq + 4``, i.e. behind a wait for the peers' ``seq + 2``), (c) the heap has not
    been re-allocated and (d) no collective call has used it since the forward."""
    return bool(GATHER_REUSE and which == 3 and gather_tag is not None and
                tuple(gather_tag) == (generation, seq))


def flash_attention_backward(k4, q4, v4, g4, neg_lse, nds, mask, scale, offset, dk4, dq4, dv4, bits=None,
                             which=3, gather_tag=None):
    """Backward of ``flash_attention_forward`` without any ``[T/N, T]`` slab: two deterministic kernels recompute
    the probability tiles from the forward's log-sum-exp (``neg_lse = -L``, ``nds = -scale * rowsum(dO * O)``,
    both fp32 ``[B, H, stat_ld]`` with ``stat_ld`` a multiple of 64).  ``dk4`` gets the row-side gradient
    (all-gather of q / v fused in), ``dq4`` / ``dv4`` the column-side ones (all-gather AND the reduce-scatter of
    the ranks' partial sums fused in).  ``which``: bit 0 = row-side kernel, bit 1 = column-side kernel (tools).

    ``gather_tag`` = ``(heap generation, sequence number)`` of the forward call that gathered these very ``q4`` /
    ``v4``: if no collective call has touched the symmetric heap since, its workspace parity still holds the complete
    gather and both kernels read it in place (no pushes, no flag waits: 2 x (N-1)/N x |q, v| less NVLink traffic
    and no start-up stall).  Safe under the epoch protocol: the row-side call (seq + 1) only READS the forward's
    parity, the column-side call (seq + 2) owns that parity anyway, and the next call that pushes into it
    (>= seq + 4) waits for every peer's seq + 2."""
    import ctypes
    from distributed_dot_product.parallel import runtime
    lib = _native.load()
    world = runtime.get_world_size()
    b, h, tn, _ = k4.shape
    total = world * q4.size(-2)
    stat_ld = neg_lse.size(-1)
    ldw = sb1w = ldwt = sb1wt = 0
    bits_t = None
    if mask is not None:
        if bits is None:
            bits, ldw, sb1w = _pack_mask(mask, b, h, tn, total, k4.device)
        else:
            ldw, sb1w = (total + 31) // 32, bits.numel() // b
        bits_t, ldwt, sb1wt = _pack_mask(mask, b, h, tn, total, k4.device, transposed=True)
    va, vb, vv, vg = _g.as_view(k4), _g.as_view(q4), _g.as_view(v4), _g.as_view(g4)
    vda, vdb, vdv = _g.as_view(dk4), _g.as_view(dq4), _g.as_view(dv4)
    nl, nd = ctypes.c_void_p(neg_lse.data_ptr()), ctypes.c_void_p(nds.data_ptr())
    chunk = _g.chunk_rows_from_offset(offset, q4.size(-2), 2 * b * h)

    reuse_ws = None
    if world > 1:
        # both kernels' needs first: growing the heap re-allocates it (and would drop a gather we mean to reuse)
        hp = runtime.heap()
        need_ws = need_flags = 0
        for bit, ws_fn in ((1, lib.ddp_attn_fwd_workspace), (2, lib.ddp_attn_bwd_cols_workspace)):
            if which & bit:
                ws, nf = ctypes.c_size_t(), ctypes.c_size_t()
                _native.check(ws_fn(ctypes.byref(vb), ctypes.byref(vv), world, chunk, ctypes.byref(ws),
                                    ctypes.byref(nf)))
                need_ws, need_flags = max(need_ws, ws.value), max(need_flags, nf.value)
        hp.ensure(need_ws, need_flags)
        if gather_reusable(gather_tag, hp.generation, hp.seq, which):
            reuse_ws = hp.comm[hp.seq & 1].ws[runtime.get_rank()]
            global gather_reuse_count
            gather_reuse_count += 1

    def comm_for():
        if world == 1:
            return None
        cm = hp.next_comm(chunk)
        cm.gather_ws = reuse_ws
        return ctypes.byref(cm)

    if which & 1:
        _native.check(lib.ddp_attn_bwd_rows(
            ctypes.byref(va), ctypes.byref(vg), ctypes.byref(vb), ctypes.byref(vv), ctypes.byref(vda), nl, nd, stat_ld,
            ctypes.c_void_p(bits.data_ptr()) if bits is not None else None, ldw, sb1w, float(scale),
            comm_for(), 0, _g._stream()))
        _native.count_launch()
    if which & 2:
        _native.check(lib.ddp_attn_bwd_cols(
            ctypes.byref(va), ctypes.byref(vg), ctypes.byref(vb), ctypes.byref(vv), ctypes.byref(vdv), ctypes.byref(vdb),
            nl, nd, stat_ld, ctypes.c_void_p(bits_t.data_ptr()) if bits_t is not None else None, ldwt, sb1wt,
            float(scale), comm_for(), 0, _g._stream()))
        _native.count_launch()


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
        from distributed_dot_product.parallel import runtime
        world = runtime.get_world_size()
        # the fused epilogue reads P through a TMA map: P rows must be 16-byte aligned and, across ranks, every
        # rank segment (T/N columns) must start on 16 bytes (ddp_gemm_ex refuses anything else with -2)
        fused = (FUSE_SOFTMAX_BWD and (probs.stride(-2) * probs.element_size()) % 16 == 0 and
                 (world == 1 or (v4.size(-2) * probs.element_size()) % 16 == 0))
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


class _FlashAttentionCore(torch.autograd.Function):
    """Training path without the probability slab: the forward keeps only O and the row log-sum-exp, the
    backward recomputes the tiles (``flash_attention_backward``).  Memory per rank is O(T/N * H * d) instead of
    O(T/N * T * H); a fully masked row yields NaN in the forward (as the reference) and zero gradients."""

    @staticmethod
    def forward(ctx, keys, queries, values, mask, num_heads, scale, offset):
        k4, q4, v4 = _heads(keys, num_heads), _heads(queries, num_heads), _heads(values, num_heads)
        b, h, tn, _ = k4.shape
        from distributed_dot_product.parallel import runtime
        total = runtime.get_world_size() * q4.size(-2)
        bits = None
        if mask is not None:
            bits, _, _ = _pack_mask(mask, b, h, tn, total, keys.device)
        out = torch.empty(keys.shape[:-1] + (values.size(-1),), dtype=values.dtype, device=values.device)
        lse = flash_attention_forward(k4, q4, v4, mask, scale, offset, _heads(out, num_heads), want_lse=True,
                                      bits=bits)
        ctx.save_for_backward(keys, queries, values, out, lse, mask, bits)
        ctx.num_heads, ctx.scale, ctx.offset = num_heads, scale, offset
        ctx.gather_tag = None
        if runtime.get_world_size() > 1:
            hp = runtime.heap()
            ctx.gather_tag = (hp.generation, hp.seq)   # the call that gathered queries / values
        return out

    @staticmethod
    def backward(ctx, grad_out):
        keys, queries, values, out, lse, mask, bits = ctx.saved_tensors
        h, scale, offset = ctx.num_heads, ctx.scale, ctx.offset
        if not grad_out.is_contiguous():
            grad_out = grad_out.contiguous()
        import ctypes
        b, tn, hd = grad_out.shape
        # delta_i = rowsum(dO_i * O_i) per (row, head); the kernels want -L and -scale*delta as [B, H, stat_ld],
        # zero-padded to a multiple of 64 rows: one native launch (ddp_attn_bwd_stats)
        stat_ld = (tn + 63) // 64 * 64
        stats = torch.empty(2, b, h, stat_ld, dtype=torch.float32, device=keys.device)
        neg_lse, nds = stats[0], stats[1]
        vg, vo = _g.as_view(_heads(grad_out, h)), _g.as_view(_heads(out, h))
        _native.check(_native.load().ddp_attn_bwd_stats(
            ctypes.byref(vg), ctypes.byref(vo), ctypes.c_void_p(lse.data_ptr()), ctypes.c_void_p(neg_lse.data_ptr()),
            ctypes.c_void_p(nds.data_ptr()), stat_ld, float(scale), _g._stream()))
        _native.count_launch()
        d_keys, d_queries, d_values = torch.empty_like(keys), torch.empty_like(queries), torch.empty_like(values)
        flash_attention_backward(_heads(keys, h), _heads(queries, h), _heads(values, h), _heads(grad_out, h),
                                 neg_lse, nds, mask, scale, offset, _heads(d_keys, h), _heads(d_queries, h),
                                 _heads(d_values, h), bits=bits, gather_tag=ctx.gather_tag)
        return d_keys, d_queries, d_values, None, None, None, None


#: train through the slab-free kernels too (forward keeps O + log-sum-exp, backward recomputes the tiles)
FLASH_TRAIN = bool(int(__import__('os').environ.get('DDP_FLASH_TRAIN', '1')))


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
    if FLASH_TRAIN and needs_grad and scale > 0:
        from distributed_dot_product.parallel import runtime
        k4, q4, v4 = _heads(keys, num_heads), _heads(queries, num_heads), _heads(values, num_heads)
        if flash_eligible(k4, q4, v4, mask, runtime.get_world_size()) and k4.size(-2) == q4.size(-2):
            return _FlashAttentionCore.apply(keys, queries, values, mask, num_heads, scale, offset)
    return _AttentionCore.apply(keys, queries, values, mask, num_heads, scale, offset)
