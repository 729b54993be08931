# -*- coding: utf-8 -*-
"""The arithmetic of the slab-free attention kernels (csrc/attn/*.cuh), restated tile by tile in PyTorch and
checked against dense softmax / autograd on the CPU.

This pins the formulas the CUDA kernels implement - independent (row, column-half) online-softmax streams with a
LAZY running maximum (it only moves when it grew by more than 2^8), the merge of the two streams of a row, the
log2-domain log-sum-exp handed to the backward, and the backward's recomputation
``P = exp2(scale*log2e*S - L)``, ``dS = P * (scale*dP - scale*delta)`` - so a change of the kernels' math has a
CPU-side oracle (reference semantics: module.py:61-69 and its autograd)."""
import math

import pytest
import torch

LOG2E = 1.4426950408889634
TAU = 8.0


def _dense(k, q, v, mask, scale):
    s = (k @ q.transpose(-1, -2)) * scale
    if mask is not None:
        s = s.masked_fill(mask, float('-inf'))
    return torch.softmax(s, dim=-1) @ v, torch.logsumexp(s, dim=-1) * LOG2E


def _forward_streams(k, q, v, mask, scale, step=128):
    """O and L (log2 domain) of one head, computed like attn_fwd_kernel: per row two streams (column halves of
    every `step`-column block), each with its own lazily updated reference maximum, merged at the end."""
    m_rows, t = k.size(0), q.size(0)
    c = scale * LOG2E
    half = step // 2
    m_used = torch.full((2, m_rows), float('-inf'), dtype=k.dtype)
    l = torch.zeros(2, m_rows, dtype=k.dtype)
    o = torch.zeros(2, m_rows, v.size(1), dtype=k.dtype)
    for n0 in range(0, t, step):
        for h in range(2):
            lo, hi = n0 + h * half, min(n0 + (h + 1) * half, t)
            if lo >= hi:
                continue
            s = (k @ q[lo:hi].T) * c
            dead = mask[:, lo:hi] if mask is not None else torch.zeros_like(s, dtype=torch.bool)
            s = s.masked_fill(dead, float('-inf'))
            m_blk = s.max(dim=1).values
            need = m_blk > m_used[h] + TAU                      # lazy: the reference point moves rarely
            f = torch.where(need, torch.exp2(m_used[h] - m_blk), torch.ones_like(m_blk))
            f = torch.where(torch.isnan(f), torch.zeros_like(f), f)   # -inf - -inf never happens when need
            o[h] *= f[:, None]
            l[h] *= f
            m_used[h] = torch.where(need, m_blk, m_used[h])
            ref = torch.where(torch.isinf(m_used[h]), torch.zeros_like(m_blk), m_used[h])
            p = torch.exp2(s - ref[:, None])
            assert p.max() <= 2 ** TAU * 1.0001                # what keeps bf16 P in range in the kernel
            l[h] += p.sum(dim=1)
            o[h] += p @ v[lo:hi]
    m_all = torch.maximum(m_used[0], m_used[1])
    f = [torch.where(torch.isinf(m_used[h]), torch.zeros_like(m_all), torch.exp2(m_used[h] - m_all)) for h in range(2)]
    l_all = l[0] * f[0] + l[1] * f[1]
    out = (o[0] * f[0][:, None] + o[1] * f[1][:, None]) / l_all[:, None]
    return out, m_all + torch.log2(l_all)


def _backward_tiles(k, q, v, g, out, lse2, mask, scale, step=64):
    """dK, dQ, dV of one head, computed like attn_bwd_rows_kernel / attn_bwd_cols_kernel from -L and -scale*delta."""
    c = scale * LOG2E
    neg_lse = torch.nan_to_num(-lse2, nan=0.0, posinf=0.0, neginf=0.0)
    nds = torch.nan_to_num(-scale * (g * out).sum(dim=1), nan=0.0, posinf=0.0, neginf=0.0)
    dk, dq, dv = torch.zeros_like(k), torch.zeros_like(q), torch.zeros_like(v)
    for n0 in range(0, q.size(0), step):
        cols = slice(n0, min(n0 + step, q.size(0)))
        p = torch.exp2((k @ q[cols].T) * c + neg_lse[:, None])
        if mask is not None:
            p = p.masked_fill(mask[:, cols], 0.0)
        ds = p * ((g @ v[cols].T) * scale + nds[:, None])
        dk += ds @ q[cols]            # row side, accumulated over column blocks
        dq[cols] += ds.T @ k          # column side, accumulated over row blocks (and ranks)
        dv[cols] += p.T @ g
    return dk, dq, dv


@pytest.mark.parametrize('masked,ramp', [(False, False), (True, False), (True, True)])
def test_forward_streams_and_backward_tiles_match_dense(masked, ramp):
    torch.manual_seed(0)
    m_rows, t, d = 96, 520, 16
    k = torch.randn(m_rows, d, dtype=torch.float64)
    q = torch.randn(t, d, dtype=torch.float64)
    v = torch.randn(t, d, dtype=torch.float64)
    g = torch.randn(m_rows, d, dtype=torch.float64)
    if ramp:  # later columns carry much larger scores: forces the lazy maximum to move several times
        q = q * torch.linspace(1.0, 40.0, t, dtype=torch.float64)[:, None]
    mask = None
    if masked:
        mask = torch.rand(m_rows, t) < 0.3
        mask[7, :260] = True        # one stream of the row sees nothing for a long time
        mask[9, 64:128] = True      # a dead half block
    scale = 1.0 / math.sqrt(d)
    kk, qq, vv = (x.clone().requires_grad_(True) for x in (k, q, v))
    ref_out, ref_lse2 = _dense(kk, qq, vv, mask, scale)
    ref_out.backward(g)
    out, lse2 = _forward_streams(k, q, v, mask, scale)
    assert torch.allclose(out, ref_out.detach(), atol=1e-10)
    assert torch.allclose(lse2, ref_lse2.detach(), atol=1e-9)
    dk, dq, dv = _backward_tiles(k, q, v, g, out, lse2, mask, scale)
    assert torch.allclose(dk, kk.grad, atol=1e-9)
    assert torch.allclose(dq, qq.grad, atol=1e-9)
    assert torch.allclose(dv, vv.grad, atol=1e-9)


def test_fully_masked_row_is_nan_forward_and_zero_gradient():
    torch.manual_seed(1)
    m_rows, t, d = 8, 128, 8
    k, q, v, g = (torch.randn(n, d, dtype=torch.float64) for n in (m_rows, t, t, m_rows))
    mask = torch.zeros(m_rows, t, dtype=torch.bool)
    mask[3] = True
    scale = 1.0 / math.sqrt(d)
    out, lse2 = _forward_streams(k, q, v, mask, scale)
    assert torch.isnan(out[3]).all() and torch.isfinite(out[[0, 1, 2, 4, 5, 6, 7]]).all()
    dk, dq, dv = _backward_tiles(k, q, v, g, out, lse2, mask, scale)
    assert torch.equal(dk[3], torch.zeros(d, dtype=torch.float64))
    assert torch.isfinite(dq).all() and torch.isfinite(dv).all()
