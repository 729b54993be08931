# -*- coding: utf-8 -*-
"""Slab-free attention forward (``ddp_attn_fwd``) against a plain PyTorch fp32 reference (one GPU).

Reference semantics: module.py:61-69 (scale, masked_fill(-inf), softmax over the gathered axis, attn @ V).
"""
import math
import os
import subprocess

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference(k, q, v, mask, heads, scale):
    b, tn, hd = k.shape
    d = hd // heads
    k4 = k.float().view(b, tn, heads, d).transpose(1, 2)
    q4 = q.float().view(b, q.size(1), heads, d).transpose(1, 2)
    v4 = v.float().view(b, v.size(1), heads, -1).transpose(1, 2)
    s = (k4 @ q4.transpose(-1, -2)) * scale
    if mask is not None:
        s = s.masked_fill(mask.unsqueeze(1), float('-inf'))
    p = torch.softmax(s, dim=-1)
    return (p @ v4).transpose(1, 2).reshape(b, tn, -1)


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
@pytest.mark.parametrize('b,heads,tn,t,d,masked', [
    (1, 1, 256, 256, 64, False),
    (2, 4, 300, 300, 32, True),     # ragged rows / columns, random mask
    (1, 16, 512, 1024, 64, True),   # the flagship head shape, more columns than rows
    (1, 2, 100, 72, 40, True),      # head dim not a power of two
])
def test_flash_forward_matches_reference(cuda_device, dtype, b, heads, tn, t, d, masked):
    from distributed_dot_product.ops.attention import flash_attention_forward, flash_eligible, _heads
    torch.manual_seed(1)
    k = torch.randn(b, tn, heads * d, device=cuda_device).to(dtype)
    q = torch.randn(b, t, heads * d, device=cuda_device).to(dtype)
    v = torch.randn(b, t, heads * d, device=cuda_device).to(dtype)
    mask = None
    if masked:
        mask = torch.rand(b, tn, t, device=cuda_device) < 0.3
        mask[:, 3, :] = True          # a fully masked row -> NaN, as softmax over -inf in the reference
        mask[:, 5, : t - 1] = True    # a single live column
    scale = 1.0 / math.sqrt(d)
    k4, q4, v4 = _heads(k, heads), _heads(q, heads), _heads(v, heads)
    assert flash_eligible(k4, q4, v4, mask, 1)
    out = torch.empty(b, tn, heads * d, device=cuda_device, dtype=dtype)
    lse = flash_attention_forward(k4, q4, v4, mask, scale, 64, _heads(out, heads), want_lse=True)
    ref = _reference(k, q, v, mask, heads, scale)
    got = out.float()
    if masked:
        assert torch.isnan(got[:, 3]).all() and torch.isnan(ref[:, 3]).all()
        got = got.clone()
        ref = ref.clone()
        got[:, 3] = 0
        ref[:, 3] = 0
    err = (got - ref).abs().max().item()
    assert err <= 2e-2 * max(1.0, ref.abs().max().item()), err
    # log-sum-exp (log2 domain) of the scaled, masked scores
    s = (_heads(k, heads).float() @ _heads(q, heads).float().transpose(-1, -2)) * scale
    if mask is not None:
        s = s.masked_fill(mask.unsqueeze(1), float('-inf'))
    lse_ref = torch.logsumexp(s, dim=-1) * math.log2(math.e)
    ok = torch.isfinite(lse_ref)
    assert (lse[ok] - lse_ref[ok]).abs().max().item() <= 2e-2


def test_module_inference_uses_flash_and_matches_training_path(cuda_device):
    """``DistributedDotProductAttn`` under ``torch.no_grad()`` runs the slab-free kernel; its output must agree
    with the training path (three kernels, probability slab kept) and with the dense twin."""
    from distributed_dot_product import _native
    from distributed_dot_product.module import DistributedDotProductAttn
    torch.manual_seed(2)
    m = DistributedDotProductAttn(256, num_heads=4, add_bias=True, offset=64).to(cuda_device).to(torch.bfloat16)
    x = torch.randn(2, 384, 256, device=cuda_device, dtype=torch.bfloat16)
    mask = torch.rand(2, 384, 384, device=cuda_device) < 0.2
    xg = x.clone().requires_grad_(True)
    train_out = m(xg, xg, xg, mask)
    before = _native.launch_count
    with torch.no_grad():
        infer_out = m(x, x, x, mask)
    launches = _native.launch_count - before
    # 4 projections + ONE attention kernel (the slab path needs 3 kernels for the core); the mask was packed by
    # the training call above and is served from the packed-mask cache (else one more launch)
    assert launches in (5, 6), launches
    err = (infer_out.float() - train_out.float()).abs().max().item()
    assert err <= 3e-2 * max(1.0, train_out.float().abs().max().item()), err


def test_attention_harness_emulated_ranks(cuda_device):
    """The standalone CUDA harness validates the fused all-gather of the attention kernel with emulated ranks on
    one GPU (produce phase of every rank, then consume phase) - see csrc/tests/attn_harness.cu."""
    exe = os.path.join(ROOT, 'build', 'attn_harness')
    if not os.path.exists(exe):
        subprocess.check_call(['python', os.path.join(ROOT, 'build.py'), 'harness'], cwd=ROOT)
    r = subprocess.run([exe, 'quick'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    tail = '\n'.join(r.stdout.splitlines()[-20:])
    assert r.returncode == 0, tail
    assert 'fail=0' in r.stdout, tail


@pytest.mark.parametrize('heads,tn,d,masked', [(4, 384, 64, True), (2, 200, 32, True), (16, 256, 64, False)])
def test_flash_training_matches_slab_path(cuda_device, heads, tn, d, masked):
    """Gradients of the slab-free training path (forward keeps O + log-sum-exp, two backward kernels recompute
    the tiles) against the slab path of the same layer and against an fp32 autograd reference."""
    import distributed_dot_product.ops.attention as att
    torch.manual_seed(3)
    b = 2
    k = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    q = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    v = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    g = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    mask = (torch.rand(b, tn, tn, device=cuda_device) < 0.3) if masked else None
    scale = 1.0 / math.sqrt(d)

    def run(flash):
        old = att.FLASH_TRAIN
        att.FLASH_TRAIN = flash
        try:
            kk, qq, vv = (t.clone().requires_grad_(True) for t in (k, q, v))
            out = att.attention_core(kk, qq, vv, mask, heads, scale, 64)
            out.backward(g)
            return out.detach().float(), kk.grad.float(), qq.grad.float(), vv.grad.float()
        finally:
            att.FLASH_TRAIN = old

    kf, qf, vf = (t.float().requires_grad_(True) for t in (k, q, v))
    ref_out = _reference(kf, qf, vf, mask, heads, scale)
    ref_out.backward(g.float())
    ref = (ref_out.detach(), kf.grad, qf.grad, vf.grad)
    flash, slab = run(True), run(False)
    for name, f, s, r in zip(('out', 'dK', 'dQ', 'dV'), flash, slab, ref):
        bound = 3e-2 * max(1.0, r.abs().max().item())
        assert (f - r).abs().max().item() <= bound, (name, 'flash vs fp32', (f - r).abs().max().item(), bound)
        assert (s - r).abs().max().item() <= bound, (name, 'slab vs fp32', (s - r).abs().max().item(), bound)


def test_attention_backward_harness_emulated_ranks(cuda_device):
    """Backward kernels incl. the fused reduce-scatter with emulated ranks (csrc/tests/attn_harness.cu, mode bwd)."""
    exe = os.path.join(ROOT, 'build', 'attn_harness')
    if not os.path.exists(exe):
        subprocess.check_call(['python', os.path.join(ROOT, 'build.py'), 'harness'], cwd=ROOT)
    r = subprocess.run([exe, 'bwd'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    tail = '\n'.join(r.stdout.splitlines()[-20:])
    assert r.returncode == 0, tail
    assert 'fail=0' in r.stdout, tail


def test_packed_mask_cache_follows_mask_contents(cuda_device):
    """The packed-mask cache is keyed by tensor identity + version counter: an in-place update of the mask and a
    brand-new mask tensor must both be seen by the next call."""
    from distributed_dot_product.ops.attention import attention_core
    torch.manual_seed(4)
    b, heads, tn, d = 1, 2, 128, 32
    k, q, v = (torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16) for _ in range(3))
    scale = 1.0 / math.sqrt(d)
    mask = torch.zeros(b, tn, tn, dtype=torch.bool, device=cuda_device)
    with torch.no_grad():
        o1 = attention_core(k, q, v, mask, heads, scale, 64).float()
        o1b = attention_core(k, q, v, mask, heads, scale, 64).float()     # cache hit
        mask[:, :, tn // 2:] = True                                      # in place: version counter moves
        o2 = attention_core(k, q, v, mask, heads, scale, 64).float()
        del mask
        mask3 = torch.zeros(b, tn, tn, dtype=torch.bool, device=cuda_device)   # may reuse the freed storage
        o3 = attention_core(k, q, v, mask3, heads, scale, 64).float()
    assert torch.equal(o1, o1b)
    ref2 = _reference(k, q, v, torch.cat([torch.zeros(b, tn, tn // 2, dtype=torch.bool, device=cuda_device),
                                          torch.ones(b, tn, tn // 2, dtype=torch.bool, device=cuda_device)], -1),
                      heads, scale)
    assert (o2 - ref2).abs().max().item() <= 2e-2 * max(1.0, ref2.abs().max().item())
    assert torch.equal(o3, o1)


def test_attention_harness_more_items_than_ctas(cuda_device):
    """160 work items on 148 CTAs (csrc/tests/attn_harness.cu, mode multi): every barrier phase, ring slot and
    step counter of the three attention kernels is carried from one work item into the next and checked element
    by element against a double-precision CPU reference."""
    exe = os.path.join(ROOT, 'build', 'attn_harness')
    if not os.path.exists(exe):
        subprocess.check_call(['python', os.path.join(ROOT, 'build.py'), 'harness'], cwd=ROOT)
    r = subprocess.run([exe, 'multi'], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    tail = '\n'.join(r.stdout.splitlines()[-20:])
    assert r.returncode == 0, tail
    assert 'fail=0' in r.stdout, tail


@pytest.mark.parametrize('mask_kind', ['random', 'causal'])
def test_flash_training_many_items_per_cta(cuda_device, mask_kind):
    """The flagship head shape at 256 work items per kernel (16 heads x 4096 rows = 256 row-block pairs on 148
    CTAs, so most CTAs run a second item): output and the three input gradients of the slab-free path, element
    by element, against a plain PyTorch fp32 reference (reference semantics: module.py:61-69 + autograd)."""
    import distributed_dot_product.ops.attention as att
    torch.manual_seed(6)
    b, heads, tn, d = 1, 16, 4096, 64
    k = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    q = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    v = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    g = torch.randn(b, tn, heads * d, device=cuda_device).to(torch.bfloat16)
    if mask_kind == 'random':
        mask = torch.rand(b, tn, tn, device=cuda_device) < 0.3
    else:
        mask = torch.triu(torch.ones(tn, tn, dtype=torch.bool, device=cuda_device), diagonal=1).unsqueeze(0)
    scale = 1.0 / math.sqrt(d)
    assert att.FLASH_TRAIN
    kk, qq, vv = (t.clone().requires_grad_(True) for t in (k, q, v))
    from distributed_dot_product import _native
    before = _native.launch_count
    out = att.attention_core(kk, qq, vv, mask, heads, scale, 64)
    out.backward(g)
    assert _native.launch_count - before <= 6  # pack x2, fwd, rowdot, bwd rows, bwd cols: the slab-free path ran
    got = (out.detach().float(), kk.grad.float(), qq.grad.float(), vv.grad.float())
    # fp32 reference, one head at a time (a [16, 4096, 4096] fp32 slab per tensor would also fit, but why)
    ref = [torch.zeros(b, tn, heads * d, device=cuda_device) for _ in range(4)]
    for h in range(heads):
        sl = slice(h * d, (h + 1) * d)
        kf, qf, vf = (t[..., sl].float().requires_grad_(True) for t in (k, q, v))
        s_ = (kf @ qf.transpose(-1, -2)) * scale
        s_ = s_.masked_fill(mask, float('-inf'))
        o = torch.softmax(s_, dim=-1) @ vf
        o.backward(g[..., sl].float())
        for dst, src in zip(ref, (o.detach(), kf.grad, qf.grad, vf.grad)):
            dst[..., sl] = src
    for name, f, r in zip(('out', 'dK', 'dQ', 'dV'), got, ref):
        err = (f - r).abs().max().item()
        bound = 3e-2 * max(1.0, r.abs().max().item())
        assert err <= bound, (name, mask_kind, err, bound)
        # a wrong 128-row block would leave the max error small relative to a large tensor: also check per block
        blk = (f - r).abs().view(b, tn // 128, 128, heads, d).amax(dim=(2, 4))
        assert blk.max().item() <= bound, (name, mask_kind, 'block', blk.max().item())
