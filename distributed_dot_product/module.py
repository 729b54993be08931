# -*- coding: utf-8 -*-
"""``DistributedDotProductAttn``: multi-head dot-product attention over a sequence split across ranks.

Same constructor, ``forward(keys, queries, values, attn_mask)`` signature, ``distributed`` switch and
``state_dict`` layout (``keys`` / ``queries`` / ``values`` / ``composition`` Linear layers, in that
order) as the reference (reference: distributed_dot_product/module.py:22-76), including its
orientation: the score slab is ``K Q^T`` - rows are this rank's *keys*, columns are all *queries*.

On CUDA every stage is a hand-written sm_100a kernel: the four projections run on the tcgen05 GEMM with
the bias in the epilogue, the head split / merge is expressed in the TMA tensor maps (strided views, no
copies), ``K Q^T`` and ``attn V`` are the fused all-gather->GEMM kernels, and scale + mask + softmax is
one in-place pass over the slab.
"""
import math

import torch
import torch.nn as nn
from torch import Tensor

from distributed_dot_product.multiplication.ops import (
    FullMultiplication, RightTransposeMultiplication)
from distributed_dot_product.ops.attention import attention_core
from distributed_dot_product.ops.linear import linear


class DistributedDotProductAttn(nn.Module):
    def __init__(self, key_dim: int, value_dim: int = None, query_dim: int = None, num_heads: int = 1,
                 add_bias: bool = False, offset: int = 32, distributed: bool = True,
                 compute_dtype: torch.dtype = None):
        super(DistributedDotProductAttn, self).__init__()
        assert key_dim % num_heads == 0
        value_dim = value_dim if value_dim is not None else key_dim
        query_dim = query_dim if query_dim is not None else key_dim
        if num_heads > 1 and value_dim != key_dim:
            # the reference views V with key_dim // num_heads and fails later with a cryptic error
            # (module.py:54); fail here with a clear one.
            raise ValueError('value_dim must equal key_dim when num_heads > 1 (got %d and %d)'
                             % (value_dim, key_dim))
        self.num_heads = num_heads
        self.value_dim = value_dim
        self.offset = offset
        self.distributed = distributed
        # Mixed precision (not in the reference, whose ops fail under autocast): fp32 parameters and activations,
        # bf16 / fp16 tensor-core math.  The inputs and the four weights are cast once per call and everything
        # between them runs on the slab-free 16-bit kernels; the output comes back in the input dtype and the
        # parameter gradients in fp32.  ``None`` follows ``torch.autocast('cuda', ...)`` when that is active.
        if compute_dtype not in (None, torch.bfloat16, torch.float16):
            raise ValueError('compute_dtype must be None, torch.bfloat16 or torch.float16')
        self.compute_dtype = compute_dtype
        self.dim = key_dim // self.num_heads
        self.keys = nn.Linear(key_dim, key_dim, bias=add_bias)
        self.queries = nn.Linear(query_dim, key_dim, bias=add_bias)
        self.values = nn.Linear(value_dim, value_dim, bias=add_bias)
        self.composition = nn.Linear(value_dim, value_dim, bias=add_bias)

    def _native_path(self, x):
        return (self.distributed and x.is_cuda and
                x.dtype in (torch.float16, torch.bfloat16, torch.float32))

    def forward(self, keys: Tensor, queries: Tensor, values: Tensor, attn_mask: Tensor = None) -> Tensor:
        if attn_mask is not None and attn_mask.dtype != torch.bool:
            raise RuntimeError('masked_fill only supports boolean masks: attn_mask must be a bool tensor '
                               '(True = masked out)')
        scale = 1.0 / math.sqrt(self.dim)
        if self._native_path(keys) and keys.dim() == 3:
            # ---- B200 path: every stage is a hand-written sm_100a kernel
            cd = self.compute_dtype
            if cd is None and torch.is_autocast_enabled('cuda'):
                cd = torch.get_autocast_dtype('cuda')
                cd = cd if cd in (torch.bfloat16, torch.float16) else None
            if cd is not None and keys.dtype == torch.float32:
                seen = {}

                def low(t):  # one cast per distinct tensor (self-attention passes the same one three times)
                    if t is None:
                        return None
                    if id(t) not in seen:
                        seen[id(t)] = t.to(cd)
                    return seen[id(t)]

                k = linear(low(keys), low(self.keys.weight), low(self.keys.bias))
                q = linear(low(queries), low(self.queries.weight), low(self.queries.bias))
                v = linear(low(values), low(self.values.weight), low(self.values.bias))
                o = attention_core(k, q, v, attn_mask, self.num_heads, scale, self.offset)
                o = linear(o, low(self.composition.weight), low(self.composition.bias))
                return o if self.compute_dtype is None else o.to(keys.dtype)   # autocast keeps the low dtype
            k = linear(keys, self.keys.weight, self.keys.bias)
            q = linear(queries, self.queries.weight, self.queries.bias)
            v = linear(values, self.values.weight, self.values.bias)
            o = attention_core(k, q, v, attn_mask, self.num_heads, scale, self.offset)
            return linear(o, self.composition.weight, self.composition.bias)

        # ---- reference composition (CPU / gloo plumbing, and the dense oracle when distributed=False)
        keys = self.keys(keys)
        queries = self.queries(queries)
        values = self.values(values)
        if self.num_heads > 1:
            if attn_mask is not None:
                attn_mask = attn_mask.unsqueeze(1)
                attn_mask = attn_mask.expand(attn_mask.size(0), self.num_heads, *attn_mask.size()[2:])
            keys = keys.view(*keys.size()[:-1], self.num_heads, self.dim).transpose(-2, -3)
            queries = queries.view(*queries.size()[:-1], self.num_heads, self.dim).transpose(-2, -3)
            values = values.view(*values.size()[:-1], self.num_heads, self.dim).transpose(-2, -3)
        if self.distributed:
            projection = RightTransposeMultiplication.apply(keys, queries, self.offset)
        else:
            projection = torch.matmul(keys, queries.transpose(-1, -2))
        projection = projection * scale
        if attn_mask is not None:
            projection = projection.masked_fill(attn_mask, -float('inf'))
        attn = torch.softmax(projection, dim=-1)
        if self.distributed:
            outputs = FullMultiplication.apply(attn, values, self.offset)
        else:
            outputs = torch.matmul(attn, values)
        if self.num_heads > 1:
            outputs = outputs.transpose(-3, -2)
            outputs = outputs.reshape(*outputs.size()[:-2], self.value_dim)
        return self.composition(outputs)
