# -*- coding: utf-8 -*-
"""Sequence-parallel Transformer encoder layers on ``DistributedDotProductAttn``.

Everything except attention is position-wise, so with the time axis split as ``T/N`` rows per GPU only the
attention operator communicates (reference motivation: README.md:7-9; BASELINE.json config 4: 16 heads,
d_model = 1024, T = 65536 on 8 GPUs).  Weights are replicated; call
``distributed_dot_product.parallel.allreduce_gradients(model)`` after ``backward()``.
"""
import torch
import torch.nn as nn

from distributed_dot_product.module import DistributedDotProductAttn
from distributed_dot_product.ops.linear import linear


class _MLP(nn.Module):
    def __init__(self, d_model, d_ff):
        super().__init__()
        self.up = nn.Linear(d_model, d_ff)
        self.down = nn.Linear(d_ff, d_model)

    def forward(self, x):
        h = torch.nn.functional.gelu(linear(x, self.up.weight, self.up.bias))
        return linear(h, self.down.weight, self.down.bias)


class SequenceParallelTransformerLayer(nn.Module):
    """Pre-norm encoder layer: ``x + Attn(LN(x))`` then ``x + MLP(LN(x))`` on ``[B, T/N, d_model]``."""

    def __init__(self, d_model: int = 1024, num_heads: int = 16, d_ff: int = None, offset: int = 1024,
                 distributed: bool = True):
        super().__init__()
        self.norm1 = nn.LayerNorm(d_model)
        self.attn = DistributedDotProductAttn(d_model, num_heads=num_heads, offset=offset, distributed=distributed)
        self.norm2 = nn.LayerNorm(d_model)
        self.mlp = _MLP(d_model, d_ff or 4 * d_model)

    def forward(self, x: torch.Tensor, attn_mask: torch.Tensor = None) -> torch.Tensor:
        y = self.norm1(x)
        x = x + self.attn(y, y, y, attn_mask)
        return x + self.mlp(self.norm2(x))


class SequenceParallelTransformerEncoder(nn.Module):
    def __init__(self, num_layers: int, d_model: int = 1024, num_heads: int = 16, d_ff: int = None,
                 offset: int = 1024, distributed: bool = True):
        super().__init__()
        self.layers = nn.ModuleList([
            SequenceParallelTransformerLayer(d_model, num_heads, d_ff, offset, distributed) for _ in range(num_layers)])
        self.norm = nn.LayerNorm(d_model)

    def forward(self, x: torch.Tensor, attn_mask: torch.Tensor = None) -> torch.Tensor:
        for layer in self.layers:
            x = layer(x, attn_mask)
        return self.norm(x)
