# -*- coding: utf-8 -*-
"""Non-local block (embedded Gaussian) over positions split across GPUs.

The reference README motivates the operator with Non-local neural networks (reference: README.md:7-9;
BASELINE.json config 5: video features with T = 16 x 56 x 56 = 50176 positions, d = 512).  The block flattens
the space-time positions of a feature map into the "time" axis; every rank holds a contiguous slice of the
positions, the pairwise term softmax(theta(x_i) . phi(x_j)) g(x_j) is ``DistributedDotProductAttn`` with one
head, and the output projection W_z plus the residual connection are local.
"""
import torch
import torch.nn as nn

from distributed_dot_product.module import DistributedDotProductAttn


class NonLocalBlock(nn.Module):
    """``z = x + W_z( softmax(theta(x) phi(x)^T / sqrt(d)) g(x) )`` on ``[B, P/N, C]`` position slices.

    ``channels``: feature dim C;  ``offset``: communication chunking of the underlying operator;
    ``zero_init``: start as identity (W_z = 0), the usual initialisation of non-local blocks.
    """

    def __init__(self, channels: int, offset: int = 1024, distributed: bool = True, zero_init: bool = True):
        super().__init__()
        self.attn = DistributedDotProductAttn(channels, num_heads=1, add_bias=False, offset=offset,
                                              distributed=distributed)
        self.norm = nn.LayerNorm(channels)
        if zero_init:
            nn.init.zeros_(self.attn.composition.weight)

    def forward(self, x: torch.Tensor, attn_mask: torch.Tensor = None) -> torch.Tensor:
        """``x``: ``[B, P/N, C]`` (this rank's positions, channels last).  A 5-D ``[B, C, t, h, w]`` local
        slab is also accepted and returned in the same layout."""
        five_d = x.dim() == 5
        if five_d:
            b, c, t, h, w = x.shape
            x = x.flatten(2).transpose(1, 2).contiguous()
        y = self.norm(x)
        z = x + self.attn(y, y, y, attn_mask)
        if five_d:
            z = z.transpose(1, 2).reshape(b, c, t, h, w)
        return z
