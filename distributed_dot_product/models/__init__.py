# -*- coding: utf-8 -*-
"""Model families built on the sequence-parallel attention operator.

The reference ships one module and motivates it with Transformers and Non-local blocks
(reference: README.md:7-9); these are those two users, plus the layer itself.
"""
from distributed_dot_product.module import DistributedDotProductAttn  # noqa: F401
from distributed_dot_product.models.nonlocal_block import NonLocalBlock  # noqa: F401
from distributed_dot_product.models.transformer import (  # noqa: F401
    SequenceParallelTransformerLayer, SequenceParallelTransformerEncoder)
