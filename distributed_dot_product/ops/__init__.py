# -*- coding: utf-8 -*-
"""Kernel-level operators (thin autograd/ctypes wrappers over libddp_b200.so).

Submodules: ``gemm`` (fused collective GEMM front end), ``softmax``, ``linear``, ``attention``.
Function re-exports use names that do not shadow the submodules.
"""
from distributed_dot_product.ops.gemm import gemm as fused_gemm  # noqa: F401
from distributed_dot_product.ops.softmax import masked_softmax, masked_softmax_  # noqa: F401
from distributed_dot_product.ops.linear import linear as linear_fn  # noqa: F401
from distributed_dot_product.ops.attention import (  # noqa: F401
    attention_core, flash_attention_backward, flash_attention_forward, flash_eligible)
