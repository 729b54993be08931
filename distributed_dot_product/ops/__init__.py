# -*- coding: utf-8 -*-
"""Kernel-level operators (thin autograd/ctypes wrappers over libddp_b200.so)."""
from distributed_dot_product.ops.gemm import gemm  # noqa: F401
from distributed_dot_product.ops.softmax import masked_softmax, masked_softmax_  # noqa: F401
from distributed_dot_product.ops.linear import linear  # noqa: F401
