# -*- coding: utf-8 -*-
"""Blackwell-native sequence-parallel dot-product attention operators.

Same public surface as the reference package (reference: distributed_dot_product/__init__.py:9-10):
``DistributedDotProductAttn``, the three autograd ops, the three functional products and
``utils.comm``; the implementation underneath is hand-written sm_100a CUDA (tcgen05 / TMEM / TMA)
with the collectives fused into the GEMM kernels over NVLink peer memory.
"""

VERSION_INFO = (0, 1, 0, 'b200')
__version__ = '.'.join(map(str, VERSION_INFO))
