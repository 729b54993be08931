# -*- coding: utf-8 -*-
"""Sequence-parallel runtime: bootstrap, symmetric heap, gradient helpers."""
from distributed_dot_product.parallel.runtime import (  # noqa: F401
    barrier, check_status, get_local_rank, get_rank, get_world_size, heap, init, shutdown)
from distributed_dot_product.parallel.grads import (  # noqa: F401
    allreduce_gradients, broadcast_parameters)
