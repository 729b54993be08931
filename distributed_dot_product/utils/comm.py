# -*- coding: utf-8 -*-
"""Rank / world helpers with the reference's names (reference: utils/comm.py:13-30).

The reference initialises Horovod and MPI at import time; here nothing happens until the first call,
and the launch model is torchrun's (``RANK`` / ``WORLD_SIZE`` / ``LOCAL_RANK`` in the environment).
"""
from distributed_dot_product.parallel import runtime


def init(backend=None):
    """Optional explicit initialisation (idempotent); every helper below calls it lazily."""
    runtime.init(backend=backend)


def get_world_size():
    return runtime.get_world_size()


def get_rank():
    return runtime.get_rank()


def get_local_rank():
    return runtime.get_local_rank()


def is_main_process():
    return runtime.get_rank() == 0


def synchronize():
    """Host barrier across ranks (reference: utils/comm.py:25-30).  The GPU kernels do not need it:
    they synchronise through device-side flags in the symmetric heap."""
    runtime.barrier()


def shutdown():
    runtime.shutdown()
