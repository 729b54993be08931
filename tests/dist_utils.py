# -*- coding: utf-8 -*-
"""Spawn-based multi-process harness: N real processes on one host (gloo on CPU, NCCL + CUDA IPC on GPU),
the torchrun-style replacement of the reference's ``horovodrun -np N --mpi pytest`` (README.md:173-176)."""
import os
import socket
import sys
import traceback

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, fn, args, use_cuda, errq):
    os.environ['RANK'] = str(rank)
    os.environ['LOCAL_RANK'] = str(rank)
    os.environ['WORLD_SIZE'] = str(world)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    os.environ.setdefault('OMP_NUM_THREADS', '1')
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    if not use_cuda:
        os.environ['CUDA_VISIBLE_DEVICES'] = ''
    torch.set_num_threads(1)
    try:
        from distributed_dot_product.utils import comm
        comm.init()
        fn(rank, world, *args)
        comm.synchronize()
        comm.shutdown()
    except Exception:
        errq.put((rank, traceback.format_exc()))
        raise


def run_distributed(fn, world, *args, use_cuda=False, timeout=240):
    """Runs ``fn(rank, world, *args)`` in ``world`` processes and fails if any of them raised."""
    ctx = mp.get_context('spawn')
    errq = ctx.SimpleQueue()
    port = _free_port()
    procs = []
    for r in range(world):
        p = ctx.Process(target=_worker, args=(r, world, port, fn, args, use_cuda, errq))
        p.start()
        procs.append(p)
    for p in procs:
        p.join(timeout)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.terminate()
    errors = []
    while not errq.empty():
        errors.append(errq.get())
    if errors:
        raise AssertionError('\n'.join('rank %d:\n%s' % e for e in errors))
    if alive:
        raise AssertionError('distributed test timed out')
    bad = [p.exitcode for p in procs if p.exitcode != 0]
    if bad:
        raise AssertionError('worker exit codes: %s' % bad)
