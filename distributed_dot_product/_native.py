# -*- coding: utf-8 -*-
"""ctypes binding of libddp_b200.so (C ABI in include/ddp_b200.h).

The library is built in-tree by ``build.py`` (nvcc, sm_100a) and loaded lazily the first time a CUDA
tensor reaches one of the ops.  On a machine with a GPU a missing library is a hard error: the ops
never fall back to PyTorch kernels silently.
"""
import ctypes
import os
import threading

MAX_WORLD = 16
F16, BF16, F32 = 0, 1, 2
OP_NT, OP_NN, OP_TN = 0, 1, 2

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_lib', 'libddp_b200.so')
_lib = None
_lock = threading.Lock()


class Tensor(ctypes.Structure):
    _fields_ = [('ptr', ctypes.c_void_p), ('rows', ctypes.c_int64), ('cols', ctypes.c_int64),
                ('ld', ctypes.c_int64), ('b0', ctypes.c_int64), ('b1', ctypes.c_int64),
                ('sb0', ctypes.c_int64), ('sb1', ctypes.c_int64), ('dtype', ctypes.c_int32),
                ('_pad', ctypes.c_int32)]


class Comm(ctypes.Structure):
    _fields_ = [('world', ctypes.c_int32), ('rank', ctypes.c_int32), ('seq', ctypes.c_uint32),
                ('phase_mask', ctypes.c_int32), ('chunk_rows', ctypes.c_int32), ('_pad', ctypes.c_int32),
                ('ws', ctypes.c_void_p * MAX_WORLD), ('flags', ctypes.c_void_p * MAX_WORLD),
                ('done', ctypes.c_void_p * MAX_WORLD), ('exit_counter', ctypes.c_void_p),
                ('status', ctypes.c_void_p), ('mc_ws', ctypes.c_void_p), ('gather_ws', ctypes.c_void_p)]


class Epilogue(ctypes.Structure):
    _fields_ = [('mode', ctypes.c_int32), ('scale', ctypes.c_float), ('P', ctypes.POINTER(Tensor)),
                ('delta', ctypes.c_void_p), ('mask_bits', ctypes.c_void_p), ('mask_ld_words', ctypes.c_int64),
                ('mask_sb1_words', ctypes.c_int64), ('stats', ctypes.c_void_p), ('rowstat', ctypes.c_void_p)]


EPI_NONE, EPI_SOFTMAX_BWD, EPI_SOFTMAX_STATS, EPI_SOFTMAX_APPLY = 0, 1, 2, 3


class NativeError(RuntimeError):
    pass


def lib_path():
    return _LIB_PATH


def available():
    return os.path.exists(_LIB_PATH)


def load():
    """Returns the loaded library, building nothing: raises if the .so is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(_LIB_PATH):
            raise NativeError(
                'libddp_b200.so not found at %s - run `python build.py` (nvcc, sm_100a) first; '
                'CUDA tensors are never routed to a PyTorch fallback' % _LIB_PATH)
        lib = ctypes.CDLL(_LIB_PATH)
        c = ctypes
        tp = c.POINTER(Tensor)
        lib.ddp_last_error.restype = c.c_char_p
        lib.ddp_version.restype = c.c_int
        lib.ddp_num_sms.restype = c.c_int
        lib.ddp_gemm_workspace.argtypes = [c.c_int, tp, tp, c.c_int, c.c_int, c.POINTER(c.c_size_t),
                                           c.POINTER(c.c_size_t)]
        lib.ddp_gemm.argtypes = [c.c_int, tp, tp, tp, c.c_float, c.c_void_p, c.POINTER(Comm), c.c_int,
                                 c.c_void_p]
        lib.ddp_gemm_ex.argtypes = [c.c_int, tp, tp, tp, c.c_float, c.c_void_p, c.POINTER(Comm),
                                    c.POINTER(Epilogue), c.c_int, c.c_void_p]
        lib.ddp_gemm_stats_tiles.argtypes = [tp, tp, c.c_int]
        lib.ddp_pack_mask.argtypes = [c.c_void_p, c.c_int64, c.c_int64, c.c_int64, c.c_int64, c.c_int64, c.c_void_p,
                                      c.c_int64, c.c_void_p]
        lib.ddp_softmax_combine.argtypes = [c.c_void_p, c.c_int64, c.c_int32, c.c_void_p, c.c_void_p]
        lib.ddp_softmax_fwd.argtypes = [tp, tp, c.c_void_p, c.c_int64, c.c_int64, c.c_float, c.c_void_p]
        lib.ddp_softmax_bwd.argtypes = [tp, tp, tp, c.c_void_p, c.c_float, c.c_void_p]
        lib.ddp_rowdot.argtypes = [tp, tp, c.c_void_p, c.c_void_p]
        lib.ddp_colsum_scratch_bytes.argtypes = [c.c_int64, c.c_int64]
        lib.ddp_colsum_scratch_bytes.restype = c.c_size_t
        lib.ddp_colsum.argtypes = [tp, c.c_void_p, c.c_void_p, c.c_size_t, c.c_void_p]
        lib.ddp_sum_partials.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_int, c.c_int64, c.c_void_p]
        lib.ddp_heap_alloc.argtypes = [c.c_size_t, c.POINTER(c.c_void_p), c.c_void_p]
        lib.ddp_heap_open.argtypes = [c.c_void_p, c.POINTER(c.c_void_p)]
        lib.ddp_heap_close.argtypes = [c.c_void_p]
        lib.ddp_heap_free.argtypes = [c.c_void_p]
        lib.ddp_read_u32_async.argtypes = [c.c_void_p, c.c_void_p, c.c_void_p]
        lib.ddp_vmm_query.argtypes = [c.POINTER(c.c_int), c.POINTER(c.c_int), c.POINTER(c.c_size_t)]
        lib.ddp_vmm_alloc.argtypes = [c.c_size_t, c.POINTER(c.c_void_p), c.POINTER(c.c_int)]
        lib.ddp_vmm_open.argtypes = [c.c_int, c.c_size_t, c.POINTER(c.c_void_p)]
        lib.ddp_vmm_release.argtypes = [c.c_void_p]
        lib.ddp_mc_create.argtypes = [c.c_size_t, c.c_int, c.POINTER(c.c_ulonglong), c.POINTER(c.c_int)]
        lib.ddp_mc_import.argtypes = [c.c_int, c.POINTER(c.c_ulonglong)]
        lib.ddp_mc_add_device.argtypes = [c.c_ulonglong]
        lib.ddp_mc_bind.argtypes = [c.c_ulonglong, c.c_void_p, c.c_size_t, c.POINTER(c.c_void_p)]
        lib.ddp_attn_fwd_workspace.argtypes = [tp, tp, c.c_int, c.c_int, c.POINTER(c.c_size_t), c.POINTER(c.c_size_t)]
        lib.ddp_attn_fwd.argtypes = [tp, tp, tp, tp, c.c_void_p, c.c_void_p, c.c_int64, c.c_int64, c.c_float,
                                     c.POINTER(Comm), c.c_int, c.c_void_p]
        lib.ddp_debug_timeout_site_attn.restype = c.c_int
        lib.ddp_attn_bwd_stats.argtypes = [tp, tp, c.c_void_p, c.c_void_p, c.c_void_p, c.c_int64, c.c_float, c.c_void_p]
        lib.ddp_attn_bwd_rows.argtypes = [tp, tp, tp, tp, tp, c.c_void_p, c.c_void_p, c.c_int64, c.c_void_p, c.c_int64,
                                          c.c_int64, c.c_float, c.POINTER(Comm), c.c_int, c.c_void_p]
        lib.ddp_attn_bwd_cols_workspace.argtypes = [tp, tp, c.c_int, c.c_int, c.POINTER(c.c_size_t),
                                                    c.POINTER(c.c_size_t)]
        lib.ddp_attn_bwd_cols.argtypes = [tp, tp, tp, tp, tp, tp, c.c_void_p, c.c_void_p, c.c_int64, c.c_void_p,
                                          c.c_int64, c.c_int64, c.c_float, c.POINTER(Comm), c.c_int, c.c_void_p]
        lib.ddp_attn_mask_words.argtypes = [c.c_int64, c.c_int64, c.c_int64, c.c_int]
        lib.ddp_attn_mask_words.restype = c.c_size_t
        lib.ddp_attn_pack_mask.argtypes = [c.c_void_p, c.c_int64, c.c_int64, c.c_int64, c.c_int64, c.c_int64, c.c_void_p,
                                           c.c_int, c.c_void_p]
        lib.ddp_debug_timeout_site.restype = c.c_int
        lib.ddp_debug_timeout_site_rowops.restype = c.c_int
        lib.ddp_read_u32.argtypes = [c.c_void_p, c.POINTER(c.c_uint32)]
        lib.ddp_memset_async.argtypes = [c.c_void_p, c.c_int, c.c_size_t, c.c_void_p]
        _lib = lib
    return _lib


def last_error():
    return load().ddp_last_error().decode()


def check(code):
    if code != 0:
        raise NativeError('libddp_b200 error %d: %s' % (code, load().ddp_last_error().decode()))


#: number of kernels launched through this binding (bench.py reports it as ``gpu_launches``)
launch_count = 0


def count_launch(n=1):
    global launch_count
    launch_count += n
