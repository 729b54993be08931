# -*- coding: utf-8 -*-
"""Public surface parity with the reference package."""
import inspect

import pytest
import torch


def test_import_paths_and_version():
    import distributed_dot_product as pkg
    from distributed_dot_product.module import DistributedDotProductAttn  # noqa: F401
    from distributed_dot_product.multiplication.functions import (  # noqa: F401
        distributed_matmul_all, distributed_matmul_nt, distributed_matmul_tn, measure, DEBUG)
    from distributed_dot_product.multiplication.ops import (  # noqa: F401
        FullMultiplication, LeftTransposeMultiplication, RightTransposeMultiplication)
    from distributed_dot_product.utils.comm import (  # noqa: F401
        get_rank, get_world_size, is_main_process, synchronize)
    assert pkg.VERSION_INFO[:3] == (0, 1, 0)
    assert pkg.__version__.startswith('0.1.0')


def test_signatures_match_reference():
    from distributed_dot_product.module import DistributedDotProductAttn
    from distributed_dot_product.multiplication import functions as F
    sig = inspect.signature(DistributedDotProductAttn.__init__)
    # the reference's seven parameters, in its order; anything this framework adds comes after them with a default
    assert list(sig.parameters)[1:8] == ['key_dim', 'value_dim', 'query_dim', 'num_heads', 'add_bias', 'offset',
                                         'distributed']
    assert all(p.default is not inspect.Parameter.empty for p in list(sig.parameters.values())[8:])
    assert sig.parameters['offset'].default == 32 and sig.parameters['distributed'].default is True
    fwd = inspect.signature(DistributedDotProductAttn.forward)
    assert list(fwd.parameters)[1:] == ['keys', 'queries', 'values', 'attn_mask']
    assert list(inspect.signature(F.distributed_matmul_nt).parameters)[:3] == ['left', 'right', 'offset']
    assert list(inspect.signature(F.distributed_matmul_all).parameters)[:3] == ['left', 'right', 'offset']
    assert list(inspect.signature(F.distributed_matmul_tn).parameters)[:2] == ['left', 'right']
    assert inspect.signature(F.distributed_matmul_nt).parameters['offset'].default == 32


@pytest.mark.parametrize('bias', [False, True])
def test_state_dict_layout(bias):
    from distributed_dot_product.module import DistributedDotProductAttn
    m = DistributedDotProductAttn(8, value_dim=16, query_dim=12, add_bias=bias)
    keys = list(m.state_dict().keys())
    if bias:
        assert keys == ['keys.weight', 'keys.bias', 'queries.weight', 'queries.bias', 'values.weight',
                        'values.bias', 'composition.weight', 'composition.bias']
    else:
        assert keys == ['keys.weight', 'queries.weight', 'values.weight', 'composition.weight']
    sd = m.state_dict()
    assert tuple(sd['keys.weight'].shape) == (8, 8)
    assert tuple(sd['queries.weight'].shape) == (8, 12)
    assert tuple(sd['values.weight'].shape) == (16, 16)
    assert tuple(sd['composition.weight'].shape) == (16, 16)
    # a reference-shaped state_dict loads
    m2 = DistributedDotProductAttn(8, value_dim=16, query_dim=12, add_bias=bias, distributed=False)
    m2.load_state_dict({k: v.clone() for k, v in sd.items()})


def test_constructor_constraints():
    from distributed_dot_product.module import DistributedDotProductAttn
    with pytest.raises(AssertionError):
        DistributedDotProductAttn(10, num_heads=4)
    with pytest.raises(ValueError):
        DistributedDotProductAttn(8, value_dim=16, num_heads=2)
    m = DistributedDotProductAttn(8, num_heads=2, offset=7)
    assert m.dim == 4 and m.offset == 7 and m.num_heads == 2 and m.value_dim == 8


def test_float_mask_raises_and_none_mask_ok():
    from distributed_dot_product.module import DistributedDotProductAttn
    m = DistributedDotProductAttn(8)
    x = torch.rand(1, 5, 8)
    with pytest.raises(RuntimeError):
        m(x, x, x, torch.zeros(1, 5, 5))
    assert m(x, x, x, None).shape == (1, 5, 8)


def test_fully_masked_row_is_nan_like_reference():
    from distributed_dot_product.module import DistributedDotProductAttn
    m = DistributedDotProductAttn(8)
    x = torch.rand(1, 4, 8)
    mask = torch.zeros(1, 4, 4, dtype=torch.bool)
    mask[0, 2] = True
    out = m(x, x, x, mask)
    assert torch.isnan(out[0, 2]).all() and not torch.isnan(out[0, 0]).any()


def test_comm_single_process():
    from distributed_dot_product.utils import comm
    assert comm.get_world_size() == 1 and comm.get_rank() == 0 and comm.is_main_process()
    comm.synchronize()


def test_single_process_functions_match_dense():
    from distributed_dot_product.multiplication.functions import (
        distributed_matmul_all, distributed_matmul_nt, distributed_matmul_tn)
    a = torch.arange(2 * 5 * 6).view(2, 5, 6)
    b = torch.arange(2 * 5 * 6).view(2, 5, 6) - 7
    s = torch.arange(2 * 5 * 5).view(2, 5, 5)
    assert (distributed_matmul_nt(a, b, offset=2) == a.matmul(b.transpose(-1, -2))).all()
    assert (distributed_matmul_all(s, b, offset=4) == s.matmul(b)).all()
    assert (distributed_matmul_tn(s, b) == s.transpose(-1, -2).matmul(b)).all()
    # input dtype is preserved (the reference returns fp32 for nt/all regardless: SURVEY Q3)
    for dt in (torch.float64, torch.float32, torch.bfloat16, torch.int64):
        assert distributed_matmul_nt(a.to(dt), b.to(dt)).dtype == dt
        assert distributed_matmul_all(s.to(dt), b.to(dt)).dtype == dt
        assert distributed_matmul_tn(s.to(dt), b.to(dt)).dtype == dt


def test_debug_measure_prints(capsys, monkeypatch):
    from distributed_dot_product.multiplication import functions as F
    monkeypatch.setattr(F, 'DEBUG', True)
    a = torch.rand(1, 4, 3)
    F.distributed_matmul_nt(a, a)
    out = capsys.readouterr().out
    assert 'distributed_matmul_nt - Left' in out and 'elapsed time' in out


def test_native_library_contract():
    from distributed_dot_product import _native
    assert _native.lib_path().endswith('libddp_b200.so')


def test_example_script_runs_single_process():
    """example.py (reference: example.py:1-33) end to end in one CPU process."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES='', MASTER_ADDR='127.0.0.1')
    for k in ('RANK', 'WORLD_SIZE', 'LOCAL_RANK'):
        env.pop(k, None)
    r = subprocess.run([sys.executable, os.path.join(root, 'example.py')], cwd=root, env=env, timeout=600,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    assert 'loss' in r.stdout and '(1, 4096, 768)' in r.stdout, r.stdout[-500:]


def test_distributed_matmul_block_single_process():
    """The reference's dead helper (functions.py:151-157) works here: local product (+ sum over ranks)."""
    from distributed_dot_product.multiplication.functions import distributed_matmul_block
    a, b = torch.rand(2, 5, 7), torch.rand(2, 7, 3)
    assert torch.allclose(distributed_matmul_block(a, b), a @ b)
    assert torch.allclose(distributed_matmul_block(a, b, transpose=True), (a @ b).transpose(-1, -2))


def test_flash_eligibility_rules():
    """Which operands the slab-free attention kernels accept (everything else takes the slab path)."""
    from distributed_dot_product.ops.attention import _heads, flash_eligible
    def ops(dtype, hd, heads, tn=64):
        x = torch.zeros(1, tn, hd, dtype=dtype)
        return _heads(x, heads), _heads(x, heads), _heads(x, heads)
    assert flash_eligible(*ops(torch.bfloat16, 1024, 16), None, 1)
    assert flash_eligible(*ops(torch.float16, 64, 2), None, 1)
    assert not flash_eligible(*ops(torch.float32, 1024, 16), None, 1)          # tf32 tensors: slab path
    assert not flash_eligible(*ops(torch.bfloat16, 1024, 4), None, 1)          # head dim 256 > 64
    assert not flash_eligible(*ops(torch.bfloat16, 24, 2), None, 1)            # head dim 12: not a multiple of 8
    mask = torch.zeros(1, 48, 96, dtype=torch.bool)
    assert flash_eligible(*ops(torch.bfloat16, 128, 2, tn=64), mask, 1)
    assert not flash_eligible(*ops(torch.bfloat16, 128, 2, tn=48), mask, 2)    # masked, world > 1: T/N % 32 != 0
    assert not flash_eligible(*ops(torch.bfloat16, 128, 2), mask.float(), 1)   # only bool masks


def test_compute_dtype_argument():
    """``compute_dtype`` (mixed precision on the CUDA path) is validated, keeps the reference's ``state_dict`` and
    leaves the CPU composition (fp32 parameters, fp32 math) alone."""
    import pytest
    from distributed_dot_product.module import DistributedDotProductAttn
    with pytest.raises(ValueError):
        DistributedDotProductAttn(32, compute_dtype=torch.float64)
    torch.manual_seed(0)
    plain = DistributedDotProductAttn(32, num_heads=2, distributed=False)
    mixed = DistributedDotProductAttn(32, num_heads=2, distributed=False, compute_dtype=torch.bfloat16)
    assert list(mixed.state_dict().keys()) == list(plain.state_dict().keys())
    mixed.load_state_dict(plain.state_dict())
    x = torch.rand(1, 6, 32)
    assert torch.equal(mixed(x, x, x), plain(x, x, x))


def test_graphed_step_is_importable_without_cuda():
    from distributed_dot_product.utils.graphs import GraphedStep
    assert callable(GraphedStep)


def test_gather_reuse_rules(monkeypatch):
    """When the attention backward may read the forward's all-gather in place (ops/attention.py)."""
    import distributed_dot_product.ops.attention as att
    monkeypatch.setattr(att, 'GATHER_REUSE', True)
    assert att.gather_reusable((3, 7), 3, 7, 3)
    assert not att.gather_reusable((3, 7), 3, 8, 3)      # another collective call used the heap since the forward
    assert not att.gather_reusable((3, 7), 4, 7, 3)      # the heap was re-allocated
    assert not att.gather_reusable((3, 7), 3, 7, 1)      # one kernel only: the parity argument needs both
    assert not att.gather_reusable((3, 7), 3, 7, 2)
    assert not att.gather_reusable(None, 3, 7, 3)        # forward ran without a heap (one process) or outside autograd
    monkeypatch.setattr(att, 'GATHER_REUSE', False)
    assert not att.gather_reusable((3, 7), 3, 7, 3)
