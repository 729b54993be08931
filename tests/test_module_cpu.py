# -*- coding: utf-8 -*-
"""``DistributedDotProductAttn`` against its dense twin (``distributed=False``) over gloo.

Reference recipe (tests/test_gradient.py:73-121): same weights, gather the inputs, run the dense module on
the whole sequence, compare input gradients and summed parameter gradients.  Extended with the forward
output, causal masks, biases, distinct query/value dims and offsets that do not divide.
This is BASELINE.json config 1 (CPU/gloo world_size=2, T=64, key_dim=32, heads=1, offset=32).
"""
import pytest
import torch

from dist_utils import run_distributed


def _worker(rank, world, length, dim, heads, bias, causal, offset, qdim):
    import torch.distributed as dist
    from distributed_dot_product.module import DistributedDotProductAttn
    from distributed_dot_product.parallel import allreduce_gradients, broadcast_parameters
    torch.manual_seed(100 + rank)
    qdim = qdim or dim
    model = DistributedDotProductAttn(dim, query_dim=qdim, num_heads=heads, add_bias=bias, offset=offset).double()
    twin = DistributedDotProductAttn(dim, query_dim=qdim, num_heads=heads, add_bias=bias, distributed=False).double()
    broadcast_parameters(model)
    twin.load_state_dict(model.state_dict())
    k = torch.rand(1, length, dim, dtype=torch.float64, requires_grad=True)
    q = torch.rand(1, length, qdim, dtype=torch.float64, requires_grad=True)
    t = length * world
    gmask = torch.zeros(1, t, t, dtype=torch.bool)
    if causal:
        gmask = torch.triu(torch.ones(t, t, dtype=torch.bool), diagonal=1).unsqueeze(0)
    mask = gmask[:, rank * length:(rank + 1) * length]
    out = model(k, q, k, mask)
    (out * out).sum().backward()

    def gather(x):
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x.detach().contiguous())
        return torch.cat(parts, dim=1)

    kg = gather(k).requires_grad_(True)
    qg = gather(q).requires_grad_(True)
    gt = twin(kg, qg, kg, gmask)
    (gt * gt).sum().backward()
    rows = slice(rank * length, (rank + 1) * length)
    assert torch.allclose(out, gt[:, rows], atol=1e-9)
    assert torch.allclose(k.grad, kg.grad[:, rows], atol=1e-9)
    assert torch.allclose(q.grad, qg.grad[:, rows], atol=1e-9)
    allreduce_gradients(model)
    for (n1, p1), (n2, p2) in zip(model.named_parameters(), twin.named_parameters()):
        assert n1 == n2
        assert torch.allclose(p1.grad, p2.grad, atol=1e-8), n1


@pytest.mark.parametrize('heads', [1, 4])
def test_module_vs_dense_twin(heads):
    run_distributed(_worker, 2, 18, 64, heads, False, False, 32, None)


def test_baseline_config1():
    # BASELINE.json config 1: world 2, T=64 (32 rows/rank), key_dim=32, heads=1, offset=32
    run_distributed(_worker, 2, 32, 32, 1, False, False, 32, None)


def test_module_causal_bias_ragged_offset_three_ranks():
    run_distributed(_worker, 3, 10, 32, 2, True, True, 7, None)


def test_module_distinct_query_dim():
    run_distributed(_worker, 2, 6, 16, 1, True, True, 4, 24)
