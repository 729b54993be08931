# -*- coding: utf-8 -*-
"""Model families on the operator: Non-local block and Transformer encoder layer vs their dense twins."""
import torch

from dist_utils import run_distributed


def _worker(rank, world):
    import torch.distributed as dist
    from distributed_dot_product.models import NonLocalBlock, SequenceParallelTransformerLayer
    from distributed_dot_product.parallel import broadcast_parameters
    torch.manual_seed(50 + rank)
    tn, c = 6, 16
    for make in (lambda d: NonLocalBlock(c, offset=4, distributed=d, zero_init=False),
                 lambda d: SequenceParallelTransformerLayer(c, num_heads=2, d_ff=32, offset=4, distributed=d)):
        model, twin = make(True).double(), make(False).double()
        broadcast_parameters(model)
        twin.load_state_dict(model.state_dict())
        x = torch.rand(1, tn, c, dtype=torch.float64, requires_grad=True)
        out = model(x)
        out.sum().backward()
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x.detach())
        xg = torch.cat(parts, 1).requires_grad_(True)
        ref = twin(xg)
        ref.sum().backward()
        rows = slice(rank * tn, (rank + 1) * tn)
        assert torch.allclose(out, ref[:, rows], atol=1e-9)
        assert torch.allclose(x.grad, xg.grad[:, rows], atol=1e-9)


def test_models_two_ranks():
    run_distributed(_worker, 2)


def test_nonlocal_block_5d_identity_init():
    from distributed_dot_product.models import NonLocalBlock
    blk = NonLocalBlock(8)
    x = torch.rand(2, 8, 2, 3, 3)
    assert torch.allclose(blk(x), x)
