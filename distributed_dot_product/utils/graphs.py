# -*- coding: utf-8 -*-
"""CUDA-graph capture of a whole training step (not in the reference, which launches op by op).

One fwd+bwd step of the flagship layer is ~26 kernel launches, half of them 20 us projections: on one GPU the host
cannot enqueue them as fast as the device retires them (about 0.2 ms of a 2.0 ms step are launch gaps).  A captured
step replays as ONE launch.

Single-process only: the fused collective kernels take their sequence number (the epoch of the symmetric-heap flags)
as a launch parameter, so a replay would present the flags of the captured call.  ``GraphedStep`` refuses to capture
when the world size is above 1.
"""
import torch


class GraphedStep(object):
    """``GraphedStep(fn)`` runs ``fn()`` a few times on a side stream (lazy initialisations, packed-mask cache,
    allocator warm-up), captures one more call, and replays the capture on every ``__call__``; the value ``fn``
    returned during capture (tensors in the graph's private pool, overwritten by every replay) is returned each
    time.  ``fn`` must read its inputs from tensors that keep their address (copy new data INTO them) and must not
    synchronise."""

    def __init__(self, fn, warmup=3):
        from distributed_dot_product.parallel import runtime
        if runtime.get_world_size() > 1:
            raise RuntimeError('GraphedStep: the collective kernels cannot be replayed from a CUDA graph '
                               '(their sequence number is a launch parameter)')
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(int(warmup), 1)):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        # captured on the stream the warm-up ran on: caches keyed by stream (packed masks) hit during capture
        with torch.cuda.graph(self.graph, stream=side):
            self.result = fn()

    def __call__(self):
        self.graph.replay()
        return self.result
