# -*- coding: utf-8 -*-
"""Microbenchmark of the in-place masked softmax at long rows on one GPU: [1, 16, rows, T] bf16."""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_dot_product.ops.softmax import softmax_forward  # noqa: E402
dev = torch.device('cuda')
for T, rows in ((8192, 8192), (32768, 2048), (65536, 1024)):
    x = torch.randn(1, 16, rows, T, device=dev, dtype=torch.bfloat16)
    mask = torch.zeros(1, rows, T, device=dev).bool()
    for _ in range(2):
        softmax_forward(x, mask, 0.125, inplace=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        softmax_forward(x, mask, 0.125, inplace=True)
    e1.record(); e1.synchronize()
    ms = e0.elapsed_time(e1) / 5
    gb = x.numel() * 2 * 2 / 1e9
    print('T=%d rows=%d: %.3f ms  %.2f TB/s (read+write)' % (T, rows, ms, gb / ms))
