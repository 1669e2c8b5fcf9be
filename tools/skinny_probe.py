#!/usr/bin/env python
"""Few-row Linear: cuBLAS (F.linear) against msda_skinny_linear_f32, per call inside a CUDA graph of 64 calls."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.ops.functions import skinny_linear  # noqa: E402


def graph_us(fn, reps=64):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                fn()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (20 * reps)


for M, N, K in [(12, 512, 512), (12, 16, 1024), (12, 1024, 16), (4, 512, 512), (30, 512, 512), (12, 2048, 1024), (12, 1608, 512),
                (100, 512, 512), (375, 512, 512), (100, 256, 512), (375, 2048, 512)]:
    x = torch.randn(M, K, device="cuda")
    w = torch.randn(N, K, device="cuda")
    b = torch.randn(N, device="cuda")
    t_ref = graph_us(lambda: torch.nn.functional.linear(x, w, b))
    t_new = graph_us(lambda: skinny_linear(x, w, b))
    print(f"M={M:3d} N={N:5d} K={K:5d}: cuBLAS {t_ref:6.2f} us   skinny {t_new:6.2f} us")
