#!/usr/bin/env python
"""Does cutting one backward call into batch chunks (memset + kernel [+ bf16 convert] per chunk, so that the fp32
accumulation image of a chunk stays in the 126 MB L2) pay?  Times fused backward on N=64 as 1, 2, 4, 8 calls."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402

for dtype in (torch.float32, torch.bfloat16):
    w = bench.make_workload(64, 1000, 8, 64, 4, 4, "encoder", "cuda:0", dtype=dtype, seed=100)
    for parts in (1, 2, 4, 8):
        nb = 64 // parts
        ops = []
        for i in range(parts):
            sub = dict(w)
            for k in ("value", "ref", "off", "logits", "grad_out"):
                sub[k] = w[k][i * nb:(i + 1) * nb]
            sub["N"] = nb
            ops.append(bench.FusedOp(sub))
        for _ in range(5):
            for op in ops:
                op.backward()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(30):
            for op in ops:
                op.backward()
        e.record()
        torch.cuda.synchronize()
        print(f"{str(dtype):16s} {parts} call(s) of N={nb}: backward {s.elapsed_time(e) / 30:.4f} ms")
        del ops
        torch.cuda.empty_cache()
