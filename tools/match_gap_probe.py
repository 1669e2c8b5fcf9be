#!/usr/bin/env python
"""What the host round trip of the Hungarian matching costs the PDVC step, interleaved A/B on one box:
StagedStep(device_match=False) -- cost matrices read back, scipy on the host, indices uploaded (pdvc/matcher.py:113-117) --
against device_match=True (csrc/msda_match.cuh inside stage 1's graph; the host never waits inside a step).  Also the
kernel's own time on the step's matrices.
    python tools/match_gap_probe.py [config] [steps]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.ops.functions import linear_sum_assignment_cuda  # noqa: E402
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402


def ms_per_step(name, steps, device_match):
    tr = Trainer(name, "cuda:0", mode="graphs")
    tr.net.device_match = device_match
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(15):
        tr.step(video)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = tr.step(video)
    float(loss)
    torch.cuda.synchronize()
    out = (time.perf_counter() - t0) / steps * 1e3
    shape = (tr.cfg.dec_layers, tr.cfg.num_queries, tr.n_gt)
    del tr
    torch.cuda.empty_cache()
    return out, shape


def kernel_us(shape, reps=200):
    costs = torch.randn(*shape, device="cuda")
    for _ in range(10):
        linear_sum_assignment_cuda(costs)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        linear_sum_assignment_cuda(costs)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "yc2_tsp_pdvc_yc2shape"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 150
    rows = {False: [], True: []}
    shape = None
    for _ in range(3):
        for flag in (False, True):
            ms, shape = ms_per_step(name, steps, flag)
            rows[flag].append(ms)
    for flag in (False, True):
        print(f"{name} device_match={flag}: " + " ".join(f"{v:.3f}" for v in rows[flag]) + f"  ms/step (min {min(rows[flag]):.3f})")
    for shp in (shape, (2, 10, 4), (4, 100, 30), (6, 300, 100), (1, 1024, 64)):
        print(f"msda_lsap_f32 {shp} (matrices, queries, events): {kernel_us(shp):.1f} us per call (back-to-back launches, random costs)")


if __name__ == "__main__":
    main()
