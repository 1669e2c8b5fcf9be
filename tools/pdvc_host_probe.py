#!/usr/bin/env python
"""Host time of one PDVC training step: each phase of Trainer.step timed on the host with the GPU idle at the start of the
step (so nothing blocks on a full launch queue): what the step costs the CPU, i.e. the floor below which a faster GPU step
would not show.    python tools/pdvc_host_probe.py [config]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402
from dense_video_captioning_b200.pdvc.train import to_device  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "yc2_tsp_pdvc_yc2shape"
    tr = Trainer(name, "cuda:0", mode="graphs")
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(10):
        tr.step(video)
    torch.cuda.synchronize()
    acc = {"to_device": 0.0, "forward (2 graph launches + matching)": 0.0, "backward (2 graph launches)": 0.0, "optimizer": 0.0,
           "whole step() call": 0.0, "device time of the step": 0.0}
    n = 30
    for _ in range(n):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        dt = to_device(video, tr.device)
        t1 = time.perf_counter()
        total, losses, _ = tr.net(dt)
        t2 = time.perf_counter()
        total.backward()
        t3 = time.perf_counter()
        tr.optimizer.step()
        t4 = time.perf_counter()
        b.record()
        torch.cuda.synchronize()
        acc["to_device"] += t1 - t0
        acc["forward (2 graph launches + matching)"] += t2 - t1
        acc["backward (2 graph launches)"] += t3 - t2
        acc["optimizer"] += t4 - t3
        acc["device time of the step"] += a.elapsed_time(b) * 1e-3
    for _ in range(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tr.step(video)
        acc["whole step() call"] += time.perf_counter() - t0
    for k, v in acc.items():
        print(f"{k:42s} {v / n * 1e3:7.3f} ms")


if __name__ == "__main__":
    main()
