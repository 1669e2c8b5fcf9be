#!/usr/bin/env python
"""Several trainers of one configuration built one after the other in ONE process: step time and device time of each of the
four graph replays -- the capture is not deterministic in speed (stream / allocation state of the process at capture time).
    python tools/seq_probe.py [config ...]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402


def one(name, steps=100):
    tr = Trainer(name, "cuda:0", mode="graphs")
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(15):
        tr.step(video)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = tr.step(video)
    float(loss)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / steps * 1e3
    marks, replay = [], torch.cuda.CUDAGraph.replay

    def timed(self):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        replay(self)
        b.record()
        marks.append((a, b))
    torch.cuda.CUDAGraph.replay = timed
    try:
        for _ in range(5):
            tr.step(video)
        torch.cuda.synchronize()
    finally:
        torch.cuda.CUDAGraph.replay = replay
    per = [sum(marks[i + 4 * k][0].elapsed_time(marks[i + 4 * k][1]) for k in range(5)) / 5 for i in range(4)]
    del tr
    torch.cuda.empty_cache()
    return ms, per


def main():
    names = sys.argv[1:] or ["yc2_tsp_pdvc"] * 5
    for n in names:
        ms, per = one(n)
        print(f"{n:28s} {ms:.3f} ms/step   graph replays " + " ".join(f"{v:.2f}" for v in per), flush=True)


if __name__ == "__main__":
    main()
