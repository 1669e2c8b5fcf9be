#!/usr/bin/env python
"""Does running the per-layer caption loops on separate streams shorten the captured stage-2 graphs?"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402
from dense_video_captioning_b200.pdvc import graphed  # noqa: E402

for par in (False, True):
    orig = graphed.LossStage.__init__

    def init(self, *a, _par=par, **k):
        orig(self, *a, **k)
        self.parallel_layers = _par

    graphed.LossStage.__init__ = init
    tr = Trainer("yc2_tsp_pdvc_yc2shape", "cuda:0", mode="graphs")
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(6):
        tr.step(video)
    torch.cuda.synchronize()
    from dense_video_captioning_b200.pdvc.train import to_device
    dt = to_device(video, tr.device)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    fw, bw = [], []
    for _ in range(10):
        tr.optimizer.zero_grad(set_to_none=True)
        torch.cuda.synchronize()
        ev[0].record()
        total, losses, _ = tr.net(dt)
        ev[1].record()
        total.backward()
        ev[2].record()
        torch.cuda.synchronize()
        fw.append(ev[0].elapsed_time(ev[1]))
        bw.append(ev[1].elapsed_time(ev[2]))
    def replay_ms(g):
        g.replay(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            g.replay()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / 10
    s1 = next(iter(tr.net._stage1.values())).graphs
    s2 = next(iter(tr.net._stage2.values())).graphs
    print(f"parallel_layers={par}: forward-graph replays: stage1 {replay_ms(s1[0]):.2f} ms | stage2 {replay_ms(s2[0]):.2f} ms")
    print(f"parallel_layers={par}: forward (stage1 + match + stage2) {sorted(fw)[5]:.2f} ms, backward {sorted(bw)[5]:.2f} ms")
    graphed.LossStage.__init__ = orig
    del tr
    torch.cuda.empty_cache()
