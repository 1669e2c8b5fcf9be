#!/usr/bin/env python
"""A bf16 Trainer as the first thing a process does (capture robustness): python tools/bf16_first_probe.py [SWITCH=0 ...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.pdvc import Trainer, model, synthetic_video  # noqa: E402

for arg in sys.argv[1:]:
    k, v = arg.split("=")
    setattr(model, k, bool(int(v)))
tr = Trainer("yc2_newmodel_core_bf16", "cuda:0", mode="graphs")
video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
for _ in range(5):
    loss = tr.step(video)
print("ok", sys.argv[1:], float(loss))
