#!/usr/bin/env python
"""A/B of a module-level switch of the PDVC trainer on one box, interleaved:
    python tools/pdvc_ab.py dense_video_captioning_b200.pdvc.model:USE_STEP_LINEAR [config] [steps]"""
import importlib
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402


def ms_per_step(name, steps):
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
    out = (time.perf_counter() - t0) / steps * 1e3
    del tr
    torch.cuda.empty_cache()
    return out


def main():
    mod_name, attr = sys.argv[1].split(":")
    name = sys.argv[2] if len(sys.argv) > 2 else "yc2_tsp_pdvc_yc2shape"
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 150
    mod = importlib.import_module(mod_name)
    rows = {False: [], True: []}
    for _ in range(3):
        for flag in (False, True):
            setattr(mod, attr, flag)
            rows[flag].append(ms_per_step(name, steps))
    for flag in (False, True):
        print(f"{attr}={flag}: " + " ".join(f"{v:.3f}" for v in rows[flag]) + f"  ms/step (min {min(rows[flag]):.3f})")


if __name__ == "__main__":
    main()
