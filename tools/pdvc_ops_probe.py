#!/usr/bin/env python
"""Which autograd nodes / aten ops make up the PDVC step's kernel launches: one staged step (the two stages run eagerly, same
ops as the captured graphs) under torch.profiler, aggregated by operator name.
    python tools/pdvc_ops_probe.py [config]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "yc2_tsp_pdvc_yc2shape"
    tr = Trainer(name, "cuda:0", mode="staged")
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(3):
        tr.step(video)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
        tr.step(video)
        torch.cuda.synchronize()
    rows = []
    for e in prof.key_averages():
        rows.append((e.count, e.key, getattr(e, "self_device_time_total", getattr(e, "self_cuda_time_total", 0.0))))
    rows.sort(key=lambda r: -r[0])
    print(f"{'calls':>6}  {'self device us':>14}  name")
    for count, key, dev in rows[:90]:
        print(f"{count:6d}  {dev:14.1f}  {key[:110]}")


    print("\nby input shape (elementwise / reduction ops with >= 4 calls):")
    rows = []
    for e in prof.key_averages(group_by_input_shape=True):
        if e.key in ("aten::add_", "aten::add", "aten::mul", "aten::copy_", "aten::cat", "aten::sum", "aten::div", "aten::fill_",
                     "aten::sub", "aten::neg", "aten::mm", "aten::addmm", "aten::clamp", "aten::index", "aten::stack") and e.count >= 4:
            rows.append((e.count, e.key, str(e.input_shapes)[:110]))
    rows.sort(key=lambda r: -r[0])
    for count, key, shapes in rows[:70]:
        print(f"{count:6d}  {key:14s} {shapes}")


if __name__ == "__main__":
    main()
