#!/usr/bin/env python
"""Where a PDVC training step spends its time on the GPU box: wall / device ms per phase and kernel-launch counts.
    python tools/pdvc_profile.py [config]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dense_video_captioning_b200.pdvc import Trainer, synthetic_video  # noqa: E402
from dense_video_captioning_b200.pdvc.train import to_device  # noqa: E402


def graph_mode_profile(tr):
    """Two-graph step: device time by kernel under replay, plus the host-visible phases."""
    from torch.profiler import ProfilerActivity, profile
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(6):
        tr.step(video)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        float(tr.step(video))
    torch.cuda.synchronize()
    print(f"{tr.mode}: {(time.perf_counter() - t0) / 20 * 1e3:.2f} ms/step")
    # device time of each graph replay, in issue order (stage 1 fwd, stage 2 fwd, stage 2 bwd, stage 1 bwd)
    marks, replay = [], torch.cuda.CUDAGraph.replay

    def timed_replay(self):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        replay(self)
        b.record()
        marks.append((a, b))

    torch.cuda.CUDAGraph.replay = timed_replay
    try:
        first = torch.cuda.Event(enable_timing=True)
        last = torch.cuda.Event(enable_timing=True)
        first.record()
        float(tr.step(video))
        last.record()
        torch.cuda.synchronize()
    finally:
        torch.cuda.CUDAGraph.replay = replay
    print("graph replays (device ms, issue order):", " ".join(f"{a.elapsed_time(b):.2f}" for a, b in marks),
          f"| whole step {first.elapsed_time(last):.2f}")
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            float(tr.step(video))
        torch.cuda.synchronize()
    ev = prof.key_averages()
    kernels = [e for e in ev if e.device_type == torch.autograd.DeviceType.CUDA]
    n_launch = sum(e.count for e in kernels) / 3
    dev_ms = sum(e.device_time_total for e in kernels) / 3e3
    print(f"kernel nodes per step: {n_launch:.0f}, summed device time {dev_ms:.2f} ms/step")
    for e in sorted(kernels, key=lambda e: -e.device_time_total)[:40]:
        print(f"  {e.count / 3:7.1f} x {e.device_time_total / 3e3:8.3f} ms  {e.key[:120]}")
    # kernel tables of single graphs replayed alone (forward graphs only: the backward graphs gather with buffers that are
    # only valid inside a live step)
    s1 = next(iter(tr.net._stage1.values())).graphs
    s2 = next(iter(tr.net._stage2.values())).graphs
    for label, g in (("stage 1 forward", s1[0]), ("stage 2 forward", s2[0])):
        with profile(activities=[ProfilerActivity.CUDA]) as gp:
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
        ks = [e for e in gp.key_averages() if e.device_type == torch.autograd.DeviceType.CUDA]
        print(f"{label}: {sum(e.count for e in ks) / 3:.0f} nodes, summed device time {sum(e.device_time_total for e in ks) / 3e3:.2f} ms")
        for e in sorted(ks, key=lambda e: -e.device_time_total)[:22]:
            print(f"  {e.count / 3:7.1f} x {e.device_time_total / 3e3:8.3f} ms  {e.key[:120]}")
    cpu_ops = sorted([e for e in ev if e.device_type == torch.autograd.DeviceType.CPU], key=lambda e: -e.self_cpu_time_total)[:15]
    print("top host ops (self cpu, per step):")
    for e in cpu_ops:
        print(f"  {e.count / 3:7.1f} x {e.self_cpu_time_total / 3e3:8.3f} ms  {e.key[:100]}")


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "yc2_tsp_pdvc_yc2shape"
    mode = sys.argv[2] if len(sys.argv) > 2 else "eager"
    tr = Trainer(name, "cuda:0", mode=mode)
    if mode != "eager":
        return graph_mode_profile(tr)
    video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
    for _ in range(5):
        tr.step(video)
    torch.cuda.synchronize()

    def phase(fn):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        return r, 1e3 * (t1 - t0), 1e3 * (t2 - t0)

    rows = []
    for _ in range(3):
        dt = to_device(video, tr.device)
        tr.optimizer.zero_grad(set_to_none=True)
        losses, h, d = phase(lambda: tr.net(dt))
        rows.append(("forward (model+criterion+captions)", h, d))
        total = sum(losses[k].float() * w for k, w in tr.weight_dict.items() if k in losses)
        _, h, d = phase(lambda: total.backward())
        rows.append(("backward", h, d))
        _, h, d = phase(lambda: (torch.nn.utils.clip_grad_norm_(tr.params, tr.grad_clip), tr.optimizer.step()))
        rows.append(("clip + adam", h, d))
    for r in rows[-3:]:
        print(f"{r[0]:40s} host {r[1]:7.2f} ms   host+drain {r[2]:7.2f} ms")

    # finer forward split
    m = tr.model
    dt = to_device(video, tr.device)
    with torch.no_grad():
        def enc():
            srcs, masks, pos = m.base_encoder(dt['video_tensor'], ~dt['video_mask'], dt['video_length'][:, 1])
            a = m.transformer.prepare_encoder_inputs(srcs, masks, pos)
            return a, m.transformer.forward_encoder(*a)
        (a, memory), h, d = phase(enc)
        print(f"{'  base encoder + encoder (no grad)':40s} host {h:7.2f} ms   host+drain {d:7.2f} ms")

    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        tr.step(video)
        torch.cuda.synchronize()
    ev = prof.key_averages()
    kernels = [e for e in ev if e.device_type == torch.autograd.DeviceType.CUDA]
    n_launch = sum(e.count for e in kernels)
    dev_ms = sum(e.device_time_total for e in kernels) / 1e3
    print(f"kernel launches per step: {n_launch}, summed device time {dev_ms:.2f} ms")
    for e in sorted(kernels, key=lambda e: -e.device_time_total)[:25]:
        print(f"  {e.count:5d} x {e.device_time_total / 1e3:8.3f} ms  {e.key[:110]}")
    cpu_ops = sorted([e for e in ev if e.device_type == torch.autograd.DeviceType.CPU], key=lambda e: -e.self_cpu_time_total)[:25]
    print("top host ops (self cpu):")
    for e in cpu_ops:
        print(f"  {e.count:5d} x {e.self_cpu_time_total / 1e3:8.3f} ms  {e.key[:100]}")


if __name__ == "__main__":
    main()
