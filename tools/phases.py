#!/usr/bin/env python
"""Per-phase clock shares of the lean kernels on the headline workload (development build only):
   make -C dense-video-captioning_b200/csrc dev && MSDA_LIB=$PWD/dense-video-captioning_b200/lib/libmsda_dev.so python tools/phases.py"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
cfg = dict(bench.HEADLINE)
w = bench.make_workload(cfg["N"], cfg["T0"], cfg["M"], cfg["D"], cfg["L"], cfg["P"], cfg["kind"], "cuda:0")
op = bench.FusedOp(w)
lib = op.lib
buf = (ctypes.c_ulonglong * 32)()
lib.msda_dev_phase_clocks.argtypes = [ctypes.c_void_p, ctypes.c_int]
for _ in range(3):
    op.forward(); op.backward()
lib.msda_dev_phase_clocks(buf, 1)
reps = 10
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(reps):
    op.forward()
e.record(); torch.cuda.synchronize(); tf = s.elapsed_time(e) / reps
s.record()
for _ in range(reps):
    op.backward()
e.record(); torch.cuda.synchronize(); tb = s.elapsed_time(e) / reps
lib.msda_dev_phase_clocks(buf, 1)
names = [["init", "raw+decode", "window calc", "bulk issue+recs", "mbar wait", "gather+store"],
         ["init", "raw+decode", "G cp.async+zero", "histogram", "scan a", "scan b", "scatter+cp wait", "row sweep", "sample grads"]]
for k, (nm, t) in enumerate(zip(names, (tf, tb))):
    v = [buf[k * 16 + i] for i in range(16)]
    tot = sum(v)
    print(("forward" if k == 0 else "backward") + f": {t*1e3:.0f} us per launch; CTA-time shares (thread-0 clocks):")
    for i, x in enumerate(v):
        if x:
            print(f"   {nm[i] if i < len(nm) else i:20s} {100.0 * x / tot:5.1f}%   {x / reps / (cfg['N'] * cfg['M']):10.0f} clk per (n,m)")
