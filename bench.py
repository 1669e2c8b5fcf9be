#!/usr/bin/env python
"""bench.py -- MSDeformAttn fwd+bwd throughput on B200 (BASELINE.json metric: queries/s and HBM GB/s).

A "step" is one fused forward + one fused backward of the 1-D temporal multi-scale deformable
attention over one batch of synthetic encoder-shaped inputs (a query = one (n, q) across all heads).

  python bench.py [--gpus N] [--steps K] [--warmup W]          our CUDA path (one JSON line)
  python bench.py --impl reference ...                         the reference's CPU path (oracle port)
  torchrun --nproc-per-node N bench.py --gpus N ...            one rank per GPU, weak scaling
  python bench.py --sweep out.json                             config-5 sweep (not a bench line)

Workload (config.workload): BASELINE.json configs[4] sweep point N=64, T0=1000 (levels 1000/500/250/125,
S=1875), M=8, D=64, L=4, P=4, Lq=S -- PDVC encoder self-attention geometry (reference points on the
level grids, deformable_transformer.py:198-208; offsets = the module's initial bias pattern,
ms_deform_attn.py:62-71, plus N(0,1) jitter; logits ~ N(0,1); value, grad_out ~ N(0,1)), fp32.
Its tensors (1.3 GB) are ~10x the 126 MB L2, so consecutive steps cannot be served from cache.
The model's real shapes (configs[0]: N=1, S=375) are launch-latency bound and are reported
in `extra` as microseconds per call.
"""
import argparse
import ctypes
import json
import math
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

HEADLINE = dict(N=64, T0=1000, M=8, D=64, L=4, P=4, kind="encoder")


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic():
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the headline kernels from the latest committed
    `ncu --set full` capture (profiles/r2_traffic.json, else r1_traffic.json; written by tools/ncu_traffic.py); {} if absent."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                return json.load(f)
        except Exception:
            continue
    return {}


def level_lengths(t0, n_levels=4):
    ts = [int(t0)]
    for _ in range(n_levels - 1):
        ts.append((ts[-1] - 1) // 2 + 1)   # Conv1d(k3,s2,p1), pdvc/base_encoder.py:37-41
    return ts


def offset_bias(M, L, P):
    """The reference's initial sampling-offset bias (ms_deform_attn.py:62-71)."""
    thetas = torch.arange(M, dtype=torch.float32) * (2.0 * math.pi / M)
    grid = torch.stack([thetas.cos(), thetas.sin()], -1)
    grid = (grid / grid.abs().max(-1, keepdim=True)[0])[:, 0]
    return grid.view(M, 1, 1) * torch.arange(1, P + 1, dtype=torch.float32).view(1, 1, P).expand(M, L, P)


def make_workload(N, T0, M, D, L, P, kind, device, dtype=torch.float32, seed=0, Lq_dec=100, jitter=1.0):
    """Synthetic fused-boundary inputs.  kind: 'encoder' (Lq = S, grid reference points) or
    'decoder' (Lq = Lq_dec, sigmoid(N(0,1)) reference points)."""
    T = level_lengths(T0, L)
    S = sum(T)
    g = torch.Generator(device="cpu").manual_seed(seed)
    if kind == "encoder":
        Lq = S
        ref1 = torch.cat([(torch.arange(t, dtype=torch.float32) + 0.5) / t for t in T])
        ref = ref1.view(1, S, 1, 1).expand(N, S, L, 1).contiguous()
    else:
        Lq = Lq_dec
        ref = torch.sigmoid(torch.randn(N, Lq, 1, 1, generator=g)).expand(N, Lq, L, 1).contiguous()
    dev = torch.device(device)
    gd = torch.Generator(device=dev).manual_seed(seed + 1) if dev.type == "cuda" else g
    value = torch.randn(N, S, M, D, generator=gd, device=dev, dtype=torch.float32).to(dtype)
    off = offset_bias(M, L, P).to(dev).view(1, 1, M, L, P) + jitter * torch.randn(N, Lq, M, L, P, generator=gd, device=dev)
    logits = torch.randn(N, Lq, M, L, P, generator=gd, device=dev)
    grad_out = torch.randn(N, Lq, M * D, generator=gd, device=dev, dtype=torch.float32).to(dtype)
    shapes = torch.tensor(T, dtype=torch.long, device=dev)
    lsi = torch.cat((shapes.new_zeros((1,)), shapes.cumsum(0)[:-1]))
    return dict(value=value.contiguous(), ref=ref.to(dev), off=off.contiguous(), logits=logits, grad_out=grad_out,
                shapes=shapes, lsi=lsi, T=T, N=N, S=S, M=M, D=D, L=L, P=P, Lq=Lq, R=1)


def algorithmic_bytes(w, elem, with_grad_ref=False):
    """SURVEY 8(d), fused boundary: every tensor touched once, no memsets, no atomic RMW."""
    N, S, M, D, L, P, Lq, R = (w[k] for k in ("N", "S", "M", "D", "L", "P", "Lq", "R"))
    val = elem * N * S * M * D
    io = elem * N * Lq * M * D
    idx = 4 * (2 * N * Lq * M * L * P + N * Lq * L * R)
    fwd = val + idx + io
    bwd_r = io + val + idx
    bwd_w = val + 4 * 2 * N * Lq * M * L * P + (4 * N * Lq * L * R if with_grad_ref else 0)
    return fwd, bwd_r + bwd_w


class FusedOp:
    """Device-resident fused fwd+bwd through the raw C ABI with preallocated outputs."""

    def __init__(self, w, pad=0):
        import dense_video_captioning_b200 as dvc
        self.lib = dvc._lib.lib()
        self.w = w
        v = w["value"]
        self.code = {torch.float32: 0, torch.bfloat16: 1}[v.dtype]
        self.out = torch.empty(w["N"], w["Lq"], w["M"] * w["D"], dtype=v.dtype, device=v.device)
        self.gv = torch.empty_like(v)
        self.goff = torch.empty_like(w["off"])
        self.glog = torch.empty_like(w["logits"])
        nb = self.lib.msda_backward_workspace_bytes(w["N"], w["S"], w["M"], w["D"], self.code)
        self.ws = torch.empty(max(nb, 16), dtype=torch.uint8, device=v.device)
        self.ws_bytes = nb
        self.pad = pad
        self.dims = (w["N"], w["S"], w["M"], w["D"], w["L"], w["Lq"], w["P"], w["R"])
        self.launches = 0

    def forward(self):
        w = self.w
        st = torch.cuda.current_stream().cuda_stream
        rc = self.lib.msda_fused_forward(w["value"].data_ptr(), w["shapes"].data_ptr(), w["lsi"].data_ptr(),
                                         w["ref"].data_ptr(), w["off"].data_ptr(), w["logits"].data_ptr(),
                                         self.out.data_ptr(), *self.dims, self.code, self.pad, 0, st)
        if rc:
            raise RuntimeError(self.lib.msda_error_string(rc).decode())
        self.launches += self.lib.msda_last_launch_count()

    def backward(self):
        w = self.w
        st = torch.cuda.current_stream().cuda_stream
        rc = self.lib.msda_fused_backward(w["value"].data_ptr(), w["shapes"].data_ptr(), w["lsi"].data_ptr(),
                                          w["ref"].data_ptr(), w["off"].data_ptr(), w["logits"].data_ptr(),
                                          w["grad_out"].data_ptr(), self.gv.data_ptr(), None, self.goff.data_ptr(),
                                          self.glog.data_ptr(), self.ws.data_ptr() if self.ws_bytes else None,
                                          self.ws_bytes, *self.dims, self.code, self.pad, 0, st)
        if rc:
            raise RuntimeError(self.lib.msda_error_string(rc).decode())
        self.launches += self.lib.msda_last_launch_count()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons during the timed region (pynvml)."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.stop_flag = threading.Event()
        self.sm, self.reasons, self.max_mhz, self.err = [], set(), None, None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
                getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
            }
            while not self.stop_flag.is_set():
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
                time.sleep(self.period)
        except Exception as e:   # pragma: no cover
            self.err = repr(e)

    def result(self):
        self.stop_flag.set()
        self.join(timeout=2)
        sm = sorted(self.sm)
        return {"sm_mhz": (sm[len(sm) // 2] if sm else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(sm), **({"error": self.err} if self.err else {})}


def dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {"device_id": torch.device(f"cuda:{local}")} if backend == "nccl" else {}
        dist.init_process_group(backend, **kw)
    return world, rank, local


def max_over_ranks(x, world, device):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def shard_batch(global_batch, world, rank):
    """Batch elements are independent: contiguous shards, no data-path collective."""
    base, rem = divmod(global_batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


# ------------------------------------------------------------------------------------------------
# CPU reference arm (the oracle's torch port = the same ATen ops the reference executes on CPU)
# ------------------------------------------------------------------------------------------------
def cpu_step_factory(cfg, n_sub, lq_sub, seed=0):
    from oracle import core_pytorch
    w = make_workload(n_sub, cfg["T0"], cfg["M"], cfg["D"], cfg["L"], cfg["P"], cfg["kind"], "cpu", seed=seed)
    lq_sub = min(lq_sub, w["Lq"])
    # a contiguous run of queries from the middle of the sequence
    q0 = (w["Lq"] - lq_sub) // 2
    ref = w["ref"][:, q0:q0 + lq_sub].contiguous()
    off0 = w["off"][:, q0:q0 + lq_sub].contiguous()
    logits0 = w["logits"][:, q0:q0 + lq_sub].contiguous()
    gout = w["grad_out"][:, q0:q0 + lq_sub].contiguous()
    shapes2 = [(1, t) for t in w["T"]]
    Tt = torch.tensor(w["T"], dtype=torch.float32)

    def step():
        value = w["value"].clone().requires_grad_(True)
        off = off0.clone().requires_grad_(True)
        logits = logits0.clone().requires_grad_(True)
        # module front-end (ms_deform_attn.py:100-116), then the reference's CPU core (func.py:41-68)
        attn = torch.softmax(logits.flatten(-2), -1).view_as(logits)
        x = ref[:, :, None, :, None, 0] + off / Tt[None, None, None, :, None]
        loc = torch.stack((x, 0.5 * torch.ones_like(x)), -1)
        out = core_pytorch.core(value, shapes2, loc, attn)
        out.backward(gout)
        return float(out.detach()[0, 0, 0])

    return step, n_sub * lq_sub


def run_reference(args, world, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = HEADLINE
    # calibrate a bounded sample so that steps+warmup fit in ~150 s
    probe, nq = cpu_step_factory(cfg, 1, 128)
    probe()
    t0 = time.perf_counter()
    probe()
    per_query = (time.perf_counter() - t0) / nq
    budget = 150.0 / max(1, args.steps + args.warmup)
    S = sum(level_lengths(cfg["T0"], cfg["L"]))
    lq_sub = int(max(32, min(S, budget / max(per_query, 1e-9))))
    step, nq = cpu_step_factory(cfg, 1, lq_sub)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    qps = nq * args.steps / dt
    line = {
        "impl": "reference", "metric": "msda_fwd_bwd_queries_per_s", "value": qps, "unit": "queries/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(cfg, world=1),
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"1 video, {lq_sub} consecutive queries of {S} (full value tensor), "
                                   "oracle/core_pytorch.py fwd+bwd incl. softmax/location front-end"},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(cfg, world):
    T = level_lengths(cfg["T0"], cfg["L"])
    return {"workload": f"msda_{cfg['kind']}_fused_fwd_bwd N={cfg['N']}/gpu T={T} S={sum(T)} Lq={sum(T)} M={cfg['M']} "
                        f"D={cfg['D']} L={cfg['L']} P={cfg['P']} (BASELINE configs[4] sweep point; PDVC encoder geometry)",
            "global_batch": cfg["N"] * world, "per_gpu_batch": cfg["N"], "parallelism": f"dp{world} (batch shards, no data-path collective)",
            "l2_policy": "inputs (1.3 GB/step) exceed the 126 MB L2; no explicit flush",
            "padding_mode": "border", "boundary": "fused (ref points + raw offsets + raw logits in)"}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def time_device(op, steps, warmup, world):
    """K steps of fwd+bwd, CUDA events on the launch stream, barrier + synchronize on both sides."""
    for _ in range(warmup):
        op.forward()
        op.backward()
    torch.cuda.synchronize()
    op.launches = 0
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
    barrier(world)
    torch.cuda.synchronize()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for i in range(steps):
        ev[i][0].record()
        op.forward()
        ev[i][1].record()
        op.backward()
        ev[i][2].record()
    end.record()
    torch.cuda.synchronize()
    barrier(world)
    total_ms = start.elapsed_time(end)
    fwd_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / steps
    bwd_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / steps
    return total_ms, fwd_ms, bwd_ms


def time_e2e(w, steps, warmup, world, chunk=8):
    """Same step through the public API with HOST (pinned) inputs and outputs: HostPipelinedFusedMSDA runs
    MSDeformAttnFusedFunction.apply + backward per batch chunk, overlapping the copy-in of the next chunk and the
    copy-out of the previous one with the kernels.  Every byte of the step's inputs and results crosses the host link
    inside the timed region."""
    from dense_video_captioning_b200.ops.host_pipeline import HostPipelinedFusedMSDA, bind_host_to_gpu_numa_node
    dev = w["value"].device
    time_e2e.numa = bind_host_to_gpu_numa_node(dev)        # pinned buffers below are first-touched on the GPU's node
    host_in = [w[k].cpu().pin_memory() for k in ("value", "ref", "off", "logits", "grad_out")]
    h2d = sum(t.numel() * t.element_size() for t in host_in)
    pipe = HostPipelinedFusedMSDA(w["shapes"], w["lsi"], dev, chunk=chunk)
    outs = None
    for _ in range(warmup):
        outs = pipe.run(*host_in)
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(steps):
        outs = pipe.run(*host_in)
    end.record()
    torch.cuda.synchronize()
    barrier(world)
    d2h = sum(t.numel() * t.element_size() for t in outs)
    return start.elapsed_time(end), h2d, d2h


def small_shape_latency(device):
    """configs[0] shapes (N=1, YouCook2): microseconds per fused fwd+bwd call, plain launches and CUDA graph."""
    res = {}
    for kind in ("decoder", "encoder"):
        w = make_workload(1, 200, 8, 64, 4, 4, kind, device, seed=3)
        op = FusedOp(w)
        for _ in range(10):
            op.forward(); op.backward()
        torch.cuda.synchronize()
        n = 200
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            op.forward(); op.backward()
        e.record()
        torch.cuda.synchronize()
        plain = 1e3 * s.elapsed_time(e) / n
        graph_us = None
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                op.forward(); op.backward()
            torch.cuda.current_stream().wait_stream(side)
            with torch.cuda.graph(g):
                for _ in range(20):
                    op.forward(); op.backward()
            g.replay()
            torch.cuda.synchronize()
            s.record()
            for _ in range(10):
                g.replay()
            e.record()
            torch.cuda.synchronize()
            graph_us = 1e3 * s.elapsed_time(e) / 200
        except Exception as ex:   # pragma: no cover
            graph_us = f"graph capture failed: {ex!r}"
        fb, bb = algorithmic_bytes(w, 4)
        res[f"yc2_{kind}_N1"] = {"us_per_fwd_bwd_plain": round(plain, 2), "us_per_fwd_bwd_graph": graph_us,
                                 "algorithmic_bytes": fb + bb, "queries": w["Lq"]}
    return res


def reference_cuda_same_box(w, steps=20):
    """The reference's own CUDA kernels (pdvc/ops/src, compiled unmodified for sm_100 by tools/build_ref_cuda.py into
    baseline/_ref/ref_cuda/) on the same workload and box: unfused boundary (locations and softmaxed weights
    precomputed, as ms_deform_attn.py:100-122 hands them over), ZEROS padding, its own allocations and memsets
    (ms_deform_attn_cuda.cu:54, :121-123).  Our unfused entry points are timed beside it on the same tensors."""
    import importlib.util
    so = os.path.join(ROOT, "baseline", "_ref", "ref_cuda", "MultiScaleDeformableAttention_refcuda.so")
    if not os.path.exists(so):
        return {"unavailable": "baseline/_ref/ref_cuda not built (tools/build_ref_cuda.py needs /root/reference)"}
    spec = importlib.util.spec_from_file_location("MultiScaleDeformableAttention_refcuda", so)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    import dense_video_captioning_b200 as dvc
    N, Lq, M, D, L, P = (w[k] for k in ("N", "Lq", "M", "D", "L", "P"))
    Tt = w["shapes"].float()
    attn = torch.softmax(w["logits"].flatten(-2), -1).view_as(w["logits"]).contiguous()
    x = w["ref"][:, :, None, :, None, 0] + w["off"] / Tt[None, None, None, :, None]
    loc = torch.stack((x, torch.full_like(x, 0.5)), -1).contiguous()
    shapes2 = torch.stack((torch.ones_like(w["shapes"]), w["shapes"]), -1).contiguous()
    value, gout = w["value"].float().contiguous(), w["grad_out"].float().contiguous()

    def timed_n(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(n):
            fn()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / n

    def timed(fn):
        return timed_n(fn, steps)

    r_fwd = timed(lambda: ref.ms_deform_attn_forward(value, shapes2, w["lsi"], loc, attn, 64))
    r_bwd = timed(lambda: ref.ms_deform_attn_backward(value, shapes2, w["lsi"], loc, attn, gout, 64))
    compat_dir = os.path.join(os.path.dirname(os.path.abspath(dvc.__file__)), "compat")
    if compat_dir not in sys.path:
        sys.path.insert(0, compat_dir)
    import MultiScaleDeformableAttention as ours   # the drop-in for the reference's extension module
    ours.padding_mode = "zeros"      # the module's default: what the extension it replaces does
    o_fwd = timed(lambda: ours.ms_deform_attn_forward(value, shapes2, w["lsi"], loc, attn, 64))
    o_bwd = timed(lambda: ours.ms_deform_attn_backward(value, shapes2, w["lsi"], loc, attn, gout, 64))
    q = N * Lq
    res = {"boundary": "unfused (sampling_locations + attention_weights in), ZEROS padding, fp32",
           "reference_cuda": {"fwd_ms": r_fwd, "bwd_ms": r_bwd, "queries_per_s": q / ((r_fwd + r_bwd) * 1e-3)},
           "ours_unfused": {"fwd_ms": o_fwd, "bwd_ms": o_bwd, "queries_per_s": q / ((o_fwd + o_bwd) * 1e-3)},
           "speedup_fwd_bwd": (r_fwd + r_bwd) / (o_fwd + o_bwd)}
    # configs[0] shapes (N = 1: where PDVC actually runs the op), microseconds per call, plain launches and CUDA-graph
    # replay (20 calls per graph), the reference's kernels and ours on the same tensors
    ours.ASSUME_1D = True        # PDVC's inputs (ms_deform_attn.py:114-117): no per-call inspection, capturable
    small = {}
    for kind in ("decoder", "encoder"):
        ws = make_workload(1, 200, 8, 64, 4, 4, kind, w["value"].device, seed=3)
        attn_s = torch.softmax(ws["logits"].flatten(-2), -1).view_as(ws["logits"]).contiguous()
        xs = ws["ref"][:, :, None, :, None, 0] + ws["off"] / ws["shapes"].float()[None, None, None, :, None]
        loc_s = torch.stack((xs, torch.full_like(xs, 0.5)), -1).contiguous()
        sh2 = torch.stack((torch.ones_like(ws["shapes"]), ws["shapes"]), -1).contiguous()
        val_s, gout_s = ws["value"].contiguous(), ws["grad_out"].contiguous()
        row = {}
        for who, mod in (("reference_cuda", ref), ("ours_unfused", ours)):
            calls = {"fwd": lambda m=mod: m.ms_deform_attn_forward(val_s, sh2, ws["lsi"], loc_s, attn_s, 64),
                     "bwd": lambda m=mod: m.ms_deform_attn_backward(val_s, sh2, ws["lsi"], loc_s, attn_s, gout_s, 64)}
            for leg, fn in calls.items():
                row[f"{who}_{leg}_us_plain"] = round(1e3 * timed_n(fn, 200), 2)
                try:
                    g = torch.cuda.CUDAGraph()
                    side = torch.cuda.Stream()
                    side.wait_stream(torch.cuda.current_stream())
                    with torch.cuda.stream(side):
                        fn()
                    torch.cuda.current_stream().wait_stream(side)
                    with torch.cuda.graph(g):
                        for _ in range(20):
                            fn()
                    row[f"{who}_{leg}_us_graph"] = round(1e3 * timed_n(g.replay, 20) / 20, 2)
                except Exception as ex:   # pragma: no cover
                    row[f"{who}_{leg}_us_graph"] = f"capture failed: {ex!r}"[:120]
        small[f"yc2_{kind}_N1"] = row
    res["small_shapes_us"] = small
    return res


def sweep_max_point(peak, steps=5):
    """BASELINE configs[4] maximum that fits one GPU: N=256, T0=2000 (S=Lq=3750), P=16, encoder geometry, fp32
    (27.5 GB of algorithmic tensors per fwd+bwd).  Timed like the headline (CUDA events, device-resident inputs)."""
    cfg = dict(N=256, T0=2000, M=8, D=64, L=4, P=16, kind="encoder")
    w = make_workload(cfg["N"], cfg["T0"], cfg["M"], cfg["D"], cfg["L"], cfg["P"], cfg["kind"], "cuda", seed=17)
    op = FusedOp(w)
    total_ms, fwd_ms, bwd_ms = time_device(op, steps, 3, 1)
    fb, bb = algorithmic_bytes(w, 4)
    out = {"workload": f"N=256 T0=2000 S=Lq=3750 M=8 D=64 L=4 P=16 encoder fp32 (largest configs[4] point)",
           "fwd_ms": fwd_ms, "bwd_ms": bwd_ms, "queries_per_s": cfg["N"] * w["Lq"] / ((fwd_ms + bwd_ms) * 1e-3),
           "forward_frac": fb / (fwd_ms * 1e-3) / 1e9 / peak, "backward_frac": bb / (bwd_ms * 1e-3) / 1e9 / peak,
           "fwd_bwd_frac": (fb + bb) / ((fwd_ms + bwd_ms) * 1e-3) / 1e9 / peak,
           "algorithmic_bytes_fwd_bwd": fb + bb}
    del op, w
    torch.cuda.empty_cache()
    return out


def pdvc_train_leg(args, world, rank, device):
    """BASELINE metric 2: PDVC training videos/s (one video per rank per step, DDP gradient all-reduce over NCCL when
    world > 1).  Every step copies its video from pinned host memory and reads the loss back (e2e by construction)."""
    from dense_video_captioning_b200.pdvc import CONFIGS, Trainer, synthetic_video
    from dense_video_captioning_b200.pdvc.train import input_bytes
    names = args.pdvc_configs.split(",") if args.pdvc_configs else list(CONFIGS)
    res = {}
    for name in names:
        tr = Trainer(name, device, world_size=world, sync=args.pdvc_sync)
        videos = [synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=7919 * i + rank) for i in range(8)]
        for i in range(args.pdvc_warmup):
            tr.step(videos[i % len(videos)])
        torch.cuda.synchronize()
        barrier(world)
        torch.cuda.synchronize()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        start.record()
        last, pending = None, None
        for i in range(args.pdvc_steps):
            loss = tr.step(videos[i % len(videos)])
            if pending is not None:
                last = float(pending)       # device->host read of every step's loss, one step late: the host issues
            pending = loss                  # step i+1 (input copy, stage-1 graph launch) while the GPU finishes step i
        last = float(pending)
        end.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier(world)
        ms = max_over_ranks(start.elapsed_time(end), world, device)
        n_params = sum(p.numel() for p in tr.params)
        res[name] = {"videos_per_s": world * args.pdvc_steps / (ms * 1e-3), "ms_per_step": ms / args.pdvc_steps,
                     "steps": args.pdvc_steps, "warmup": args.pdvc_warmup, "videos_per_step": world,
                     "wall_s_rank0": wall, "final_loss": last, "trainable_params": n_params,
                     "allreduce_bytes_per_step": 4 * n_params if world > 1 else 0,
                     "h2d_bytes_per_step": input_bytes(videos[0]), "d2h_bytes_per_step": 4,
                     "dtype": "bf16 autocast (MSDA kernels bf16 value, fp32 index side)" if tr.amp else "f32",
                     "shape": {"T": tr.cfg.frame_embedding_num, "queries": tr.cfg.num_queries,
                               "enc_layers": tr.cfg.enc_layers, "dec_layers": tr.cfg.dec_layers,
                               "feature_dim": tr.cfg.feature_dim, "vocab": tr.cfg.vocab_size, "n_gt": tr.n_gt,
                               "caption_words": tr.cap_len}}
        del tr
        torch.cuda.empty_cache()
    return res


def pdvc_cpu_baseline(name, seconds=20.0):
    """The same training step on the host cores with the reference's CPU path for the two attention modules
    (oracle/core_pytorch.module_forward = the F.grid_sample route of ms_deform_attn.py:123-124), all threads."""
    from dense_video_captioning_b200.ops.modules import MSDeformAttn, MSDeformAttnCap
    from dense_video_captioning_b200.pdvc import Trainer, synthetic_video
    from oracle import core_pytorch
    saved = MSDeformAttn.forward, MSDeformAttnCap.forward

    def attn_forward(self, query, ref, src, shapes, starts, mask=None):
        return core_pytorch.module_forward(dict(self.named_parameters()), query, ref, src, shapes, mask, self.n_heads,
                                           self.n_levels, self.n_points)

    def cap_forward(self, query, ref, src, shapes, starts, mask=None, projected_value=None):
        return core_pytorch.module_forward(dict(self.named_parameters()), query, ref, src, shapes, mask, self.n_heads,
                                           self.n_levels, self.n_points, caption=True)

    MSDeformAttn.forward, MSDeformAttnCap.forward = attn_forward, cap_forward
    try:
        torch.set_num_threads(os.cpu_count() or 1)
        tr = Trainer(name, "cpu", amp=False)
        video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=3)
        video = {k: v.cpu() for k, v in video.items()}
        tr.step(video)
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds:
            float(tr.step(video))
            n += 1
        dt = time.perf_counter() - t0
        return {"value": n / dt, "unit": "videos/s", "cores": torch.get_num_threads(), "kind": "port",
                "sample": f"{n} training steps of config {name} in {dt:.1f} s on the host cores (fp32)"}
    finally:
        MSDeformAttn.forward, MSDeformAttnCap.forward = saved


def run_ours(args, world, rank, local):
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (our arm) needs a CUDA device: there is no CPU fallback for the MSDA kernels")
    torch.cuda.set_device(local)
    device = f"cuda:{local}"
    cfg = dict(HEADLINE)
    if args.batch:
        cfg["N"] = args.batch
    if args.points:
        cfg["P"] = args.points
    peak, peak_src = load_peaks()
    # weak scaling: every rank owns its own shard of cfg.N * world videos (seeded per rank)
    lo, hi = shard_batch(cfg["N"] * world, world, rank)
    w = make_workload(hi - lo, cfg["T0"], cfg["M"], cfg["D"], cfg["L"], cfg["P"], cfg["kind"], device,
                      dtype=torch.bfloat16 if args.dtype == "bf16" else torch.float32, seed=100 + rank)
    op = FusedOp(w)
    sampler = ClockSampler(local)
    sampler.start()
    total_ms, fwd_ms, bwd_ms = time_device(op, args.steps, args.warmup, world)
    clocks = sampler.result()
    launches = op.launches
    total_ms = max_over_ranks(total_ms, world, device)
    queries_per_step = cfg["N"] * world * w["Lq"]
    qps = queries_per_step * args.steps / (total_ms * 1e-3)
    elem = 2 if args.dtype == "bf16" else 4
    fb, bb = algorithmic_bytes(w, elem)
    # dominant launch: the backward call = cudaMemsetAsync(grad_value) + msda_bwd_lean_kernel (csrc/msda_lean.cuh)
    traffic = load_traffic()
    roof = {"bound": "hbm", "kernel": "msda_fused_backward (cudaMemsetAsync(grad_value) + msda_bwd_lean_kernel)",
            "achieved": bb / (bwd_ms * 1e-3) / 1e9, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
            "frac": bb / (bwd_ms * 1e-3) / 1e9 / peak, "traffic": traffic.get("backward_bytes_per_launch"),
            "traffic_source": traffic.get("source"),
            "algorithmic_bytes_per_launch": bb, "ms_per_launch": bwd_ms,
            "on_chip_note": "both kernels are bound by shared-memory wavefronts / issue slots, not HBM: the gather reads "
                            "2*D*4 B of shared memory per sample (12.8x the HBM bytes); see DESIGN.md section 4",
            "forward": {"kernel": "msda_fwd_lean_kernel", "achieved": fb / (fwd_ms * 1e-3) / 1e9,
                        "frac": fb / (fwd_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": fb, "ms_per_launch": fwd_ms,
                        "traffic": traffic.get("forward_bytes_per_launch")},
            "fwd_bwd": {"achieved": (fb + bb) / ((fwd_ms + bwd_ms) * 1e-3) / 1e9,
                        "frac": (fb + bb) / ((fwd_ms + bwd_ms) * 1e-3) / 1e9 / peak}}
    if world == 1 and not args.no_extra:
        try:
            del op
            torch.cuda.empty_cache()
            roof["sweep_max"] = sweep_max_point(peak)
        except Exception as ex:   # must never cost the bench line
            roof["sweep_max"] = {"error": repr(ex)[:200]}
        op = None
    e2e_steps = max(3, min(args.steps, 20))
    e2e_ms, h2d, d2h = time_e2e(w, e2e_steps, 3, world)
    e2e_ms = max_over_ranks(e2e_ms, world, device)
    e2e_qps = queries_per_step * e2e_steps / (e2e_ms * 1e-3)
    line = {
        "metric": "msda_fwd_bwd_queries_per_s", "value": qps, "unit": "queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if args.dtype == "bf16" else "f32", "data": "synthetic",
        "config": workload_config(cfg, world), "roofline": roof, "clocks": clocks,
        "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps, "host_numa": getattr(time_e2e, "numa", None),
                "api": "ops.host_pipeline.HostPipelinedFusedMSDA (MSDeformAttnFusedFunction.apply + backward per 8-video chunk; copy-in, kernels and copy-out on three streams), pinned host tensors in and out"},
        "gpu_launches": launches,
        "hbm_gbs_fwd_bwd": roof["fwd_bwd"]["achieved"],
    }
    op = None
    torch.cuda.empty_cache()
    if not args.no_pdvc:
        try:
            line["pdvc_train"] = {"metric": "pdvc_train_videos_per_s", "n_gpus": world, "data": "synthetic",
                                  "parallelism": f"dp{world} (NCCL gradient all-reduce: {args.pdvc_sync}; 1 video per rank per step)" if world > 1
                                  else "single process", "configs": pdvc_train_leg(args, world, rank, device)}
        except Exception as ex:
            if world > 1:
                raise
            line["pdvc_train"] = {"error": repr(ex)}
    if rank == 0:
        if world == 1 and not args.no_cpu and not args.no_pdvc and "configs" in line.get("pdvc_train", {}):
            try:
                line["pdvc_train"]["cpu_baseline"] = pdvc_cpu_baseline("yc2_tsp_pdvc_yc2shape")
            except Exception as ex:
                line["pdvc_train"]["cpu_baseline"] = {"error": repr(ex)}
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            step, nq = cpu_step_factory(cfg, 1, 256)
            step()
            t0 = time.perf_counter()
            reps = 0
            while time.perf_counter() - t0 < 10.0:
                step()
                reps += 1
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": nq * reps / dt, "unit": "queries/s", "cores": torch.get_num_threads(),
                                    "kind": "port",
                                    "sample": f"1 video, 256 consecutive queries of {w['Lq']} (full value tensor), "
                                              f"{reps} reps in {dt:.1f} s; oracle/core_pytorch.py fwd+bwd"}
        if world == 1 and not args.no_extra:
            line["extra"] = {"small_shapes": small_shape_latency(device)}
            try:
                line["extra"]["reference_cuda_same_box"] = reference_cuda_same_box(w)
            except Exception as ex:   # the comparison must never cost the bench line
                line["extra"]["reference_cuda_same_box"] = {"error": repr(ex)}
        # compact PDVC summary LAST (the driver keeps the tail of the line): videos/s per config, the exchange mode, and
        # the weak-scaling efficiency against this box's N=1 run when one was recorded (the driver runs N=1,2,4,8 back to
        # back on one box; the N=1 run leaves its figures in a scratch file)
        cfgs = line.get("pdvc_train", {}).get("configs")
        if cfgs:
            vps = {k: round(v["videos_per_s"], 2) for k, v in cfgs.items()}
            scratch = os.path.join(os.environ.get("TMPDIR", "/tmp"), "msda_bench_pdvc_n1.json")
            eff = None
            try:
                if world == 1:
                    with open(scratch, "w") as f:
                        json.dump(vps, f)
                else:
                    with open(scratch) as f:
                        n1 = json.load(f)
                    eff = {k: round(vps[k] / (world * n1[k]), 4) for k in vps if n1.get(k)}
            except Exception:
                pass
            line["pdvc_sync"] = args.pdvc_sync if world > 1 else "single process"
            line["pdvc_weak_efficiency_vs_n1"] = eff
            line["pdvc_videos_per_s"] = vps
        print(json.dumps(line), flush=True)


def run_sweep(args):
    """configs[4]: batch x temporal length x points x {encoder, decoder} x dtype; writes a JSON file."""
    torch.cuda.set_device(0)
    peak, _ = load_peaks()
    rows = []
    for dtype in ("f32", "bf16"):
        for kind in ("encoder", "decoder"):
            for N in (1, 4, 16, 64, 256):
                for T0 in (100, 200, 500, 1000, 2000):
                    for P in (4, 8, 16):
                        S = sum(level_lengths(T0))
                        Lq = S if kind == "encoder" else 100
                        elem = 2 if dtype == "bf16" else 4
                        need = N * (3 * S * 512 * elem + 2 * Lq * 512 * elem + 6 * Lq * 8 * 4 * P * 4) + N * S * 512 * 4
                        if need > 60e9:
                            continue
                        w = make_workload(N, T0, 8, 64, 4, P, kind, "cuda:0",
                                          dtype=torch.bfloat16 if dtype == "bf16" else torch.float32, seed=5)
                        op = FusedOp(w)
                        steps = 20 if need > 1e9 else 100
                        total_ms, fwd_ms, bwd_ms = time_device(op, steps, 3, 1)
                        fb, bb = algorithmic_bytes(w, elem)
                        rows.append(dict(dtype=dtype, kind=kind, N=N, T0=T0, P=P, S=S, Lq=Lq, fwd_us=1e3 * fwd_ms,
                                         bwd_us=1e3 * bwd_ms, queries_per_s=N * Lq / ((fwd_ms + bwd_ms) * 1e-3),
                                         fwd_gbs=fb / fwd_ms / 1e6, bwd_gbs=bb / bwd_ms / 1e6,
                                         fwd_bwd_frac_hbm=(fb + bb) / (fwd_ms + bwd_ms) / 1e6 / peak))
                        del op, w
                        torch.cuda.empty_cache()
    with open(args.sweep, "w") as f:
        json.dump({"peak_gbs": peak, "rows": rows}, f, indent=1)
    print(f"wrote {len(rows)} rows to {args.sweep}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--batch", type=int, default=0, help="videos per GPU (default: headline config)")
    ap.add_argument("--points", type=int, default=0, help="sampling points per level (default: headline config; experiments)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--sweep", default="")
    ap.add_argument("--no-pdvc", action="store_true", help="skip the PDVC training leg")
    ap.add_argument("--pdvc-steps", type=int, default=30)
    ap.add_argument("--pdvc-warmup", type=int, default=5)
    ap.add_argument("--pdvc-configs", default="", help="comma-separated subset of pdvc.CONFIGS (default: all)")
    ap.add_argument("--pdvc-sync", default="overlap", choices=["overlap", "flat", "ddp"],
                    help="gradient exchange at N>1: two buckets with the caption-head bucket reduced during the stage-1 "
                         "backward graph (default), one flat all-reduce after backward, or torch DDP")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.sweep:
        return run_sweep(args)
    if args.impl == "reference":
        # rank 0 alone runs and prints; the other ranks exit 0 without work (no process group needed)
        return run_reference(args, int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")))
    world, rank, local = dist_setup(args.gpus)
    try:
        run_ours(args, world, rank, local)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
