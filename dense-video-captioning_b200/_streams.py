"""Side streams for independent sub-chains of a step (parallel branches of a captured CUDA graph).

torch.cuda.Stream() hands out streams from a pool of 32 per device and priority, round-robin: two Stream objects created at
different times can be the SAME CUDA stream, and a capture stream (torch.cuda.graph takes a fresh default-priority one per
capture) can coincide with a "side" stream created earlier in the process -- the fork then silently serialises (measured: the
second trainer built in a process ran its loss stage 0.9 + 1.25 ms instead of 0.63 + 1.05, tools/seq_probe.py).  The branch
streams here come from the high-priority pool (never a capture stream, which is default priority), one per (role, index),
checked to be distinct CUDA streams, and live for the process.
"""
import torch

_POOL = {}
PRIORITY = -1     # high-priority pool: disjoint from the default-priority streams torch hands to captures


def branch_stream(device, role, index=0):
    device = torch.device(device)
    ordinal = device.index if device.index is not None else torch.cuda.current_device()
    pool = _POOL.setdefault(ordinal, {"streams": {}, "handles": set()})
    key = (role, index)
    stream = pool["streams"].get(key)
    if stream is None:
        for _ in range(64):
            stream = torch.cuda.Stream(ordinal, priority=PRIORITY)
            if stream.cuda_stream not in pool["handles"]:
                break
        else:
            raise RuntimeError("no distinct CUDA stream left for a parallel branch (32 per device and priority)")
        pool["handles"].add(stream.cuda_stream)
        pool["streams"][key] = stream
    return stream
