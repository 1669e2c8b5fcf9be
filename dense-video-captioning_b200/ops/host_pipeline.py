"""Host-resident fused MSDeformAttn forward+backward, pipelined over batch chunks.

When the tensors of a call live in (pinned) host memory -- the end-to-end boundary bench.py's `e2e` measures -- the
step is bound by the host link, not by the kernels: the headline workload moves 616 MB in and 614 MB out per step for
1.3 ms of kernel time.  Batch elements are independent, so the batch is cut into chunks and three things
overlap: the host->device copy of chunk c+1 (copy-in stream), the kernels of chunk c (main stream), and the device->host copy of chunk
c-1's results (copy-out stream) -- PCIe is full duplex, the two directions use separate copy engines.  Same arithmetic, same
results as one call on the whole batch (tests/test_parity_gpu.py::test_host_pipeline_equals_single_call).
"""
import torch

from .functions.ms_deform_attn_func import MSDeformAttnFusedFunction


def bind_host_to_gpu_numa_node(device):
    """Run this process (and with it the first-touch placement of the pinned buffers it allocates next) on the CPUs of the
    NUMA node the GPU hangs off, so that host<->device DMA does not cross the socket interconnect.  Returns a dict saying
    what was found and done; never raises (containers may hide sysfs or forbid sched_setaffinity)."""
    import os
    info = {"numa_node": None, "bound": False}
    try:
        p = torch.cuda.get_device_properties(torch.device(device))
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        info["pci"] = bdf
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        info["numa_node"] = node
        if node < 0:
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        info["node_cpus"], info["allowed_before"] = len(cpus), len(os.sched_getaffinity(0))
        if allowed and allowed != os.sched_getaffinity(0):
            os.sched_setaffinity(0, allowed)
            info["bound"] = True
    except Exception as ex:                                 # noqa: BLE001 -- diagnostic only
        info["error"] = repr(ex)[:120]
    return info


class HostPipelinedFusedMSDA:
    """fwd+bwd of MSDeformAttnFusedFunction with pinned host tensors in and out.

    run(value, reference_points, sampling_offsets, attn_logits, grad_output) -> (output, grad_value, grad_offsets,
    grad_logits), all pinned host tensors owned by this object (overwritten by the next run)."""

    def __init__(self, temporal_shapes, level_start_index, device, chunk=8, padding_mode="border"):
        self.shapes = temporal_shapes.to(device)
        self.lsi = level_start_index.to(device)
        self.device = torch.device(device)
        self.chunk = chunk
        self.padding_mode = padding_mode
        self.copy_in = torch.cuda.Stream(self.device)
        self.main = torch.cuda.Stream(self.device)
        self.copy_out = torch.cuda.Stream(self.device)
        self._dev = None
        self._out = None

    def _buffers(self, host_inputs):
        key = tuple((t.shape, t.dtype) for t in host_inputs)
        if self._dev is None or self._key != key:
            self._key = key
            c = self.chunk
            # two device input sets (double buffer), pinned host outputs for the whole batch
            self._dev = [[torch.empty((c,) + tuple(t.shape[1:]), dtype=t.dtype, device=self.device) for t in host_inputs]
                         for _ in range(2)]
            value, ref, off, logits, gout = host_inputs
            self._out = [torch.empty(gout.shape, dtype=gout.dtype).pin_memory(), torch.empty(value.shape, dtype=value.dtype).pin_memory(),
                         torch.empty(off.shape, dtype=off.dtype).pin_memory(), torch.empty(logits.shape, dtype=logits.dtype).pin_memory()]
        return self._dev, self._out

    def run(self, value, reference_points, sampling_offsets, attn_logits, grad_output):
        host_inputs = (value, reference_points, sampling_offsets, attn_logits, grad_output)
        for t in host_inputs:
            if t.is_cuda or not t.is_pinned():
                raise RuntimeError("HostPipelinedFusedMSDA takes pinned host tensors")
        dev, outs = self._buffers(host_inputs)
        n = value.shape[0]
        bounds = [(lo, min(lo + self.chunk, n)) for lo in range(0, n, self.chunk)]
        caller = torch.cuda.current_stream(self.device)
        self.copy_in.wait_stream(caller)
        self.main.wait_stream(caller)
        loaded = [torch.cuda.Event() for _ in bounds]
        consumed = [torch.cuda.Event() for _ in bounds]
        for c, (lo, hi) in enumerate(bounds):
            slot = dev[c % 2]
            with torch.cuda.stream(self.copy_in):
                if c >= 2:
                    self.copy_in.wait_event(consumed[c - 2])           # the slot's previous chunk has been consumed
                for d, h in zip(slot, host_inputs):
                    d[:hi - lo].copy_(h[lo:hi], non_blocking=True)
                loaded[c].record(self.copy_in)
            with torch.cuda.stream(self.main):
                self.main.wait_event(loaded[c])
                v, r, o, lg, g = (d[:hi - lo] for d in slot)
                v = v.detach().requires_grad_(True)
                o = o.detach().requires_grad_(True)
                lg = lg.detach().requires_grad_(True)
                out = MSDeformAttnFusedFunction.apply(v, self.shapes, self.lsi, r, o, lg, self.padding_mode)
                out.backward(g)
                consumed[c].record(self.main)
                results = (out.detach(), v.grad, o.grad, lg.grad)
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(consumed[c])
                for h, d in zip(outs, results):
                    d.record_stream(self.copy_out)                     # allocated on `main`, read here
                    h[lo:hi].copy_(d, non_blocking=True)
        caller.wait_stream(self.main)
        caller.wait_stream(self.copy_out)
        # the returned tensors are HOST buffers filled by asynchronous copies: block the host until they have landed
        # (stream order alone only protects later GPU work)
        done = torch.cuda.Event()
        done.record(self.copy_out)
        done.synchronize()
        return tuple(outs)
