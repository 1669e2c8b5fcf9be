"""Gradient-norm clipping + Adam over flat gradient buckets in two kernel launches (csrc/msda_optim.cuh).

The reference ends a step with `torch.nn.utils.clip_grad_norm_(model.parameters(), opt.grad_clip)` and
`optimizer.step()` of `optim.Adam(model.parameters(), lr=opt.lr, weight_decay=opt.weight_decay)` (train.py:185-187, :123).
With the gradients already laid out in the trainer's flat all-reduce buckets (pdvc/train.py, sync="overlap") the same
arithmetic needs no per-tensor launches: one pass for the global norm, one for the update of every parameter.
`FlatClipAdam` keeps torch.optim.Adam's public surface -- `param_groups` (lr, betas, eps, weight_decay are read from
group 0 at every step, so lr schedulers work), `state_dict()` / `load_state_dict()` in Adam's own format (a checkpoint
written by torch.optim.Adam loads here and vice versa) -- and needs CUDA: there is no CPU path.
"""
import numpy as np
import torch

from .. import _lib

CHUNK = 8192
_SEGMENT = np.dtype([("g", "<u8"), ("n", "<i8")])
_CHUNK = np.dtype([("param", "<u8"), ("grad", "<u8"), ("exp_avg", "<u8"), ("exp_avg_sq", "<u8"), ("n", "<i4"), ("pad", "<i4")])


class FlatClipAdam:
    def __init__(self, buckets, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_norm=0.0, order=None):
        """buckets: list of (flat_grad, params, grad_views) as built by Trainer._setup_buckets -- `grad_views[i]` is the
        view of `flat_grad` that holds params[i]'s gradient when step() is called.  max_norm <= 0: no clipping.
        order: the parameter list whose order numbers the entries of state_dict() (what torch.optim.Adam(order) would
        write); default: bucket order."""
        assert _SEGMENT.itemsize == 16 and _CHUNK.itemsize == 40
        self.buckets = buckets
        self.params = [p for _, group, _ in buckets for p in group]
        if order is not None:
            assert {id(p) for p in order} == {id(p) for p in self.params}
            self.params = list(order)
        self.param_groups = [{"params": self.params, "lr": lr, "betas": tuple(betas), "eps": eps,
                              "weight_decay": weight_decay, "amsgrad": False, "maximize": False}]
        self.max_norm = float(max_norm)
        self.t = 0
        device = buckets[0][0].device
        if device.type != "cuda":
            raise RuntimeError("FlatClipAdam runs on CUDA only (use torch.optim.Adam + clip_grad_norm_ elsewhere)")
        self.device = device
        self.exp_avg = [torch.zeros_like(flat) for flat, _, _ in buckets]
        self.exp_avg_sq = [torch.zeros_like(flat) for flat, _, _ in buckets]
        segments = np.zeros(len(buckets), dtype=_SEGMENT)
        chunks = []
        for b, (flat, group, views) in enumerate(buckets):
            assert flat.dtype == torch.float32 and flat.is_contiguous()
            segments[b] = (flat.data_ptr(), flat.numel())
            for p, view in zip(group, views):
                assert p.dtype == torch.float32 and p.is_contiguous() and view.numel() == p.numel()
                offset = (view.data_ptr() - flat.data_ptr()) // 4
                for c in range(0, p.numel(), CHUNK):
                    n = min(CHUNK, p.numel() - c)
                    chunks.append((p.data_ptr() + 4 * c, flat.data_ptr() + 4 * (offset + c),
                                   self.exp_avg[b].data_ptr() + 4 * (offset + c),
                                   self.exp_avg_sq[b].data_ptr() + 4 * (offset + c), n, 0))
        self.n_segments, self.n_chunks = len(buckets), len(chunks)
        self._segments = torch.from_numpy(segments.view(np.uint8).copy()).to(device)
        self._chunks = torch.from_numpy(np.array(chunks, dtype=_CHUNK).view(np.uint8).copy()).to(device)
        self._partials = torch.zeros(1024, dtype=torch.float32, device=device)
        self.grad_norm = torch.zeros((), dtype=torch.float32, device=device)    # total norm of the last step (device scalar)
        self._offsets = {id(p): (b, (view.data_ptr() - flat.data_ptr()) // 4)
                         for b, (flat, group, views) in enumerate(buckets) for p, view in zip(group, views)}

    def zero_grad(self, set_to_none=True):
        """Nothing to do: the gradients are views of the buckets and every backward rewrites them completely."""

    @torch.no_grad()
    def step(self):
        group = self.param_groups[0]
        self.t += 1
        beta1, beta2 = group["betas"]
        with torch.cuda.device(self.device):
            rc = _lib.lib().msda_adam_clip_step(
                self._segments.data_ptr(), self.n_segments, self._chunks.data_ptr(), self.n_chunks,
                self._partials.data_ptr(), self._partials.numel(), self.grad_norm.data_ptr(), float(group["lr"]), float(beta1),
                float(beta2), float(group["eps"]), float(group["weight_decay"]), self.t, self.max_norm,
                torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(rc, "msda_adam_clip_step")

    # ---- torch.optim.Adam's checkpoint format ---------------------------------------------------------------------
    def _state_views(self, p):
        b, off = self._offsets[id(p)]
        return (self.exp_avg[b][off:off + p.numel()].view_as(p), self.exp_avg_sq[b][off:off + p.numel()].view_as(p))

    def state_dict(self):
        state = {}
        if self.t > 0:
            for i, p in enumerate(self.params):
                m, v = self._state_views(p)
                state[i] = {"step": torch.tensor(float(self.t)), "exp_avg": m.clone(), "exp_avg_sq": v.clone()}
        group = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        group["params"] = list(range(len(self.params)))
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        group = sd["param_groups"][0]
        for k in ("lr", "betas", "eps", "weight_decay"):
            if k in group:
                self.param_groups[0][k] = tuple(group[k]) if k == "betas" else group[k]
        steps = set()
        for i, p in enumerate(self.params):
            entry = sd["state"].get(i)
            m, v = self._state_views(p)
            if entry is None:
                m.zero_()
                v.zero_()
                continue
            m.copy_(entry["exp_avg"])
            v.copy_(entry["exp_avg_sq"])
            steps.add(int(entry["step"]))
        if len(steps) > 1:
            raise ValueError("FlatClipAdam keeps one step count for all parameters; the checkpoint has several")
        self.t = steps.pop() if steps else 0
