"""Autograd functions over the C-ABI CUDA library -- mirror of the reference's
`pdvc/ops/functions/ms_deform_attn_func.py`.

  MSDeformAttnFunction          same name, argument order and outputs as the reference class
                                (ms_deform_attn_func.py:20-38); calls msda_ms_deform_attn_forward /
                                _backward instead of the pybind11 module (func.py:25-26, :34-36).
  ms_deform_attn_core_pytorch   same signature as the reference helper (func.py:41-68) but executed
                                by the CUDA library (there is no PyTorch/CPU path in this package).
  MSDeformAttnFusedFunction     new: softmax + sampling locations + core op in one kernel each way
                                (replaces ms_deform_attn.py:100-122 for the 1-D temporal case).
  MSDeformSampleFunction        new: the return_value=True branch (func.py:64-65) as a CUDA op.

Error behaviour follows the reference's host wrapper (ms_deform_attn_cuda.cu:28-38): non-contiguous
or non-CUDA tensors raise RuntimeError; so do unsupported dtypes.
"""
import contextlib
import weakref

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from ... import _lib
from ..._streams import branch_stream

_DTYPES = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16, torch.float64: _lib.F64}
_PAD = {"border": _lib.PAD_BORDER, "zeros": _lib.PAD_ZEROS}
_default_padding_mode = "border"


def set_default_padding_mode(mode):
    """'border' (default) = the reference's PyTorch path / parity oracle (func.py:58-59);
    'zeros' = the reference's CUDA kernel (ms_deform_im2col_cuda.cuh:286-289), for checkpoints trained
    with that extension."""
    global _default_padding_mode
    if mode not in _PAD:
        raise ValueError(f"padding mode must be 'border' or 'zeros', got {mode!r}")
    _default_padding_mode = mode


def get_default_padding_mode():
    return _default_padding_mode


def _pad_code(mode):
    mode = _default_padding_mode if mode is None else mode
    if mode not in _PAD:
        raise ValueError(f"padding mode must be 'border' or 'zeros', got {mode!r}")
    return _PAD[mode]


def _require(cond, msg):
    if not cond:
        raise RuntimeError(msg)


def _check_cuda_contig(**tensors):
    # same wording as AT_ASSERTM in ms_deform_attn_cuda.cu:28-38
    for name, t in tensors.items():
        _require(isinstance(t, torch.Tensor), f"{name} must be a tensor")
        _require(t.is_cuda, f"{name} must be a CUDA tensor")
        _require(t.is_contiguous(), f"{name} tensor has to be contiguous")


def _ptr(t):
    return t.data_ptr() if t is not None else None


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


# spatial_shapes tensors whose H column has already been inspected (one device->host read per tensor)
_shape_cache = {}


def _shapes_are_1d(shapes):
    """True when every level has H == 1.  Costs one tiny D2H read the first time a given shapes
    tensor is seen (keyed on tensor identity + version); the nn.Module front-ends never need it
    because they build the (1, T_l) shapes themselves."""
    key = id(shapes)
    hit = _shape_cache.get(key)
    if hit is not None and hit[0]() is shapes and hit[1] == shapes._version:
        return hit[2]
    flag = bool((shapes[:, 0] == 1).all().item())
    if len(_shape_cache) > 256:
        _shape_cache.clear()
    _shape_cache[key] = (weakref.ref(shapes), shapes._version, flag)
    return flag


_layout_cache = {}


def _check_level_layout(shapes, lsi, S):
    """The reference asserts sum_l H_l*W_l == Len_in (ms_deform_attn.py:89); the kernels index value rows as
    level_start_index[l] + i and tile self-attention by sum(T_l) == Lq, so an inconsistent pair of level tensors
    would read out of bounds or leave output rows unwritten.  Checked on the host once per (shapes, lsi) tensor pair
    (keyed on identity + version: one small device->host read the first time a pair is seen, none afterwards; skipped
    while a CUDA graph is being captured -- the warm-up runs before a capture have already seen the tensors)."""
    key = (id(shapes), id(lsi), int(S))
    hit = _layout_cache.get(key)
    if hit is not None and hit[0]() is shapes and hit[1]() is lsi and hit[2] == (shapes._version, lsi._version):
        return
    if torch.cuda.is_current_stream_capturing():
        return
    sz = (shapes if shapes.dim() == 1 else shapes[:, 0] * shapes[:, 1]).cpu()
    want = torch.cat((sz.new_zeros(1), sz.cumsum(0)[:-1]))
    _require(int(sz.sum()) == int(S), f"spatial_shapes cover {int(sz.sum())} rows but value has {int(S)} "
                                      "(the reference asserts their equality, ms_deform_attn.py:89)")
    _require(bool((sz > 0).all()), "every level must have at least one row")
    _require(torch.equal(lsi.cpu(), want), "level_start_index is not the exclusive prefix sum of the level sizes")
    if len(_layout_cache) > 256:
        _layout_cache.clear()
    _layout_cache[key] = (weakref.ref(shapes), weakref.ref(lsi), (shapes._version, lsi._version))


def _auto_1d(shapes, is1d, loc, pad):
    """Decide whether the 1-D temporal kernels may take a call whose caller did not say.
    H == 1 everywhere is enough under BORDER (the clamped y has no effect at all); under ZEROS the
    row weight depends on y, so the dummy y = 0.5 of ms_deform_attn.py:114-115 is verified too
    (one more tiny D2H read -- the nn.Modules pass assume_1d=True and never get here)."""
    flat = is1d if is1d is not None else _shapes_are_1d(shapes)
    if flat and pad == _lib.PAD_ZEROS:
        flat = bool((loc[..., 1] == 0.5).all().item())
    return flat


def _normalise_shapes(shapes):
    """Accept (L,2) [H,W] (reference layout) or (L,) temporal lengths."""
    _require(shapes.dtype == torch.int64, "spatial_shapes must be int64")
    if shapes.dim() == 1:
        return torch.stack([torch.ones_like(shapes), shapes], -1).contiguous(), True
    _require(shapes.dim() == 2 and shapes.shape[1] == 2, "spatial_shapes must be (L,2) or (L,)")
    return shapes, None


def _index_dtype_ok(value, *idx_tensors):
    """loc / attn dtype rule: equal to value's, except bf16 values, which take fp32 (or bf16,
    up-converted) index-side tensors."""
    for t in idx_tensors:
        if value.dtype == torch.bfloat16:
            _require(t.dtype in (torch.float32, torch.bfloat16), "bf16 value needs fp32/bf16 locations and weights")
        else:
            _require(t.dtype == value.dtype, "value, sampling_locations and attention_weights must share a dtype")


class MSDeformAttnFunction(Function):
    """Reference signature: apply(value, value_spatial_shapes, value_level_start_index,
    sampling_locations, attention_weights, im2col_step) -> (N, Lq, M*D).

    Optional trailing arguments (not in the reference): padding_mode ('border'|'zeros'|None) and
    assume_1d (True skips the H==1 inspection; the modules pass it)."""

    @staticmethod
    def forward(ctx, value, value_spatial_shapes, value_level_start_index, sampling_locations,
                attention_weights, im2col_step=64, padding_mode=None, assume_1d=None, force_generic=False):
        ctx.n_inputs = 9
        _check_cuda_contig(value=value, spatial_shapes=value_spatial_shapes,
                           level_start_index=value_level_start_index, sampling_loc=sampling_locations,
                           attn_weight=attention_weights)
        _require(value.dtype in _DTYPES, f"unsupported dtype {value.dtype} (float32, bfloat16, float64)")
        _require(value.dim() == 4, "value must be (N, S, M, D)")
        _require(sampling_locations.dim() == 6 and sampling_locations.shape[-1] == 2,
                 "sampling_locations must be (N, Lq, M, L, P, 2)")
        _index_dtype_ok(value, sampling_locations, attention_weights)
        shapes, is1d = _normalise_shapes(value_spatial_shapes)
        _require(value_level_start_index.dtype == torch.int64, "level_start_index must be int64")
        N, S, M, D = value.shape
        _, Lq, M2, L, P, _ = sampling_locations.shape
        _require(M2 == M and sampling_locations.shape[0] == N, "sampling_locations does not match value")
        _require(tuple(attention_weights.shape) == (N, Lq, M, L, P), "attention_weights must be (N, Lq, M, L, P)")
        _require(shapes.shape[0] == L and value_level_start_index.numel() == L, "level tensors do not match L")
        _check_level_layout(value_spatial_shapes, value_level_start_index, S)
        pad = _pad_code(padding_mode)
        if assume_1d is None:
            assume_1d = _auto_1d(shapes, is1d, sampling_locations, pad)
        loc, attn = sampling_locations, attention_weights
        if value.dtype == torch.bfloat16:
            loc, attn = loc.float(), attn.float()
        # force_generic: True -> the shape-generic kernels; "no_small" -> never the single-phase small-problem kernels (tests)
        flags = (_lib.FLAG_ASSUME_1D if assume_1d else 0) | (_lib.FLAG_FORCE_GENERIC if force_generic is True else 0) | \
                (_lib.FLAG_NO_SMALL if force_generic == "no_small" else 0)
        out = torch.empty((N, Lq, M * D), dtype=value.dtype, device=value.device)
        with torch.cuda.device(value.device):
            rc = _lib.lib().msda_ms_deform_attn_forward(
                _ptr(value), _ptr(shapes), _ptr(value_level_start_index), _ptr(loc), _ptr(attn), _ptr(out),
                N, S, M, D, L, Lq, P, _DTYPES[value.dtype], pad, flags, _stream(value.device))
        _lib.check(rc, "msda_ms_deform_attn_forward")
        ctx.im2col_step = im2col_step
        ctx.pad, ctx.flags = pad, flags
        ctx.idx_dtypes = (sampling_locations.dtype, attention_weights.dtype)
        ctx.save_for_backward(value, shapes, value_level_start_index, loc, attn)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        value, shapes, lsi, loc, attn = ctx.saved_tensors
        N, S, M, D = value.shape
        _, Lq, _, L, P, _ = loc.shape
        grad_output = grad_output.contiguous()
        _require(grad_output.dtype == value.dtype, "grad_output dtype must match value")
        grad_value = torch.empty_like(value)
        grad_loc = torch.empty_like(loc)
        grad_attn = torch.empty_like(attn)
        lib = _lib.lib()
        code = _DTYPES[value.dtype]
        ws_bytes = lib.msda_backward_workspace_bytes(N, S, M, D, code)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=value.device) if ws_bytes else None
        with torch.cuda.device(value.device):
            rc = lib.msda_ms_deform_attn_backward(
                _ptr(value), _ptr(shapes), _ptr(lsi), _ptr(loc), _ptr(attn), _ptr(grad_output),
                _ptr(grad_value), _ptr(grad_loc), _ptr(grad_attn), _ptr(ws), ws_bytes,
                N, S, M, D, L, Lq, P, code, ctx.pad, ctx.flags, _stream(value.device))
        _lib.check(rc, "msda_ms_deform_attn_backward")
        if grad_loc.dtype != ctx.idx_dtypes[0]:
            grad_loc = grad_loc.to(ctx.idx_dtypes[0])
        if grad_attn.dtype != ctx.idx_dtypes[1]:
            grad_attn = grad_attn.to(ctx.idx_dtypes[1])
        # (grad_value, None, None, grad_sampling_loc, grad_attn_weight, None) as func.py:38, plus
        # None for the optional extras
        return (grad_value, None, None, grad_loc, grad_attn, None, None, None, None)


class MSDeformAttnFusedFunction(Function):
    """apply(value, temporal_shapes, level_start_index, reference_points, sampling_offsets,
    attn_logits, padding_mode=None) -> (N, Lq, M*D)

    value (N,S,M,D) f32|bf16; temporal_shapes (L,) int64; reference_points (N,Lq,L,1|2) f32;
    sampling_offsets, attn_logits (N,Lq,M,L,P) f32 -- the raw Linear outputs of the module
    (ms_deform_attn.py:99-100).  Nothing but the inputs is saved for backward."""

    @staticmethod
    def forward(ctx, value, temporal_shapes, level_start_index, reference_points, sampling_offsets,
                attn_logits, padding_mode=None, no_small=False):
        _check_cuda_contig(value=value, temporal_shapes=temporal_shapes, level_start_index=level_start_index,
                           reference_points=reference_points, sampling_offsets=sampling_offsets,
                           attn_logits=attn_logits)
        _require(value.dtype in (torch.float32, torch.bfloat16), "fused path takes float32 or bfloat16 values")
        _require(temporal_shapes.dtype == torch.int64 and temporal_shapes.dim() == 1, "temporal_shapes must be (L,) int64")
        _require(level_start_index.dtype == torch.int64, "level_start_index must be int64")
        for name, t in (("reference_points", reference_points), ("sampling_offsets", sampling_offsets),
                        ("attn_logits", attn_logits)):
            _require(t.dtype == torch.float32, f"{name} must be float32")
        N, S, M, D = value.shape
        _require(sampling_offsets.dim() == 5, "sampling_offsets must be (N, Lq, M, L, P)")
        _, Lq, M2, L, P = sampling_offsets.shape
        R = reference_points.shape[-1]
        _require(M2 == M and sampling_offsets.shape[0] == N, "sampling_offsets does not match value")
        _require(tuple(attn_logits.shape) == (N, Lq, M, L, P), "attn_logits must be (N, Lq, M, L, P)")
        _require(tuple(reference_points.shape) == (N, Lq, L, R), "reference_points must be (N, Lq, L, R)")
        if R not in (1, 2):
            raise ValueError(f"Last dim of reference_points must be 1 or 2, but get {R} instead.")
        _require(temporal_shapes.numel() == L and level_start_index.numel() == L, "level tensors do not match L")
        _check_level_layout(temporal_shapes, level_start_index, S)
        pad = _pad_code(padding_mode)
        out = torch.empty((N, Lq, M * D), dtype=value.dtype, device=value.device)
        with torch.cuda.device(value.device):
            rc = _lib.lib().msda_fused_forward(
                _ptr(value), _ptr(temporal_shapes), _ptr(level_start_index), _ptr(reference_points),
                _ptr(sampling_offsets), _ptr(attn_logits), _ptr(out), N, S, M, D, L, Lq, P, R,
                _DTYPES[value.dtype], pad, _lib.FLAG_NO_SMALL if no_small else 0, _stream(value.device))
        _lib.check(rc, "msda_fused_forward")
        ctx.pad = pad
        ctx.flags = _lib.FLAG_NO_SMALL if no_small else 0
        ctx.save_for_backward(value, temporal_shapes, level_start_index, reference_points, sampling_offsets,
                              attn_logits)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        value, shapes, lsi, ref, off, logits = ctx.saved_tensors
        N, S, M, D = value.shape
        _, Lq, _, L, P = off.shape
        R = ref.shape[-1]
        grad_output = grad_output.contiguous()
        _require(grad_output.dtype == value.dtype, "grad_output dtype must match value")
        grad_value = torch.empty_like(value)
        grad_off = torch.empty_like(off)
        grad_logits = torch.empty_like(logits)
        grad_ref = torch.empty_like(ref) if ctx.needs_input_grad[3] else None
        lib = _lib.lib()
        code = _DTYPES[value.dtype]
        ws_bytes = lib.msda_backward_workspace_bytes(N, S, M, D, code)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=value.device) if ws_bytes else None
        with torch.cuda.device(value.device):
            rc = lib.msda_fused_backward(
                _ptr(value), _ptr(shapes), _ptr(lsi), _ptr(ref), _ptr(off), _ptr(logits), _ptr(grad_output),
                _ptr(grad_value), _ptr(grad_ref), _ptr(grad_off), _ptr(grad_logits), _ptr(ws), ws_bytes,
                N, S, M, D, L, Lq, P, R, code, ctx.pad, ctx.flags, _stream(value.device))
        _lib.check(rc, "msda_fused_backward")
        return grad_value, None, None, grad_ref, grad_off, grad_logits, None, None


class MSDeformSampleFunction(Function):
    """apply(value, spatial_shapes, level_start_index, sampling_locations, padding_mode=None,
    assume_1d=None) -> un-weighted samples (N*M, D, Lq, L, P)  (func.py:64-65)."""

    @staticmethod
    def forward(ctx, value, value_spatial_shapes, value_level_start_index, sampling_locations,
                padding_mode=None, assume_1d=None):
        _check_cuda_contig(value=value, spatial_shapes=value_spatial_shapes,
                           level_start_index=value_level_start_index, sampling_loc=sampling_locations)
        _require(value.dtype in _DTYPES, f"unsupported dtype {value.dtype}")
        _index_dtype_ok(value, sampling_locations)
        shapes, is1d = _normalise_shapes(value_spatial_shapes)
        N, S, M, D = value.shape
        _, Lq, _, L, P, _ = sampling_locations.shape
        _check_level_layout(value_spatial_shapes, value_level_start_index, S)
        pad = _pad_code(padding_mode)
        if assume_1d is None:
            assume_1d = _auto_1d(shapes, is1d, sampling_locations, pad)
        loc = sampling_locations.float() if value.dtype == torch.bfloat16 else sampling_locations
        flags = _lib.FLAG_ASSUME_1D if assume_1d else 0
        samples = torch.empty((N * M, D, Lq, L, P), dtype=value.dtype, device=value.device)
        with torch.cuda.device(value.device):
            rc = _lib.lib().msda_sample_forward(
                _ptr(value), _ptr(shapes), _ptr(value_level_start_index), _ptr(loc), _ptr(samples),
                N, S, M, D, L, Lq, P, _DTYPES[value.dtype], pad, flags, _stream(value.device))
        _lib.check(rc, "msda_sample_forward")
        ctx.pad, ctx.flags = pad, flags
        ctx.loc_dtype = sampling_locations.dtype
        ctx.save_for_backward(value, shapes, value_level_start_index, loc)
        return samples

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_samples):
        value, shapes, lsi, loc = ctx.saved_tensors
        N, S, M, D = value.shape
        _, Lq, _, L, P, _ = loc.shape
        grad_samples = grad_samples.contiguous()
        _require(grad_samples.dtype == value.dtype, "grad_samples dtype must match value")
        grad_value = torch.empty_like(value)
        grad_loc = torch.empty_like(loc)
        lib = _lib.lib()
        code = _DTYPES[value.dtype]
        ws_bytes = lib.msda_backward_workspace_bytes(N, S, M, D, code)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=value.device) if ws_bytes else None
        with torch.cuda.device(value.device):
            rc = lib.msda_sample_backward(
                _ptr(value), _ptr(shapes), _ptr(lsi), _ptr(loc), _ptr(grad_samples), _ptr(grad_value),
                _ptr(grad_loc), _ptr(ws), ws_bytes, N, S, M, D, L, Lq, P, code, ctx.pad, ctx.flags,
                _stream(value.device))
        _lib.check(rc, "msda_sample_backward")
        if grad_loc.dtype != ctx.loc_dtype:
            grad_loc = grad_loc.to(ctx.loc_dtype)
        return grad_value, None, None, grad_loc, None, None


class MSDeformCapSampleFunction(Function):
    """apply(value, temporal_shapes, level_start_index, reference_points, sampling_offsets, padding_mode=None)
    -> un-weighted samples in the kernel's layout (N, Lq, M, L*P, D).

    Fused front-end of MSDeformAttnCap (ms_deform_attn_for_caption.py:100-123) for the 1-D temporal case: locations
    from reference points + raw offsets, dummy y, sampling -- one kernel forward, one backward.  The reference's
    (N*M, D, Lq, L, P) tensor is `out.view(N, Lq, M, L, P, D).permute(0, 2, 5, 1, 3, 4).reshape(N*M, D, Lq, L, P)`."""

    @staticmethod
    def supported(value):
        ch = 8 if value.dtype == torch.bfloat16 else 4
        return value.dtype in (torch.float32, torch.bfloat16) and value.shape[-1] % ch == 0

    @staticmethod
    def forward(ctx, value, temporal_shapes, level_start_index, reference_points, sampling_offsets, padding_mode=None,
                run=None):
        """run: a StepSampleRun when this call is one word step of a decoding run (same value / reference points at
        every step): the steps' backward passes then add their grad_value / grad_reference_points into ONE buffer."""
        _check_cuda_contig(value=value, temporal_shapes=temporal_shapes, level_start_index=level_start_index,
                           reference_points=reference_points, sampling_offsets=sampling_offsets)
        _require(value.dtype in (torch.float32, torch.bfloat16), "fused sampler takes float32 or bfloat16 values")
        _require(temporal_shapes.dtype == torch.int64 and temporal_shapes.dim() == 1, "temporal_shapes must be (L,) int64")
        _require(reference_points.dtype == torch.float32 and sampling_offsets.dtype == torch.float32,
                 "reference_points and sampling_offsets must be float32")
        N, S, M, D = value.shape
        _require(sampling_offsets.dim() == 5, "sampling_offsets must be (N, Lq, M, L, P)")
        _, Lq, M2, L, P = sampling_offsets.shape
        R = reference_points.shape[-1]
        _require(M2 == M and sampling_offsets.shape[0] == N, "sampling_offsets does not match value")
        _require(tuple(reference_points.shape) == (N, Lq, L, R), "reference_points must be (N, Lq, L, R)")
        if R not in (1, 2):
            raise ValueError(f"Last dim of reference_points must be 1 or 2, but get {R} instead.")
        _check_level_layout(temporal_shapes, level_start_index, S)
        pad = _pad_code(padding_mode)
        out = torch.empty((N, Lq, M, L * P, D), dtype=value.dtype, device=value.device)
        with torch.cuda.device(value.device):
            rc = _lib.lib().msda_cap_sample_forward(
                _ptr(value), _ptr(temporal_shapes), _ptr(level_start_index), _ptr(reference_points),
                _ptr(sampling_offsets), _ptr(out), N, S, M, D, L, Lq, P, R, _DTYPES[value.dtype], pad, 0,
                _stream(value.device))
        _lib.check(rc, "msda_cap_sample_forward")
        ctx.pad = pad
        ctx.run, ctx.index = None, 0
        if run is not None and value.dtype == torch.float32:
            ctx.run, ctx.index = run, run.enter(value, reference_points)
        ctx.save_for_backward(value, temporal_shapes, level_start_index, reference_points, sampling_offsets)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_samples):
        value, shapes, lsi, ref, off = ctx.saved_tensors
        N, S, M, D = value.shape
        _, Lq, _, L, P = off.shape
        R = ref.shape[-1]
        grad_samples = grad_samples.contiguous()
        _require(grad_samples.dtype == value.dtype, "grad_samples dtype must match value")
        run, flags = ctx.run, 0
        if run is not None:
            # one zero-filled buffer per decoding run; every step's kernel adds into it (MSDA_FLAG_ACCUMULATE) and step 0,
            # whose backward runs last, hands it to autograd -- instead of a zero-fill and an accumulation per step
            grad_value, grad_ref = run.buffers(value, ref, ctx.needs_input_grad[3])
            flags = _lib.FLAG_ACCUMULATE
        else:
            grad_value = torch.empty_like(value)
            grad_ref = torch.empty_like(ref) if ctx.needs_input_grad[3] else None
        grad_off = torch.empty_like(off)
        lib = _lib.lib()
        code = _DTYPES[value.dtype]
        ws_bytes = lib.msda_backward_workspace_bytes(N, S, M, D, code)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=value.device) if ws_bytes else None
        with torch.cuda.device(value.device):
            rc = lib.msda_cap_sample_backward(
                _ptr(value), _ptr(shapes), _ptr(lsi), _ptr(ref), _ptr(off), _ptr(grad_samples), _ptr(grad_value),
                _ptr(grad_ref), _ptr(grad_off), _ptr(ws), ws_bytes, N, S, M, D, L, Lq, P, R, code, ctx.pad, flags,
                _stream(value.device))
        _lib.check(rc, "msda_cap_sample_backward")
        if run is not None and not run.leave(ctx.index):
            return None, None, None, None, grad_off, None, None        # the sums travel with step 0
        return grad_value, None, None, grad_ref, grad_off, None, None


class StepSampleRun:
    """The word steps of ONE teacher-forced decoding run through MSDeformCapSampleFunction (LSTM_DSA.py:74-96 calls the
    sampler once per word on the same memory and reference points): their gradients with respect to the value tensor and
    the reference points are sums over the steps, formed here in one buffer by the kernels themselves.  Steps are entered
    in order 0, 1, 2, ...; their backward passes run last-to-first (every step depends on the previous one through the
    LSTM state), which `leave` checks."""

    def __init__(self):
        self.calls = 0
        self._value = self._ref = None
        self._grad_value = self._grad_ref = None
        self._left = None

    def enter(self, value, ref):
        if self.calls == 0:
            self._value, self._ref = value, ref
        _require(value is self._value or value.data_ptr() == self._value.data_ptr(), "StepSampleRun: one value tensor per run")
        _require(ref is self._ref or ref.data_ptr() == self._ref.data_ptr(), "StepSampleRun: one reference tensor per run")
        index, self.calls = self.calls, self.calls + 1
        return index

    def buffers(self, value, ref, need_ref):
        if self._left is None:
            self._left = self.calls
            self._grad_value = torch.zeros_like(value)
            self._grad_ref = torch.zeros_like(ref) if need_ref else None
        return self._grad_value, self._grad_ref

    def leave(self, index):
        """True for the step that hands the sums to autograd (step 0, the last one to run)."""
        self._left -= 1
        _require(index == self._left, "StepSampleRun: backward passes must run from the last step to the first, each once")
        if index != 0:
            return False
        self.calls, self._left, self._value, self._ref = 0, None, None, None
        return True


PROJ_WGRAD_SIDE_STREAM = True   # ProjLinearFunction.backward: weight gradient on a side stream beside the data gradient


class ProjLinearFunction(Function):
    """The Linear projections of MSDeformAttn on the tcgen05 tensor cores (csrc/msda_proj.cuh).

    apply(x, w0, b0, w1, b1, row_mask) -> y0                      (w1 is None: one layer)
                                       -> (y0, y1)                (two layers that share x: sampling_offsets and
                                                                   attention_weights, ms_deform_attn.py:99-100)
    x (..., K); w_s (N_s, K); b_s (N_s) or None; row_mask (...) bool or None -- True rows of y0 are zero
    (the masked_fill of ms_deform_attn.py:97-98).  Under torch.autocast(bf16) the operands are taken in bf16
    (bf16 x bf16 -> fp32 accumulate); otherwise fp32 operands run as 3xTF32 split products (fp32-level accuracy).
    Weight and bias gradients are produced in fp32."""

    @staticmethod
    def supported(x, w0, w1=None):
        if not (x.is_cuda and x.dtype in (torch.float32, torch.bfloat16)):
            return False
        K, N0 = x.shape[-1], w0.shape[0]
        N1 = w1.shape[0] if w1 is not None else 0
        return K % 8 == 0 and N0 % 8 == 0 and N1 % 8 == 0 and (N1 == 0 or N0 % 128 == 0) and x.numel() > 0

    @staticmethod
    def forward(ctx, x, w0, b0, w1, b1, row_mask):
        amp = torch.is_autocast_enabled("cuda")
        dt = torch.get_autocast_dtype("cuda") if amp else x.dtype
        _require(dt in (torch.float32, torch.bfloat16), "ProjLinearFunction takes float32 or bfloat16 operands")
        K = x.shape[-1]
        x2 = x.reshape(-1, K).to(dt).contiguous()
        rows = x2.shape[0]
        ws = [w.to(dt).contiguous() if w is not None else None for w in (w0, w1)]
        bs = [b.float().contiguous() if b is not None else None for b in (b0, b1)]
        N0, N1 = w0.shape[0], (w1.shape[0] if w1 is not None else 0)
        mask = None
        if row_mask is not None:
            mask = row_mask.reshape(-1).contiguous().view(torch.uint8)
            _require(mask.numel() == rows, "row_mask does not match the rows of x")
        y0 = torch.empty((rows, N0), dtype=dt, device=x.device)
        y1 = torch.empty((rows, N1), dtype=dt, device=x.device) if N1 else None
        with torch.cuda.device(x.device):
            rc = _lib.lib().msda_proj_forward(_ptr(x2), _ptr(ws[0]), _ptr(bs[0]), _ptr(ws[1]), _ptr(bs[1]), _ptr(mask),
                                              _ptr(y0), _ptr(y1), rows, K, N0, N1, _DTYPES[dt], _stream(x.device))
        _lib.check(rc, "msda_proj_forward")
        ctx.save_for_backward(x2, ws[0], ws[1], mask)
        ctx.meta = (x.shape, x.dtype, K, N0, N1, dt, b0 is not None, b1 is not None,
                    w0.dtype, w1.dtype if w1 is not None else None)
        lead = x.shape[:-1]
        if N1:
            return y0.view(*lead, N0), y1.view(*lead, N1)
        return y0.view(*lead, N0)

    @staticmethod
    @once_differentiable
    def backward(ctx, *grads):
        x2, w0, w1, mask = ctx.saved_tensors
        xshape, xdtype, K, N0, N1, dt, has_b0, has_b1, w0dt, w1dt = ctx.meta
        rows = x2.shape[0]
        dev = x2.device
        dy0 = grads[0]
        dy0 = (dy0 if dy0 is not None else torch.zeros((rows, N0), dtype=dt, device=dev)).reshape(rows, N0).to(dt).contiguous()
        dy1 = None
        if N1:
            dy1 = grads[1]
            dy1 = (dy1 if dy1 is not None else torch.zeros((rows, N1), dtype=dt, device=dev)).reshape(rows, N1).to(dt).contiguous()
        lib = _lib.lib()
        code = _DTYPES[dt]
        dx = None
        with torch.cuda.device(dev):
            dw0 = torch.empty((N0, K), dtype=torch.float32, device=dev)
            db0 = torch.empty((N0,), dtype=torch.float32, device=dev) if has_b0 else None
            dw1 = torch.empty((N1, K), dtype=torch.float32, device=dev) if N1 else None
            db1 = torch.empty((N1,), dtype=torch.float32, device=dev) if (N1 and has_b1) else None
            # the weight gradient needs dy and x, the data gradient dy and W: independent, both latency-bound on a fraction
            # of the SMs -> the weight gradient runs on a side stream beside the data gradient (a parallel branch of a
            # captured backward graph); joined before returning, so autograd sees every result on the caller's stream
            side = main = None
            if PROJ_WGRAD_SIDE_STREAM and ctx.needs_input_grad[0]:
                main = torch.cuda.current_stream(dev)
                side = branch_stream(dev, "proj_wgrad")
                side.wait_stream(main)
            with (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
                rc = lib.msda_proj_wgrad(_ptr(dy0), _ptr(dy1), _ptr(x2), _ptr(mask), _ptr(dw0), _ptr(db0), _ptr(dw1), _ptr(db1),
                                         rows, K, N0, N1, code, _stream(dev))
            _lib.check(rc, "msda_proj_wgrad")
            if ctx.needs_input_grad[0]:
                dx = torch.empty((rows, K), dtype=dt, device=dev)
                rc = lib.msda_proj_dgrad(_ptr(dy0), _ptr(dy1), _ptr(w0), _ptr(w1), _ptr(mask), _ptr(dx), rows, K, N0, N1, code, _stream(dev))
                _lib.check(rc, "msda_proj_dgrad")
                dx = dx.view(xshape).to(xdtype)
            if side is not None:
                main.wait_stream(side)
        return (dx, dw0.to(w0dt), db0, dw1.to(w1dt) if dw1 is not None else None, db1, None)


def ms_deform_attn_core_pytorch(value, value_spatial_shapes, sampling_locations, attention_weights,
                                return_value=False):
    """Same call as the reference helper (func.py:41-68): no level_start_index argument, optional
    return_value.  Executed by the CUDA library; CPU tensors raise RuntimeError (the reference runs
    F.grid_sample here -- this package ships no CPU path)."""
    _require(value.is_cuda, "value must be a CUDA tensor (this package has no CPU path)")
    shapes = torch.as_tensor(value_spatial_shapes, dtype=torch.int64, device=value.device)
    shapes2, _ = _normalise_shapes(shapes.contiguous())
    sizes = shapes2[:, 0] * shapes2[:, 1]
    lsi = torch.cat((sizes.new_zeros((1,)), sizes.cumsum(0)[:-1])).contiguous()
    value = value.contiguous()
    loc = sampling_locations.contiguous()
    if return_value:
        return MSDeformSampleFunction.apply(value, shapes2, lsi, loc)
    return MSDeformAttnFunction.apply(value, shapes2, lsi, loc, attention_weights.contiguous(), 64)


def skinny_linear(x, weight, bias=None):
    """y = x @ weight.T (+ bias) for an input with few rows: x (M,K), weight (N,K) fp32 CUDA tensors, K % 4 == 0
    (csrc/msda_steplinear.cuh; C ABI msda_skinny_linear_f32)."""
    _check_cuda_contig(x=x, weight=weight)
    _require(x.dtype == torch.float32 and weight.dtype == torch.float32, "skinny_linear takes float32 tensors")
    _require(x.dim() == 2 and weight.dim() == 2 and x.shape[1] == weight.shape[1], "x (M,K) and weight (N,K) expected")
    _require(x.shape[1] % 4 == 0, "K must be a multiple of 4")
    if bias is not None:
        _check_cuda_contig(bias=bias)
        _require(bias.dtype == torch.float32 and bias.numel() == weight.shape[0], "bias must be (N,) float32")
    M, K = x.shape
    N = weight.shape[0]
    y = torch.empty((M, N), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        rc = _lib.lib().msda_skinny_linear_f32(_ptr(x), _ptr(weight), _ptr(bias), _ptr(y), M, N, K, _stream(x.device))
    _lib.check(rc, "msda_skinny_linear_f32")
    return y


class StepLinear:
    """One nn.Linear applied once per word step of a teacher-forced decoding run (`sampling_offsets` of MSDeformAttnCap,
    ms_deform_attn_for_caption.py:100; `h2att`, LSTM_DSA.py:249), for ONE forward/backward pass:

      * forward and data gradient through the few-row kernel (skinny_linear), fp32 also under autocast;
      * the weight and bias gradients of all steps in ONE GEMM / ONE reduction: each step's backward only records its
        (input, grad_output); step 0 -- whose backward runs last, every later step depends on it through the LSTM state --
        computes  dW = [dY_0; ...; dY_n]^T [X_0; ...; X_n]  instead of n GEMMs and 2n gradient accumulations.

    Steps must be applied in order 0, 1, 2, ... and each exactly once per pass (checked when the gradients are formed)."""

    def __init__(self, linear):
        self.weight, self.bias = linear.weight, linear.bias
        self.calls = 0
        self._pending = []
        self._weight_t = None

    def __call__(self, x):
        index, self.calls = self.calls, self.calls + 1
        return _StepLinearFunction.apply(x, self.weight, self.bias, self, index)

    def _transposed(self, weight):
        if self._weight_t is None:
            self._weight_t = weight.t().contiguous()          # once per backward pass, shared by all steps
        return self._weight_t


class _StepLinearFunction(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, x, weight, bias, owner, index):
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        ctx.save_for_backward(x2, weight)
        ctx.owner, ctx.index, ctx.lead, ctx.has_bias = owner, index, lead, bias is not None
        return skinny_linear(x2, weight.contiguous(), bias).view(*lead, weight.shape[0])

    @staticmethod
    @once_differentiable
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad_y):
        x2, weight = ctx.saved_tensors
        owner = ctx.owner
        g2 = grad_y.reshape(-1, grad_y.shape[-1]).contiguous().float()
        grad_x = None
        if ctx.needs_input_grad[0]:
            grad_x = skinny_linear(g2, owner._transposed(weight), None).view(*ctx.lead, weight.shape[1])
        owner._pending.append((ctx.index, x2, g2))
        if ctx.index != 0:
            return grad_x, None, None, None, None
        pending, owner._pending, owner._weight_t = owner._pending, [], None
        _require(sorted(i for i, _, _ in pending) == list(range(owner.calls)),
                 "StepLinear: every step must take part in the backward pass exactly once")
        owner.calls = 0
        xs = torch.cat([x for _, x, _ in pending])
        gs = torch.cat([g for _, _, g in pending])
        grad_w = gs.t().mm(xs) if ctx.needs_input_grad[1] else None
        grad_b = gs.sum(0) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return grad_x, grad_w, grad_b, None, None


class LSTMCellFunction(Function):
    """apply(pre, rec, c_prev) -> (h, c): one step of a one-layer nn.LSTM (LSTM_DSA.py:206, :261) given its gate
    pre-activations split as pre + rec, both (B, 4H) float32 in torch's gate order i,f,g,o; `pre` may be a row-strided
    view (one word step of a (B, steps, 4H) tensor).  One kernel each way (csrc/msda_steplinear.cuh)."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, pre, rec, c_prev):
        _check_cuda_contig(rec=rec, c_prev=c_prev)
        _require(pre.is_cuda and pre.dim() == 2 and pre.stride(1) == 1 and pre.shape == rec.shape, "pre must be (B,4H), unit inner stride")
        _require(pre.dtype == rec.dtype == c_prev.dtype == torch.float32, "LSTM cell takes float32 tensors")
        B, H = c_prev.shape
        _require(rec.shape[1] == 4 * H, "gates must be (B, 4H)")
        c, h, act = torch.empty_like(c_prev), torch.empty_like(c_prev), torch.empty_like(rec)
        with torch.cuda.device(rec.device):
            rc = _lib.lib().msda_lstm_cell_forward_f32(_ptr(pre), pre.stride(0), _ptr(rec), _ptr(c_prev), _ptr(c), _ptr(h),
                                                       _ptr(act), B, H, _stream(rec.device))
        _lib.check(rc, "msda_lstm_cell_forward_f32")
        ctx.save_for_backward(act, c, c_prev)
        return h, c

    @staticmethod
    @once_differentiable
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad_h, grad_c):
        act, c, c_prev = ctx.saved_tensors
        B, H = c_prev.shape
        grad_h = grad_h.contiguous().float() if grad_h is not None else None
        grad_c = grad_c.contiguous().float() if grad_c is not None else None
        grad_gates, grad_c_prev = torch.empty_like(act), torch.empty_like(c_prev)
        with torch.cuda.device(act.device):
            rc = _lib.lib().msda_lstm_cell_backward_f32(_ptr(grad_h), _ptr(grad_c), _ptr(act), _ptr(c), _ptr(c_prev),
                                                        _ptr(grad_gates), _ptr(grad_c_prev), B, H, _stream(act.device))
        _lib.check(rc, "msda_lstm_cell_backward_f32")
        return grad_gates, grad_gates, grad_c_prev


class CaptionAttendFunction(Function):
    """apply(samples, hproj, alpha_weight, alpha_bias, D) -> attended (BM, D).

    Soft attention over the deformable samples of the caption head (LSTM_DSA.py:247-257) in one kernel each way
    (csrc/msda_capattend.cuh).  samples (BM, K, D + A) float32: per sample the clip feature (D) followed by ctx2att of it
    (A); hproj (BM // M, A) = h2att(h); alpha_weight (A,) / alpha_bias (1,) = alpha_net."""

    @staticmethod
    def supported(samples, k, d):
        a = samples.shape[-1] - d
        return (samples.is_cuda and samples.dtype == torch.float32 and k <= 64 and d % 4 == 0 and a % 4 == 0
                and d <= 3072 and 0 < a <= 1536)

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, samples, hproj, alpha_weight, alpha_bias, D, run=None):
        """run: a StepAttendRun when this call is one word step of a decoding run: the steps' alpha_net gradients are then
        reduced once for the whole run instead of once per word (plus two accumulations per word in autograd)."""
        samples, hproj = samples.contiguous(), hproj.contiguous()
        ctx.run, ctx.index = run, (run.enter() if run is not None else 0)
        ctx.alpha_shape = alpha_weight.shape
        alpha_weight, alpha_bias = alpha_weight.contiguous().view(-1), alpha_bias.contiguous().view(-1)
        _check_cuda_contig(samples=samples, hproj=hproj, alpha_weight=alpha_weight, alpha_bias=alpha_bias)
        BM, K, row = samples.shape
        A = row - D
        _require(A > 0 and alpha_weight.numel() == A and hproj.shape[-1] == A and BM % hproj.shape[0] == 0,
                 "samples (BM,K,D+A), hproj (B,A), alpha (A,) expected")
        M = BM // hproj.shape[0]
        attended = torch.empty((BM, D), dtype=torch.float32, device=samples.device)
        weight = torch.empty((BM, K), dtype=torch.float32, device=samples.device)
        with torch.cuda.device(samples.device):
            rc = _lib.lib().msda_cap_attend_forward_f32(_ptr(samples), _ptr(hproj), _ptr(alpha_weight), _ptr(alpha_bias),
                                                        _ptr(attended), _ptr(weight), BM, M, K, D, A, _stream(samples.device))
        _lib.check(rc, "msda_cap_attend_forward_f32")
        ctx.save_for_backward(samples, hproj, alpha_weight, weight)
        ctx.dims = (BM, M, K, D, A)
        return attended

    @staticmethod
    @once_differentiable
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad_attended):
        samples, hproj, alpha_weight, weight = ctx.saved_tensors
        BM, M, K, D, A = ctx.dims
        grad_attended = grad_attended.contiguous().float()
        grad_samples = torch.empty_like(samples)
        grad_hproj = torch.empty((BM, A), dtype=torch.float32, device=samples.device)
        run = ctx.run
        if run is not None:
            grad_alpha = run.rows(ctx.index, BM, A + 1, samples.device)     # this step's slot of the run's (steps, BM, A+1)
        else:
            grad_alpha = torch.empty((BM, A + 1), dtype=torch.float32, device=samples.device)
        with torch.cuda.device(samples.device):
            rc = _lib.lib().msda_cap_attend_backward_f32(_ptr(samples), _ptr(hproj), _ptr(alpha_weight), _ptr(weight),
                                                         _ptr(grad_attended), _ptr(grad_samples), _ptr(grad_hproj),
                                                         _ptr(grad_alpha), BM, M, K, D, A, _stream(samples.device))
        _lib.check(rc, "msda_cap_attend_backward_f32")
        if M > 1:
            grad_hproj = grad_hproj.view(BM // M, M, A).sum(1)
        if run is not None:
            total = run.leave(ctx.index)
            if total is None:
                return grad_samples, grad_hproj, None, None, None, None     # alpha_net's gradients travel with step 0
        else:
            total = grad_alpha.sum(0)
        return grad_samples, grad_hproj, total[:A].view(ctx.alpha_shape), total[A:], None, None


class StepAttendRun:
    """The word steps of ONE decoding run through CaptionAttendFunction: every step's backward writes its per-row alpha_net
    gradients into its slot of one (steps, BM, A+1) buffer; step 0 -- whose backward runs last -- reduces the buffer once."""

    def __init__(self):
        self.calls = 0
        self._buf = None
        self._left = None

    def enter(self):
        index, self.calls = self.calls, self.calls + 1
        return index

    def rows(self, index, bm, width, device):
        if self._left is None:
            self._left = self.calls
            self._buf = torch.empty((self.calls, bm, width), dtype=torch.float32, device=device)
        return self._buf[index]

    def leave(self, index):
        self._left -= 1
        _require(index == self._left, "StepAttendRun: backward passes must run from the last step to the first, each once")
        if index != 0:
            return None
        total = self._buf.view(-1, self._buf.shape[-1]).sum(0)
        self.calls, self._left, self._buf = 0, None, None
        return total
