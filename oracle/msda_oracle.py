"""oracle/msda_oracle.py -- TEST INFRASTRUCTURE ONLY (checker, never the product path).

numpy/ctypes front-end of the plain-C oracle (oracle/msda_oracle.c), plus numpy restatements
of the host-side arithmetic of the reference module (softmax over levels x points and the
sampling-location formula, pdvc/ops/modules/ms_deform_attn.py:99-112).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Parity is pinned by tests/golden/*.npz, which were produced by running the
reference's own ms_deform_attn_core_pytorch (imported unmodified from /root/reference) in
fp64 -- see tests/golden/make_golden.py.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libmsda_oracle.so")
_lib = None

BORDER = 0  # pdvc/ops/functions/ms_deform_attn_func.py:58-59 (padding_mode='border')
ZEROS = 1   # pdvc/ops/src/cuda/ms_deform_im2col_cuda.cuh:286-289 (reference CUDA kernel)


def build(force=False):
    """Compile oracle/msda_oracle.c with gcc (oracle/Makefile)."""
    if force or not os.path.exists(_LIB_PATH):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _prep(dtype, *arrs):
    return [None if a is None else np.ascontiguousarray(a, dtype=dtype) for a in arrs]


def _shapes2(shapes):
    """Accept (L,) temporal lengths or (L,2) [H,W]; return int64 (L,2)."""
    s = np.asarray(shapes, dtype=np.int64)
    if s.ndim == 1:
        s = np.stack([np.ones_like(s), s], -1)
    return np.ascontiguousarray(s)


def _suffix(dtype):
    return {np.dtype(np.float64): "f64", np.dtype(np.float32): "f32"}[np.dtype(dtype)]


def forward(value, shapes, loc, attn, pad_mode=BORDER, dtype=np.float64, return_samples=False):
    """out (N,Lq,M*D) of ms_deform_attn_core_pytorch (func.py:41-68).

    return_samples=True additionally returns the return_value=True tensor (N*M, D, Lq, L, P)
    (func.py:64-65).
    """
    value, loc, attn = _prep(dtype, value, loc, attn)
    shapes = _shapes2(shapes)
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = loc.shape
    assert int((shapes[:, 0] * shapes[:, 1]).sum()) == S
    out = np.empty((N, Lq, M * D), dtype=dtype)
    samples = np.empty((N * M, D, Lq, L, P), dtype=dtype) if return_samples else None
    fn = getattr(lib(), "msda_oracle_forward_" + _suffix(dtype))
    fn(_ptr(value), _ptr(shapes), _ptr(loc), _ptr(attn), N, S, M, D, L, Lq, P, int(pad_mode),
       _ptr(out), _ptr(samples))
    return (out, samples) if return_samples else out


def backward(value, shapes, loc, attn, grad_out, pad_mode=BORDER, dtype=np.float64):
    """(grad_value, grad_loc, grad_attn) -- autograd through func.py:41-68."""
    value, loc, attn, grad_out = _prep(dtype, value, loc, attn, grad_out)
    shapes = _shapes2(shapes)
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = loc.shape
    gv = np.empty_like(value)
    gl = np.empty_like(loc)
    ga = np.empty_like(attn)
    fn = getattr(lib(), "msda_oracle_backward_" + _suffix(dtype))
    fn(_ptr(value), _ptr(shapes), _ptr(loc), _ptr(attn), _ptr(grad_out), N, S, M, D, L, Lq, P,
       int(pad_mode), _ptr(gv), _ptr(gl), _ptr(ga))
    return gv, gl, ga


def sample_backward(value, shapes, loc, grad_samples, pad_mode=BORDER, dtype=np.float64):
    """(grad_value, grad_loc) of the return_value=True branch (func.py:64-65)."""
    value, loc, grad_samples = _prep(dtype, value, loc, grad_samples)
    shapes = _shapes2(shapes)
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = loc.shape
    gv = np.empty_like(value)
    gl = np.empty_like(loc)
    fn = getattr(lib(), "msda_oracle_sample_backward_" + _suffix(dtype))
    fn(_ptr(value), _ptr(shapes), _ptr(loc), _ptr(grad_samples), N, S, M, D, L, Lq, P,
       int(pad_mode), _ptr(gv), _ptr(gl))
    return gv, gl


# ---------------------------------------------------------------------------------------------
# Host-side arithmetic of the module (the "fused front-end" of the CUDA path)
# ---------------------------------------------------------------------------------------------

def softmax_lp(logits):
    """softmax over the flattened (L*P) axis -- ms_deform_attn.py:100-101. logits (N,Lq,M,L,P)."""
    x = np.asarray(logits, dtype=np.float64)
    sh = x.shape
    x = x.reshape(sh[0], sh[1], sh[2], -1)
    x = x - x.max(-1, keepdims=True)
    e = np.exp(x)
    return (e / e.sum(-1, keepdims=True)).reshape(sh)


def sampling_locations(ref, offsets, temporal_shapes, n_points):
    """1-D sampling locations x -- ms_deform_attn.py:103-112.

    ref (N,Lq,L,1|2), offsets (N,Lq,M,L,P), temporal_shapes (L,).  Returns x (N,Lq,M,L,P).
    """
    ref = np.asarray(ref, dtype=np.float64)
    off = np.asarray(offsets, dtype=np.float64)
    T = np.asarray(temporal_shapes, dtype=np.float64)
    if ref.shape[-1] == 1:
        return ref[:, :, None, :, None, 0] + off / T[None, None, None, :, None]
    if ref.shape[-1] == 2:
        return ref[:, :, None, :, None, 0] + off / n_points * ref[:, :, None, :, None, 1] * 0.5
    raise ValueError("Last dim of reference_points must be 1 or 2")


def fused_forward(value, temporal_shapes, ref, offsets, logits, pad_mode=BORDER):
    """out of the module core given raw offsets / logits / reference points (fp64)."""
    P = np.asarray(offsets).shape[-1]
    x = sampling_locations(ref, offsets, temporal_shapes, P)
    loc = np.stack([x, np.full_like(x, 0.5)], -1)  # ms_deform_attn.py:114-116
    return forward(value, temporal_shapes, loc, softmax_lp(logits), pad_mode)


def fused_backward(value, temporal_shapes, ref, offsets, logits, grad_out, pad_mode=BORDER):
    """(grad_value, grad_ref, grad_offsets, grad_logits) of fused_forward (fp64, chain rule by hand)."""
    off = np.asarray(offsets, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    T = np.asarray(temporal_shapes, dtype=np.float64)
    P = off.shape[-1]
    A = softmax_lp(logits)
    x = sampling_locations(ref, off, temporal_shapes, P)
    loc = np.stack([x, np.full_like(x, 0.5)], -1)
    gv, gl, ga = backward(value, temporal_shapes, loc, A, grad_out, pad_mode)
    gx = gl[..., 0]
    # softmax Jacobian over the flattened L*P axis
    dot = (A * ga).sum((-1, -2), keepdims=True)
    g_logits = A * (ga - dot)
    g_ref = np.zeros_like(ref)
    if ref.shape[-1] == 1:
        g_off = gx / T[None, None, None, :, None]
        g_ref[..., 0] = gx.sum((2, 4))
    else:
        scale = ref[:, :, None, :, None, 1] * 0.5 / P
        g_off = gx * scale
        g_ref[..., 0] = gx.sum((2, 4))
        g_ref[..., 1] = (gx * off * 0.5 / P).sum((2, 4))
    return gv, g_ref, g_off, g_logits
