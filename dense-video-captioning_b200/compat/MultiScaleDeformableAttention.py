"""Stand-in for the reference's compiled extension module `MultiScaleDeformableAttention`
(pdvc/ops/src/vision.cpp:13-16).  Put this directory on PYTHONPATH and the reference's own
`pdvc/ops/functions/ms_deform_attn_func.py` (`import MultiScaleDeformableAttention as MSDA`, func.py:18)
runs unmodified on top of the B200 library:

    ms_deform_attn_forward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, im2col_step)
        -> Tensor (N, Lq, M*D)
    ms_deform_attn_backward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, grad_output,
                            im2col_step) -> [grad_value, grad_sampling_loc, grad_attn_weight]

Module-level switches:
  padding_mode  'zeros' (default): what the extension this module replaces does (ms_deform_im2col_cuda.cuh:286-289);
                'border': the reference's PyTorch path; None: follow functions.get_default_padding_mode().
  ASSUME_1D     None (default): decide per call -- every level has H == 1 (cached per shapes tensor) and, under 'zeros',
                every sampling_loc[..., 1] == 0.5, which costs ONE blocking device->host read per call (the row weight
                depends on y there), so such calls are not CUDA-graph capturable;
                True: the caller guarantees both (PDVC always: ms_deform_attn.py:114-117) -- no host read, capturable;
                False: always take the shape-generic 2-D kernels.
"""
import os
import sys

import torch

_root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from dense_video_captioning_b200 import _lib  # noqa: E402
from dense_video_captioning_b200.ops.functions import ms_deform_attn_func as _f  # noqa: E402

padding_mode = "zeros"
ASSUME_1D = None


def _common(value, spatial_shapes, level_start_index, sampling_loc, attn_weight):
    _f._check_cuda_contig(value=value, spatial_shapes=spatial_shapes, level_start_index=level_start_index,
                          sampling_loc=sampling_loc, attn_weight=attn_weight)
    _f._require(value.dtype in (torch.float32, torch.float64), "value must be float32 or float64")
    _f._require(sampling_loc.dtype == value.dtype and attn_weight.dtype == value.dtype, "dtype mismatch")
    N, S, M, D = value.shape
    _, Lq, _, L, P, _ = sampling_loc.shape
    pad = _f._pad_code(padding_mode)
    one_d = _f._auto_1d(spatial_shapes, None, sampling_loc, pad) if ASSUME_1D is None else bool(ASSUME_1D)
    flags = _lib.FLAG_ASSUME_1D if one_d else 0
    return N, S, M, D, L, Lq, P, _f._DTYPES[value.dtype], pad, flags


def ms_deform_attn_forward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, im2col_step):
    N, S, M, D, L, Lq, P, code, pad, flags = _common(value, spatial_shapes, level_start_index, sampling_loc, attn_weight)
    out = torch.empty((N, Lq, M * D), dtype=value.dtype, device=value.device)
    with torch.cuda.device(value.device):
        rc = _lib.lib().msda_ms_deform_attn_forward(
            value.data_ptr(), spatial_shapes.data_ptr(), level_start_index.data_ptr(), sampling_loc.data_ptr(),
            attn_weight.data_ptr(), out.data_ptr(), N, S, M, D, L, Lq, P, code, pad, flags,
            torch.cuda.current_stream(value.device).cuda_stream)
    _lib.check(rc, "ms_deform_attn_forward")
    return out


def ms_deform_attn_backward(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, grad_output,
                            im2col_step):
    N, S, M, D, L, Lq, P, code, pad, flags = _common(value, spatial_shapes, level_start_index, sampling_loc, attn_weight)
    _f._check_cuda_contig(grad_output=grad_output)
    gv, gl, ga = torch.empty_like(value), torch.empty_like(sampling_loc), torch.empty_like(attn_weight)
    with torch.cuda.device(value.device):
        rc = _lib.lib().msda_ms_deform_attn_backward(
            value.data_ptr(), spatial_shapes.data_ptr(), level_start_index.data_ptr(), sampling_loc.data_ptr(),
            attn_weight.data_ptr(), grad_output.data_ptr(), gv.data_ptr(), gl.data_ptr(), ga.data_ptr(), None, 0,
            N, S, M, D, L, Lq, P, code, pad, flags, torch.cuda.current_stream(value.device).cuda_stream)
    _lib.check(rc, "ms_deform_attn_backward")
    return [gv, gl, ga]
