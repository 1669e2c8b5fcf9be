"""MSDeformAttn -- drop-in for the reference module `pdvc/ops/modules/ms_deform_attn.py:30-126`.

Same constructor, same parameter names (sampling_offsets, attention_weights, value_proj, output_proj
-- checkpoints load unchanged), same initialisation, same forward signature.  The body between the
Linear projections (softmax over levels x points, sampling-location arithmetic, dummy-y stacking and
the core op, ms_deform_attn.py:100-122) runs as ONE fused CUDA kernel forward and one backward.
"""
import math
import warnings

import torch
import torch.nn.functional as F
from torch import nn
from torch.nn.init import constant_, xavier_uniform_

from ... import _lib
from ..._streams import branch_stream
from ..functions.ms_deform_attn_func import MSDeformAttnFunction, MSDeformAttnFusedFunction, ProjLinearFunction

# The module's Linear projections on the hand-written tcgen05 GEMMs (csrc/msda_proj.cuh; value_proj with its padding
# mask in the epilogue, sampling_offsets and attention_weights as ONE GEMM, output_proj).  False: torch's library GEMMs.
USE_PROJ_GEMM = True


PARALLEL_VALUE_PROJ = True    # value projection on a side stream beside the offsets / logits projection


def _proj(x, lin0, lin1=None, row_mask=None):
    """lin0(x) [, lin1(x)] with the row mask applied to lin0's output; tcgen05 path when the shapes allow it."""
    w1 = lin1.weight if lin1 is not None else None
    if USE_PROJ_GEMM and ProjLinearFunction.supported(x, lin0.weight, w1):
        return ProjLinearFunction.apply(x, lin0.weight, lin0.bias, w1, lin1.bias if lin1 is not None else None, row_mask)
    y0 = lin0(x)
    if row_mask is not None:
        y0 = y0.masked_fill(row_mask[..., None], float(0))
    return (y0, lin1(x)) if lin1 is not None else y0


def _is_power_of_2(n):
    if (not isinstance(n, int)) or (n < 0):
        raise ValueError("invalid input for _is_power_of_2: {} (type: {})".format(n, type(n)))
    return (n & (n - 1) == 0) and n != 0


def _fused_ok(head_dim, dtype):
    """Mirror of fast_lpg() in csrc/msda_api.cu: does the vectorised kernel take this head?"""
    ch = 8 if dtype == torch.bfloat16 else 4
    if dtype not in (torch.float32, torch.bfloat16) or head_dim % ch:
        return False
    lanes = head_dim // ch
    return lanes >= 4 and (lanes < 32 or head_dim % (32 * ch) == 0)


class _MSDeformFrontEnd(nn.Module):
    """Parameters, init and projections shared by MSDeformAttn and MSDeformAttnCap."""

    query_scale = 1          # MSDeformAttnCap projects a 2*d_model query
    centre_offsets = False   # MSDeformAttnCap subtracts the per-level mean of the offset bias

    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        if d_model % n_heads != 0:
            raise ValueError('d_model must be divisible by n_heads, but got {} and {}'.format(d_model, n_heads))
        _d_per_head = d_model // n_heads
        if not _is_power_of_2(_d_per_head):
            warnings.warn("You'd better set d_model in MSDeformAttn to make the dimension of each attention head a "
                          "power of 2 which is more efficient in our CUDA implementation.")
        self.im2col_step = 64   # kept for API compatibility (ms_deform_attn.py:48); one launch covers any batch
        self.d_model = d_model
        self.n_levels = n_levels
        self.n_heads = n_heads
        self.n_points = n_points
        self.padding_mode = None   # None -> functions.get_default_padding_mode() ('border')

        self.sampling_offsets = nn.Linear(self.query_scale * d_model, n_heads * n_levels * n_points)
        self.attention_weights = nn.Linear(self.query_scale * d_model, n_heads * n_levels * n_points)
        self.value_proj = nn.Linear(d_model, d_model)
        self.output_proj = nn.Linear(d_model, d_model)
        self._reset_parameters()

    def _reset_parameters(self):
        # ms_deform_attn.py:62-77 / ms_deform_attn_for_caption.py:60-76: zero offset weights, offset bias
        # = sign-normalised cos(2*pi*h/M) * (p+1) per head, zero attention weights, xavier value/out.
        constant_(self.sampling_offsets.weight.data, 0.)
        thetas = torch.arange(self.n_heads, dtype=torch.float32) * (2.0 * math.pi / self.n_heads)
        grid_init = torch.stack([thetas.cos(), thetas.sin()], -1)
        grid_init = (grid_init / grid_init.abs().max(-1, keepdim=True)[0]).view(self.n_heads, 1, 1, 2)
        grid_init = grid_init[..., 0].repeat(1, self.n_levels, self.n_points)
        for i in range(self.n_points):
            grid_init[:, :, i] *= i + 1
        if self.centre_offsets:
            grid_init = grid_init - grid_init.mean(2, keepdim=True)
        with torch.no_grad():
            self.sampling_offsets.bias = nn.Parameter(grid_init.view(-1))
        constant_(self.attention_weights.weight.data, 0.)
        constant_(self.attention_weights.bias.data, 0.)
        xavier_uniform_(self.value_proj.weight.data)
        constant_(self.value_proj.bias.data, 0.)
        xavier_uniform_(self.output_proj.weight.data)
        constant_(self.output_proj.bias.data, 0.)

    def project_value(self, input_flatten, input_padding_mask=None):
        """value_proj + padding mask (ms_deform_attn.py:95-98) -> (N,S,M,D).  Callers that attend to the same memory
        many times (the caption LSTM, one call per word) may compute it once and pass it as `projected_value`."""
        N, Len_in, _ = input_flatten.shape
        value = _proj(input_flatten, self.value_proj, row_mask=input_padding_mask)
        return value.view(N, Len_in, self.n_heads, self.d_model // self.n_heads)

    def _project(self, query, reference_points, input_flatten, input_spatial_shapes, input_padding_mask,
                 projected_value=None, need_logits=True, sampling_offsets=None):
        """value (N,S,M,D), raw offsets and logits (N,Lq,M,L,P) -- ms_deform_attn.py:95-100.  `sampling_offsets`: the
        output of self.sampling_offsets(query) when the caller has already computed it (functions.StepLinear)."""
        if not query.is_cuda:
            raise RuntimeError("MSDeformAttn (B200 build) runs on CUDA tensors only: there is no CPU path "
                               "(the reference falls back to F.grid_sample here, ms_deform_attn.py:123-124)")
        _lib.lib()   # fail loudly if the CUDA library is missing
        N, Len_q, _ = query.shape
        N, Len_in, _ = input_flatten.shape
        if reference_points.shape[-1] not in (1, 2):
            raise ValueError('Last dim of reference_points must be 1 or 2, but get {} instead.'.format(
                reference_points.shape[-1]))
        # value_proj(input) and the offsets / logits projection of the query are independent GEMMs, each latency-bound on a
        # fraction of the SMs: the value projection runs on a side stream (a parallel branch of a captured graph, forward and
        # -- autograd replays a node on the stream its forward ran on -- backward)
        side = main = None
        if projected_value is not None:
            value = projected_value
        elif PARALLEL_VALUE_PROJ and USE_PROJ_GEMM:
            main = torch.cuda.current_stream(query.device)
            side = branch_stream(query.device, "value_proj")
            side.wait_stream(main)
            with torch.cuda.stream(side):
                value = self.project_value(input_flatten, input_padding_mask)
        else:
            value = self.project_value(input_flatten, input_padding_mask)
        logits = None
        if sampling_offsets is None and need_logits:
            sampling_offsets, logits = _proj(query, self.sampling_offsets, self.attention_weights)   # one GEMM
        elif sampling_offsets is None:
            sampling_offsets = _proj(query, self.sampling_offsets)
        elif need_logits:
            logits = _proj(query, self.attention_weights)
        offsets = sampling_offsets.view(N, Len_q, self.n_heads, self.n_levels, self.n_points)
        if logits is not None:
            logits = logits.view(N, Len_q, self.n_heads, self.n_levels, self.n_points)
        if side is not None:
            main.wait_stream(side)
            if not torch.cuda.is_current_stream_capturing():
                value.record_stream(main)
        return value, offsets, logits

    def _locations_and_weights(self, reference_points, offsets, logits, input_spatial_shapes):
        """Unfused front-end (only for head dims the vectorised kernel does not take):
        ms_deform_attn.py:101-117 in torch ops on the GPU."""
        N, Len_q = offsets.shape[:2]
        attn = None
        if logits is not None:
            attn = F.softmax(logits.reshape(N, Len_q, self.n_heads, -1).float(), -1).view_as(logits).contiguous()
        offsets = offsets.float()
        if reference_points.shape[-1] == 1:
            x = reference_points[:, :, None, :, None, 0] + offsets / input_spatial_shapes[None, None, None, :, None]
        else:
            x = reference_points[:, :, None, :, None, 0] + offsets / self.n_points * reference_points[:, :, None, :, None, 1] * 0.5
        loc = torch.stack((x, 0.5 * x.new_ones(x.shape)), -1)
        shapes2 = torch.stack([input_spatial_shapes.new_ones(input_spatial_shapes.shape), input_spatial_shapes], -1)
        return loc.contiguous(), attn, shapes2.contiguous()


class MSDeformAttn(_MSDeformFrontEnd):
    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4):
        """
        Multi-Scale Deformable Attention Module (1-D temporal)
        :param d_model      hidden dimension
        :param n_levels     number of feature levels
        :param n_heads      number of attention heads
        :param n_points     number of sampling points per attention head per feature level
        """
        super().__init__(d_model, n_levels, n_heads, n_points)

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes, input_level_start_index,
                input_padding_mask=None):
        """
        :param query                       (N, Length_{query}, C)
        :param reference_points            (N, Length_{query}, n_levels, 1), range in [0, 1], including padding area
                                        or (N, Length_{query}, n_levels, 2), add additional (c, l) to form reference boxes
        :param input_flatten               (N, sum_l T_l, C)
        :param input_spatial_shapes        (n_levels,), [T_0, T_1, ..., T_{L-1}]
        :param input_level_start_index     (n_levels,), [0, T_0, T_0+T_1, ...]
        :param input_padding_mask          (N, sum_l T_l), True for padding elements, False for non-padding elements

        :return output                     (N, Length_{query}, C)
        """
        value, offsets, logits = self._project(query, reference_points, input_flatten, input_spatial_shapes,
                                               input_padding_mask)
        value = value.contiguous()
        if _fused_ok(value.shape[-1], value.dtype):
            output = MSDeformAttnFusedFunction.apply(
                value, input_spatial_shapes.contiguous(), input_level_start_index.contiguous(),
                reference_points.float().contiguous(), offsets.float().contiguous(), logits.float().contiguous(),
                self.padding_mode)
        else:
            loc, attn, shapes2 = self._locations_and_weights(reference_points, offsets, logits, input_spatial_shapes)
            if value.dtype not in (torch.bfloat16,):
                loc, attn = loc.to(value.dtype), attn.to(value.dtype)
            output = MSDeformAttnFunction.apply(value, shapes2, input_level_start_index.contiguous(), loc, attn,
                                                self.im2col_step, self.padding_mode, True)
        output = _proj(output, self.output_proj)
        return output
