"""MSDeformAttnCap -- drop-in for `pdvc/ops/modules/ms_deform_attn_for_caption.py:30-123`.

Same front-end as MSDeformAttn on a 2*d_model query (hidden state + event feature of the caption
LSTM, pdvc/CaptioningHead/LSTM_DSA.py:239-240); returns the raw sampled values (N*M, D, Lq, L, P) with
no weighting and no output projection (ms_deform_attn_for_caption.py:120-123).  The reference always
takes the F.grid_sample path here, even on CUDA; this module runs the sampler CUDA kernel."""
from ..functions.ms_deform_attn_func import MSDeformCapSampleFunction, MSDeformSampleFunction
from .ms_deform_attn import _MSDeformFrontEnd


class MSDeformAttnCap(_MSDeformFrontEnd):
    query_scale = 2
    centre_offsets = True   # ms_deform_attn_for_caption.py:68

    @property
    def supports_rows_layout(self):
        """forward(..., rows_layout=True) exists -- False when `forward` has been replaced (the CPU tests and the CPU
        baseline of bench.py swap in the oracle, which returns the reference's layout)."""
        return type(self).forward is _CUDA_FORWARD

    def __init__(self, d_model=256, n_levels=4, n_heads=8, n_points=4):
        super().__init__(d_model, n_levels, n_heads, n_points)
        # The reference's caption sampler ALWAYS samples with border padding (its grid_sample route), also on CUDA
        # (ms_deform_attn_for_caption.py:120-121): functions.set_default_padding_mode('zeros') -- the switch for
        # checkpoints trained with the reference CUDA extension -- must not reach it.
        self.padding_mode = "border"

    def forward(self, query, reference_points, input_flatten, input_spatial_shapes, input_level_start_index,
                input_padding_mask=None, projected_value=None, sampling_offsets=None, rows_layout=False,
                sample_run=None):
        """rows_layout=False: the reference's (N*M, D, Lq, L, P) (ms_deform_attn_for_caption.py:120-121 via func.py:64-65).
        rows_layout=True: the same samples as (N, Lq, M, L*P, D) -- the layout the kernel writes and the soft attention over
        the samples reads (LSTM_DSA.py:239-257 permutes to it): saves a transposing copy each way, forward and backward.
        sample_run: functions.StepSampleRun when this is one word step of a decoding run (gradient sums formed in place)."""
        value, offsets, logits = self._project(query, reference_points, input_flatten, input_spatial_shapes,
                                               input_padding_mask, projected_value, need_logits=False,
                                               sampling_offsets=sampling_offsets)
        value = value.contiguous()
        if MSDeformCapSampleFunction.supported(value):
            # one fused kernel each way (csrc/msda_capsample.cuh); the reference's (N*M, D, Lq, L, P) is returned as a
            # permuted view of the kernel's (N, Lq, M, L*P, D) buffer: same shape, same values, D-contiguous memory
            N, Lq = offsets.shape[:2]
            M, L, P, D = self.n_heads, self.n_levels, self.n_points, value.shape[-1]
            samples = MSDeformCapSampleFunction.apply(
                value, input_spatial_shapes.contiguous(), input_level_start_index.contiguous(),
                reference_points.float().contiguous(), offsets.float().contiguous(), self.padding_mode, sample_run)
            if rows_layout:
                return samples.view(N, Lq, M, L * P, D)
            return samples.view(N, Lq, M, L, P, D).permute(0, 2, 5, 1, 3, 4).reshape(N * M, D, Lq, L, P)
        # head dims the vectorised kernel does not take: torch front-end + generic sampler kernel.
        # the reference computes attention weights here as well (:102-103); they never reach the output, so the
        # projection is skipped (attention_weights / output_proj stay in the state_dict for checkpoint compatibility)
        loc, _, shapes2 = self._locations_and_weights(reference_points, offsets, logits, input_spatial_shapes)
        value = value.contiguous()
        if value.dtype != loc.dtype and value.dtype.is_floating_point and value.dtype.itemsize >= 4:
            loc = loc.to(value.dtype)
        out = MSDeformSampleFunction.apply(value, shapes2, input_level_start_index.contiguous(), loc,
                                           self.padding_mode, True)
        if rows_layout:
            N, Lq = offsets.shape[:2]
            out = out.reshape(N, self.n_heads, -1, Lq, self.n_levels * self.n_points).permute(0, 3, 1, 4, 2)
        return out


_CUDA_FORWARD = MSDeformAttnCap.forward
