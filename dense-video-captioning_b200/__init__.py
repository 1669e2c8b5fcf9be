"""B200-native (sm_100a) multi-scale deformable attention for PDVC dense video captioning.

Drop-in for the hot path of trtuananh/Dense-Video-Captioning: `pdvc/ops` (MSDeformAttnFunction,
MSDeformAttn, MSDeformAttnCap and the `MultiScaleDeformableAttention` extension module).  Host code
is Python/PyTorch; all arithmetic runs in the hand-written CUDA library `lib/libmsda_sm100.so`
behind the C ABI of include/msda.h.  There is no CPU path: calling any op without the library or
with CPU tensors raises.
"""
from . import _lib  # noqa: F401
from .ops.functions import (MSDeformAttnFunction, MSDeformAttnFusedFunction, MSDeformSampleFunction,
                            ms_deform_attn_core_pytorch)
from .ops.modules import MSDeformAttn, MSDeformAttnCap

__all__ = ["MSDeformAttnFunction", "MSDeformAttnFusedFunction", "MSDeformSampleFunction",
           "ms_deform_attn_core_pytorch", "MSDeformAttn", "MSDeformAttnCap"]
__version__ = "0.1.0"
