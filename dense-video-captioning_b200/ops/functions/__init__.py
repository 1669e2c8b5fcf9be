from .ms_deform_attn_func import (CaptionAttendFunction, LSTMCellFunction, MSDeformAttnFunction, MSDeformAttnFusedFunction, MSDeformCapSampleFunction,
                                  MSDeformSampleFunction, ProjLinearFunction, StepAttendRun, StepLinear, StepSampleRun,
                                  get_default_padding_mode, ms_deform_attn_core_pytorch,
                                  set_default_padding_mode, skinny_linear)
from .matching import linear_sum_assignment_cuda, lsap_supported, pdvc_match_cuda

__all__ = ["linear_sum_assignment_cuda", "lsap_supported", "pdvc_match_cuda", "MSDeformAttnFunction", "MSDeformAttnFusedFunction", "MSDeformSampleFunction", "MSDeformCapSampleFunction",
           "ms_deform_attn_core_pytorch", "set_default_padding_mode", "get_default_padding_mode", "StepLinear", "StepSampleRun", "StepAttendRun", "skinny_linear", "LSTMCellFunction", "CaptionAttendFunction", "ProjLinearFunction"]
