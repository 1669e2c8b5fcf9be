from .ms_deform_attn_func import (MSDeformAttnFunction, MSDeformAttnFusedFunction, MSDeformCapSampleFunction,
                                  MSDeformSampleFunction,
                                  get_default_padding_mode, ms_deform_attn_core_pytorch,
                                  set_default_padding_mode)

__all__ = ["MSDeformAttnFunction", "MSDeformAttnFusedFunction", "MSDeformSampleFunction", "MSDeformCapSampleFunction",
           "ms_deform_attn_core_pytorch", "set_default_padding_mode", "get_default_padding_mode"]
