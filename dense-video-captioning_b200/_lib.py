"""ctypes binding of libmsda_sm100.so (include/msda.h).

This is the only place the package touches native code.  Loading is lazy (so that CPU-only host
logic -- parameter init, state_dict handling -- can be imported and tested without a GPU), but every
compute entry point goes through `lib()`, which raises if the library has not been built: there is
no fallback implementation.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# MSDA_LIB selects another build of the same library (development: lib/libmsda_dev.so)
LIB_PATH = os.environ.get("MSDA_LIB") or os.path.join(_HERE, "lib", "libmsda_sm100.so")
CSRC = os.path.join(_HERE, "csrc")

F32, BF16, F64 = 0, 1, 2
PAD_BORDER, PAD_ZEROS = 0, 1
FLAG_ASSUME_1D, FLAG_FORCE_GENERIC, FLAG_NO_SMALL, FLAG_ACCUMULATE = 1, 2, 4, 8
STATIC_SCRATCH_BYTES = 16 << 20    # MSDA_STATIC_SCRATCH_BYTES
ERR_STATIC_SCRATCH = -8
PATH_NONE, PATH_LEAN, PATH_TILE, PATH_FAST, PATH_GENERIC, PATH_SMALL, PATH_PROJ, PATH_MATCH = range(8)
LSAP_MAX_DIM = 1024

_vp, _i, _u, _sz, _ll = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint, ctypes.c_size_t, ctypes.c_longlong

# name -> (restype, argtypes); mirrors include/msda.h one to one
SIGNATURES = {
    "msda_version": (_i, []),
    "msda_arch": (ctypes.c_char_p, []),
    "msda_error_string": (ctypes.c_char_p, [_i]),
    "msda_last_launch_count": (_i, []),
    "msda_last_launch_path": (_i, []),
    "msda_backward_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "msda_ms_deform_attn_forward": (_i, [_vp] * 6 + [_i] * 7 + [_i, _i, _u, _vp]),
    "msda_ms_deform_attn_backward": (_i, [_vp] * 9 + [_vp, _sz] + [_i] * 7 + [_i, _i, _u, _vp]),
    "msda_fused_forward": (_i, [_vp] * 7 + [_i] * 8 + [_i, _i, _u, _vp]),
    "msda_fused_backward": (_i, [_vp] * 11 + [_vp, _sz] + [_i] * 8 + [_i, _i, _u, _vp]),
    "msda_fused_backward_bf16_static": (_i, [_vp] * 11 + [_i] * 8 + [_i, _u, _vp]),
    "msda_sample_forward": (_i, [_vp] * 5 + [_i] * 7 + [_i, _i, _u, _vp]),
    "msda_sample_backward": (_i, [_vp] * 7 + [_vp, _sz] + [_i] * 7 + [_i, _i, _u, _vp]),
    "msda_cap_sample_forward": (_i, [_vp] * 6 + [_i] * 8 + [_i, _i, _u, _vp]),
    "msda_cap_sample_backward": (_i, [_vp] * 9 + [_vp, _sz] + [_i] * 8 + [_i, _i, _u, _vp]),
    "msda_cap_attend_forward_f32": (_i, [_vp] * 6 + [_i] * 5 + [_vp]),
    "msda_cap_attend_backward_f32": (_i, [_vp] * 8 + [_i] * 5 + [_vp]),
    "msda_skinny_linear_f32": (_i, [_vp] * 4 + [_i] * 3 + [_vp]),
    "msda_lstm_cell_forward_f32": (_i, [_vp, _ll] + [_vp] * 5 + [_i, _i, _vp]),
    "msda_lstm_cell_backward_f32": (_i, [_vp] * 7 + [_i, _i, _vp]),
    "msda_proj_forward": (_i, [_vp] * 8 + [_i] * 5 + [_vp]),
    "msda_proj_dgrad": (_i, [_vp] * 6 + [_i] * 5 + [_vp]),
    "msda_proj_wgrad": (_i, [_vp] * 8 + [_i] * 5 + [_vp]),
    "msda_lsap_f32": (_i, [_vp] * 4 + [_i] * 3 + [_vp]),
    "msda_pdvc_match_f32": (_i, [_vp] * 8 + [_i] * 4 + [ctypes.c_double] * 5 + [_vp]),
    "msda_adam_clip_step": (_i, [_vp, _i, _vp, _i, _vp, _i, _vp] + [ctypes.c_double] * 5 + [_i, ctypes.c_double, _vp]),
}

_lib = None


def build(verbose=False):
    """Compile csrc/ for sm_100a with nvcc (csrc/Makefile) into lib/libmsda_sm100.so."""
    cmd = ["make", "-C", CSRC] + ([] if verbose else ["-s"])
    subprocess.check_call(cmd)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the MSDeformAttn CUDA library has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'` or `make -C "
                f"{CSRC}`). There is no CPU or PyTorch fallback for this op.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            if not hasattr(handle, name) and os.environ.get("MSDA_LIB"):
                continue      # development A/B against an older build of the library (tools/r2_ab.sh)
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(code, what):
    if code != 0:
        msg = lib().msda_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")
