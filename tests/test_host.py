"""CPU tests of the host side: the C-ABI library loads and exports everything include/msda.h declares,
argument validation works without a GPU, the drop-in modules keep the reference's parameter names and
initialisation, and nothing silently falls back to a CPU implementation."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden

import dense_video_captioning_b200 as dvc
from dense_video_captioning_b200 import _lib


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "msda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(msda_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 11
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/msda.h but not exported"
    # and the ctypes signature table covers the header one to one
    assert sorted(_lib.SIGNATURES) == declared


def test_library_is_sm100a_only():
    lib = _lib.lib()
    assert lib.msda_arch() == b"sm_100a"
    assert lib.msda_version() == 100
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode == 0:
        archs = set(re.findall(r"sm_\d+a?", out.stdout))
        assert archs == {"sm_100a"}, archs


def test_argument_validation_needs_no_gpu():
    lib = _lib.lib()
    nul = None
    buf = (ctypes.c_char * 64)()
    p = ctypes.cast(buf, ctypes.c_void_p).value
    p16 = (p + 15) // 16 * 16
    dims = (1, 8, 2, 32, 2, 3, 4)
    # unknown dtype / pad mode / bad dims / NULL / misaligned
    assert lib.msda_ms_deform_attn_forward(p16, p16, p16, p16, p16, p16, *dims, 7, 0, 0, nul) == -1
    assert lib.msda_ms_deform_attn_forward(p16, p16, p16, p16, p16, p16, *dims, 0, 5, 0, nul) == -6
    assert lib.msda_ms_deform_attn_forward(p16, p16, p16, p16, p16, p16, 1, 8, 2, 0, 2, 3, 4, 0, 0, 0, nul) == -2
    assert lib.msda_ms_deform_attn_forward(p16, p16, p16, p16, p16, p16, 1, 8, 2, 32, 64, 3, 4, 0, 0, 0, nul) == -2
    assert lib.msda_ms_deform_attn_forward(nul, p16, p16, p16, p16, p16, *dims, 0, 0, 0, nul) == -3
    assert lib.msda_ms_deform_attn_forward(p16 + 4, p16, p16, p16, p16, p16, *dims, 0, 0, 0, nul) == -4
    # bf16 backward without workspace
    assert lib.msda_ms_deform_attn_backward(*([p16] * 9), nul, 0, *dims, 1, 0, 0, nul) == -5
    assert lib.msda_fused_forward(*([p16] * 7), *dims, 3, 0, 0, 0, nul) == -2      # R must be 1 or 2
    assert lib.msda_fused_forward(*([p16] * 7), 1, 8, 2, 30, 2, 3, 4, 1, 0, 0, 0, nul) == -7   # D=30: no vector path
    assert b"workspace" in lib.msda_error_string(-5)
    assert lib.msda_backward_workspace_bytes(2, 10, 8, 64, _lib.BF16) == 2 * 10 * 8 * 64 * 4
    assert lib.msda_backward_workspace_bytes(2, 10, 8, 64, _lib.F32) == 0


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.lib()


def test_cpu_tensors_raise_not_fallback():
    m = dvc.MSDeformAttn(64, 4, 4, 4)
    T = torch.tensor([8, 4, 2, 1])
    lsi = torch.tensor([0, 8, 12, 14])
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 3, 64), torch.zeros(1, 3, 4, 1), torch.zeros(1, 15, 64), T, lsi)
    with pytest.raises(RuntimeError, match="CUDA"):
        dvc.MSDeformAttnFunction.apply(torch.zeros(1, 15, 4, 16), torch.stack([torch.ones_like(T), T], -1), lsi,
                                       torch.zeros(1, 3, 4, 4, 4, 2), torch.zeros(1, 3, 4, 4, 4), 64)
    with pytest.raises(RuntimeError):
        dvc.ms_deform_attn_core_pytorch(torch.zeros(1, 15, 4, 16), [(1, 8), (1, 4), (1, 2), (1, 1)],
                                        torch.zeros(1, 3, 4, 4, 4, 2), torch.zeros(1, 3, 4, 4, 4))


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: no file of the package may mention it."""
    pkg = os.path.join(ROOT, "dense-video-captioning_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "grid_sample(" not in text, f


@pytest.mark.parametrize("cls_name,tag", [("MSDeformAttn", "attn_ref1_nomask"), ("MSDeformAttnCap", "cap_ref1_nomask")])
def test_module_parameters_match_reference(cls_name, tag):
    g = load_golden("module_" + tag)
    cls = getattr(dvc, cls_name)
    m = cls(int(g["d_model"]), int(g["n_levels"]), int(g["n_heads"]), int(g["n_points"]))
    ref_keys = sorted(k[len("param."):] for k in g if k.startswith("param."))
    assert sorted(m.state_dict().keys()) == ref_keys
    for k in ref_keys:
        assert tuple(m.state_dict()[k].shape) == g["param." + k].shape, k
    # _reset_parameters reproduces the reference's offset-bias pattern bit for bit (ms_deform_attn.py:62-71)
    assert np.array_equal(m.sampling_offsets.bias.detach().numpy().astype(np.float64), g["init_offset_bias"])
    assert float(m.sampling_offsets.weight.abs().max()) == 0.0
    assert float(m.attention_weights.weight.abs().max()) == 0.0 and float(m.attention_weights.bias.abs().max()) == 0.0
    assert float(m.value_proj.bias.abs().max()) == 0.0 and float(m.output_proj.bias.abs().max()) == 0.0
    assert m.im2col_step == 64
    m.load_state_dict({k: torch.from_numpy(g["param." + k]).float() for k in ref_keys}, strict=True)


def test_module_rejects_bad_arguments():
    with pytest.raises(ValueError, match="divisible"):
        dvc.MSDeformAttn(30, 4, 8, 4)
    with pytest.warns(UserWarning, match="power of 2"):
        dvc.MSDeformAttn(48, 4, 4, 4)


def test_padding_mode_switch():
    from dense_video_captioning_b200.ops import functions as F
    assert F.get_default_padding_mode() == "border"
    F.set_default_padding_mode("zeros")
    assert F.get_default_padding_mode() == "zeros"
    F.set_default_padding_mode("border")
    with pytest.raises(ValueError):
        F.set_default_padding_mode("reflect")


def test_compat_extension_module_surface():
    sys.path.insert(0, os.path.join(ROOT, "dense-video-captioning_b200", "compat"))
    import MultiScaleDeformableAttention as MSDA   # the name the reference imports (func.py:18)
    assert callable(MSDA.ms_deform_attn_forward) and callable(MSDA.ms_deform_attn_backward)
    with pytest.raises(RuntimeError, match="CUDA"):
        MSDA.ms_deform_attn_forward(torch.zeros(1, 3, 1, 4), torch.tensor([[1, 3]]), torch.tensor([0]),
                                    torch.zeros(1, 1, 1, 1, 1, 2), torch.zeros(1, 1, 1, 1, 1), 64)


def test_bench_helpers():
    import bench
    assert bench.level_lengths(200) == [200, 100, 50, 25]
    assert bench.level_lengths(100) == [100, 50, 25, 13]
    # SURVEY Appendix B, unfused fp32 decoder row is 3 174 400 B; the fused boundary drops one idx tensor
    w = dict(N=1, S=375, M=8, D=64, L=4, P=4, Lq=100, R=1)
    fwd, bwd = bench.algorithmic_bytes(w, 4)
    assert fwd == 4 * (375 * 512 + 2 * 100 * 8 * 16 + 100 * 4 + 100 * 512)
    assert bwd == 4 * (100 * 512 + 375 * 512 + 2 * 12800 + 400) + 4 * (375 * 512 + 2 * 12800)
    # shards partition the batch
    for world in (1, 2, 3, 8):
        spans = [bench.shard_batch(64, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == 64
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    b = bench.offset_bias(8, 4, 4)
    assert torch.equal(b.reshape(-1), dvc.MSDeformAttn(512, 4, 8, 4).sampling_offsets.bias.detach())


def test_training_harness_kernels_have_no_cpu_path():
    """The device-side matcher and the fused optimizer refuse CPU tensors instead of falling back (scipy / torch.optim.Adam
    are the host-side tools; the trainer picks them itself when it runs on the CPU)."""
    import pytest
    import torch
    from dense_video_captioning_b200.ops.functions import linear_sum_assignment_cuda, lsap_supported, pdvc_match_cuda
    from dense_video_captioning_b200.pdvc.optim import FlatClipAdam
    costs = torch.zeros(2, 5, 3)
    assert not lsap_supported(costs)
    with pytest.raises(RuntimeError):
        linear_sum_assignment_cuda(costs)
    with pytest.raises(RuntimeError):
        pdvc_match_cuda(torch.zeros(2, 5, 1), torch.zeros(2, 5, 2), torch.zeros(3, dtype=torch.long), torch.zeros(3, 2), 1, 1, 1, 0.25, 2.0)
    p = torch.zeros(4, requires_grad=True)
    flat = torch.zeros(4)
    with pytest.raises(RuntimeError):
        FlatClipAdam([(flat, [p], [flat.view_as(p)])])
