"""Same-box comparison against the reference's OWN CUDA kernels: the unmodified
pdvc/ops/src sources compiled for sm_100 by tools/build_ref_cuda.py into baseline/_ref/ref_cuda/ (a
built .so only; git-ignored, shipped to the GPU box).  The reference kernel zero-pads
(ms_deform_im2col_cuda.cuh:286-289), so it pins this library's ZEROS mode -- forward and all three
gradients, including the dummy-y location gradient the reference kernel writes (cuh:158-159).

Skipped when the .so is absent (e.g. /root/reference was not mounted at build time)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "baseline", "_ref", "ref_cuda", "MultiScaleDeformableAttention_refcuda.so")


@pytest.fixture(scope="module")
def refcuda():
    if not os.path.exists(SO):
        pytest.skip("reference CUDA extension not built (tools/build_ref_cuda.py)")
    spec = importlib.util.spec_from_file_location("MultiScaleDeformableAttention_refcuda", SO)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def dvc():
    import dense_video_captioning_b200 as pkg
    return pkg


def _inputs(seed, N, M, D, shapes, Lq, P, two_d, dtype):
    inp = cases.make_core_inputs(seed, N, M, D, shapes, Lq, P, lo=-0.1, hi=1.1, two_d=two_d)
    dev = "cuda"
    t = {k: torch.from_numpy(np.ascontiguousarray(inp[k])).to(dtype).to(dev) for k in ("value", "loc", "attn", "grad_out")}
    shapes2 = cases.as_shapes2(shapes)
    t["shapes"] = torch.from_numpy(shapes2).to(dev)
    t["lsi"] = torch.from_numpy(cases.level_start_index(shapes2)).to(dev)
    return t


CASES = [
    # name, N, M, D, shapes, Lq, P, two_d, dtype
    ("yc2_dec", 1, 8, 64, cases.level_lengths(200), 100, 4, False, torch.float32),
    ("yc2_enc_n4", 4, 8, 64, cases.level_lengths(200), 375, 4, False, torch.float32),
    ("anet_dec", 2, 8, 64, cases.level_lengths(100), 10, 4, False, torch.float32),
    ("cap_m1_d512", 1, 1, 512, cases.level_lengths(200), 6, 4, False, torch.float32),
    ("p8_d32", 2, 4, 32, cases.level_lengths(64), 50, 8, False, torch.float32),
    ("p16_l4", 2, 8, 64, cases.level_lengths(120), 37, 16, False, torch.float32),     # L*P = 64 (lean kernels, 2 samples/lane)
    ("p12_l4", 1, 4, 64, cases.level_lengths(75), 21, 12, False, torch.float32),      # L*P = 48
    ("ref_test_2d_f64", 1, 2, 2, [(6, 4), (3, 2)], 2, 2, True, torch.float64),      # pdvc/ops/test.py:21-25
    ("ragged_2d", 2, 4, 16, [(5, 7), (3, 4), (1, 9)], 13, 3, True, torch.float32),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_zeros_mode_matches_reference_cuda_kernels(refcuda, dvc, case):
    _, N, M, D, shapes, Lq, P, two_d, dtype = case
    t = _inputs(11, N, M, D, shapes, Lq, P, two_d, dtype)
    want_out = refcuda.ms_deform_attn_forward(t["value"], t["shapes"], t["lsi"], t["loc"], t["attn"], 64)
    want_gv, want_gl, want_ga = refcuda.ms_deform_attn_backward(
        t["value"], t["shapes"], t["lsi"], t["loc"], t["attn"], t["grad_out"].view(N, Lq, M * D).contiguous(), 64)

    value = t["value"].clone().requires_grad_(True)
    loc = t["loc"].clone().requires_grad_(True)
    attn = t["attn"].clone().requires_grad_(True)
    out = dvc.MSDeformAttnFunction.apply(value, t["shapes"], t["lsi"], loc, attn, 64, "zeros")
    out.backward(t["grad_out"].view_as(out))
    torch.cuda.synchronize()

    rel = 1e-11 if dtype == torch.float64 else 2e-5   # two fp32 kernels with different summation order

    def close(got, want, what):
        scale = float(want.abs().max())
        err = float((got - want).abs().max())
        assert err <= rel * scale + 1e-6 * (dtype != torch.float64) + 1e-13, f"{what}: {err:.3e} vs scale {scale:.3e}"

    close(out, want_out, "output")
    close(value.grad, want_gv, "grad_value")
    close(attn.grad, want_ga, "grad_attn_weight")
    # grad_sampling_loc is piecewise constant in the location: an fp32 floor flip at a knot changes one element by O(1)
    d = (loc.grad - want_gl).abs()
    bad = float((d > rel * float(want_gl.abs().max()) + 1e-6).float().mean())
    assert bad <= 1e-3, f"grad_sampling_loc: {bad:.2e} of elements differ"
