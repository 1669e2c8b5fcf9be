"""GPU parity tests (run on the B200 box with -m gpu): the CUDA library, called through the C ABI by
the autograd functions, against the CPU oracle (oracle/) on the same seeded inputs and against the
committed golden vectors produced by the unmodified reference.

Stated tolerances:
  fp32 kernel vs fp64 oracle on identical (fp32-representable) inputs:  |err| <= 2e-4 * max|ref| + 1e-6
  bf16 kernel vs fp64 oracle on the bf16-rounded inputs:                |err| <= 1.6e-2 * max|ref| + 1e-6
  fp64 (generic kernels) vs fp64 oracle:                                |err| <= 1e-10 * max|ref| + 1e-12
"""
import numpy as np
import pytest
import torch

import cases
from conftest import load_golden
from oracle import msda_oracle as O

pytestmark = pytest.mark.gpu

TOL = {torch.float32: (2e-4, 1e-6), torch.bfloat16: (1.6e-2, 1e-6), torch.float64: (1e-10, 1e-12)}
PAD = {"border": O.BORDER, "zeros": O.ZEROS}


@pytest.fixture(scope="module")
def dvc():
    import dense_video_captioning_b200 as pkg
    assert torch.cuda.is_available()
    return pkg


def _round_inputs(inp, dtype):
    """Round the seeded fp64 inputs to what the kernel will see, so oracle and kernel share bits."""
    idx_dt = torch.float64 if dtype == torch.float64 else torch.float32
    out = {}
    for k in ("value", "grad_out"):
        out[k] = torch.from_numpy(np.ascontiguousarray(inp[k])).to(dtype)
    for k in ("loc", "attn"):
        out[k] = torch.from_numpy(np.ascontiguousarray(inp[k])).to(idx_dt)
    return out


def _np64(t):
    return t.detach().double().cpu().numpy()


def _assert_close(got, want, dtype, what, rel_scale=1.0):
    rel, ab = TOL[dtype]
    want = np.asarray(want, dtype=np.float64)
    got = np.asarray(got, dtype=np.float64)
    assert got.shape == want.shape, f"{what}: shape {got.shape} vs {want.shape}"
    scale = np.abs(want).max()
    err = np.abs(got - want).max()
    assert np.isfinite(got).all(), f"{what}: non-finite values"
    assert err <= rel * rel_scale * scale + ab, f"{what}: max|err| {err:.3e} > {rel * rel_scale:.1e} * {scale:.3e}"


def _run_core(dvc, t, shapes, pad, **kw):
    dev = "cuda"
    value = t["value"].to(dev).requires_grad_(True)
    loc = t["loc"].to(dev).requires_grad_(True)
    attn = t["attn"].to(dev).requires_grad_(True)
    shp = torch.from_numpy(shapes).to(dev)
    lsi = torch.from_numpy(cases.level_start_index(shapes)).to(dev)
    out = dvc.MSDeformAttnFunction.apply(value, shp, lsi, loc, attn, 64, pad, *kw.get("extra", ()))
    out.backward(t["grad_out"].to(dev))
    torch.cuda.synchronize()
    return out, value.grad, loc.grad, attn.grad


def _oracle(t, shapes, pad):
    v, l, a, g = (_np64(t[k]) for k in ("value", "loc", "attn", "grad_out"))
    out = O.forward(v, shapes, l, a, PAD[pad])
    gv, gl, ga = O.backward(v, shapes, l, a, g, PAD[pad])
    return out, gv, gl, ga


ONE_D = [c for c in cases.ALL_CORE_CASES if c not in ("ref_test_2d", "ragged_2d", "vec_2d")]


@pytest.mark.parametrize("path", ["auto", "no_small"])     # auto: small problems take csrc/msda_small.cuh; no_small: the staged kernels
@pytest.mark.parametrize("pad", ["border", "zeros"])
@pytest.mark.parametrize("name", ONE_D)
def test_core_fp32_vs_oracle(dvc, name, pad, path):
    inp = cases.core_inputs(name)
    t = _round_inputs(inp, torch.float32)
    got = _run_core(dvc, t, inp["shapes"], pad, extra=(None, "no_small") if path == "no_small" else ())
    want = _oracle(t, inp["shapes"], pad)
    for g, w, what in zip(got, want, ("out", "grad_value", "grad_loc", "grad_attn")):
        _assert_close(_np64(g), w, torch.float32, f"{name}/{pad}/{what}")
    if pad == "border":
        assert float(got[2][..., 1].abs().max()) == 0.0   # dummy-y gradient is exactly 0


@pytest.mark.parametrize("name", ["yc2_dec", "anet_enc", "p8_d32", "edges", "wide_range", "p16_d128_l1"])
def test_core_fp32_generic_kernels_vs_oracle(dvc, name):
    """Same problems through the shape-generic kernels (MSDA_FLAG_FORCE_GENERIC)."""
    inp = cases.core_inputs(name)
    t = _round_inputs(inp, torch.float32)
    got = _run_core(dvc, t, inp["shapes"], "border", extra=(None, True))
    want = _oracle(t, inp["shapes"], "border")
    for g, w, what in zip(got, want, ("out", "grad_value", "grad_loc", "grad_attn")):
        _assert_close(_np64(g), w, torch.float32, f"{name}/generic/{what}")


@pytest.mark.parametrize("pad", ["border", "zeros"])
@pytest.mark.parametrize("name", ["ref_test_2d", "ragged_2d", "vec_2d", "odd_d71"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_core_2d_and_fp64_vs_oracle(dvc, name, pad, dtype):
    """True 2-D levels and fp64 (what pdvc/ops/test.py exercises) run on the generic kernels."""
    inp = cases.core_inputs(name)
    t = _round_inputs(inp, dtype)
    got = _run_core(dvc, t, inp["shapes"], pad)
    want = _oracle(t, inp["shapes"], pad)
    for g, w, what in zip(got, want, ("out", "grad_value", "grad_loc", "grad_attn")):
        _assert_close(_np64(g), w, dtype, f"{name}/{pad}/{what}")


@pytest.mark.parametrize("path", ["auto", "no_small"])
@pytest.mark.parametrize("pad", ["border", "zeros"])
@pytest.mark.parametrize("name", ["yc2_dec", "anet_enc", "p8_d32", "p16_d128_l1", "edges", "uniform01"])
def test_core_bf16_vs_oracle(dvc, name, pad, path):
    inp = cases.core_inputs(name)
    if inp["value"].shape[-1] % 8:
        pytest.skip("bf16 vector path needs D % 8 == 0")
    t = _round_inputs(inp, torch.bfloat16)
    got = _run_core(dvc, t, inp["shapes"], pad, extra=(None, "no_small") if path == "no_small" else ())
    want = _oracle(t, inp["shapes"], pad)
    for g, w, what in zip(got, want, ("out", "grad_value", "grad_loc", "grad_attn")):
        _assert_close(_np64(g), w, torch.bfloat16, f"{name}/{pad}/bf16/{what}")


@pytest.mark.parametrize("name", ["yc2_dec", "anet_enc", "edges", "cap_m1_d512", "uniform01"])
def test_core_fp32_vs_reference_golden(dvc, name):
    """Directly against the numbers the unmodified reference produced (tests/golden)."""
    inp = cases.core_inputs(name)
    g = load_golden("core_border_" + name)
    t = _round_inputs(inp, torch.float32)
    got = _run_core(dvc, t, inp["shapes"], "border")
    # the golden run used the un-rounded fp64 inputs: allow the input-rounding error on top
    for arr, key in zip(got, ("out", "grad_value", "grad_loc", "grad_attn")):
        _assert_close(_np64(arr), g[key], torch.float32, f"{name}/golden/{key}", rel_scale=4.0)


def test_zeros_mode_vs_reference_golden(dvc):
    inp = cases.core_inputs("yc2_dec")
    g = load_golden("core_zeros_yc2_dec")
    t = _round_inputs(inp, torch.float32)
    got = _run_core(dvc, t, inp["shapes"], "zeros")
    for arr, key in zip(got, ("out", "grad_value", "grad_loc", "grad_attn")):
        _assert_close(_np64(arr), g[key], torch.float32, f"zeros/golden/{key}", rel_scale=4.0)


# ------------------------------------------------------------------------------------------------
# fused front-end
# ------------------------------------------------------------------------------------------------
def _fused_inputs(seed, N, M, D, T, Lq, P, R):
    rs = np.random.RandomState(seed)
    L = len(T)
    S = sum(T)
    value = rs.standard_normal((N, S, M, D))
    if R == 1:
        ref = rs.uniform(0.0, 1.0, (N, Lq, L, 1))
    else:
        ref = np.stack([rs.uniform(0.05, 0.95, (N, Lq, L)), rs.uniform(0.02, 0.5, (N, Lq, L))], -1)
    off = rs.standard_normal((N, Lq, M, L, P)) * 2.5
    logits = rs.standard_normal((N, Lq, M, L, P)) * 1.5
    grad_out = rs.standard_normal((N, Lq, M * D))
    return dict(value=value, ref=ref, off=off, logits=logits, grad_out=grad_out, T=np.asarray(T, dtype=np.int64))


def _nudge_off_knots(inp, P):
    """Shift offsets so that no sample sits within 2e-4 px of an interpolation knot."""
    x = O.sampling_locations(inp["ref"].astype(np.float32).astype(np.float64),
                             inp["off"].astype(np.float32).astype(np.float64), inp["T"], P)
    T = inp["T"].astype(np.float64)[None, None, None, :, None]
    pix = x * T - 0.5
    frac = pix - np.floor(pix)
    bad = (frac < 5e-4) | (frac > 1 - 5e-4)
    inp["off"] = np.where(bad, inp["off"] + 0.0137, inp["off"])
    return inp


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("pad", ["border", "zeros"])
@pytest.mark.parametrize("cfg", [
    dict(seed=21, N=1, M=8, D=64, T=cases.level_lengths(200), Lq=100, P=4, R=1),
    dict(seed=22, N=2, M=8, D=64, T=cases.level_lengths(100), Lq=188, P=4, R=2),
    dict(seed=23, N=1, M=4, D=32, T=cases.level_lengths(40), Lq=19, P=8, R=1),
    dict(seed=24, N=1, M=2, D=128, T=[30, 15], Lq=9, P=16, R=2),
    dict(seed=25, N=2, M=3, D=16, T=[12], Lq=5, P=3, R=1),
    dict(seed=26, N=2, M=8, D=64, T=cases.level_lengths(120), Lq=70, P=16, R=2),   # L*P = 64, decoder tiling, boxes
    dict(seed=27, N=1, M=4, D=64, T=cases.level_lengths(90), Lq=33, P=10, R=1),    # L*P = 40
])
@pytest.mark.parametrize("path", ["auto", "no_small"])
def test_fused_vs_oracle(dvc, cfg, pad, dtype, path):
    if dtype == torch.bfloat16 and cfg["D"] < 32:
        pytest.skip("bf16 vector path needs D >= 32")
    inp = _nudge_off_knots(_fused_inputs(**cfg), cfg["P"])
    dev = "cuda"
    value = torch.from_numpy(inp["value"]).to(dtype).to(dev).requires_grad_(True)
    ref = torch.from_numpy(inp["ref"]).float().to(dev).requires_grad_(True)
    off = torch.from_numpy(inp["off"]).float().to(dev).requires_grad_(True)
    logits = torch.from_numpy(inp["logits"]).float().to(dev).requires_grad_(True)
    gout = torch.from_numpy(inp["grad_out"]).to(dtype).to(dev)
    T = torch.from_numpy(inp["T"]).to(dev)
    lsi = torch.from_numpy(cases.level_start_index(cases.as_shapes2(inp["T"]))).to(dev)
    out = dvc.MSDeformAttnFusedFunction.apply(value, T, lsi, ref, off, logits, pad, path == "no_small")
    took = dvc._lib.lib().msda_last_launch_path()
    assert (took != dvc._lib.PATH_SMALL) if path == "no_small" else True
    out.backward(gout)
    torch.cuda.synchronize()
    v64, r64, o64, l64, g64 = (_np64(x) for x in (value, ref, off, logits, gout))
    want_out = O.fused_forward(v64, inp["T"], r64, o64, l64, PAD[pad])
    gv, g_ref, g_off, g_logits = O.fused_backward(v64, inp["T"], r64, o64, l64, g64, PAD[pad])
    _assert_close(_np64(out), want_out, dtype, "fused/out")
    _assert_close(_np64(value.grad), gv, dtype, "fused/grad_value")
    _assert_close(_np64(off.grad), g_off, dtype, "fused/grad_offsets")
    _assert_close(_np64(logits.grad), g_logits, dtype, "fused/grad_logits")
    _assert_close(_np64(ref.grad), g_ref, dtype, "fused/grad_ref")


# ------------------------------------------------------------------------------------------------
# sampler (return_value=True branch)
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["cap_m1_d512", "edges", "p8_d32"])
def test_sampler_vs_reference_golden(dvc, name):
    inp = cases.core_inputs(name)
    g = load_golden("core_border_" + name)
    dev = "cuda"
    value = torch.from_numpy(inp["value"]).float().to(dev).requires_grad_(True)
    loc = torch.from_numpy(inp["loc"]).float().to(dev).requires_grad_(True)
    shp = torch.from_numpy(inp["shapes"]).to(dev)
    lsi = torch.from_numpy(inp["lsi"]).to(dev)
    samples = dvc.MSDeformSampleFunction.apply(value, shp, lsi, loc)
    gs = np.random.RandomState(1234).standard_normal(tuple(samples.shape))
    samples.backward(torch.from_numpy(gs).float().to(dev))
    torch.cuda.synchronize()
    _assert_close(_np64(samples), g["samples"], torch.float32, "samples", rel_scale=4.0)
    _assert_close(_np64(value.grad), g["sgrad_value"], torch.float32, "sgrad_value", rel_scale=4.0)
    _assert_close(_np64(loc.grad), g["sgrad_loc"], torch.float32, "sgrad_loc", rel_scale=4.0)


# ------------------------------------------------------------------------------------------------
# module drop-in: reference state_dict -> new module, output and gradients vs the reference's
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["attn_ref1_nomask", "attn_ref1_mask", "attn_ref2_nomask", "attn_ref2_mask",
                                 "cap_ref1_nomask", "cap_ref1_mask", "cap_ref2_nomask", "cap_ref2_mask"])
def test_module_dropin_vs_reference_golden(dvc, tag):
    g = load_golden("module_" + tag)
    cls = dvc.MSDeformAttnCap if tag.startswith("cap") else dvc.MSDeformAttn
    m = cls(int(g["d_model"]), int(g["n_levels"]), int(g["n_heads"]), int(g["n_points"]))
    sd = {k[len("param."):]: torch.from_numpy(v).float() for k, v in g.items() if k.startswith("param.")}
    m.load_state_dict(sd, strict=True)
    m = m.cuda()
    dev = "cuda"
    query = torch.from_numpy(g["query"]).float().to(dev).requires_grad_(True)
    src = torch.from_numpy(g["src"]).float().to(dev).requires_grad_(True)
    ref = torch.from_numpy(g["ref"]).float().to(dev).requires_grad_(True)
    T = torch.from_numpy(g["T"]).to(dev)
    lsi = torch.cat((T.new_zeros((1,)), T.cumsum(0)[:-1]))
    mask = torch.from_numpy(g["mask"]).to(dev) if g["mask"].size else None
    out = m(query, ref, src, T, lsi, mask)
    out.backward(torch.from_numpy(g["grad_out"]).float().to(dev))
    torch.cuda.synchronize()
    _assert_close(_np64(out), g["out"], torch.float32, tag + "/out", rel_scale=4.0)
    _assert_close(_np64(query.grad), g["grad_query"], torch.float32, tag + "/grad_query", rel_scale=4.0)
    _assert_close(_np64(src.grad), g["grad_src"], torch.float32, tag + "/grad_src", rel_scale=4.0)
    _assert_close(_np64(ref.grad), g["grad_ref"], torch.float32, tag + "/grad_ref", rel_scale=4.0)
    for name, p in m.named_parameters():
        want = g["gparam." + name]
        if p.grad is None:
            assert np.abs(want).max() == 0.0, f"{tag}: missing grad for {name}"
            continue
        _assert_close(_np64(p.grad), want, torch.float32, f"{tag}/grad {name}", rel_scale=4.0)


# ------------------------------------------------------------------------------------------------
# API contract (ms_deform_attn_cuda.cu:28-52)
# ------------------------------------------------------------------------------------------------
def test_api_contract_errors(dvc):
    inp = cases.core_inputs("p8_d32")
    t = _round_inputs(inp, torch.float32)
    dev = "cuda"
    shp = torch.from_numpy(inp["shapes"]).to(dev)
    lsi = torch.from_numpy(inp["lsi"]).to(dev)
    v, l, a = t["value"].to(dev), t["loc"].to(dev), t["attn"].to(dev)
    with pytest.raises(RuntimeError, match="contiguous"):
        dvc.MSDeformAttnFunction.apply(v.transpose(1, 2).contiguous().transpose(1, 2), shp, lsi, l, a, 64)
    with pytest.raises(RuntimeError, match="CUDA"):
        dvc.MSDeformAttnFunction.apply(v.cpu(), shp, lsi, l, a, 64)
    with pytest.raises(RuntimeError, match="dtype"):
        dvc.MSDeformAttnFunction.apply(v.half(), shp, lsi, l.half(), a.half(), 64)
    with pytest.raises(RuntimeError, match="dtype"):
        dvc.MSDeformAttnFunction.apply(v, shp, lsi, l.double(), a, 64)
    with pytest.raises(RuntimeError):
        dvc.ms_deform_attn_core_pytorch(v.cpu(), inp["shapes"], l.cpu(), a.cpu())
    m = dvc.MSDeformAttn(64, 4, 4, 4)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.zeros(1, 3, 64), torch.zeros(1, 3, 4, 1), torch.zeros(1, 15, 64),
          torch.tensor([8, 4, 2, 1]), torch.tensor([0, 8, 12, 14]))


@pytest.mark.parametrize("N", [65, 130])
def test_batch_not_multiple_of_im2col_step(dvc, N):
    """The reference asserts batch % min(batch, 64) == 0 (ms_deform_attn_cuda.cu:50-52); one launch here
    covers any N -- check every sample against the oracle of sample 0 replicated."""
    base = cases.make_core_inputs(seed=31, N=1, M=2, D=32, shapes=cases.level_lengths(16), Lq=5, P=4)
    t = _round_inputs(base, torch.float32)
    rep = {k: v.repeat(N, *([1] * (v.dim() - 1))).contiguous() for k, v in t.items()}
    got = _run_core(dvc, rep, base["shapes"], "border")
    want = _oracle(t, base["shapes"], "border")
    for g, w, what in zip(got, want, ("out", "grad_value", "grad_loc", "grad_attn")):
        g = _np64(g)
        _assert_close(g[:1], w, torch.float32, f"N={N}/{what}[0]")
        _assert_close(g[-1:], w, torch.float32, f"N={N}/{what}[-1]")


def test_core_pytorch_name_runs_on_cuda(dvc):
    """ms_deform_attn_core_pytorch keeps its reference signature but is CUDA-backed."""
    inp = cases.core_inputs("p8_d32")
    t = _round_inputs(inp, torch.float32)
    out = dvc.ms_deform_attn_core_pytorch(t["value"].cuda(), inp["shapes"], t["loc"].cuda(), t["attn"].cuda())
    want = O.forward(_np64(t["value"]), inp["shapes"], _np64(t["loc"]), _np64(t["attn"]))
    _assert_close(_np64(out), want, torch.float32, "core_pytorch/out")
    samples = dvc.ms_deform_attn_core_pytorch(t["value"].cuda(), inp["shapes"], t["loc"].cuda(), t["attn"].cuda(),
                                              return_value=True)
    _, want_s = O.forward(_np64(t["value"]), inp["shapes"], _np64(t["loc"]), _np64(t["attn"]), return_samples=True)
    _assert_close(_np64(samples), want_s, torch.float32, "core_pytorch/samples")


def test_backward_run_to_run_variation_is_bounded(dvc):
    """grad_value uses fp32 atomics (like the reference, ms_deform_im2col_cuda.cuh:126-153): summation
    order may vary, the result may not by more than rounding."""
    inp = cases.core_inputs("anet_enc")
    t = _round_inputs(inp, torch.float32)
    a = _np64(_run_core(dvc, t, inp["shapes"], "border")[1])
    b = _np64(_run_core(dvc, t, inp["shapes"], "border")[1])
    assert np.abs(a - b).max() <= 1e-5 * np.abs(a).max()


# ------------------------------------------------------------------------------------------------
# encoder self-attention geometry (Lq == S): level-grid tiling + staged value windows in both kernels
# ------------------------------------------------------------------------------------------------
def _encoder_inputs(seed, N, M, D, T0, P, jitter):
    """Reference points on the level grids (deformable_transformer.py:198-208), offsets = the module's
    initial bias (ms_deform_attn.py:62-71) + N(0, jitter) rows."""
    import bench
    rs = np.random.RandomState(seed)
    T = cases.level_lengths(T0)
    L, S = len(T), sum(T)
    ref1 = np.concatenate([(np.arange(t) + 0.5) / t for t in T])
    ref = np.broadcast_to(ref1[None, :, None, None], (N, S, L, 1)).copy()
    bias = bench.offset_bias(M, L, P).numpy().astype(np.float64)
    off = bias[None, None] + jitter * rs.standard_normal((N, S, M, L, P))
    return dict(value=rs.standard_normal((N, S, M, D)), ref=ref, off=off,
                logits=rs.standard_normal((N, S, M, L, P)), grad_out=rs.standard_normal((N, S, M * D)),
                T=np.asarray(T, dtype=np.int64))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("pad", ["border", "zeros"])
@pytest.mark.parametrize("cfg", [
    dict(seed=31, N=2, M=8, D=64, T0=300, P=4, jitter=1.0),    # 10 time spans per (n, m): several tiles, halos
    dict(seed=32, N=1, M=4, D=64, T0=521, P=4, jitter=6.0),    # odd lengths, wide windows
    dict(seed=33, N=1, M=2, D=32, T0=160, P=8, jitter=40.0),   # windows that do not fit: L1 gather fallback
    dict(seed=34, N=2, M=8, D=64, T0=300, P=16, jitter=2.0),   # L*P = 64: two samples per lane (lean kernels, SPL = 2)
    dict(seed=35, N=1, M=4, D=64, T0=200, P=12, jitter=3.0),   # L*P = 48: SPL = 2 with dead lanes, P not a power of two
    dict(seed=36, N=3, M=2, D=128, T0=97, P=9, jitter=1.5),    # L*P = 36, D = 128, odd everything
])
@pytest.mark.parametrize("path", ["auto", "no_small"])
def test_encoder_geometry_vs_oracle(dvc, cfg, pad, dtype, path):
    inp = _nudge_off_knots(_encoder_inputs(**cfg), cfg["P"])
    dev = "cuda"
    value = torch.from_numpy(inp["value"]).to(dtype).to(dev).requires_grad_(True)
    ref = torch.from_numpy(inp["ref"]).float().to(dev).requires_grad_(True)
    off = torch.from_numpy(inp["off"]).float().to(dev).requires_grad_(True)
    logits = torch.from_numpy(inp["logits"]).float().to(dev).requires_grad_(True)
    gout = torch.from_numpy(inp["grad_out"]).to(dtype).to(dev)
    T = torch.from_numpy(inp["T"]).to(dev)
    lsi = torch.from_numpy(cases.level_start_index(cases.as_shapes2(inp["T"]))).to(dev)
    out = dvc.MSDeformAttnFusedFunction.apply(value, T, lsi, ref, off, logits, pad, path == "no_small")
    out.backward(gout)
    torch.cuda.synchronize()
    v64, r64, o64, l64, g64 = (_np64(x) for x in (value, ref, off, logits, gout))
    want_out = O.fused_forward(v64, inp["T"], r64, o64, l64, PAD[pad])
    gv, g_ref, g_off, g_logits = O.fused_backward(v64, inp["T"], r64, o64, l64, g64, PAD[pad])
    _assert_close(_np64(out), want_out, dtype, "enc/out")
    _assert_close(_np64(value.grad), gv, dtype, "enc/grad_value")
    _assert_close(_np64(logits.grad), g_logits, dtype, "enc/grad_logits")
    _assert_close(_np64(ref.grad), g_ref, dtype, "enc/grad_ref", rel_scale=4.0)
    # grad_offsets is piecewise constant in the location: allow a 1e-3 fraction of knot-flip outliers
    rel, ab = TOL[dtype]
    d = np.abs(_np64(off.grad) - g_off)
    assert (d > rel * np.abs(g_off).max() + ab).mean() < 1e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_full_size_properties(dvc, dtype):
    """BASELINE.json-sized problem (bench.py's headline workload) where the oracle is too slow: size-independent
    properties.  (1) out is linear in value, so <out, g> == <value, grad_value> (adjointness);
    (2) softmax Jacobian: sum over (l, p) of grad_logits is 0 for every (n, q, m);
    (3) the fused entry points agree with the unfused ones fed torch-computed locations / weights;
    (4) the first video of the batch agrees with the fp64 oracle on a slice of its queries."""
    import bench
    cfg = dict(bench.HEADLINE)
    w = bench.make_workload(cfg["N"], cfg["T0"], cfg["M"], cfg["D"], cfg["L"], cfg["P"], cfg["kind"], "cuda", dtype=dtype, seed=7)
    value = w["value"].clone().requires_grad_(True)
    off = w["off"].clone().requires_grad_(True)
    logits = w["logits"].clone().requires_grad_(True)
    out = dvc.MSDeformAttnFusedFunction.apply(value, w["shapes"], w["lsi"], w["ref"], off, logits, "border")
    out.backward(w["grad_out"])
    torch.cuda.synchronize()
    assert torch.isfinite(out).all() and torch.isfinite(value.grad).all()
    lhs = float((out.detach().double() * w["grad_out"].double()).sum())
    rhs = float((value.detach().double() * value.grad.double()).sum())
    scale = float((out.detach().double().abs() * w["grad_out"].double().abs()).sum())
    assert abs(lhs - rhs) <= (3e-3 if dtype == torch.bfloat16 else 1e-5) * scale, (lhs, rhs, scale)
    gl = logits.grad.double().flatten(-2).sum(-1)
    assert float(gl.abs().max()) <= 1e-4 * float(logits.grad.abs().max()) * (50 if dtype == torch.bfloat16 else 1) + 1e-6
    # (3) unfused path on torch-computed loc / attn (a 4-video slice keeps the temporaries small)
    nb = 4
    Tt = w["shapes"].float()
    attn = torch.softmax(w["logits"][:nb].flatten(-2), -1).view_as(w["logits"][:nb]).contiguous()
    x = w["ref"][:nb, :, None, :, None, 0] + w["off"][:nb] / Tt[None, None, None, :, None]
    loc = torch.stack((x, torch.full_like(x, 0.5)), -1).contiguous()
    shapes2 = torch.stack([torch.ones_like(w["shapes"]), w["shapes"]], -1).contiguous()
    v2 = w["value"][:nb].clone().requires_grad_(True)
    out2 = dvc.MSDeformAttnFunction.apply(v2, shapes2, w["lsi"], loc, attn, 64, "border", True)
    out2.backward(w["grad_out"][:nb].contiguous())
    rel = 2e-2 if dtype == torch.bfloat16 else 3e-4   # x recomputed in fp32 outside the kernel: a few knot flips
    d = (out2.detach().double() - out[:nb].detach().double()).abs()
    assert float((d > rel * float(out.detach().abs().max())).double().mean()) < 1e-3
    dgv = (v2.grad.double() - value.grad[:nb].double()).abs()
    assert float((dgv > rel * float(value.grad.abs().max())).double().mean()) < 1e-3
    # (4) oracle on a query slice of video 0 (queries only matter through their own rows of the inputs)
    q0, q1 = 900, 1100
    T = np.asarray(w["T"], dtype=np.int64)
    v64 = _np64(w["value"][:1])
    want = O.fused_forward(v64, T, _np64(w["ref"][:1, q0:q1]), _np64(w["off"][:1, q0:q1]), _np64(w["logits"][:1, q0:q1]), O.BORDER)
    _assert_close(_np64(out[:1, q0:q1]), want, dtype, "full/out-slice")


def test_host_pipeline_equals_single_call(dvc):
    """ops/host_pipeline.HostPipelinedFusedMSDA (chunks on three streams, pinned host tensors in and out) returns what one
    fused call on the whole batch returns: chunking the batch is pure scheduling.  Uneven last chunk included."""
    import bench
    from dense_video_captioning_b200.ops.host_pipeline import HostPipelinedFusedMSDA
    w = bench.make_workload(11, 64, 8, 64, 4, 4, "encoder", "cuda", seed=21)
    v = w["value"].clone().requires_grad_(True)
    o = w["off"].clone().requires_grad_(True)
    lg = w["logits"].clone().requires_grad_(True)
    out = dvc.MSDeformAttnFusedFunction.apply(v, w["shapes"], w["lsi"], w["ref"], o, lg, "border")
    out.backward(w["grad_out"])
    torch.cuda.synchronize()
    host = [w[k].cpu().pin_memory() for k in ("value", "ref", "off", "logits", "grad_out")]
    pipe = HostPipelinedFusedMSDA(w["shapes"], w["lsi"], "cuda", chunk=4)
    for _ in range(2):      # second run reuses the buffers
        got = pipe.run(*host)
        torch.cuda.synchronize()
    # chunks of 4 videos are small problems (csrc/msda_small.cuh), the whole batch takes the staged kernels: same results up
    # to fp32 rounding (different summation order of the 16 samples of a query)
    for a, b in ((got[0], out.detach()), (got[2], o.grad), (got[3], lg.grad)):
        assert float((a - b.cpu()).abs().max()) <= 1e-5 * float(b.abs().max())
    # grad_value is accumulated with float atomics: equal up to summation order
    assert float((got[1] - v.grad.cpu()).abs().max()) <= 1e-5 * float(v.grad.abs().max())
    with pytest.raises(RuntimeError):
        pipe.run(*[t.cuda() for t in host])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("pad", ["border", "zeros"])
@pytest.mark.parametrize("shape", [(1, 1, 512, 6, 4, 1), (1, 1, 512, 6, 4, 2), (3, 4, 32, 17, 8, 1), (2, 2, 24, 5, 3, 2)],
                         ids=["cap_R1", "cap_R2", "n3_m4_d32_p8", "n2_m2_d24_p3_R2"])
def test_fused_caption_sampler_vs_oracle(dvc, shape, pad, dtype):
    """msda_cap_sample_forward/backward (csrc/msda_capsample.cuh): samples, grad_value, grad_offsets and
    grad_reference_points against the oracle's sampler + the chain rule of ms_deform_attn_for_caption.py:104-112."""
    from dense_video_captioning_b200.ops.functions import MSDeformCapSampleFunction
    N, M, D, Lq, P, R = shape
    if dtype == torch.bfloat16 and D % 8:
        pytest.skip("bf16 vector path needs D % 8 == 0")
    T = cases.level_lengths(200)
    L, S = len(T), sum(T)
    rs = np.random.RandomState(3)
    value = torch.from_numpy(rs.standard_normal((N, S, M, D))).to(dtype)
    ref = rs.uniform(0.05, 0.95, (N, Lq, L, R)).astype(np.float32)
    if R == 2:
        ref[..., 1] = rs.uniform(0.05, 0.3, (N, Lq, L))
    off = (rs.standard_normal((N, Lq, M, L, P)) * 3).astype(np.float32)
    x = O.sampling_locations(ref, off, T, P)
    x = cases.keep_off_knots(x, T)                       # move the offsets so that locations stay off the knots
    Tl = np.asarray(T, dtype=np.float64)[None, None, None, :, None]
    scale = (1.0 / Tl) if R == 1 else (ref[:, :, None, :, None, 1].astype(np.float64) * 0.5 / P)
    off = ((x - ref[:, :, None, :, None, 0]) / scale).astype(np.float32)
    gs = torch.from_numpy(rs.standard_normal((N, Lq, M, L * P, D))).to(dtype)
    dev = "cuda"
    tv = value.to(dev).requires_grad_(True)
    tr = torch.from_numpy(ref).to(dev).requires_grad_(True)
    to = torch.from_numpy(off).to(dev).requires_grad_(True)
    shp = torch.tensor(T, dtype=torch.long, device=dev)
    lsi = torch.cat((shp.new_zeros((1,)), shp.cumsum(0)[:-1]))
    out = MSDeformCapSampleFunction.apply(tv, shp, lsi, tr, to, pad)
    out.backward(gs.to(dev))
    torch.cuda.synchronize()
    # oracle: exact fp64 on the values the kernel saw
    v64 = _np64(value)
    x64 = O.sampling_locations(ref, off, T, P)
    loc = np.stack([x64, np.full_like(x64, 0.5)], -1)
    attn = np.full((N, Lq, M, L, P), 1.0)
    _, want = O.forward(v64, T, loc, attn, PAD[pad], return_samples=True)       # (N*M, D, Lq, L, P)
    want_k = want.reshape(N, M, D, Lq, L * P).transpose(0, 3, 1, 4, 2)           # kernel layout (N, Lq, M, LP, D)
    gs_ref = _np64(gs).reshape(N, Lq, M, L, P, D).transpose(0, 2, 5, 1, 3, 4).reshape(N * M, D, Lq, L, P)
    gv, gl = O.sample_backward(v64, T, loc, np.ascontiguousarray(gs_ref), PAD[pad])
    gx = gl[..., 0]
    g_off = gx * scale
    g_ref = np.zeros(ref.shape)
    g_ref[..., 0] = gx.sum(axis=(2, 4))
    if R == 2:
        g_ref[..., 1] = (gx * off.astype(np.float64) * 0.5 / P).sum(axis=(2, 4))
    _assert_close(_np64(out), want_k, dtype, "samples")
    _assert_close(_np64(tv.grad), gv, dtype, "grad_value", rel_scale=2.0)
    for name, got, ref_arr in (("grad_offsets", to.grad, g_off), ("grad_reference_points", tr.grad, g_ref)):
        d = np.abs(_np64(got) - ref_arr)
        rel, ab = TOL[dtype]
        bad = float((d > 4 * rel * np.abs(ref_arr).max() + ab).mean())
        assert bad <= 2e-3, f"{name}: {bad:.2e} of elements differ"


def test_caption_module_returns_reference_shape_as_view(dvc):
    """MSDeformAttnCap keeps the reference's output shape (N*M, D, Lq, L, P); for the caption head (N = M = 1) it is a
    view of the kernel's D-contiguous buffer, and the head's own permute (LSTM_DSA.py:241-242) is contiguous again."""
    m = dvc.MSDeformAttnCap(512, 4, 1, 4).cuda()
    T = torch.tensor(cases.level_lengths(100), device="cuda")
    lsi = torch.cat((T.new_zeros((1,)), T.cumsum(0)[:-1]))
    q = torch.randn(1, 6, 1024, device="cuda")
    src = torch.randn(1, int(T.sum()), 512, device="cuda")
    ref = torch.rand(1, 6, 4, 1, device="cuda")
    out = m(q, ref, src, T, lsi)
    assert tuple(out.shape) == (1, 512, 6, 4, 4)
    back = out.reshape(1, 1, 512, 6, 16).permute(0, 3, 1, 4, 2)
    assert back.is_contiguous()


@pytest.mark.parametrize("M,N,K", [(12, 512, 512), (12, 16, 1024), (1, 7, 4), (9, 33, 516), (30, 1608, 512), (17, 2048, 1536)])
def test_skinny_linear_matches_torch_fp32(M, N, K):
    """Floating-point kernel: compared with F.linear in float64 at fp32 rounding level (rel 2e-6 of the row scale)."""
    from dense_video_captioning_b200.ops.functions import skinny_linear
    g = torch.Generator().manual_seed(5)
    x = torch.randn(M, K, generator=g).cuda()
    w = torch.randn(N, K, generator=g).cuda()
    b = torch.randn(N, generator=g).cuda()
    want = torch.nn.functional.linear(x.double(), w.double(), b.double())
    got = skinny_linear(x, w, b)
    scale = float((x.double().abs() @ w.double().abs().t()).max())
    assert float((got.double() - want).abs().max()) <= 2e-6 * scale
    got = skinny_linear(x, w)
    assert float((got.double() - (want - b.double())).abs().max()) <= 2e-6 * scale


def test_step_linear_gradients_match_linear_in_a_recurrence():
    """StepLinear inside a recurrence (each step's input depends on the previous step's output): outputs, input
    gradients and the deferred weight / bias gradients equal nn.Linear's."""
    from dense_video_captioning_b200.ops.functions import StepLinear
    torch.manual_seed(2)
    lin = torch.nn.Linear(64, 64).cuda()
    x0 = torch.randn(10, 64, device="cuda", requires_grad=True)
    drive = torch.randn(7, 10, 64, device="cuda")

    def run(apply):
        h, outs = x0, []
        for i in range(7):
            h = torch.tanh(apply(h) + drive[i])
            outs.append(h)
        loss = (torch.stack(outs) ** 2).sum()
        grads = torch.autograd.grad(loss, [x0, lin.weight, lin.bias])
        return [loss.detach()] + list(grads)

    want = run(lin)
    step = StepLinear(lin)
    got = run(step)
    assert step.calls == 0 and not step._pending          # the pass was closed by step 0's backward
    for a, b in zip(got, want):
        assert float((a - b).abs().max()) <= 2e-5 * float(b.abs().max())
    got2 = run(step)                                         # reusable after a complete pass
    for a, b in zip(got2, want):
        assert float((a - b).abs().max()) <= 2e-5 * float(b.abs().max())


def test_step_lstm_matches_nn_lstm_float64():
    """StepLSTM (hoisted input projections + few-row kernel + fused cell kernel) against nn.LSTM run step by step in
    float64 on the CPU: outputs, final state, gradients of both weight matrices and of every input (rel 2e-5)."""
    from dense_video_captioning_b200.pdvc.model import StepLSTM
    torch.manual_seed(4)
    B, E, C, H, steps = 10, 32, 48, 64, 6
    rnn = torch.nn.LSTM(E + 2 * C, H, 1, bias=False)
    ref = torch.nn.LSTM(E + 2 * C, H, 1, bias=False).double()
    ref.load_state_dict({k: v.double() for k, v in rnn.state_dict().items()})
    rnn = rnn.cuda()
    assert StepLSTM.supported(rnn)
    embedded = torch.randn(B, steps, E)
    query = torch.randn(1, B, C)
    attended = torch.randn(steps, B, C)
    mix = torch.randn(H, C) * 0.3            # makes each step's `attended` depend on the previous hidden state

    def run(module, dev, dtype, stepper):
        emb = embedded.to(dev, dtype).requires_grad_(True)
        q = query.to(dev, dtype).requires_grad_(True)
        att = attended.to(dev, dtype).requires_grad_(True)
        m = mix.to(dev, dtype)
        state = (torch.zeros(1, B, H, device=dev, dtype=dtype), torch.zeros(1, B, H, device=dev, dtype=dtype))
        step = stepper(module, emb, q)
        outs = []
        for i in range(steps):
            h_prev = state[0] if state[0].dim() == 2 else state[0][-1]     # StepLSTM hands its state on as (B, H) tensors
            a = att[i] + h_prev @ m
            out, state = step(i, emb, q, a, state)
            outs.append(out)
        outs = torch.stack(outs)
        loss = (outs ** 2).sum() + (state[1] ** 2).sum()
        grads = torch.autograd.grad(loss, [emb, q, att, module.weight_ih_l0, module.weight_hh_l0])
        return [outs.detach(), state[1].detach().reshape(1, B, H)] + [g for g in grads]

    def ref_stepper(module, emb, q):
        def step(i, emb, q, a, state):
            out, state = module(torch.cat((emb[:, i].unsqueeze(0), a.unsqueeze(0), q), 2), state)
            return out.squeeze(0), state
        return step

    def new_stepper(module, emb, q):
        cell = StepLSTM(module, emb, q)
        return lambda i, emb, q, a, state: cell(a, state)

    want = run(ref, "cpu", torch.float64, ref_stepper)
    got = run(rnn, "cuda", torch.float32, new_stepper)
    for g, w in zip(got, want):
        assert float((g.double().cpu() - w).abs().max()) <= 2e-5 * float(w.abs().max())


@pytest.mark.parametrize("B,M,K,D,A", [(12, 1, 16, 512, 512), (5, 2, 12, 64, 48), (3, 1, 64, 32, 20), (1, 1, 1, 8, 4)])
def test_caption_attend_matches_torch_float64(B, M, K, D, A):
    """Fused soft attention over the samples (LSTM_DSA.py:247-257) against the torch expression in float64: output and
    all four gradients (rel 2e-5)."""
    from dense_video_captioning_b200.ops.functions import CaptionAttendFunction
    g = torch.Generator().manual_seed(9)
    samples = torch.randn(B * M, K, D + A, generator=g)
    hproj = torch.randn(B, A, generator=g)
    alpha_w = torch.randn(1, A, generator=g) * 0.3
    alpha_b = torch.randn(1, generator=g)
    grad = torch.randn(B * M, D, generator=g)

    def run(dev, dtype, fused):
        s, h, w, b = (t.to(dev, dtype).requires_grad_(True) for t in (samples, hproj, alpha_w, alpha_b))
        if fused:
            out = CaptionAttendFunction.apply(s, h, w, b, D)
        else:
            clip, ctx = s[..., :D], s[..., D:]
            scores = torch.tanh(ctx + h.repeat_interleave(M, 0)[:, None, :]) @ w.t() + b          # (BM, K, 1)
            out = (torch.softmax(scores.squeeze(-1), 1)[:, :, None] * clip).sum(1)
        grads = torch.autograd.grad(out, [s, h, w, b], grad.to(dev, dtype))
        return [out.detach()] + list(grads)

    want = run("cpu", torch.float64, False)
    got = run("cuda", torch.float32, True)
    for name, a, b in zip(("attended", "grad_samples", "grad_hproj", "grad_alpha", "grad_alpha_bias"), got, want):
        scale = max(float(b.abs().max()), 1.0 if name == "grad_alpha_bias" else 0.0)
        assert a.shape == b.shape, name
        assert float((a.double().cpu() - b).abs().max()) <= 2e-5 * scale, name


def test_sample_run_accumulates_like_autograd():
    """MSDA_FLAG_ACCUMULATE / StepSampleRun: three sampler calls on one value tensor and one set of reference points, their
    backward passes adding into one buffer, against the same three calls with autograd's own accumulation (2e-6 of max|g|:
    the two differ only in the order of fp32 additions)."""
    from dense_video_captioning_b200.ops.functions import MSDeformCapSampleFunction, StepSampleRun
    g = torch.Generator().manual_seed(11)
    N, S, M, D, Lq, P = 1, 375, 1, 64, 7, 4
    T = torch.tensor([200, 100, 50, 25], device="cuda")
    lsi = torch.cat((T.new_zeros(1), T.cumsum(0)[:-1]))
    res = {}
    for use_run in (False, True):
        value = torch.randn(N, S, M, D, generator=g.manual_seed(11)).cuda().requires_grad_(True)
        ref = torch.stack([torch.rand(N, Lq, 4, generator=g), torch.rand(N, Lq, 4, generator=g) * 0.3 + 0.05], -1).cuda().requires_grad_(True)
        run = StepSampleRun() if use_run else None
        total, offs = 0, []
        for step in range(3):
            off = (torch.randn(N, Lq, M, 4, P, generator=g) * 2).cuda().requires_grad_(True)
            offs.append(off)
            out = MSDeformCapSampleFunction.apply(value, T, lsi, ref, off, "border", run)
            total = total + (out * (step + 1)).sum() + total * 0.5     # later steps depend on earlier ones, as in the word loop
        total.backward()
        res[use_run] = [value.grad, ref.grad] + [o.grad for o in offs]
        if use_run:
            assert run.calls == 0 and run._left is None               # the run closed itself
    for a, b in zip(res[False], res[True]):
        assert float((a - b).abs().max()) <= 2e-6 * float(a.abs().max())
