"""CPU tests: the oracle (plain-C restatement and the torch port) against the golden vectors
produced by the unmodified reference (tests/golden/make_golden.py).  No GPU, no /root/reference."""
import numpy as np
import pytest
import torch

import cases
from conftest import golden_tol, load_golden
from oracle import core_pytorch
from oracle import msda_oracle as O

ZERO_CASES = ["yc2_dec", "edges", "wide_range", "ragged_2d", "p8_d32"]
SAMPLE_CASES = ["cap_m1_d512", "edges", "p8_d32"]


def _close(got, want, what):
    want64 = np.asarray(want, dtype=np.float64)
    tol = golden_tol(want)
    scale = max(np.abs(want64).max(), 1e-30)
    err = np.abs(np.asarray(got, dtype=np.float64) - want64).max()
    assert err <= tol * scale + 1e-300, f"{what}: max err {err:.3e} vs scale {scale:.3e}"


def _check_inputs(inp, g):
    cs = cases.checksum({k: inp[k] for k in ("value", "loc", "attn", "grad_out")})
    assert abs(cs - float(g["input_checksum"])) <= 1e-9 * max(1.0, abs(cs)), "seeded inputs drifted from the fixtures"


@pytest.mark.parametrize("name", cases.ALL_CORE_CASES)
def test_c_oracle_matches_reference_border(name):
    inp = cases.core_inputs(name)
    g = load_golden("core_border_" + name)
    _check_inputs(inp, g)
    out = O.forward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], O.BORDER)
    gv, gl, ga = O.backward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], inp["grad_out"], O.BORDER)
    _close(out, g["out"], "out")
    _close(gv, g["grad_value"], "grad_value")
    _close(gl, g["grad_loc"], "grad_loc")
    _close(ga, g["grad_attn"], "grad_attn")
    if "out__sum" in g:  # big tensors were stored as fp32: check the fp64 checksums too
        assert abs(out.sum() - float(g["out__sum"])) <= 1e-9 * float(g["out__abssum"])
        assert abs(gv.sum() - float(g["grad_value__sum"])) <= 1e-9 * float(g["grad_value__abssum"])


@pytest.mark.parametrize("name", ZERO_CASES)
def test_c_oracle_matches_grid_sample_zeros(name):
    inp = cases.core_inputs(name)
    g = load_golden("core_zeros_" + name)
    out = O.forward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], O.ZEROS)
    gv, gl, ga = O.backward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], inp["grad_out"], O.ZEROS)
    _close(out, g["out"], "out")
    _close(gv, g["grad_value"], "grad_value")
    _close(gl, g["grad_loc"], "grad_loc")
    _close(ga, g["grad_attn"], "grad_attn")


@pytest.mark.parametrize("name", SAMPLE_CASES)
def test_c_oracle_return_value_branch(name):
    inp = cases.core_inputs(name)
    g = load_golden("core_border_" + name)
    _, samples = O.forward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], O.BORDER, return_samples=True)
    _close(samples, g["samples"], "samples")
    gs = np.random.RandomState(1234).standard_normal(samples.shape)
    gv, gl = O.sample_backward(inp["value"], inp["shapes"], inp["loc"], gs, O.BORDER)
    _close(gv, g["sgrad_value"], "sgrad_value")
    _close(gl, g["sgrad_loc"], "sgrad_loc")


def test_border_grad_y_is_zero_in_1d():
    inp = cases.core_inputs("yc2_dec")
    _, gl, _ = O.backward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], inp["grad_out"], O.BORDER)
    assert np.all(gl[..., 1] == 0.0)


def test_f32_oracle_close_to_f64():
    inp = cases.core_inputs("p8_d32")
    o64 = O.forward(inp["value"], inp["shapes"], inp["loc"], inp["attn"])
    o32 = O.forward(inp["value"], inp["shapes"], inp["loc"], inp["attn"], dtype=np.float32)
    assert np.abs(o32 - o64).max() <= 2e-5 * np.abs(o64).max()


@pytest.mark.parametrize("name", ["yc2_dec", "edges", "ragged_2d", "odd_d71"])
def test_torch_port_matches_reference(name):
    inp = cases.core_inputs(name)
    g = load_golden("core_border_" + name)
    t = {k: torch.from_numpy(np.ascontiguousarray(inp[k])).double() for k in ("value", "loc", "attn", "grad_out")}
    for k in ("value", "loc", "attn"):
        t[k].requires_grad_(True)
    out = core_pytorch.core(t["value"], inp["shapes"], t["loc"], t["attn"])
    out.backward(t["grad_out"])
    _close(out.detach().numpy(), g["out"], "out")
    _close(t["value"].grad.numpy(), g["grad_value"], "grad_value")
    _close(t["loc"].grad.numpy(), g["grad_loc"], "grad_loc")
    _close(t["attn"].grad.numpy(), g["grad_attn"], "grad_attn")


@pytest.mark.parametrize("tag", ["attn_ref1_nomask", "attn_ref1_mask", "attn_ref2_nomask", "attn_ref2_mask",
                                 "cap_ref1_mask", "cap_ref2_nomask"])
def test_torch_port_module_matches_reference(tag):
    g = load_golden("module_" + tag)
    params = {k[len("param."):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("param.")}
    mask = torch.from_numpy(g["mask"]) if g["mask"].size else None
    out = core_pytorch.module_forward(params, torch.from_numpy(g["query"]), torch.from_numpy(g["ref"]),
                                      torch.from_numpy(g["src"]), torch.from_numpy(g["T"]), mask,
                                      n_heads=int(g["n_heads"]), n_levels=int(g["n_levels"]),
                                      n_points=int(g["n_points"]), caption=tag.startswith("cap"))
    _close(out.numpy(), g["out"], "module out")


@pytest.mark.parametrize("tag", ["attn_ref1_mask", "attn_ref2_nomask"])
def test_fused_numpy_front_end_matches_reference_module(tag):
    """softmax + sampling-location arithmetic + chain rule (oracle.msda_oracle.fused_*) against the
    reference module's autograd, through the projections recomputed in numpy."""
    g = load_golden("module_" + tag)
    P = {k[len("param."):]: v for k, v in g.items() if k.startswith("param.")}
    N, Lq, C = g["query"].shape
    M, L, Pn = int(g["n_heads"]), int(g["n_levels"]), int(g["n_points"])
    S = g["src"].shape[1]
    value = g["src"] @ P["value_proj.weight"].T + P["value_proj.bias"]
    if g["mask"].size:
        value = np.where(g["mask"][..., None], 0.0, value)
    value = value.reshape(N, S, M, C // M)
    off = (g["query"] @ P["sampling_offsets.weight"].T + P["sampling_offsets.bias"]).reshape(N, Lq, M, L, Pn)
    logits = (g["query"] @ P["attention_weights.weight"].T + P["attention_weights.bias"]).reshape(N, Lq, M, L, Pn)
    core_out = O.fused_forward(value, g["T"], g["ref"], off, logits)
    out = core_out @ P["output_proj.weight"].T + P["output_proj.bias"]
    _close(out, g["out"], "module out via fused numpy front-end")
    g_core = g["grad_out"] @ P["output_proj.weight"]
    gv, g_ref, g_off, g_logits = O.fused_backward(value, g["T"], g["ref"], off, logits, g_core)
    _close(g_ref, g["grad_ref"], "grad_ref")
    g_query = g_off.reshape(N, Lq, -1) @ P["sampling_offsets.weight"] + g_logits.reshape(N, Lq, -1) @ P["attention_weights.weight"]
    _close(g_query, g["grad_query"], "grad_query")
    gv = gv.reshape(N, S, C)
    if g["mask"].size:
        gv = np.where(g["mask"][..., None], 0.0, gv)
    _close(gv @ P["value_proj.weight"], g["grad_src"], "grad_src")
