"""GPU parity of the tcgen05 projection GEMMs (csrc/msda_proj.cuh) against float64 torch.

What is replaced: the four nn.Linear layers of MSDeformAttn (pdvc/ops/modules/ms_deform_attn.py:95-100, :125) -- value_proj with
the padding-mask zeroing of :97-98, sampling_offsets and attention_weights as one GEMM, output_proj -- forward, data gradient,
weight gradient and bias gradient.

Stated tolerances (max |err| relative to max |reference|, reference = float64 F.linear on the operands the kernel sees):
  fp32 operands (3xTF32 split products, fp32 accumulation in TMEM):  2e-6  -- fp32 level; plain TF32 would be ~5e-4
  bf16 operands (bf16 x bf16 -> fp32 accumulate, bf16 outputs):      6e-3  (one bf16 rounding of the output)
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TOL = {torch.float32: 2e-6, torch.bfloat16: 6e-3}


@pytest.fixture(scope="module")
def dvc():
    import dense_video_captioning_b200 as pkg
    assert torch.cuda.is_available()
    return pkg


def _rel(got, want):
    want = want.double()
    return float((got.double() - want).abs().max() / want.abs().max().clamp_min(1e-30))


def _case(rows, K, N0, N1, masked, dtype, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    lead = (2, rows // 2) if rows % 2 == 0 and rows > 2 else (rows,)
    x = torch.randn(*lead, K, generator=g)
    w0 = torch.randn(N0, K, generator=g) / K ** 0.5
    b0 = torch.randn(N0, generator=g)
    w1 = torch.randn(N1, K, generator=g) / K ** 0.5 if N1 else None
    b1 = torch.randn(N1, generator=g) if N1 else None
    mask = (torch.rand(*lead, generator=g) < 0.2) if masked else None
    g0 = torch.randn(*lead, N0, generator=g)
    g1 = torch.randn(*lead, N1, generator=g) if N1 else None
    rd = (lambda t: t.to(dtype).float()) if dtype == torch.bfloat16 else (lambda t: t)   # operands as the kernel sees them
    return dict(x=rd(x), w0=rd(w0), b0=b0, w1=rd(w1) if N1 else None, b1=b1, mask=mask, g0=rd(g0), g1=rd(g1) if N1 else None)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,K,N0,N1,masked", [
    (375, 512, 512, 0, True),      # value_proj, YouCook2 encoder memory, padding mask
    (375, 512, 128, 128, False),   # sampling_offsets || attention_weights on the encoder queries (M*L*P = 128 each)
    (100, 512, 128, 128, False),   # ... on the decoder queries
    (100, 512, 512, 0, False),     # output_proj, decoder
    (188, 512, 512, 0, True),      # ActivityNet
    (10, 512, 128, 128, False),    # ActivityNet decoder: 10 queries
    (1, 64, 8, 0, False),          # smallest legal problem
    (37, 264, 72, 0, True),        # nothing a multiple of the tile sizes
    (300, 768, 384, 0, True),      # 768-d model (configs[3]), 3 M-tiles, 6 N-tiles
    (129, 1024, 256, 256, True),   # two segments, long reduction, rows just over one tile
], ids=["value375", "offlog375", "offlog100", "out100", "anet188", "anet10", "tiny", "ragged", "d768", "seg129"])
def test_proj_linear_vs_float64(dvc, rows, K, N0, N1, masked, dtype):
    from dense_video_captioning_b200.ops.functions import ProjLinearFunction
    c = _case(rows, K, N0, N1, masked, dtype, seed=rows + K + N0)
    dev = "cuda"
    x = c["x"].to(dtype).to(dev).requires_grad_(True)
    w0 = c["w0"].to(dev).requires_grad_(True)          # master weights stay fp32, as nn.Linear parameters are
    b0 = c["b0"].to(dev).requires_grad_(True)
    w1 = c["w1"].to(dev).requires_grad_(True) if N1 else None
    b1 = c["b1"].to(dev).requires_grad_(True) if N1 else None
    mask = c["mask"].to(dev) if masked else None
    assert ProjLinearFunction.supported(x, w0, w1)
    ctx = torch.autocast("cuda", dtype=torch.bfloat16) if dtype == torch.bfloat16 else torch.autocast("cuda", enabled=False)
    with ctx:
        out = ProjLinearFunction.apply(x, w0, b0, w1, b1, mask)
    assert dvc._lib.lib().msda_last_launch_path() == dvc._lib.PATH_PROJ
    outs = out if N1 else (out,)
    gs = [c["g0"].to(dtype).to(dev)] + ([c["g1"].to(dtype).to(dev)] if N1 else [])
    torch.autograd.backward(outs, gs)
    torch.cuda.synchronize()

    # float64 reference on the same operands
    X = c["x"].double().requires_grad_(True)
    W0, B0 = c["w0"].double().requires_grad_(True), c["b0"].double().requires_grad_(True)
    ref0 = F.linear(X, W0, B0)
    if masked:
        ref0 = ref0.masked_fill(c["mask"][..., None], 0.0)
    refs, params = [ref0], [W0, B0]
    if N1:
        W1, B1 = c["w1"].double().requires_grad_(True), c["b1"].double().requires_grad_(True)
        refs.append(F.linear(X, W1, B1))
        params += [W1, B1]
    torch.autograd.backward(refs, [c["g0"].double()] + ([c["g1"].double()] if N1 else []))

    tol = TOL[dtype]
    for i, (o, r) in enumerate(zip(outs, refs)):
        assert o.dtype == dtype and o.shape == r.shape
        assert _rel(o.cpu(), r.detach()) <= tol, f"y{i}: {_rel(o.cpu(), r.detach()):.2e}"
    assert _rel(x.grad.cpu(), X.grad) <= tol, f"dx: {_rel(x.grad.cpu(), X.grad):.2e}"
    got = [w0.grad, b0.grad] + ([w1.grad, b1.grad] if N1 else [])
    for name, gt, p in zip(("dw0", "db0", "dw1", "db1"), got, params):
        assert gt.dtype == torch.float32
        # weight / bias gradients leave the kernel in fp32 in both modes (fp32 accumulator, no output rounding)
        assert _rel(gt.cpu(), p.grad) <= (2e-6 if dtype == torch.float32 else 1e-5), f"{name}: {_rel(gt.cpu(), p.grad):.2e}"


def test_proj_accuracy_is_fp32_level_not_tf32(dvc):
    """The 3xTF32 split against what the same products give in plain TF32 (torch's allow_tf32 matmul) and in cuBLAS fp32:
    the hand-written kernel must sit with fp32, orders of magnitude below TF32."""
    from dense_video_captioning_b200.ops.functions import ProjLinearFunction
    g = torch.Generator(device="cpu").manual_seed(5)
    x = torch.randn(375, 512, generator=g).cuda()
    w = (torch.randn(512, 512, generator=g) / 512 ** 0.5).cuda()
    ref = x.double() @ w.double().t()
    ours = ProjLinearFunction.apply(x, w, None, None, None, None)
    fp32 = x @ w.t()
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        tf32 = x @ w.t()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    e_ours, e_fp32, e_tf32 = _rel(ours, ref), _rel(fp32, ref), _rel(tf32, ref)
    print(f"relative error vs float64: ours (3xTF32) {e_ours:.2e}, cuBLAS fp32 {e_fp32:.2e}, cuBLAS TF32 {e_tf32:.2e}")
    assert e_ours <= 2e-6
    assert e_ours <= 4 * e_fp32 + 1e-7
    assert e_tf32 > 20 * e_ours


def test_module_with_and_without_proj_gemm_agree(dvc):
    """MSDeformAttn end to end (forward + every gradient) with the tcgen05 projections against the same module on torch's
    library GEMMs; padding mask and both reference-point branches."""
    from dense_video_captioning_b200.ops.modules import ms_deform_attn as M
    torch.manual_seed(0)
    T = [200, 100, 50, 25]
    S = sum(T)
    mod = M.MSDeformAttn(512, 4, 8, 4).cuda()
    with torch.no_grad():
        mod.sampling_offsets.weight.normal_(0, 0.02)
        mod.attention_weights.weight.normal_(0, 0.05)
    shapes = torch.tensor(T, device="cuda")
    lsi = torch.cat((shapes.new_zeros(1), shapes.cumsum(0)[:-1]))
    res = {}
    for R in (1, 2):
        query = torch.randn(2, 100, 512, device="cuda")
        src = torch.randn(2, S, 512, device="cuda")
        ref = torch.rand(2, 100, 4, R, device="cuda") * (0.9 if R == 1 else 0.4) + 0.05
        mask = torch.zeros(2, S, dtype=torch.bool, device="cuda")
        mask[1, 150:200] = True
        gout = torch.randn(2, 100, 512, device="cuda")
        for flag in (True, False):
            M.USE_PROJ_GEMM = flag
            q, s_ = query.clone().requires_grad_(True), src.clone().requires_grad_(True)
            mod.zero_grad(set_to_none=True)
            out = mod(q, ref, s_, shapes, lsi, mask)
            out.backward(gout)
            res[flag] = [out.detach(), q.grad, s_.grad] + [p.grad.clone() for p in mod.parameters()]
        M.USE_PROJ_GEMM = True
        names = ["out", "dquery", "dsrc"] + [n for n, _ in mod.named_parameters()]
        for n, a, b in zip(names, res[True], res[False]):
            # grad of the offsets path is piecewise constant in the locations: both runs see bit-different offsets near knots
            tol = 2e-3 if ("sampling_offsets" in n or n in ("dquery",)) else 2e-5
            assert _rel(a, b) <= tol, f"R={R} {n}: {_rel(a, b):.2e}"


def test_proj_unsupported_shapes_fall_back(dvc):
    from dense_video_captioning_b200.ops.functions import ProjLinearFunction
    from dense_video_captioning_b200.ops.modules import ms_deform_attn as M
    x = torch.randn(5, 36, device="cuda")                 # K not a multiple of 8
    lin = torch.nn.Linear(36, 16).cuda()
    assert not ProjLinearFunction.supported(x, lin.weight)
    assert torch.equal(M._proj(x, lin), lin(x))
    lib = dvc._lib.lib()
    p = x.data_ptr()
    assert lib.msda_proj_forward(p, p, None, None, None, None, p, None, 5, 36, 16, 0, 0, None) == -7
