"""Direct probe of the three projection GEMM entry points (fp32 / bf16): relative error of each output vs float64."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import dense_video_captioning_b200 as dvc
lib = dvc._lib.lib()
st = lambda: torch.cuda.current_stream().cuda_stream
def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())
for dt, code in ((torch.float32, 0), (torch.bfloat16, 1)):
    for rows, K, N0, N1 in ((375, 512, 512, 0), (100, 512, 128, 128), (37, 264, 72, 0)):
        torch.manual_seed(1)
        x = torch.randn(rows, K, device="cuda").to(dt)
        w0 = (torch.randn(N0, K, device="cuda") / K ** 0.5).to(dt)
        w1 = (torch.randn(N1, K, device="cuda") / K ** 0.5).to(dt) if N1 else None
        b0 = torch.randn(N0, device="cuda")
        b1 = torch.randn(N1, device="cuda") if N1 else None
        dy0 = torch.randn(rows, N0, device="cuda").to(dt)
        dy1 = torch.randn(rows, N1, device="cuda").to(dt) if N1 else None
        p = lambda t: t.data_ptr() if t is not None else None
        y0 = torch.full((rows, N0), 7.0, device="cuda", dtype=dt); y1 = torch.full((rows, N1), 7.0, device="cuda", dtype=dt) if N1 else None
        rc = lib.msda_proj_forward(p(x), p(w0), p(b0), p(w1), p(b1), None, p(y0), p(y1), rows, K, N0, N1, code, st())
        torch.cuda.synchronize()
        r = [rel(y0, x.double() @ w0.double().t() + b0.double())] + ([rel(y1, x.double() @ w1.double().t() + b1.double())] if N1 else [])
        print(dt, (rows, K, N0, N1), "fwd rc", rc, r)
        dx = torch.full((rows, K), 7.0, device="cuda", dtype=dt)
        rc = lib.msda_proj_dgrad(p(dy0), p(dy1), p(w0), p(w1), None, p(dx), rows, K, N0, N1, code, st())
        torch.cuda.synchronize()
        want = dy0.double() @ w0.double() + (dy1.double() @ w1.double() if N1 else 0)
        print("   dgrad rc", rc, rel(dx, want), float(dx.float().abs().max()))
        dw0 = torch.full((N0, K), 7.0, device="cuda"); db0 = torch.full((N0,), 7.0, device="cuda")
        dw1 = torch.full((N1, K), 7.0, device="cuda") if N1 else None; db1 = torch.full((N1,), 7.0, device="cuda") if N1 else None
        rc = lib.msda_proj_wgrad(p(dy0), p(dy1), p(x), None, p(dw0), p(db0), p(dw1), p(db1), rows, K, N0, N1, code, st())
        torch.cuda.synchronize()
        r = [rel(dw0, dy0.double().t() @ x.double()), rel(db0, dy0.double().sum(0))]
        if N1: r += [rel(dw1, dy1.double().t() @ x.double()), rel(db1, dy1.double().sum(0))]
        print("   wgrad rc", rc, r)
