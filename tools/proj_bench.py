"""Microseconds per call, CUDA-graph replayed (20 calls per graph): the tcgen05 projection GEMMs vs torch's library calls."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import dense_video_captioning_b200 as dvc
lib = dvc._lib.lib()

def graph_us(fn, reps=20, iters=20):
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): fn()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): g.replay()
    e.record(); torch.cuda.synchronize()
    return 1e3 * s.elapsed_time(e) / (reps * iters)

p = lambda t: t.data_ptr() if t is not None else None
for dt, code in ((torch.float32, 0), (torch.bfloat16, 1)):
    for rows, K, N0, N1 in ((375, 512, 512, 0), (375, 512, 128, 128), (100, 512, 512, 0), (100, 512, 128, 128), (375, 768, 768, 0)):
        x = torch.randn(rows, K, device="cuda").to(dt); w0 = torch.randn(N0, K, device="cuda").to(dt) * 0.04
        w1 = torch.randn(N1, K, device="cuda").to(dt) * 0.04 if N1 else None
        b0 = torch.randn(N0, device="cuda"); b1 = torch.randn(N1, device="cuda") if N1 else None
        b0t, b1t = b0.to(dt), (b1.to(dt) if N1 else None)
        dy0 = torch.randn(rows, N0, device="cuda").to(dt); dy1 = torch.randn(rows, N1, device="cuda").to(dt) if N1 else None
        y0 = torch.empty(rows, N0, device="cuda", dtype=dt); y1 = torch.empty(rows, N1, device="cuda", dtype=dt) if N1 else None
        dx = torch.empty(rows, K, device="cuda", dtype=dt)
        dw0 = torch.empty(N0, K, device="cuda"); db0 = torch.empty(N0, device="cuda")
        dw1 = torch.empty(N1, K, device="cuda") if N1 else None; db1 = torch.empty(N1, device="cuda") if N1 else None
        st = lambda: torch.cuda.current_stream().cuda_stream
        ours = {
            "fwd": lambda: lib.msda_proj_forward(p(x), p(w0), p(b0), p(w1), p(b1), None, p(y0), p(y1), rows, K, N0, N1, code, st()),
            "dgrad": lambda: lib.msda_proj_dgrad(p(dy0), p(dy1), p(w0), p(w1), None, p(dx), rows, K, N0, N1, code, st()),
            "wgrad": lambda: lib.msda_proj_wgrad(p(dy0), p(dy1), p(x), None, p(dw0), p(db0), p(dw1), p(db1), rows, K, N0, N1, code, st()),
        }
        def t_fwd():
            F.linear(x, w0, b0t)
            if N1: F.linear(x, w1, b1t)
        def t_dgrad():
            r = dy0 @ w0
            if N1: r = r + dy1 @ w1
        def t_wgrad():
            dy0.t() @ x; dy0.sum(0)
            if N1: dy1.t() @ x; dy1.sum(0)
        ref = {"fwd": t_fwd, "dgrad": t_dgrad, "wgrad": t_wgrad}
        row = " ".join(f"{k}: ours {graph_us(ours[k]):5.2f} torch {graph_us(ref[k]):5.2f} us |" for k in ours)
        print(f"{str(dt)[6:]:9s} rows={rows} K={K} N={N0}+{N1}  {row}")
