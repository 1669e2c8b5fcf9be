"""Phase clocks (clock64 deltas of CTA (0,0,0), thread 0) of one projection GEMM launch; needs -DMSDA_PROJ_TIMING."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
dbg = torch.zeros(16, dtype=torch.int64, device="cuda")
os.environ["MSDA_PROJ_DEBUG_PTR"] = str(dbg.data_ptr())
import dense_video_captioning_b200 as dvc
lib = dvc._lib.lib()
p = lambda t: t.data_ptr() if t is not None else None
names = ["setup", "loop", "done-wait+tmem-ld", "cluster sync 1", "dsmem + sync 2", "reduce+store"]
for dt, code in ((torch.float32, 0), (torch.bfloat16, 1)):
    rows, K, N0 = 375, 512, 512
    x = torch.randn(rows, K, device="cuda").to(dt); w0 = torch.randn(N0, K, device="cuda").to(dt); b0 = torch.randn(N0, device="cuda")
    y0 = torch.empty(rows, N0, device="cuda", dtype=dt); dy0 = torch.randn(rows, N0, device="cuda").to(dt)
    dw0 = torch.empty(N0, K, device="cuda"); db0 = torch.empty(N0, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for mode in ("fwd", "wgrad"):
        for _ in range(3):
            if mode == "fwd":
                lib.msda_proj_forward(p(x), p(w0), p(b0), None, None, None, p(y0), None, rows, K, N0, 0, code, st)
            else:
                lib.msda_proj_wgrad(p(dy0), None, p(x), None, p(dw0), p(db0), None, None, rows, K, N0, 0, code, st)
            torch.cuda.synchronize()
        t = dbg.cpu().tolist()
        print(dt, mode, {n: t[i + 1] - t[i] for i, n in enumerate(names)}, "total", t[6] - t[0])
