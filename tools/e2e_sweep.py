import sys, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__)))); sys.path.insert(0, __import__("os").path.join(__import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))), "tests"))
import bench
w = bench.make_workload(64, 1000, 8, 64, 4, 4, "encoder", "cuda:0", seed=100)
for chunk in (2, 4, 8, 16, 32, 64):
    ms, h2d, d2h = bench.time_e2e(w, 10, 3, 1, chunk=chunk)
    print(chunk, round(ms / 10, 2), "ms/step", round((h2d + d2h) / (ms / 10) / 1e6, 1), "GB/s both ways")
# pure copies for reference
hin = w["value"].cpu().pin_memory(); dv = torch.empty_like(w["value"]); hout = torch.empty_like(hin).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for name, both in (("h2d only", False), ("h2d + d2h concurrent", True)):
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    for _ in range(5):
        with torch.cuda.stream(s1): dv.copy_(hin, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): hout.copy_(w["value"], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(name, round(ms, 2), "ms per 246 MB", round(hin.numel() * 4 / ms / 1e6, 1), "GB/s per direction")
