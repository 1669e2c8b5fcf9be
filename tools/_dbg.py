import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
import cases, test_parity_gpu as tp
import dense_video_captioning_b200 as dvc
cfg = dict(seed=31, N=2, M=8, D=64, T0=300, P=4, jitter=1.0)
inp = tp._nudge_off_knots(tp._encoder_inputs(**cfg), 4)
dev = "cuda"
value = torch.from_numpy(inp["value"]).float().to(dev).requires_grad_(True)
ref = torch.from_numpy(inp["ref"]).float().to(dev)
off = torch.from_numpy(inp["off"]).float().to(dev).requires_grad_(True)
logits = torch.from_numpy(inp["logits"]).float().to(dev).requires_grad_(True)
gout = torch.from_numpy(inp["grad_out"]).float().to(dev)
T = torch.from_numpy(inp["T"]).to(dev)
lsi = torch.from_numpy(cases.level_start_index(cases.as_shapes2(inp["T"]))).to(dev)
pad = sys.argv[1]
out = dvc.MSDeformAttnFusedFunction.apply(value, T, lsi, ref, off, logits, pad)
torch.cuda.synchronize(); print("fwd ok", float(out.abs().max()))
out.backward(gout)
torch.cuda.synchronize(); print("bwd ok", float(value.grad.abs().max()))
