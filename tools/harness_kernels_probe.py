#!/usr/bin/env python
"""One call each of the training-harness kernels of round 2 on the shapes of the YouCook2-shaped PDVC step (for ncu):
msda_pdvc_match_f32 (2 layers x 100 queries x 6 events), msda_lsap_f32 (4 x 100 x 30), msda_adam_clip_step (18.2 M
parameters in two buckets)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dense_video_captioning_b200.ops.functions import linear_sum_assignment_cuda, pdvc_match_cuda  # noqa: E402
from dense_video_captioning_b200.pdvc.optim import FlatClipAdam  # noqa: E402

g = torch.Generator().manual_seed(0)
logits = torch.randn(2, 100, 1, generator=g).cuda()
boxes = torch.stack([torch.rand(2, 100, generator=g), torch.rand(2, 100, generator=g) * 0.4 + 0.02], -1).cuda()
labels = torch.zeros(6, dtype=torch.long, device="cuda")
tboxes = torch.stack([torch.rand(6, generator=g), torch.rand(6, generator=g) * 0.3 + 0.05], -1).cuda()
costs = torch.randn(4, 100, 30, generator=g).cuda()
sizes = [512 * 512] * 40 + [2048 * 512] * 6 + [1608 * 512, 512, 2048, 7]
params = [torch.randn(n, device="cuda").requires_grad_(True) for n in sizes]
buckets = []
for group in (params[:10], params[10:]):
    flat = torch.randn(sum(p.numel() for p in group), device="cuda")
    views, off = [], 0
    for p in group:
        views.append(flat[off:off + p.numel()].view_as(p))
        off += p.numel()
    buckets.append((flat, group, views))
opt = FlatClipAdam(buckets, lr=5e-5, weight_decay=1e-4, max_norm=100.0)
for _ in range(3):
    pdvc_match_cuda(logits, boxes, labels, tboxes, 2.0, 0.0, 4.0, 0.25, 2.0)
    linear_sum_assignment_cuda(costs)
    opt.step()
torch.cuda.synchronize()
print("parameters:", sum(sizes), "bytes moved by the update (7 x 4 B each):", 28 * sum(sizes))
