"""Fused clip + Adam (csrc/msda_optim.cuh, pdvc/optim.FlatClipAdam) against what it replaces:
torch.nn.utils.clip_grad_norm_ followed by torch.optim.Adam(lr, weight_decay).step() (reference train.py:185-187, :123).

Reference arithmetic: the same two torch calls in float64 on the same parameters / gradients.  Tolerance: parameters after
5 steps within 2e-6 of max|p| (fp32 arithmetic against float64; the update is lr * O(1) per step), the reported gradient norm
within 2e-6 relative.  torch's own fp32 fused Adam is held to the same bar as a cross-check of the bar itself.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _buckets(sizes, seed):
    g = torch.Generator().manual_seed(seed)
    params = [torch.randn(n, generator=g).cuda().requires_grad_(True) for n in sizes]
    half = len(params) // 2
    buckets = []
    for group in (params[:half], params[half:]):
        flat = torch.zeros(sum(p.numel() for p in group), device="cuda")
        views, off = [], 0
        for p in group:
            views.append(flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        buckets.append((flat, group, views))
    return params, buckets


@pytest.mark.parametrize("max_norm,scale", [(100.0, 50.0), (100.0, 0.01), (0.0, 1.0)], ids=["clipping", "below_max", "no_clip"])
@pytest.mark.parametrize("weight_decay", [0.0, 1e-4])
def test_flat_clip_adam_matches_torch_float64(max_norm, scale, weight_decay):
    from dense_video_captioning_b200.pdvc.optim import FlatClipAdam
    sizes = [1, 7, 512, 8191, 8192, 8193, 100_003, 3, 512 * 512, 1607 * 512 + 5]
    params, buckets = _buckets(sizes, seed=1)
    lr = 5e-2                                    # large on purpose: differences in the update are not hidden by a tiny lr
    ours = FlatClipAdam(buckets, lr=lr, weight_decay=weight_decay, max_norm=max_norm, order=params)
    ref_p = [p.detach().double().clone().requires_grad_(True) for p in params]
    ref = torch.optim.Adam(ref_p, lr=lr, weight_decay=weight_decay)
    f32_p = [p.detach().clone().requires_grad_(True) for p in params]
    f32 = torch.optim.Adam(f32_p, lr=lr, weight_decay=weight_decay, fused=True)
    g = torch.Generator().manual_seed(2)
    for step in range(5):
        grads = [(torch.randn(p.numel(), generator=g) * scale * (1 + step)).cuda() for p in params]
        for (flat, group, views) in buckets:
            for p, v in zip(group, views):
                v.copy_(grads[[id(q) for q in params].index(id(p))].view_as(p))
        for q, q32, gr in zip(ref_p, f32_p, grads):
            q.grad = gr.double().view_as(q)
            q32.grad = gr.clone().view_as(q32)
        ours.step()
        want_norm = float(torch.sqrt(sum((gr.double() ** 2).sum() for gr in grads)))
        if max_norm > 0:
            torch.nn.utils.clip_grad_norm_(ref_p, max_norm)
            torch.nn.utils.clip_grad_norm_(f32_p, max_norm)
            assert abs(float(ours.grad_norm) - want_norm) <= 2e-6 * want_norm
        ref.step()
        f32.step()
    for i, (p, q, q32) in enumerate(zip(params, ref_p, f32_p)):
        scale_p = float(q.abs().max())
        err = float((p.detach().double() - q.detach()).abs().max())
        err32 = float((q32.detach().double() - q.detach()).abs().max())
        assert err <= 2e-6 * max(scale_p, 1.0), (i, sizes[i], err)
        assert err <= 4 * err32 + 1e-7 * max(scale_p, 1.0), (i, sizes[i], err, err32)       # no worse than torch's fp32 kernel


def test_flat_clip_adam_state_dict_round_trip_with_torch_adam():
    from dense_video_captioning_b200.pdvc.optim import FlatClipAdam
    sizes = [5, 300, 9000, 64]
    params, buckets = _buckets(sizes, seed=3)
    ours = FlatClipAdam(buckets, lr=1e-3, weight_decay=1e-4, max_norm=0.0, order=params)
    twin = [p.detach().clone().requires_grad_(True) for p in params]
    adam = torch.optim.Adam(twin, lr=1e-3, weight_decay=1e-4)
    g = torch.Generator().manual_seed(4)

    def both_step():
        for (flat, group, views) in buckets:
            flat.copy_(torch.randn(flat.numel(), generator=g).cuda())
        for p, q in zip(params, twin):
            b, off = ours._offsets[id(p)]
            q.grad = buckets[b][0][off:off + p.numel()].view_as(q).clone()
        ours.step()
        adam.step()

    for _ in range(3):
        both_step()
    sd = ours.state_dict()
    assert set(sd["state"].keys()) == set(range(len(params))) and float(sd["state"][0]["step"]) == 3.0
    # a torch Adam restored from OUR state continues exactly like ours, and ours restored from torch's continues like torch
    adam2 = torch.optim.Adam(twin, lr=1e-3, weight_decay=1e-4)
    adam2.load_state_dict(sd)
    ours.load_state_dict(adam.state_dict())
    adam = adam2
    for _ in range(2):
        both_step()
    for p, q in zip(params, twin):
        assert float((p - q).abs().max()) <= 2e-6 * max(float(q.abs().max()), 1.0)


def test_trainer_with_flat_optimizer_tracks_trainer_with_torch_adam():
    """Three steps of the graph trainer with the fused clip + Adam against the same trainer on clip_grad_norm_ + torch's Adam.
    Two runs of one step differ by the order of fp32 atomics in the backward; Adam turns a noise-level gradient into an update
    of +-lr, so parameters are compared to a few lr per step and the losses to 1e-3."""
    from dense_video_captioning_b200.pdvc import Trainer, synthetic_video
    out = {}
    for flag in (False, True):
        tr = Trainer("yc2_tsp_pdvc_yc2shape", "cuda:0", mode="graphs", flat_optimizer=flag, dropout=0.0)
        assert tr._flat_optimizer == flag
        video = synthetic_video(tr.cfg, tr.n_gt, tr.cap_len, seed=1)
        losses = [float(tr.step(video)) for _ in range(3)]
        out[flag] = (losses, {n: p.detach().clone() for n, p in tr.model.named_parameters() if p.requires_grad})
        del tr
    (l0, p0), (l1, p1) = out[False], out[True]
    assert l0[0] > l0[-1], "loss must decrease"
    for a, b in zip(l0, l1):
        assert abs(a - b) <= 1e-3 * abs(a), (l0, l1)
    lr, steps = 5e-5, 3
    for n in p0:
        assert float((p0[n] - p1[n]).abs().max()) <= 4 * lr * steps, n
