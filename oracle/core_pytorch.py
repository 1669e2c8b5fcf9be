"""oracle/core_pytorch.py -- TEST INFRASTRUCTURE ONLY (checker / timed CPU baseline).

A torch restatement of the reference's CPU path for the hot path, executing the same ATen
operators the reference executes (F.grid_sample per level, stack, weight, sum), so that it can
stand in for the reference on the GPU box where /root/reference does not exist:

  * core()           <- ms_deform_attn_core_pytorch, pdvc/ops/functions/ms_deform_attn_func.py:41-68
  * module_forward() <- MSDeformAttn.forward,        pdvc/ops/modules/ms_deform_attn.py:79-126
                        (and MSDeformAttnCap.forward, ms_deform_attn_for_caption.py:78-123,
                        with return_value=True)

tests/test_oracle.py checks core() against the committed golden fixtures (tests/golden/), which
come from the unmodified reference function run in this container.  bench.py times core()
forward + backward as the `cpu_baseline` / `--impl reference` arm ("kind": "port").
The product package never imports this file.
"""
import torch
import torch.nn.functional as F


def core(value, spatial_shapes, sampling_locations, attention_weights, return_value=False,
         padding_mode="border"):
    """value (N,S,M,D); spatial_shapes iterable of (H,W); sampling_locations (N,Lq,M,L,P,2) in
    [0,1] nominal; attention_weights (N,Lq,M,L,P).  Returns (N,Lq,M*D), or the raw samples
    (N*M, D, Lq, L, P) when return_value is set (func.py:64-65)."""
    n, _, m, d = value.shape
    lq, n_lvl, n_pts = sampling_locations.shape[1], sampling_locations.shape[3], sampling_locations.shape[4]
    hw = [(int(h), int(w)) for h, w in spatial_shapes]
    per_level = value.split([h * w for h, w in hw], dim=1)            # func.py:48
    grids = sampling_locations * 2 - 1                                 # func.py:49  [0,1] -> [-1,1]
    sampled = []
    for lvl, (h, w) in enumerate(hw):
        # (N, H*W, M, D) -> (N*M, D, H, W)                               func.py:53
        img = per_level[lvl].reshape(n, h * w, m * d).permute(0, 2, 1).reshape(n * m, d, h, w)
        # (N, Lq, M, P, 2) -> (N*M, Lq, P, 2)                            func.py:55
        grid = grids[:, :, :, lvl].permute(0, 2, 1, 3, 4).reshape(n * m, lq, n_pts, 2)
        sampled.append(F.grid_sample(img, grid, mode="bilinear", padding_mode=padding_mode,
                                     align_corners=False))             # func.py:58-59
    stacked = torch.stack(sampled, dim=-2)                             # (N*M, D, Lq, L, P)
    if return_value:
        return stacked
    w_attn = attention_weights.permute(0, 2, 1, 3, 4).reshape(n * m, 1, lq, n_lvl * n_pts)  # func.py:62
    out = (stacked.flatten(-2) * w_attn).sum(-1).view(n, m * d, lq)   # func.py:67
    return out.transpose(1, 2).contiguous()                            # func.py:68


def module_forward(params, query, reference_points, input_flatten, temporal_shapes,
                   padding_mask=None, n_heads=8, n_levels=4, n_points=4, caption=False,
                   padding_mode="border"):
    """Whole-module restatement (ms_deform_attn.py:95-125).  `params` is a state_dict-like
    mapping with the reference's parameter names (sampling_offsets.*, attention_weights.*,
    value_proj.*, output_proj.*).  temporal_shapes: int64 tensor (L,)."""
    n, lq, _ = query.shape
    _, s, c = input_flatten.shape
    value = F.linear(input_flatten, params["value_proj.weight"], params["value_proj.bias"])
    if padding_mask is not None:
        value = value.masked_fill(padding_mask[..., None], 0.0)
    value = value.view(n, s, n_heads, c // n_heads)
    off = F.linear(query, params["sampling_offsets.weight"], params["sampling_offsets.bias"])
    off = off.view(n, lq, n_heads, n_levels, n_points)
    attn = F.linear(query, params["attention_weights.weight"], params["attention_weights.bias"])
    attn = F.softmax(attn.view(n, lq, n_heads, n_levels * n_points), -1)
    attn = attn.view(n, lq, n_heads, n_levels, n_points)
    ref0 = reference_points[:, :, None, :, None, 0]
    if reference_points.shape[-1] == 1:
        x = ref0 + off / temporal_shapes[None, None, None, :, None]
    elif reference_points.shape[-1] == 2:
        x = ref0 + off / n_points * reference_points[:, :, None, :, None, 1] * 0.5
    else:
        raise ValueError("Last dim of reference_points must be 1 or 2")
    loc = torch.stack((x, 0.5 * torch.ones_like(x)), -1)               # ms_deform_attn.py:114-116
    shapes = [(1, int(t)) for t in temporal_shapes]
    if caption:
        return core(value, shapes, loc, attn, return_value=True, padding_mode=padding_mode)
    out = core(value, shapes, loc, attn, padding_mode=padding_mode)
    return F.linear(out, params["output_proj.weight"], params["output_proj.bias"])
