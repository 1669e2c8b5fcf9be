#!/usr/bin/env python
"""Whole-model golden fixture: runs the UNMODIFIED reference PDVC (pdvc/pdvc.py build(), criterion, matcher, LSTM-DSA
caption head, its own MSDeformAttn / MSDeformAttnCap on the grid_sample CPU path) in THIS container, in float64, on a
seeded synthetic video, and records every loss term and the gradient norm / checksum of every parameter.

Build-container only (needs /root/reference); stubs as in SURVEY Appendix C.  Weights are not stored: both sides fill
them with tests/cases.fill_parameters().  Dropout is 0 so that the training-mode forward is deterministic.

    python tests/golden/make_golden_pdvc.py        -> tests/golden/pdvc_anet_c3d.npz
"""
import collections
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402

REF = "/root/reference"


def load_reference(cfg_path, overrides):
    for name in ("MultiScaleDeformableAttention", "colorlog", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.path.insert(0, REF)
    os.chdir(REF)
    sys.argv = ["x", "--cfg_path", cfg_path, "--device", "cpu"]
    import opts
    opt = opts.parse_opts()
    for k, v in overrides.items():
        setattr(opt, k, v)
    from pdvc.pdvc import build
    model, criterion, _ = build(opt)
    return opt, model, criterion


def main():
    torch.manual_seed(0)
    torch.set_default_dtype(torch.float64)   # the reference creates helper tensors with the default dtype
    opt, model, criterion = load_reference("cfgs/anet_c3d_pdvc.yml",
                                           dict(transformer_dropout_prob=0.0, drop_prob=0.0, vocab_size=300))
    model.double().train()
    criterion.double()
    cases.fill_parameters(model)
    b = cases.pdvc_batch(5, opt.frame_embedding_num, opt.feature_dim, n_gt=4, cap_len=9, vocab_size=opt.vocab_size)
    t = {k: torch.from_numpy(v) for k, v in b.items()}
    dt = collections.defaultdict(lambda: None, {
        "video_tensor": t["video_tensor"], "video_mask": t["video_mask"], "video_length": t["video_length"],
        "video_target": [{"boxes": t["boxes"], "labels": t["labels"]}], "cap_tensor": t["cap_tensor"],
        "cap_mask": t["cap_mask"], "gt_boxes": t["boxes"][None], "gt_boxes_mask": torch.ones(1, 4, dtype=torch.bool)})
    out, loss = model(dt, criterion, "queries")
    wd = criterion.weight_dict
    total = sum(loss[k] * wd[k] for k in loss if k in wd)
    total.backward()
    rec = {"loss_names": np.array(sorted(loss.keys())), "total_loss": np.float64(total.item()),
           "loss_values": np.array([float(loss[k]) for k in sorted(loss.keys())]),
           "weight_names": np.array(sorted(wd.keys())), "weight_values": np.array([float(wd[k]) for k in sorted(wd.keys())]),
           "pred_boxes": out["pred_boxes"].detach().numpy(), "pred_logits": out["pred_logits"].detach().numpy(),
           "pred_count": out["pred_count"].detach().numpy(),
           "matched_query": out["matched_indices"][0][0][0].numpy(), "matched_target": out["matched_indices"][0][0][1].numpy(),
           "torch_version": np.array(torch.__version__), "vocab_size": np.int64(opt.vocab_size)}
    names, norms, sums = [], [], []
    seen = set()
    for key, p in model.named_parameters():
        name = cases.canonical_param_name(key)
        if name in seen or p.grad is None:
            continue
        seen.add(name)
        names.append(name)
        norms.append(float(p.grad.norm()))
        sums.append(float((p.grad * torch.arange(1, p.numel() + 1, dtype=p.dtype).view_as(p).remainder(7).add(1)).sum()))
    rec.update(grad_names=np.array(names), grad_norms=np.array(norms), grad_checksums=np.array(sums))
    path = os.path.join(HERE, "pdvc_anet_c3d.npz")
    np.savez_compressed(path, **rec)
    print("wrote", path, "total loss", total.item(), {k: round(float(v), 6) for k, v in loss.items()})
    print(len(names), "parameters with gradients")


if __name__ == "__main__":
    main()
