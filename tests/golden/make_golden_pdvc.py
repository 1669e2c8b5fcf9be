#!/usr/bin/env python
"""Whole-model golden fixture: runs the UNMODIFIED reference PDVC (pdvc/pdvc.py build(), criterion, matcher, LSTM-DSA
caption head, its own MSDeformAttn / MSDeformAttnCap on the grid_sample CPU path) in THIS container, in float64, on a
seeded synthetic video, and records every loss term and the gradient norm / checksum of every parameter.

Build-container only (needs /root/reference); stubs as in SURVEY Appendix C.  Weights are not stored: both sides fill
them with tests/cases.fill_parameters().  Dropout is 0 so that the training-mode forward is deterministic.

    python tests/golden/make_golden_pdvc.py [name]   -> tests/golden/pdvc_<name>.npz   (one name per process:
                                                        the reference parses its options from sys.argv once)
        anet_c3d       cfgs/anet_c3d_pdvc.yml (10 queries, T = 100, 2+2 layers)              BASELINE configs[2]
        yc2shape       cfgs/yc2_tsp_pdvc.yml + the YouCook2 shapes of cfgs/yc2_tsn_pdvcl.yml:26,31
                       (100 queries, T = 200, vocab 1607, max_eseq_length 20)                 BASELINE configs[1](b)
        newmodel_core  cfgs/yc2_newModel_sound.yml:8-19 (768-d features, 3+3 layers, 100 queries, T = 200): the PDVC
                       core of the MViT+HuBERT model                                          BASELINE configs[3]
        anet_c3d_refcuda   anet_c3d with the padding semantics of the reference's GPU build: MSDeformAttn (encoder,
                       decoder) samples with ZERO padding, as MSDeformAttnFunction's CUDA kernel does
                       (ms_deform_im2col_cuda.cuh:286-289; module call site ms_deform_attn.py:120-122), while the caption
                       head's MSDeformAttnCap stays on grid_sample / border (ms_deform_attn_for_caption.py:120-121 calls
                       the PyTorch core unconditionally).  The reference's extension is not built here, so the module's
                       core call is redirected to the oracle's grid_sample restatement with padding_mode="zeros" -- the
                       only patched line; everything else is the unmodified reference.
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


def load_reference(cfg_path, overrides, zero_padded_attn=False):
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
    if zero_padded_attn:
        sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
        from oracle import core_pytorch
        import pdvc.ops.modules.ms_deform_attn as ref_mod

        def zero_padded_core(value, shapes, loc, attn, return_value=False):
            return core_pytorch.core(value, shapes, loc, attn, return_value=return_value, padding_mode="zeros")
        ref_mod.ms_deform_attn_core_pytorch = zero_padded_core      # MSDeformAttn only; the caption module keeps its own import
    model, criterion, _ = build(opt)
    return opt, model, criterion


FIXTURES = {
    "anet_c3d": ("cfgs/anet_c3d_pdvc.yml", dict(vocab_size=300), dict(seed=5, n_gt=4, cap_len=9)),
    "yc2shape": ("cfgs/yc2_tsp_pdvc.yml", dict(num_queries=100, frame_embedding_num=200, vocab_size=1607, max_eseq_length=20),
                 dict(seed=6, n_gt=6, cap_len=10)),
    "newmodel_core": ("cfgs/yc2_newModel_sound.yml", dict(), dict(seed=7, n_gt=5, cap_len=8)),
    "anet_c3d_refcuda": ("cfgs/anet_c3d_pdvc.yml", dict(vocab_size=300), dict(seed=5, n_gt=4, cap_len=9)),
}


def main():
    fixture = sys.argv[1] if len(sys.argv) > 1 else "anet_c3d"
    cfg_path, overrides, batch = FIXTURES[fixture]
    torch.manual_seed(0)
    torch.set_default_dtype(torch.float64)   # the reference creates helper tensors with the default dtype
    opt, model, criterion = load_reference(cfg_path, dict(transformer_dropout_prob=0.0, drop_prob=0.0, **overrides),
                                           zero_padded_attn=fixture.endswith("_refcuda"))
    model.double().train()
    criterion.double()
    cases.fill_parameters(model)
    n_gt = batch["n_gt"]
    b = cases.pdvc_batch(batch["seed"], opt.frame_embedding_num, opt.feature_dim, n_gt=n_gt, cap_len=batch["cap_len"],
                         vocab_size=opt.vocab_size)
    t = {k: torch.from_numpy(v) for k, v in b.items()}
    dt = collections.defaultdict(lambda: None, {
        "video_tensor": t["video_tensor"], "video_mask": t["video_mask"], "video_length": t["video_length"],
        "video_target": [{"boxes": t["boxes"], "labels": t["labels"]}], "cap_tensor": t["cap_tensor"],
        "cap_mask": t["cap_mask"], "gt_boxes": t["boxes"][None], "gt_boxes_mask": torch.ones(1, n_gt, dtype=torch.bool)})
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
           "torch_version": np.array(torch.__version__), "vocab_size": np.int64(opt.vocab_size),
           "config": np.array([opt.feature_dim, opt.num_queries, opt.frame_embedding_num, opt.max_eseq_length, opt.enc_layers,
                               opt.dec_layers, opt.hidden_dim, opt.transformer_ff_dim], dtype=np.int64),
           "batch": np.array([batch["seed"], n_gt, batch["cap_len"]], dtype=np.int64)}
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
    path = os.path.join(HERE, f"pdvc_{fixture}.npz")
    np.savez_compressed(path, **rec)
    print("wrote", path, "total loss", total.item(), {k: round(float(v), 6) for k, v in loss.items()})
    print(len(names), "parameters with gradients")


if __name__ == "__main__":
    main()
