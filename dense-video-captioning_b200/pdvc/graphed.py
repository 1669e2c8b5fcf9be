"""The PDVC training step as two CUDA graphs around the one host decision it contains.

A batch-1 PDVC step is ~4 100 kernel launches for ~16 ms of device work (profiles/r1_pdvc_profile.txt): eager
execution is bound by launch overhead (68 ms/step).  The only point where the host must see device results is the
Hungarian assignment (scipy, pdvc/matcher.py:115-117).  So the step is split there:

    stage 1  (graph)  features -> pyramid -> encoder -> decoder -> heads -> matching cost matrices of all layers
                      [-> Hungarian matching of all layers on the device, csrc/msda_match.cuh: `device_match`]
    host              ONE read-back of the cost matrices, linear_sum_assignment per layer, indices uploaded
                      [skipped with `device_match`: the indices are static outputs of stage 1's graph and the host
                       never waits for the device inside a step]
    stage 2  (graph)  set losses with the matched indices + the teacher-forced caption LSTM loops -> total loss

Each stage is captured once per shape key by `graph_module` below (forward graph + backward graph; the MSDeformAttn /
MSDeformAttnCap CUDA kernels are captured like any other launch: the C ABI does no allocation and no
synchronisation).  `total.backward()` replays stage 2's backward graph, then stage 1's.  Parameter gradients leave the
graphs either through autograd's AccumulateGrad (default; what torch DDP's hooks need) or -- `grad_sinks`, used by
Trainer(sync="overlap") -- through multi-tensor copies captured at the end of each backward graph that write them
straight into the trainer's flat gradient buckets; a stage's `after_backward` callback then starts that bucket's NCCL
all-reduce, so the caption-head bucket (stage 2) is reduced while stage 1's backward graph replays.

Shape key of stage 2: (events, matched pairs, word steps per layer) -- a new key is captured on first use and cached.
`graphs=False` runs the same two stages eagerly (used by the CPU tests and as the reference for the graph tests).
"""
import contextlib
import gc

import numpy as np
import torch
from scipy.optimize import linear_sum_assignment
from torch import nn

from .._streams import branch_stream
from ..ops.functions.matching import linear_sum_assignment_cuda, pdvc_match_cuda
from .model import caption_steps


def pyramid_lengths(t0, n_levels):
    """Level lengths of the feature pyramid: Conv1d(k3,s2,p1) per extra level (pdvc/base_encoder.py:37-41)."""
    out = [int(t0)]
    for _ in range(n_levels - 1):
        out.append((out[-1] - 1) // 2 + 1)
    return out


def graph_module(module, sample_args, warmup=3, autocast=None, grad_sinks=None):
    """Capture `module(*sample_args)` as a forward CUDA graph and its autograd backward as a second graph; return a
    callable with the same signature whose result is differentiable w.r.t. the tensor arguments and the module's
    trainable parameters (the scheme of torch.cuda.make_graphed_callables, written out so that the capture runs in
    `thread_local` error mode -- NCCL's watchdog thread may poll events while a rank is capturing -- and outputs that
    carry no gradient are marked non-differentiable).

    grad_sinks: {id(parameter): tensor view} -- the backward graph ends with multi-tensor copies of those parameters'
    gradients into the given views (zero-fill where a parameter got none) and autograd receives no gradient for them:
    the caller owns `.grad` (views of a flat bucket).  The returned callable's `after_backward` attribute, when set, is
    called right after the backward graph has been enqueued."""
    params = tuple(p for p in module.parameters() if p.requires_grad)
    surface = tuple(sample_args) + params
    diff_inputs = tuple(t for t in surface if t.requires_grad)
    ctx = (lambda: torch.autocast("cuda", dtype=autocast, cache_enabled=False)) if autocast else contextlib.nullcontext

    def run():
        with ctx():
            return tuple(module(*sample_args))

    def backward_of(outs, grad_outs):
        diff_outs = tuple(o for o in outs if o.requires_grad)
        return torch.autograd.grad(diff_outs, diff_inputs, tuple(g for g in grad_outs if g is not None),
                                   only_inputs=True, allow_unused=True)

    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(warmup):
            outs = run()
            backward_of(outs, tuple(torch.ones_like(o) if o.requires_grad else None for o in outs))
        del outs
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()

    pool = torch.cuda.graph_pool_handle()
    fwd_graph, bwd_graph = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    # No cyclic garbage collection while a capture is open: a collection that happens to run mid-capture may destroy CUDA
    # graphs / pools of an earlier Trainer or of evicted shape keys, and the driver calls that makes invalidate the capture
    # ("operation failed due to a previous error during capture").  torch.cuda.graph collects once on entry only.
    gc_was_enabled = gc.isenabled()
    gc.collect()
    gc.disable()
    try:
        with torch.cuda.graph(fwd_graph, pool=pool, capture_error_mode="thread_local"):
            static_outs = run()
        static_grad_outs = tuple(torch.zeros_like(o) if o.requires_grad else None for o in static_outs)
        with torch.cuda.graph(bwd_graph, pool=pool, capture_error_mode="thread_local"):
            grads = backward_of(static_outs, static_grad_outs)
            if grad_sinks:
                grads = list(grads)
                src, dst = [], []
                for i, t in enumerate(diff_inputs):
                    sink = grad_sinks.get(id(t))
                    if sink is None:
                        continue
                    if grads[i] is None:
                        sink.zero_()
                    else:
                        src.append(grads[i].reshape(sink.shape).to(sink.dtype))
                        dst.append(sink)
                    grads[i] = None                  # nothing for AccumulateGrad: the sink IS the parameter's .grad
                if src:
                    torch._foreach_copy_(dst, src)
    finally:
        if gc_was_enabled:
            gc.enable()
    grads = list(grads)
    static_grad_inputs = tuple(grads.pop(0) if t.requires_grad else None for t in surface)
    n_user = len(sample_args)

    class Graphed(torch.autograd.Function):
        @staticmethod
        def forward(fctx, *inputs):
            for static, new in zip(surface[:n_user], inputs[:n_user]):
                if static.data_ptr() != new.data_ptr():
                    static.copy_(new)
            fwd_graph.replay()
            outs = tuple(o.detach() for o in static_outs)
            fctx.mark_non_differentiable(*(o for o, s in zip(outs, static_outs) if not s.requires_grad))
            return outs

        @staticmethod
        @torch.autograd.function.once_differentiable
        def backward(fctx, *grad_outs):
            for static, new in zip(static_grad_outs, grad_outs):
                if static is not None and new is not None and static.data_ptr() != new.data_ptr():
                    static.copy_(new)
            bwd_graph.replay()
            if call.after_backward is not None:
                call.after_backward()
            return tuple(g.detach() if g is not None else None for g in static_grad_inputs)

    def call(*user_args):
        return Graphed.apply(*(tuple(user_args) + params))

    call.graphs = (fwd_graph, bwd_graph)
    call.after_backward = None
    # the parameters' AccumulateGrad nodes were created on the capture stream and are kept alive by the captured autograd
    # graph; later backwards run on the caller's stream -- intended, autograd inserts the stream wait
    quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
    if quiet is not None:
        quiet(False)
    return call


ALIAS_STAGE_INPUTS = True     # stage 2 reads stage 1's static output buffers directly
LSAP_MAX_DIM = 1024           # largest cost matrix side msda_lsap_f32 takes (include/msda.h)
FUSED_MATCH_COST = True       # device matching forms the cost matrices inside the matching kernel (fp32 predictions)

LOSS_TERMS = ("loss_ce", "loss_counter", "loss_bbox", "loss_giou", "loss_self_iou", "cardinality_error", "loss_caption")


class FeatureStage(nn.Module):
    """Stage 1.  Tensors in: features (1,T,C), valid mask (1,T) bool, duration (1,), target segments (G,2), labels (G,)."""

    def __init__(self, model, matcher):
        super().__init__()
        self.base_encoder = model.base_encoder
        self.transformer = model.transformer
        self.query_embed = model.query_embed
        self.class_head, self.count_head, self.bbox_head = model.class_head, model.count_head, model.bbox_head
        self.__dict__["_model"] = model          # not registered: parameters are exposed through the members above
        self.__dict__["_matcher"] = matcher
        self.device_match = False                # set per call by StagedStep: assignment computed here, on the device

    def forward(self, vf, valid_mask, duration, tgt_boxes, tgt_labels):
        model = self._model
        hs, init_reference, inter_references, others = model.forward_features(vf, valid_mask, duration)
        logits, counts, boxes = model.forward_heads(hs, init_reference, inter_references)
        matched = ()
        with torch.no_grad():
            m = self._matcher
            if self.device_match and FUSED_MATCH_COST and logits.dtype == torch.float32 and boxes.dtype == torch.float32:
                # cost matrices (matcher.py:84-103) and assignment (:115) of all layers in ONE launch; one video per step
                costs, q_idx, t_idx = pdvc_match_cuda(logits[:, 0], boxes[:, 0], tgt_labels, tgt_boxes, m.cost_class,
                                                      m.cost_bbox, m.cost_giou, m.cost_alpha, m.cost_gamma)
                matched = (q_idx, t_idx)
            else:
                costs = torch.stack([m.cost_matrix(logits[l], boxes[l], tgt_labels, tgt_boxes) for l in range(hs.shape[0])])
                if self.device_match:            # (layers, Q, G); rows ascending = matcher.py:115
                    matched = linear_sum_assignment_cuda(costs)
        return (hs, init_reference, inter_references, logits, counts, boxes, others['memory'], others['mask_flatten'],
                others['valid_ratios'], costs) + tuple(matched)


class LossStage(nn.Module):
    """Stage 2 for one shape key.  Everything data-dependent arrives as a device tensor; `n_steps` (word steps per
    layer) and the pair count are fixed per instance."""

    def __init__(self, model, criterion, n_steps, level_lengths):
        super().__init__()
        self.caption_head = model.caption_head
        self.__dict__["_model"] = model
        self.__dict__["_criterion"] = criterion
        self.n_steps = tuple(n_steps)
        self.level_lengths = list(level_lengths)
        self.parallel_layers = True
        self.batch_layers = True

    def _side_streams(self, n, device):
        return [branch_stream(device, "caption_loop", i) for i in range(n)]

    def _caption_losses_per_layer(self, hs, init_reference, inter_references, others, cap_tensor, cap_mask, q_idx, t_idx):
        """Caption heads that differ per layer (share_caption_head=0) or layers with different word-step counts."""
        n_pred = hs.shape[0]
        # The caption loops of the decoder layers are independent chains of ~1 300 tiny kernels each: on the GPU every
        # layer's loop runs on its own stream (inside a capture these become parallel branches of the graph, forward and --
        # because autograd replays each node on the stream it ran on -- backward).  cuDNN's RNN calls serialise across
        # streams in host issue order (tools/graph_fork_probe.py), so the word steps of the layers are issued in lockstep:
        # the branches then wait for each other only at the LSTM cell and overlap everywhere else.
        n_side = n_pred - 1 if (hs.is_cuda and self.parallel_layers) else 0
        main = torch.cuda.current_stream() if n_side else None
        side = self._side_streams(n_side, hs.device) if n_side else []

        def on_stream(lid):
            return torch.cuda.stream(side[lid - 1]) if (n_side and lid > 0) else contextlib.nullcontext()

        for s_ in side:
            s_.wait_stream(main)
        decoders = []
        for lid in range(n_pred):
            with on_stream(lid):
                reference = init_reference if lid == 0 else inter_references[lid - 1]
                q, c = q_idx[lid], t_idx[lid]
                decoders.append(self.caption_head[lid].prepare(hs[lid][:, q], reference[:, q], others, cap_tensor[c]))
        for i in range(max(self.n_steps)):
            for lid in range(n_pred):
                if i < self.n_steps[lid]:
                    with on_stream(lid):
                        decoders[lid].step(i)
        cap = []
        for lid in range(n_pred):
            with on_stream(lid):
                c = t_idx[lid]
                cap.append(self.caption_head[lid].build_loss(decoders[lid].result(), cap_tensor[:, 1:][c], cap_mask[:, 1:][c]).mean())
        return cap, side, main

    def forward(self, hs, init_reference, inter_references, logits, counts, boxes, memory, mask_flat, valid_ratios,
                tgt_boxes, tgt_labels, cap_tensor, cap_mask, num_boxes, q_idx, t_idx):
        model, crit = self._model, self._criterion
        n_pred = hs.shape[0]
        shapes, starts = model.transformer.level_tensors(self.level_lengths, hs.device)
        others = {'memory': memory, 'mask_flatten': mask_flat, 'spatial_shapes': shapes, 'level_start_index': starts,
                  'valid_ratios': valid_ratios, 'level_lengths': self.level_lengths}
        targets = [{'boxes': tgt_boxes, 'labels': tgt_labels}]
        pairs = q_idx.shape[1]
        batch = torch.zeros(pairs, dtype=torch.long, device=hs.device)

        shared = (self.batch_layers and len(set(self.n_steps)) == 1 and n_pred > 1
                  and all(head is self.caption_head[0] for head in self.caption_head)
                  and self.caption_head[0].core.n_levels == len(self.level_lengths)
                  and inter_references.shape[-1] == 2)     # with_box_refine=0 keeps 1-coordinate references: per-layer path
        if shared:
            # One caption head shared by all decoder layers (opts.py:148, the default): the layers' matched pairs are
            # independent sentences for the SAME weights, so they are decoded as ONE batch of n_pred * pairs sentences --
            # one chain of word steps instead of n_pred.  Layer 0's single-coordinate reference (pdvc.py:262) is written in
            # the two-coordinate form with width 2P/T_l, for which the sampler's `ref0 + off/P * ref1 * 0.5`
            # (ms_deform_attn_for_caption.py:107-109) is the `ref0 + off/T_l` of :104-106.
            head = self.caption_head[0]
            side = self._side_streams(1, hs.device) if hs.is_cuda else []
            main = torch.cuda.current_stream() if side else None
            for s_ in side:
                s_.wait_stream(main)
            with (torch.cuda.stream(side[0]) if side else contextlib.nullcontext()):
                layer = torch.arange(n_pred, device=hs.device)[:, None]
                hs_pairs = hs.squeeze(1)[layer, q_idx].reshape(1, n_pred * pairs, -1)       # squeeze: a view, no select in backward
                valid = valid_ratios[0]                                                     # (L,)
                first = init_reference.squeeze(0)[q_idx[0]] * valid[None, :]                        # (pairs, L)
                width = (2.0 * head.core.n_points) / shapes.to(first.dtype)
                refs = [torch.stack([first, width[None, :].expand_as(first)], -1)]
                if n_pred > 1:
                    later = inter_references[:n_pred - 1, 0][layer[:-1], q_idx[1:]]          # (n_pred-1, pairs, 2)
                    refs.append((later[:, :, None, :] * valid[None, None, :, None]).reshape(-1, valid.shape[0], 2))
                ref = torch.cat(refs, 0)[None]                                              # (1, n_pred*pairs, L, 2)
                flat = t_idx.reshape(-1)
                decoder = head.begin(hs_pairs, ref, memory, shapes, starts, mask_flat, cap_tensor[flat].long())
                for i in range(self.n_steps[0]):
                    decoder.step(i)
                per_sentence = head.build_loss(decoder.result(), cap_tensor[:, 1:][flat], cap_mask[:, 1:][flat])
                cap = list(per_sentence.view(n_pred, pairs).mean(1).unbind(0))
        else:
            cap, side, main = self._caption_losses_per_layer(hs, init_reference, inter_references, others, cap_tensor,
                                                             cap_mask, q_idx, t_idx)
        set_losses = []
        logits_l, counts_l, boxes_l = logits.unbind(0), counts.unbind(0), boxes.unbind(0)   # one stack each in backward
        for lid in range(n_pred):
            out = {'pred_logits': logits_l[lid], 'pred_count': counts_l[lid], 'pred_boxes': boxes_l[lid]}
            set_losses.append(crit.layer_losses(out, targets, (batch, q_idx[lid], [t_idx[lid]]), [pairs], num_boxes))
        for s_ in side:
            main.wait_stream(s_)
        terms, total = [], 0
        for lid in range(n_pred):
            suffix = '' if lid == n_pred - 1 else f'_{lid}'
            losses = set_losses[lid]
            losses['loss_caption'] = cap[lid]
            for name in LOSS_TERMS:
                value = losses[name]
                terms.append(value.detach().to(torch.float64 if memory.dtype == torch.float64 else torch.float32))
                weight = crit.weight_dict.get(name + suffix, 0)
                if weight:
                    total = total + (value if value.dtype == torch.float64 else value.float()) * weight
        return total, torch.stack(terms)


class StagedStep(nn.Module):
    """forward(dt) -> (total loss, dict of loss terms, (query indices, event indices) per layer) for ONE video, as
    stage 1 -> matching -> stage 2.  The indices are device tensors with `device_match` (default on CUDA), else the host
    arrays scipy returned."""

    def __init__(self, model, criterion, graphs=True, amp=False, max_graphs=64, device_match=True):
        super().__init__()
        self.model = model
        self.criterion = criterion
        self.graphs = graphs
        self.amp = amp
        # Hungarian matching on the device (CUDA only; needs events <= queries <= 1024, else the host path below):
        # no device->host read-back inside the step, the host runs ahead of the GPU
        self.device_match = device_match
        self.max_graphs = max_graphs      # stage-2 graphs kept (one per shape key, least recently used dropped first)
        # Trainer(sync="overlap"): {id(param): flat-bucket view} and one callback per stage (see graph_module)
        self.__dict__["grad_sinks"] = None
        self.__dict__["after_backward"] = {1: None, 2: None}
        self.stage1 = FeatureStage(model, criterion.matcher)
        self.__dict__["_stage1"] = {}
        self.__dict__["_stage2"] = {}
        self.__dict__["_staging"] = {}

    # ---- host side -------------------------------------------------------------------------------------------
    def _match(self, costs, n_pred):
        """One device->host transfer for all layers, then scipy's LSAP per layer (matcher.py:115-117)."""
        host = costs.detach().float().cpu()
        q_rows, t_rows = [], []
        for lid in range(n_pred):
            rows, cols = linear_sum_assignment(host[lid])
            q_rows.append(rows)
            t_rows.append(cols)
        return np.stack(q_rows).astype(np.int64), np.stack(t_rows).astype(np.int64)

    def _upload_indices(self, q_host, t_host, device):
        """Pinned staging buffers (one pair per shape): indices go up with an async copy, no pageable-memory stall."""
        if device.type != "cuda":
            return torch.from_numpy(q_host), torch.from_numpy(t_host)
        key = q_host.shape
        if key not in self._staging:
            self._staging[key] = (torch.empty(key, dtype=torch.long).pin_memory(), torch.empty(key, dtype=torch.long).pin_memory())
        q_pin, t_pin = self._staging[key]
        q_pin.copy_(torch.from_numpy(q_host))
        t_pin.copy_(torch.from_numpy(t_host))
        return q_pin.to(device, non_blocking=True), t_pin.to(device, non_blocking=True)

    def _num_boxes(self, n_gt, device):
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            total = torch.tensor([float(n_gt)], device=device)
            torch.distributed.all_reduce(total)                      # criterion.py:169-171
            return (total / torch.distributed.get_world_size()).clamp(min=1)[0]
        return torch.full((), max(float(n_gt), 1.0), device=device)

    def _graph(self, module, args, stage):
        call = graph_module(module, args, autocast=torch.bfloat16 if self.amp else None, grad_sinks=self.grad_sinks)
        call.after_backward = self.after_backward[stage]
        return call

    def stage2_parameters(self):
        """Parameters only stage 2 (losses + caption loops) reaches: their gradients are final when stage 2's backward
        ends, i.e. before stage 1's backward starts."""
        return [p for p in self.model.caption_head.parameters() if p.requires_grad]

    def _cached(self, cache, key, make):
        """Least-recently-used cache of captured stages, at most max_graphs per stage."""
        if key in cache:
            cache[key] = cache.pop(key)                     # most recently used last
        else:
            if len(cache) >= self.max_graphs:
                cache.pop(next(iter(cache)))
            cache[key] = make()
        return cache[key]

    # ---- the step --------------------------------------------------------------------------------------------
    def forward(self, dt):
        vf, valid, duration = dt['video_tensor'], dt['video_mask'], dt['video_length'][:, 1]
        assert vf.shape[0] == 1, "one video per step (cfgs/*: batch_size 1; LSTM_DSA.py:59)"
        target = dt['video_target'][0]
        args1 = (vf, valid, duration, target['boxes'], target['labels'])
        use_graphs = self.graphs and vf.is_cuda
        n_gt, n_query = int(target['boxes'].shape[0]), int(self.model.cfg.num_queries)
        # every event is matched when there are at least as many queries, so the word-step count (the one host decision
        # that depends on WHICH captions were matched) is known without the assignment
        on_device = bool(self.device_match and vf.is_cuda and 1 <= n_gt <= n_query <= LSAP_MAX_DIM)
        stage1 = self.stage1
        stage1.device_match = on_device
        if use_graphs:      # the matching costs are part of stage 1, so its graph depends on the event count too
            key1 = (int(vf.shape[1]), n_gt, on_device)
            stage1 = self._cached(self._stage1, key1, lambda: self._graph(self.stage1, tuple(a.detach().clone() for a in args1), 1))
        # the event count is host data: its all-reduce (criterion.py:169-171) is issued before stage 1 and has long
        # finished when stage 2 reads it, instead of making the ranks meet in the middle of the step
        num_boxes = self._num_boxes(int(target['boxes'].shape[0]), vf.device)
        outs = stage1(*args1)
        hs, init_ref, inter_refs, logits, counts, boxes, memory, mask_flat, valid_ratios, costs = outs[:10]
        n_pred = hs.shape[0]
        cap_host = dt.get('cap_tensor_host')
        cap_host = dt['cap_tensor'].cpu().numpy() if cap_host is None else np.asarray(cap_host)
        if on_device:
            q_idx, t_idx = outs[10], outs[11]
            n_steps = (caption_steps(cap_host[:n_gt]),) * n_pred
            matched = (q_idx, t_idx)
        else:
            q_host, t_host = self._match(costs, n_pred)
            n_steps = tuple(caption_steps(cap_host[t_host[lid]]) for lid in range(n_pred))
            q_idx, t_idx = self._upload_indices(q_host, t_host, vf.device)
            matched = (q_host, t_host)
        args2 = (hs, init_ref, inter_refs.detach(), logits, counts, boxes, memory, mask_flat, valid_ratios.detach(),
                 target['boxes'], target['labels'], dt['cap_tensor'], dt['cap_mask'], num_boxes, q_idx, t_idx)
        key = (n_gt, int(q_idx.shape[1]), n_steps, int(vf.shape[1]), int(dt['cap_tensor'].shape[1]), on_device)

        def make_stage2():
            lengths = pyramid_lengths(vf.shape[1], self.model.cfg.num_feature_levels)
            stage2 = LossStage(self.model, self.criterion, n_steps, lengths)
            if use_graphs:
                # the first 9 arguments are static outputs of the stage-1 graph: captured in place (same storage), so the
                # replay wrapper finds identical addresses and skips its input copies; everything else is cloned
                # (with device matching the two index tensors, arguments 14 and 15, are static stage-1 outputs as well)
                static = lambda i: ALIAS_STAGE_INPUTS and (i < 9 or (on_device and i >= 14))  # noqa: E731
                sample = tuple((a.detach() if static(i) else a.detach().clone()).requires_grad_(a.requires_grad)
                               for i, a in enumerate(args2))
                stage2 = self._graph(stage2, sample, 2)
            return stage2

        total, terms = self._cached(self._stage2, key, make_stage2)(*args2)
        names = [name + ('' if lid == n_pred - 1 else f'_{lid}') for lid in range(n_pred) for name in LOSS_TERMS]
        return total, dict(zip(names, terms.unbind(0))), matched
