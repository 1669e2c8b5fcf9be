"""Synthetic-feature PDVC training harness (BASELINE.json metric 2: training videos/s at 1/2/4/8 GPUs).

What the reference does per iteration (train.py:163-194): move the batch to the device, forward with the criterion,
weighted loss sum, backward, clip_grad_norm_(100), Adam step -- batch size 1 video, with a `cuda.synchronize()` before
and after every iteration.  Same step here, minus the two forced synchronisations (and, in 'graphs' mode, minus the
host round trip of the Hungarian matching: pdvc/graphed.py), with the inputs copied from pinned host memory each step, and -- new, the reference never wraps PDVC in DDP (SURVEY finding 0.8) -- one process per GPU
with an NCCL all-reduce (mean) of the gradients.  Three exchange modes (`Trainer(sync=...)`):
  'overlap' (default)  two flat buckets; the caption-head bucket is reduced on a side stream WHILE the stage-1 backward
                       graph replays, the rest right after it; gradients reach the buckets through copies captured in
                       the backward graphs (no per-step concatenation, no AccumulateGrad traffic);
  'flat'               one concatenation + one all-reduce after both backward graphs (no overlap; round-1 default);
  'ddp'                torch DistributedDataParallel (bucket hooks on AccumulateGrad).
Ranks own disjoint videos (video i of step s goes to rank i); the learning rate is not rescaled (the reference has no
rule).

Configurations: shapes from cfgs/anet_c3d_pdvc.yml (+ cfgs/anet_c3d_pdvcl.yml), cfgs/yc2_tsp_pdvc.yml as
shipped and with the YouCook2 overrides of cfgs/yc2_tsn_pdvcl.yml:26,31, and the PDVC core of
cfgs/yc2_newModel_sound.yml on synthetic 768-d features.
"""
import contextlib
from dataclasses import replace

import numpy as np
import torch

from .._streams import branch_stream
from .model import PDVCConfig, build

CONFIGS = {
    # BASELINE configs[2]: ActivityNet C3D, 100 frame embeddings, 10 queries
    "anet_c3d_pdvc": dict(cfg=PDVCConfig(feature_dim=500, num_queries=10, frame_embedding_num=100, vocab_size=5747,
                                         max_eseq_length=10), n_gt=4, cap_len=12, amp=False),
    # BASELINE configs[1] as shipped: inherits the ActivityNet shapes, only the feature type/dim change
    "yc2_tsp_pdvc": dict(cfg=PDVCConfig(feature_dim=512, num_queries=10, frame_embedding_num=100, vocab_size=5747,
                                        max_eseq_length=10), n_gt=6, cap_len=12, amp=False),
    # BASELINE configs[1] with the YouCook2 shapes of configs[0] (100 queries, 200 frames, vocab 1607)
    "yc2_tsp_pdvc_yc2shape": dict(cfg=PDVCConfig(feature_dim=512, num_queries=100, frame_embedding_num=200,
                                                 vocab_size=1607, max_eseq_length=20), n_gt=6, cap_len=12, amp=False),
    # BASELINE configs[3]: PDVC core of yc2_newModel_sound (3+3 layers, 768-d fused features), bf16 autocast
    "yc2_newmodel_core_bf16": dict(cfg=PDVCConfig(feature_dim=768, num_queries=100, frame_embedding_num=200,
                                                  vocab_size=1607, max_eseq_length=20, enc_layers=3, dec_layers=3),
                                   n_gt=6, cap_len=12, amp=True),
}


def synthetic_video(cfg, n_gt, cap_len, seed):
    """One video's batch dict on the HOST (pinned where a GPU exists), SURVEY 8d config-2 inputs: features ~ N(0,1),
    all frames valid, duration 120 s, n_gt events with centre U(0.2,0.8) and length U(0.05,0.25), label 0, captions of
    cap_len random tokens framed by 0 (BOS/EOS) and padded to cap_len + 4."""
    rs = np.random.RandomState(seed)
    T = cfg.frame_embedding_num
    words = cap_len + 4
    feats = rs.standard_normal((1, T, cfg.feature_dim)).astype(np.float32)
    centre = rs.uniform(0.2, 0.8, n_gt)
    length = rs.uniform(0.05, 0.25, n_gt)
    boxes = np.stack([centre, length], -1).astype(np.float32)
    caps = np.zeros((n_gt, words), dtype=np.int64)
    caps[:, 1:1 + cap_len] = rs.randint(1, cfg.vocab_size, (n_gt, cap_len))
    cap_mask = np.zeros((n_gt, words), dtype=np.float32)
    cap_mask[:, :cap_len + 2] = 1.0
    host = {
        'video_tensor': torch.from_numpy(feats),
        'video_mask': torch.ones(1, T, dtype=torch.bool),
        'video_length': torch.tensor([[float(T), 120.0, float(n_gt)]]),
        'boxes': torch.from_numpy(boxes),
        'labels': torch.zeros(n_gt, dtype=torch.long),
        'cap_tensor': torch.from_numpy(caps),
        'cap_mask': torch.from_numpy(cap_mask),
    }
    if torch.cuda.is_available():
        host = {k: v.pin_memory() for k, v in host.items()}
    return host


def to_device(host, device):
    """The host->device boundary of the step (train.py:174-177)."""
    d = {k: v.to(device, non_blocking=True) for k, v in host.items()}
    return {'video_tensor': d['video_tensor'], 'video_mask': d['video_mask'], 'video_length': d['video_length'],
            'video_target': [{'boxes': d['boxes'], 'labels': d['labels']}],
            'cap_tensor': d['cap_tensor'], 'cap_mask': d['cap_mask'], 'cap_tensor_host': host['cap_tensor'].numpy()}


def input_bytes(host):
    return sum(v.numel() * v.element_size() for v in host.values())


FLAT_OPTIMIZER = True     # default of Trainer(flat_optimizer=None); module-level so tools/pdvc_ab.py can A/B it


class Trainer:
    """model + criterion + Adam(lr 5e-5, wd 1e-4) + grad clip 100 (cfgs/anet_c3d_pdvcl.yml:17,21; opts.py:104-105)."""

    def __init__(self, name_or_cfg, device, world_size=1, lr=5e-5, weight_decay=1e-4, grad_clip=100.0, amp=None,
                 seed=0, dropout=None, mode="graphs", sync="overlap", flat_optimizer=None):
        """mode: 'graphs' = two CUDA graphs with the Hungarian matching on the device between them (pdvc/graphed.py; CUDA
        only), 'staged' = the same two stages eagerly, 'eager' = the reference's control flow (model.forward with the
        criterion inside).
        sync (world_size > 1): 'overlap' | 'flat' | 'ddp', see the module docstring.
        flat_optimizer: clip + Adam as two launches over the flat gradient buckets (pdvc/optim.py) where those exist
        (mode 'graphs' with sync 'overlap'); otherwise torch's clip_grad_norm_ + fused Adam."""
        if sync not in ("overlap", "flat", "ddp"):
            raise ValueError(f"sync must be 'overlap', 'flat' or 'ddp', got {sync!r}")
        spec = CONFIGS[name_or_cfg] if isinstance(name_or_cfg, str) else dict(cfg=name_or_cfg, n_gt=4, cap_len=12, amp=False)
        cfg = spec["cfg"]
        if dropout is not None:
            cfg = replace(cfg, transformer_dropout_prob=dropout, drop_prob=dropout)
        self.cfg, self.n_gt, self.cap_len = cfg, spec["n_gt"], spec["cap_len"]
        self.amp = spec["amp"] if amp is None else amp
        self.device = torch.device(device)
        torch.manual_seed(seed)                      # identical initial weights on every rank
        self.model, self.criterion = build(cfg)
        self.model.to(self.device).train()
        self.criterion.to(self.device)
        # parameters this training path never reaches a loss through -- the proposal embedding
        # (deformable_transformer.py:136-142) and MSDeformAttnCap's attention_weights / output_proj
        # (ms_deform_attn_for_caption.py:102-103: computed, never used): they have no gradient in the reference either,
        # so Adam skips them there; freezing them keeps DDP's graph static without find_unused_parameters
        for name, p in self.model.named_parameters():
            if name.startswith(("transformer.pos_trans.", "transformer.pos_trans_norm.")) or \
                    ".core.deformable_att.attention_weights." in name or ".core.deformable_att.output_proj." in name:
                p.requires_grad_(False)
        self.params = [p for p in self.model.parameters() if p.requires_grad]
        if mode == "graphs" and self.device.type != "cuda":
            mode = "staged"
        self.mode = mode
        if self.device.type == "cuda":
            for m in self.model.modules():
                if isinstance(m, torch.nn.LSTM):
                    m.flatten_parameters()
        if mode == "eager":
            self.net = _StepModule(self.model, self.criterion)
        else:
            from .graphed import StagedStep
            self.net = StagedStep(self.model, self.criterion, graphs=(mode == "graphs"), amp=self.amp)
        self.world_size, self.sync = world_size, sync
        self._buckets = None
        if sync == "overlap" and mode != "eager":
            self._setup_buckets()
        if world_size > 1 and sync in ("flat", "overlap"):
            with torch.no_grad():                    # what DDP's constructor does: rank 0's weights and buffers everywhere
                for t in list(self.model.parameters()) + list(self.model.buffers()):
                    torch.distributed.broadcast(t, 0)
        elif world_size > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            ids = [self.device.index] if self.device.type == "cuda" else None
            if self.device.type == "cuda":
                side = torch.cuda.Stream()           # DDP built on a side stream, as graph capture inside it requires
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    self.net = DDP(self.net, device_ids=ids, gradient_as_bucket_view=True, broadcast_buffers=False)
                torch.cuda.current_stream().wait_stream(side)
            else:
                self.net = DDP(self.net, gradient_as_bucket_view=True, broadcast_buffers=False)
        # same update rule as the reference's torch.optim.Adam (train.py:123) after clip_grad_norm_ (train.py:185)
        flat_optimizer = FLAT_OPTIMIZER if flat_optimizer is None else flat_optimizer
        self._flat_optimizer = bool(flat_optimizer and self._sinks_live and self.device.type == "cuda")
        if self._flat_optimizer:
            from .optim import FlatClipAdam
            self.optimizer = FlatClipAdam(self._buckets, lr=lr, weight_decay=weight_decay, max_norm=grad_clip, order=self.params)
        else:
            self.optimizer = torch.optim.Adam(self.params, lr=lr, weight_decay=weight_decay,
                                              fused=(self.device.type == "cuda"))
        self.grad_clip = grad_clip
        self.weight_dict = self.criterion.weight_dict

    def step(self, host_batch):
        """One training iteration on one video; returns the total loss as a device scalar of its own (not a view of a
        graph's static output) without synchronising, so a caller may read step i's loss after issuing step i+1."""
        dt = to_device(host_batch, self.device)
        if self._sinks_live:
            self._pending = []                       # .grad are views of the buckets, fully rewritten by the backward graphs
        else:
            self.optimizer.zero_grad(set_to_none=True)
        if self.mode == "eager":
            ctx = torch.autocast(self.device.type, dtype=torch.bfloat16) if self.amp else contextlib.nullcontext()
            with ctx:
                losses = self.net(dt)
            total = sum(losses[k].float() * w for k, w in self.weight_dict.items() if k in losses)
        else:
            if self.amp and self.mode == "staged":
                with torch.autocast(self.device.type, dtype=torch.bfloat16):
                    total, losses, _ = self.net(dt)
            else:
                total, losses, _ = self.net(dt)       # graphs: autocast is baked into the captured stages
        total.backward()
        if self._sinks_live:
            for work in self._pending:               # both buckets were issued during backward
                work.wait()
        elif self.world_size > 1 and self.sync == "overlap" and self._buckets is not None:
            self._allreduce_buckets_after_backward()
        elif self.world_size > 1 and self.sync in ("flat", "overlap"):
            self._allreduce_gradients()
        if self._flat_optimizer:
            self.optimizer.step()                    # global-norm clipping included
        else:
            torch.nn.utils.clip_grad_norm_(self.params, self.grad_clip)
            self.optimizer.step()
        self.last_losses = losses
        return total.detach().clone()

    # ---- sync='overlap' -----------------------------------------------------------------------------------------
    def _setup_buckets(self):
        """Two flat gradient buckets: [0] parameters only stage 2 reaches (the caption head: final when stage 2's
        backward graph ends), [1] everything else (final when stage 1's ends).  With CUDA graphs the backward graphs copy
        the gradients into the buckets themselves (graph_module grad_sinks) and `.grad` is a permanent view."""
        first = {id(p) for p in self.net.stage2_parameters()}
        groups = [[p for p in self.params if id(p) in first], [p for p in self.params if id(p) not in first]]
        self._buckets = []
        for group in groups:
            # every parameter's slot starts on a 16-byte boundary (the fused optimizer's vector path, pdvc/optim.py)
            starts, offset = [], 0
            for p in group:
                starts.append(offset)
                offset += (p.numel() + 3) // 4 * 4
            flat = torch.zeros(offset, dtype=self.params[0].dtype, device=self.device)
            views = [flat[o:o + p.numel()].view_as(p) for o, p in zip(starts, group)]
            self._buckets.append((flat, group, views))
        self._sinks_live = self.mode == "graphs"
        if self._sinks_live:
            self.net.grad_sinks = {id(p): v for _, group, views in self._buckets for p, v in zip(group, views)}
            for _, group, views in self._buckets:
                for p, v in zip(group, views):
                    p.grad = v
            if self.world_size > 1:
                self._comm_stream = branch_stream(self.device, "gradient_allreduce")
                self.net.after_backward[2] = lambda: self._issue_bucket(0, side=True)
                self.net.after_backward[1] = lambda: self._issue_bucket(1, side=False)

    _sinks_live = False
    _pending = ()

    def _reduce_mean(self, flat, async_op=False):
        dist = torch.distributed
        if dist.get_backend() == "nccl":
            return dist.all_reduce(flat, op=dist.ReduceOp.AVG, async_op=async_op)
        work = dist.all_reduce(flat, async_op=False)
        flat.div_(self.world_size)
        return work

    def _issue_bucket(self, index, side):
        """Called from inside backward, right after a stage's backward graph was enqueued: start that bucket's all-reduce.
        side=True: on the communication stream, so that the collective runs beside the next backward graph."""
        flat = self._buckets[index][0]
        current = torch.cuda.current_stream(self.device)
        if side:
            self._comm_stream.wait_stream(current)
            with torch.cuda.stream(self._comm_stream):
                work = self._reduce_mean(flat, async_op=True)
            flat.record_stream(self._comm_stream)
        else:
            work = self._reduce_mean(flat, async_op=True)
        self._pending.append(work)

    def _allreduce_buckets_after_backward(self):
        """'overlap' without CUDA graphs (mode='staged', the CPU/gloo tests): same two buckets, filled by concatenation
        after backward -- the exchange pattern without the overlap."""
        for flat, group, views in self._buckets:
            grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in group]
            torch._foreach_copy_(views, grads)       # slots are 16-byte aligned: not a plain concatenation
            self._reduce_mean(flat)
            for p, view in zip(group, views):
                p.grad = view

    def _allreduce_gradients(self):
        """Mean of the gradients over the ranks: one concatenation, one all-reduce, .grad re-pointed into the flat buffer."""
        if getattr(self, "_flat", None) is None:
            self._flat = torch.empty(sum(p.numel() for p in self.params), dtype=self.params[0].dtype, device=self.device)
            self._flat_views, offset = [], 0
            for p in self.params:
                self._flat_views.append(self._flat[offset:offset + p.numel()].view_as(p))
                offset += p.numel()
        grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in self.params]
        torch.cat([g.reshape(-1) for g in grads], out=self._flat)
        if torch.distributed.get_backend() == "nccl":
            torch.distributed.all_reduce(self._flat, op=torch.distributed.ReduceOp.AVG)
        else:
            torch.distributed.all_reduce(self._flat)
            self._flat.div_(self.world_size)
        for p, view in zip(self.params, self._flat_views):
            p.grad = view


class _StepModule(torch.nn.Module):
    """forward(dt) -> loss dict; the unit DDP wraps (criterion included: it has no parameters)."""

    def __init__(self, model, criterion):
        super().__init__()
        self.model = model
        self.criterion = criterion

    def forward(self, dt):
        return self.model(dt, self.criterion)[1]
