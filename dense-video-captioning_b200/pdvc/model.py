"""PDVC around the B200 hot path: feature pyramid, deformable transformer, prediction heads, LSTM caption head
with deformable soft attention.  Written so that BASELINE.json's second metric (PDVC training videos/s) can be
measured where the reference sources do not exist; module / parameter names follow the reference, so its
checkpoints load with `load_state_dict` unchanged.

Behaviour restated from: pdvc/base_encoder.py:13-82, pdvc/position_encoding.py:11-66, pdvc/pdvc.py:33-112 (heads and
their init), :123-170 (forward), :236-323 (matched prediction), :325-445 (caption pairing),
pdvc/CaptioningHead/LSTM_DSA.py:17-109 and :195-263 (captioner), pdvc/pdvc.py:546-556 (MLP).

Differences that do not change results: host read-backs inside the step are removed (duration mask built with a
comparison instead of a per-video Python loop, level lengths taken from shapes, caption length from the host copy
of the caption tensor, one cost-matrix transfer for all decoder layers) and the caption head projects the encoder
memory once per sentence instead of once per word.
"""
import copy
import contextlib
import math

from dataclasses import dataclass, field
from types import SimpleNamespace

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from ..ops.functions import CaptionAttendFunction, LSTMCellFunction, StepAttendRun, StepLinear, StepSampleRun, get_default_padding_mode
from .._streams import branch_stream
from ..ops.modules import MSDeformAttnCap
from .transformer import DeformableTransformer, _linear, inverse_sigmoid


# Module-level switches (each an exact regrouping of the reference's expressions; tools/pdvc_ab.py A/Bs one at a time)
USE_FP32_SAMPLES = True              # bf16 autocast: caption-head samples (two projected-memory rows per sample) kept in fp32
USE_FUSED_ATTEND = True              # one kernel each way for the soft attention over the samples (csrc/msda_capattend.cuh)
USE_STEP_LSTM = True                 # hoisted input projections + fused cell kernel instead of cuDNN's single-step LSTM
USE_STEP_LINEAR = True               # few-row kernels + deferred weight gradients for the Linear layers on the caption recurrence
USE_STEP_SAMPLE_RUN = True           # word steps: sampler / alpha_net gradient sums formed once per decoding run
USE_PROJ_MLP = True                  # hidden layers of the segment heads on the projection kernels (as the FFN layers)
PARALLEL_H2ATT = True                # word step: h2att(h) on a side stream beside the offsets -> sampler chain
PARALLEL_HEADS = True                # class / count / segment heads of the decoder layers on one stream per layer
PARALLEL_POSITION_EMBEDDING = True   # masks + position embeddings of the pyramid on a side stream beside the projections


@dataclass
class PDVCConfig:
    """The subset of the reference's options (opts.py) that shapes the model and the loss."""
    feature_dim: int = 500
    hidden_dim: int = 512
    num_queries: int = 10
    num_feature_levels: int = 4
    nheads: int = 8
    enc_layers: int = 2
    dec_layers: int = 2
    enc_n_points: int = 4
    dec_n_points: int = 4
    transformer_ff_dim: int = 512
    transformer_dropout_prob: float = 0.1
    num_classes: int = 1
    max_eseq_length: int = 10
    aux_loss: bool = True
    with_box_refine: bool = True
    share_caption_head: bool = True
    # caption head ('standard' = LSTM + deformable soft attention, cfgs/anet_c3d_pdvc.yml)
    vocab_size: int = 5747
    rnn_size: int = 512
    num_layers: int = 1
    input_encoding_size: int = 512
    att_hid_size: int = 512
    drop_prob: float = 0.5
    max_caption_len: int = 30
    cap_nheads: int = 1
    cap_dec_n_points: int = 4
    cap_num_feature_levels: int = 4
    # matcher / loss (cfgs/anet_c3d_pdvcl.yml)
    set_cost_class: float = 2.0
    set_cost_bbox: float = 0.0
    set_cost_giou: float = 4.0
    cost_alpha: float = 0.25
    cost_gamma: float = 2.0
    cls_loss_coef: float = 2.0
    bbox_loss_coef: float = 0.0
    giou_loss_coef: float = 4.0
    count_loss_coef: float = 0.5
    caption_loss_coef: float = 2.0
    focal_alpha: float = 0.25
    focal_gamma: float = 2.0
    lloss_gau_mask: int = 1
    lloss_beta: float = 1.0
    frame_embedding_num: int = 100
    extras: dict = field(default_factory=dict)


class PositionEmbeddingSine(nn.Module):
    """Sine embedding of the (mask-aware, normalised) frame index concatenated with a learnt embedding of the
    video duration -- position_encoding.py:11-58."""

    def __init__(self, num_pos_feats=64, temperature=10000, scale=2 * math.pi):
        super().__init__()
        self.num_pos_feats = num_pos_feats
        self.temperature = temperature
        self.scale = scale
        self.max_duration = 256
        self.duration_embed_layer = nn.Linear(self.max_duration, self.max_duration)

    def duration_embedding(self, duration, device):
        """(N, max_duration): first int(duration) units on, the rest off (position_encoding.py:60-66) without a host loop.
        The same vector for every pyramid level of a video: BaseEncoder forms it once per forward."""
        units = torch.arange(self.max_duration, device=device)[None] < duration.int()[:, None]
        return self.duration_embed_layer(units.to(self.duration_embed_layer.weight.dtype))

    def frequencies(self, device):
        """temperature ** (2 * (i // 2) / num_pos_feats), i < num_pos_feats: constants, formed once per device."""
        key = str(device)
        if key not in self.__dict__.setdefault("_dim_t", {}):
            dim_t = torch.arange(self.num_pos_feats, dtype=torch.float32, device=device)
            self._dim_t[key] = self.temperature ** (2 * (dim_t // 2) / self.num_pos_feats)
        return self._dim_t[key]

    def forward(self, mask, duration, dur=None):
        """mask (N,T) True = padding; duration (N,) seconds -> (N, 2*num_pos_feats, T).  `dur`: duration_embedding(duration)
        when the caller has it already."""
        steps = (~mask).cumsum(1, dtype=torch.float32)
        steps = (steps - 0.5) / (steps[:, -1:] + 1e-6) * self.scale
        ang = steps[:, :, None] / self.frequencies(mask.device)
        pos_x = torch.stack((ang[:, :, 0::2].sin(), ang[:, :, 1::2].cos()), dim=3).flatten(2)
        if dur is None:
            dur = self.duration_embedding(duration, mask.device)
        dur = dur[:, None, :].expand(-1, pos_x.shape[1], -1)
        return torch.cat((pos_x, dur), dim=2).permute(0, 2, 1)


class BaseEncoder(nn.Module):
    """1x1 conv + (L-1) stride-2 convs, GroupNorm(32) each -> L temporal levels (base_encoder.py:13-82)."""

    def __init__(self, num_feature_levels, vf_dim, hidden_dim):
        super().__init__()
        assert num_feature_levels > 1, "the single-level variant of the reference uses Conv2d and is never configured"
        # hidden_dim//2 sine channels + 256 duration channels: as in the reference this adds up to hidden_dim only for
        # hidden_dim = 512, the value every reference config uses (opts.py:71)
        self.pos_embed = PositionEmbeddingSine(hidden_dim // 2)
        self.num_feature_levels = num_feature_levels
        self.hidden_dim = hidden_dim
        projs = [nn.Sequential(nn.Conv1d(vf_dim, hidden_dim, kernel_size=1), nn.GroupNorm(32, hidden_dim))]
        in_ch = vf_dim
        for _ in range(num_feature_levels - 1):
            projs.append(nn.Sequential(nn.Conv1d(in_ch, hidden_dim, kernel_size=3, stride=2, padding=1),
                                       nn.GroupNorm(32, hidden_dim)))
            in_ch = hidden_dim
        self.input_proj = nn.ModuleList(projs)
        for proj in self.input_proj:
            nn.init.xavier_uniform_(proj[0].weight, gain=1)
            nn.init.constant_(proj[0].bias, 0)

    def forward(self, vf, mask, duration):
        x = vf.transpose(1, 2)
        # level lengths of the stride-2 convolutions (k3, s2, p1): known without running them, so the masks and position
        # embeddings of all levels -- which depend on the mask and the duration only -- are formed on a side stream while
        # the projection chain runs (a parallel branch of the captured graph, forward and backward)
        lengths = [x.shape[-1]]
        for _ in range(1, self.num_feature_levels):
            lengths.append((lengths[-1] - 1) // 2 + 1)
        side = main = None
        if x.is_cuda and PARALLEL_POSITION_EMBEDDING:
            main = torch.cuda.current_stream(x.device)
            side = branch_stream(x.device, "position_embedding")
            side.wait_stream(main)
        with (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
            dur = self.pos_embed.duration_embedding(duration, mask.device)
            masks = [mask] + [F.interpolate(mask[None].float(), size=(t,)).to(torch.bool)[0] for t in lengths[1:]]
            poses = [self.pos_embed(m, duration, dur) for m in masks]
        srcs = [self.input_proj[0](x)]
        for lvl in range(1, self.num_feature_levels):
            srcs.append(self.input_proj[lvl](x if lvl == 1 else srcs[-1]))
            assert srcs[-1].shape[-1] == lengths[lvl]
        if side is not None:
            main.wait_stream(side)
            if not torch.cuda.is_current_stream_capturing():      # eager: the allocator must know the main stream reads them
                for t in poses + masks:
                    t.record_stream(main)
        poses = [p.to(src.dtype) for p, src in zip(poses, srcs)]
        return srcs, masks, poses


class MLP(nn.Module):
    def __init__(self, input_dim, hidden_dim, output_dim, num_layers):
        super().__init__()
        self.num_layers = num_layers
        dims = [input_dim] + [hidden_dim] * (num_layers - 1) + [output_dim]
        self.layers = nn.ModuleList(nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:]))

    def forward(self, x):
        for i, layer in enumerate(self.layers):
            # tcgen05 projection kernel where it applies (fp32 CUDA, widths multiples of 8), else nn.Linear
            x = _linear(x, layer) if USE_PROJ_MLP else layer(x)
            if i < self.num_layers - 1:
                x = F.relu(x)
        return x


class ShowAttendTellCore(nn.Module):
    """One word step: deformable sampling of the encoder memory around the event, soft attention over the L*P
    samples, LSTM cell (LSTM_DSA.py:195-263)."""

    def __init__(self, cfg):
        super().__init__()
        self.n_levels = cfg.cap_num_feature_levels
        self.n_heads = cfg.cap_nheads
        self.n_points = cfg.cap_dec_n_points
        self.att_feat_size = cfg.hidden_dim // cfg.cap_nheads
        self.att_hid_size = cfg.att_hid_size
        self.rnn = nn.LSTM(cfg.input_encoding_size + cfg.hidden_dim * 2, cfg.rnn_size, cfg.num_layers, bias=False,
                           dropout=cfg.drop_prob if cfg.num_layers > 1 else 0.0)
        self.att_drop = nn.Dropout(0.5)   # present (and unused) in the reference as well, LSTM_DSA.py:212
        self.deformable_att = MSDeformAttnCap(cfg.hidden_dim, self.n_levels, self.n_heads, self.n_points)
        if self.att_hid_size > 0:
            self.ctx2att = nn.Linear(self.att_feat_size, self.att_hid_size)
            self.h2att = nn.Linear(cfg.rnn_size, self.att_hid_size)
            self.alpha_net = nn.Linear(self.att_hid_size, 1)

    def forward(self, xt, state, query, reference_points, memory, temporal_shapes, level_start_index, padding_mask,
                projected_value=None, step_linears=None):
        """`step_linears`: {'h2att', 'sampling_offsets'} -> functions.StepLinear standing in for the two Linear layers on
        the recurrence (same weights; SentenceDecoder makes them per decoding run)."""
        n, lq = reference_points.shape[:2]
        k = self.n_levels * self.n_points
        # the top layer's hidden state, taken ONCE: every `state[0][-1]` is a select whose backward is a zero-fill + copy +
        # accumulation on the recurrence's critical path; StepLSTM hands the state over as 2-D tensors (no select at all)
        h_prev = state[0] if state[0].dim() == 2 else state[0][-1]
        h2att = step_linears["h2att"] if step_linears is not None else self.h2att
        # h2att(h) and the offsets -> sampler chain both hang on h only: h2att runs on a side stream beside the chain
        hproj = side = main = None
        if PARALLEL_H2ATT and step_linears is not None and h_prev.is_cuda and self.att_hid_size > 0:
            main = torch.cuda.current_stream(h_prev.device)
            side = branch_stream(h_prev.device, "h2att")
            side.wait_stream(main)
            with torch.cuda.stream(side):
                hproj = h2att(h_prev)
        joint_query = torch.cat((h_prev.unsqueeze(0), query), 2)
        extra = {}
        if step_linears is not None:
            extra["sampling_offsets"] = step_linears["sampling_offsets"](joint_query)
            if "sample_run" in step_linears and getattr(self.deformable_att, "supports_rows_layout", False):
                extra["sample_run"] = step_linears["sample_run"]
        # the reference takes (N*M, D', Lq, L, P) and permutes it to (N, Lq, M, L*P, D') (LSTM_DSA.py:239-243); the CUDA
        # module hands the samples over in that layout directly (no transposing copies)
        if getattr(self.deformable_att, "supports_rows_layout", False):
            clip = self.deformable_att(joint_query, reference_points, memory, temporal_shapes, level_start_index,
                                       padding_mask, projected_value=projected_value, rows_layout=True, **extra)
        else:
            clip = self.deformable_att(joint_query, reference_points, memory, temporal_shapes, level_start_index,
                                       padding_mask, projected_value=projected_value, **extra)   # (N*M, D', Lq, L, P)
            clip = clip.reshape(n, self.n_heads, -1, lq, k).permute(0, 3, 1, 4, 2)        # (N, Lq, M, L*P, D')
        clip = clip.reshape(n * lq, self.n_heads, k, -1)
        if side is not None:
            main.wait_stream(side)
            if not torch.cuda.is_current_stream_capturing():
                hproj.record_stream(main)
        if clip.shape[-1] == self.att_feat_size + self.att_hid_size:
            # `projected_value` carried ctx2att(value) next to value (SentenceDecoder): interpolation weights sum to 1 in
            # border mode, so sampling the projected rows IS ctx2att of the sampled rows (LSTM_DSA.py:247) without the
            # per-word GEMM, its two backward GEMMs and the per-word weight-gradient accumulation
            if USE_FUSED_ATTEND and CaptionAttendFunction.supported(clip, k, self.att_feat_size):
                attended = CaptionAttendFunction.apply(clip.reshape(n * lq * self.n_heads, k, -1),
                                                       hproj if hproj is not None else h2att(h_prev),
                                                       self.alpha_net.weight, self.alpha_net.bias, self.att_feat_size,
                                                       step_linears.get("attend_run") if step_linears else None)
                clip = None
            else:
                clip, ctx = clip[..., :self.att_feat_size], clip[..., self.att_feat_size:]
        else:
            ctx = self.ctx2att(clip)
        if clip is not None:
            scores = ctx + (hproj if hproj is not None else h2att(h_prev))[:, None, None, :]
            scores = self.alpha_net(torch.tanh(scores)).view(-1, k)
            weight = F.softmax(scores, dim=1)
            attended = torch.bmm(weight.unsqueeze(1), clip.reshape(-1, k, self.att_feat_size)).squeeze(1)
        attended = attended.reshape(n * lq, self.n_heads * self.att_feat_size)
        if step_linears is not None and "lstm" in step_linears:
            return step_linears["lstm"](attended, state)
        if state[0].dim() == 2:
            state = (state[0].unsqueeze(0), state[1].unsqueeze(0))
        step_in = torch.cat((xt.unsqueeze(0), attended.unsqueeze(0), query), 2)
        out, state = self.rnn(step_in, state)
        return out.squeeze(0), state


class StepLSTM:
    """The one-layer nn.LSTM of the word step (LSTM_DSA.py:206, :261) over ONE teacher-forced decoding run, same weights:
    the columns of W_ih that multiply the word embedding and the event query do not depend on the recurrence, so their
    products are formed for all word steps in one GEMM up front; per step only [attended, h] x [W_ih(attended) | W_hh]
    (few-row kernel, weight gradient deferred: functions.StepLinear) and the fused cell kernel (functions.LSTMCellFunction)
    remain -- fp32 throughout, where cuDNN's single-step LSTM spends ~5 kernels forward and ~10 backward per word."""

    def __init__(self, rnn, embedded, query):
        w_ih, w_hh = rnn.weight_ih_l0, rnn.weight_hh_l0
        e, c = embedded.shape[-1], query.shape[-1]
        with torch.autocast("cuda", enabled=False):
            pre = F.linear(embedded.float(), w_ih[:, :e]) + F.linear(query[0].float(), w_ih[:, e + c:])[:, None, :]
            recurrent = torch.cat((w_ih[:, e:e + c], w_hh), 1)
        self.pre = pre.unbind(1)                               # (pairs, 4H) per word step
        self.rec = StepLinear(SimpleNamespace(weight=recurrent, bias=None))
        self.i = 0

    @staticmethod
    def supported(rnn):
        return rnn.num_layers == 1 and not rnn.bias and not rnn.bidirectional and rnn.weight_ih_l0.dtype == torch.float32

    def __call__(self, attended, state):
        """state: (h, c) as (pairs, H) tensors -- or nn.LSTM's (1, pairs, H) on the first step; returned 2-D."""
        h_prev, c_prev = (state[0], state[1]) if state[0].dim() == 2 else (state[0][-1], state[1][-1])
        rec = self.rec(torch.cat((attended.float(), h_prev.float()), 1))
        h, c = LSTMCellFunction.apply(self.pre[self.i], rec, c_prev)
        self.i += 1
        return h, (h, c)


class LSTMDSACaptioner(nn.Module):
    """Teacher-forced sentence decoder for the matched (query, caption) pairs of ONE video
    (LSTM_DSA.py:17-109; the reference asserts one video per call too, :59)."""

    def __init__(self, cfg):
        super().__init__()
        self.vocab_size = cfg.vocab_size
        self.rnn_size = cfg.rnn_size
        self.num_layers = cfg.num_layers
        self.max_caption_len = cfg.max_caption_len
        self.embed = nn.Embedding(self.vocab_size + 1, cfg.input_encoding_size)
        self.logit = nn.Linear(self.rnn_size, self.vocab_size + 1)
        self.dropout = nn.Dropout(cfg.drop_prob)
        self.core = ShowAttendTellCore(cfg)
        self.embed.weight.data.uniform_(-0.1, 0.1)
        self.logit.bias.data.fill_(0)
        self.logit.weight.data.uniform_(-0.1, 0.1)

    def build_loss(self, logprobs, target, mask):
        """Masked token NLL per sentence (LSTM_DSA.py:48-52); a gather replaces the one-hot product."""
        steps = logprobs.shape[1]
        picked = logprobs.gather(2, target[:, :steps, None].long()).squeeze(2)
        return -(picked * mask[:, :steps]).sum(1) / (mask.sum(1) + 1e-6)

    def forward(self, hs, reference, others, cap_tensor, n_steps=None):
        """hs (1, pairs, C), reference (1, pairs, 1|2), cap_tensor (pairs, words).  `n_steps` = number of word steps
        to run; the reference stops at the first column >= 1 whose tokens are all 0 (:88-89) -- the caller computes
        that on the host copy of the captions so the loop does not read the device."""
        assert hs.shape[0] == 1, "the standard captioner handles one video per call (LSTM_DSA.py:59)"
        valid = others['valid_ratios']
        if reference.shape[-1] == 2:
            ref = reference[:, :, None] * torch.stack([valid, valid], -1)[:, None]
        else:
            ref = reference[:, :, None] * valid[:, None, :, None]
        memory, shapes, starts, mask = (others['memory'], others['spatial_shapes'], others['level_start_index'],
                                        others['mask_flatten'])
        core = self.core
        if core.n_levels < shapes.shape[0]:                       # LSTM_DSA.py:80-86
            lengths = others['level_lengths'][:core.n_levels]
            shapes, starts = shapes[:core.n_levels], starts[:core.n_levels]
            memory, mask, ref = memory[:, :sum(lengths)], mask[:, :sum(lengths)], ref[:, :, :core.n_levels]
        seq = cap_tensor.long()
        if n_steps is None:
            n_steps = caption_steps(seq.cpu().numpy())
        decoder = self.begin(hs, ref, memory, shapes, starts, mask, seq)
        for i in range(n_steps):
            decoder.step(i)
        return decoder.result()

    def prepare(self, hs, reference, others, cap_tensor):
        """Everything `forward` does before its word loop, as a SentenceDecoder whose steps the caller drives (the
        two-graph trainer interleaves the word steps of all decoder layers across streams, pdvc/graphed.py)."""
        assert self.core.n_levels == others['spatial_shapes'].shape[0], "level-truncated caption heads go through forward()"
        valid = others['valid_ratios']
        if reference.shape[-1] == 2:
            ref = reference[:, :, None] * torch.stack([valid, valid], -1)[:, None]
        else:
            ref = reference[:, :, None] * valid[:, None, :, None]
        return self.begin(hs, ref, others['memory'], others['spatial_shapes'], others['level_start_index'],
                          others['mask_flatten'], cap_tensor.long())

    def begin(self, hs, ref, memory, shapes, starts, mask, seq):
        return SentenceDecoder(self, hs, ref, memory, shapes, starts, mask, seq)


class SentenceDecoder:
    """State of one teacher-forced decoding run: LSTM state, projected memory, log-probabilities so far."""

    def __init__(self, head, hs, ref, memory, shapes, starts, mask, seq):
        self.head, self.hs, self.ref, self.memory, self.shapes, self.starts, self.mask, self.seq = (
            head, hs, ref, memory, shapes, starts, mask, seq)
        pairs = hs.shape[1]
        self.state = (hs.new_zeros(head.num_layers, pairs, head.rnn_size), hs.new_zeros(head.num_layers, pairs, head.rnn_size))
        core = head.core
        value = core.deformable_att.project_value(memory, mask)                           # once per sentence, not per word
        if value.is_cuda and core.att_hid_size > 0 and (core.deformable_att.padding_mode or get_default_padding_mode()) == "border":
            value = torch.cat((value, core.ctx2att(value)), -1)                           # see ShowAttendTellCore.forward
            if USE_FP32_SAMPLES and USE_FUSED_ATTEND and value.dtype == torch.bfloat16:
                value = value.float()       # under bf16 autocast: sample in fp32 so that the fused attention kernel applies
        self.value = value.contiguous()
        self.embedded = head.embed(seq)                     # teacher forcing: every input word is known up front
        self.step_linears = None
        if USE_STEP_LINEAR and value.is_cuda and core.att_hid_size > 0 and core.h2att.weight.dtype == torch.float32:
            self.step_linears = {"h2att": StepLinear(core.h2att),
                                 "sampling_offsets": StepLinear(core.deformable_att.sampling_offsets)}
            if USE_STEP_LSTM and StepLSTM.supported(core.rnn):
                self.step_linears["lstm"] = StepLSTM(core.rnn, self.embedded, hs)
            if USE_STEP_SAMPLE_RUN and self.value.dtype == torch.float32:
                self.step_linears["sample_run"] = StepSampleRun()
                self.step_linears["attend_run"] = StepAttendRun()
        self.outputs = []

    def step(self, i):
        head = self.head
        out, self.state = head.core(self.embedded[:, i], self.state, self.hs, self.ref, self.memory, self.shapes,
                                    self.starts, self.mask, projected_value=self.value, step_linears=self.step_linears)
        self.outputs.append(out)

    def result(self):
        """Log-probabilities (pairs, steps, vocab+1).  Under teacher forcing the vocabulary projection never feeds the
        recurrence, so LSTM_DSA.py:95-96 (dropout, logit, log_softmax per word) runs once over all word steps."""
        head = self.head
        return F.log_softmax(head.logit(head.dropout(torch.stack(self.outputs, 1))), dim=2)


def caption_steps(seq_host):
    """Number of word steps the reference's loop runs for these captions (LSTM_DSA.py:74-92): columns
    0 .. words-2, stopping before the first column i >= 1 that is all padding."""
    words = seq_host.shape[1]
    for i in range(1, words - 1):
        if int(seq_host[:, i].sum()) == 0:
            return i
    return words - 1


class PDVC(nn.Module):
    """pdvc/pdvc.py:33-170, :236-323 -- training path with learnt queries (`transformer_input_type='queries'`), the
    path every reference config trains with."""

    def __init__(self, cfg: PDVCConfig):
        super().__init__()
        self.cfg = cfg
        self.base_encoder = BaseEncoder(cfg.num_feature_levels, cfg.feature_dim, cfg.hidden_dim)
        self.transformer = DeformableTransformer(
            d_model=cfg.hidden_dim, nhead=cfg.nheads, num_encoder_layers=cfg.enc_layers,
            num_decoder_layers=cfg.dec_layers, dim_feedforward=cfg.transformer_ff_dim,
            dropout=cfg.transformer_dropout_prob, return_intermediate_dec=True,
            num_feature_levels=cfg.num_feature_levels, dec_n_points=cfg.dec_n_points, enc_n_points=cfg.enc_n_points)
        hidden, n_pred = cfg.hidden_dim, cfg.dec_layers
        self.query_embed = nn.Embedding(cfg.num_queries, hidden * 2)
        class_head = nn.Linear(hidden, cfg.num_classes)
        count_head = nn.Linear(hidden, cfg.max_eseq_length + 1)
        bbox_head = MLP(hidden, hidden, 2, 3)
        class_head.bias.data = torch.ones(cfg.num_classes) * (-math.log((1 - 0.01) / 0.01))   # prior prob 0.01
        nn.init.constant_(bbox_head.layers[-1].weight.data, 0)
        nn.init.constant_(bbox_head.layers[-1].bias.data, 0)
        captioner = LSTMDSACaptioner(cfg)
        if cfg.share_caption_head:
            self.caption_head = nn.ModuleList([captioner for _ in range(n_pred)])
        else:
            self.caption_head = nn.ModuleList([copy.deepcopy(captioner) for _ in range(n_pred)])
        if cfg.with_box_refine:
            self.class_head = nn.ModuleList([copy.deepcopy(class_head) for _ in range(n_pred)])
            self.count_head = nn.ModuleList([copy.deepcopy(count_head) for _ in range(n_pred)])
            self.bbox_head = nn.ModuleList([copy.deepcopy(bbox_head) for _ in range(n_pred)])
            nn.init.constant_(self.bbox_head[0].layers[-1].bias.data[1:], -2)
            self.transformer.decoder.bbox_head = self.bbox_head
        else:
            nn.init.constant_(bbox_head.layers[-1].bias.data[1:], -2)
            self.class_head = nn.ModuleList([class_head for _ in range(n_pred)])
            self.count_head = nn.ModuleList([count_head for _ in range(n_pred)])
            self.bbox_head = nn.ModuleList([bbox_head for _ in range(n_pred)])
            self.transformer.decoder.bbox_head = None

    def forward_features(self, vf, valid_mask, duration):
        """Features -> pyramid -> encoder -> decoder (pdvc.py:123-157).  Tensors in, tensors out (+ `others`)."""
        mask = ~valid_mask
        srcs, masks, pos = self.base_encoder(vf, mask, duration)
        tr = self.transformer
        src, shapes, starts, valid_ratios, pos_flat, mask_flat, lengths = tr.prepare_encoder_inputs(srcs, masks, pos)
        memory = tr.forward_encoder(src, shapes, starts, valid_ratios, pos_flat, mask_flat, lengths)
        query_mask = torch.ones(vf.shape[0], self.cfg.num_queries, dtype=torch.bool, device=vf.device)
        init_reference, tgt, reference_points, query_pos = tr.prepare_decoder_input_query(memory, self.query_embed.weight)
        hs, inter_references = tr.forward_decoder(tgt, reference_points, memory, shapes, starts, valid_ratios,
                                                  query_pos, mask_flat, query_mask, False)
        others = {'memory': memory, 'mask_flatten': mask_flat, 'spatial_shapes': shapes, 'level_start_index': starts,
                  'valid_ratios': valid_ratios, 'proposals_mask': query_mask, 'level_lengths': lengths}
        return hs, init_reference, inter_references, others

    def forward_heads(self, hs, init_reference, inter_references):
        """Class / count / segment heads on every decoder layer (pdvc.py:246-266) -> stacked (Ld, N, ...) tensors."""
        n_layers = hs.shape[0]
        logits, counts, boxes = [None] * n_layers, [None] * n_layers, [None] * n_layers
        # unbind, not hs[lid]: one stack in backward instead of a zero-fill + copy + accumulation per layer
        layers, refs = hs.unbind(0), (inter_references.unbind(0) if n_layers > 1 else ())
        # the layers' heads are independent of each other: one stream per layer (parallel branches of the captured graph)
        side, main = [], None
        if hs.is_cuda and PARALLEL_HEADS and n_layers > 1:
            main = torch.cuda.current_stream(hs.device)
            side = [branch_stream(hs.device, "heads", i) for i in range(n_layers - 1)]
            for s_ in side:
                s_.wait_stream(main)
        for lid in range(n_layers):
            with (torch.cuda.stream(side[lid - 1]) if (side and lid > 0) else contextlib.nullcontext()):
                state = layers[lid]
                reference = init_reference if lid == 0 else refs[lid - 1]
                delta = self.bbox_head[lid](state)
                unact = inverse_sigmoid(reference)
                if reference.shape[-1] == 2:
                    delta = delta + unact
                else:
                    delta = torch.cat((delta[..., :1] + unact, delta[..., 1:]), -1)
                logits[lid] = self.class_head[lid](state)
                counts[lid] = self.count_head[lid](state.max(dim=1)[0])
                boxes[lid] = delta.sigmoid()
        for s_ in side:
            main.wait_stream(s_)
        if side and not torch.cuda.is_current_stream_capturing():
            for t in logits[1:] + counts[1:] + boxes[1:]:
                t.record_stream(main)
        return torch.stack(logits), torch.stack(counts), torch.stack(boxes)

    def forward(self, dt, criterion):
        """dt: the reference's batch dict (data/video_dataset.py collate_fn): video_tensor (N,T,C), video_mask (N,T)
        True = valid, video_length (N,3) [frames, seconds, n_gt], video_target [{boxes (G,2), labels (G,)}],
        cap_tensor (G_total, words), cap_mask; optional cap_tensor_host.  Returns (outputs, loss dict)."""
        hs, init_reference, inter_references, others = self.forward_features(
            dt['video_tensor'], dt['video_mask'], dt['video_length'][:, 1])
        return self.matched_prediction(dt, criterion, hs, init_reference, inter_references, others)

    def matched_prediction(self, dt, criterion, hs, init_reference, inter_references, others):
        """Heads on every decoder layer, set losses with Hungarian matching, then the caption loss of the matched
        pairs per layer (pdvc.py:236-323)."""
        n_pred = hs.shape[0]
        logits, counts, boxes = self.forward_heads(hs, init_reference, inter_references)
        references = [init_reference if lid == 0 else inter_references[lid - 1] for lid in range(n_pred)]
        out = {'pred_logits': logits[-1], 'pred_count': counts[-1], 'pred_boxes': boxes[-1]}
        if self.cfg.aux_loss:
            out['aux_outputs'] = [{'pred_logits': logits[i], 'pred_count': counts[i], 'pred_boxes': boxes[i]}
                                  for i in range(n_pred - 1)]
        loss, last_indices, aux_indices = criterion(out, dt['video_target'])
        cap_host = dt.get('cap_tensor_host')          # numpy / host copy of the captions, if the loader kept one
        cap_host = dt['cap_tensor'].cpu().numpy() if cap_host is None else np.asarray(cap_host)
        layers = range(n_pred) if self.cfg.aux_loss else [n_pred - 1]
        for lid in layers:
            indices = last_indices if lid == n_pred - 1 else aux_indices[lid]
            cap_loss = self.caption_loss(self.caption_head[lid], dt, hs[lid], references[lid], others, indices, cap_host)
            loss['loss_caption' if lid == n_pred - 1 else f'loss_caption_{lid}'] = cap_loss
        return out, loss

    def caption_loss(self, cap_head, dt, hs, reference, others, indices, cap_host):
        """pdvc.py:325-434 with matched indices: run the captioner on the matched queries against their captions."""
        assert hs.shape[0] == 1 and len(indices) == 1
        query_ids, cap_ids = indices[0]
        q = query_ids.to(hs.device, non_blocking=True)
        c = cap_ids.to(hs.device, non_blocking=True)
        seq = dt['cap_tensor'][c]
        logprobs = cap_head(hs[:, q], reference[:, q], others, seq, n_steps=caption_steps(cap_host[cap_ids.numpy()]))
        target = dt['cap_tensor'][:, 1:][c]
        target_mask = dt['cap_mask'][:, 1:][c]
        return cap_head.build_loss(logprobs, target, target_mask).mean()


def build(cfg: PDVCConfig):
    """Counterpart of pdvc/pdvc.py:559-604: model, criterion and its loss weights."""
    from .criterion import HungarianMatcher, SetCriterion
    model = PDVC(cfg)
    weight_dict = {'loss_ce': cfg.cls_loss_coef, 'loss_bbox': cfg.bbox_loss_coef, 'loss_giou': cfg.giou_loss_coef,
                   'loss_counter': cfg.count_loss_coef, 'loss_caption': cfg.caption_loss_coef}
    if cfg.aux_loss:
        for i in range(cfg.dec_layers - 1):
            weight_dict.update({f'{k}_{i}': v for k, v in list(weight_dict.items()) if not k[-1].isdigit()})
    matcher = HungarianMatcher(cfg.set_cost_class, cfg.set_cost_bbox, cfg.set_cost_giou, cfg.cost_alpha, cfg.cost_gamma)
    criterion = SetCriterion(cfg.num_classes, matcher, weight_dict, cfg.focal_alpha, cfg.focal_gamma,
                             cfg.lloss_gau_mask, cfg.lloss_beta)
    return model, criterion
