"""The callers of the hot path: PDVC's temporal deformable transformer.

Restated from the behaviour of `pdvc/deformable_transformer.py` (encoder layer :162-189, encoder and its
level-grid reference points :192-216, decoder layer :219-271, decoder with iterative refinement :274-325,
input preparation :84-141) with the same module / parameter names, so a reference checkpoint's `transformer.*`
entries load unchanged.  Every MSDeformAttn here is the fused CUDA module of this package; the rest is
cuBLAS / ATen plumbing (out of the hot path, kept only so that the training metric of BASELINE.json can be
measured on the GPU box, where the reference sources do not exist).
"""
import copy
import math

import torch
import torch.nn.functional as F
from torch import nn

from ..ops.modules import MSDeformAttn
from ..ops.modules.ms_deform_attn import _proj

# The feed-forward Linears of the encoder / decoder layers (deformable_transformer.py:173-176, :247-250) on the same
# tcgen05 GEMM kernels as MSDeformAttn's projections (csrc/msda_proj.cuh: bias in the epilogue, bias gradient in the
# weight-gradient kernel) -- for fp32 operands, where that fusion beats cuBLAS' SIMT sgemm + separate bias kernels
# (yc2 shape: 5.87 -> 5.83 ms/step); under bf16 autocast cuBLAS' tensor-core GEMMs are faster for these plain layers
# (7.76 vs 7.82 ms/step, tools/pdvc_ab.py).  False: torch's library GEMMs always.
USE_PROJ_FFN = True


def _linear(x, lin):
    if USE_PROJ_FFN and x.is_cuda and not torch.is_autocast_enabled("cuda"):
        return _proj(x, lin)
    return lin(x)


def inverse_sigmoid(x, eps=1e-5):
    """misc/detr_utils/misc.py:540-544."""
    x = x.clamp(min=0, max=1)
    return torch.log(x.clamp(min=eps) / (1 - x).clamp(min=eps))


def _clones(module, n):
    return nn.ModuleList([copy.deepcopy(module) for _ in range(n)])


class DeformableTransformerEncoderLayer(nn.Module):
    def __init__(self, d_model=256, d_ffn=1024, dropout=0.1, n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        self.self_attn = MSDeformAttn(d_model, n_levels, n_heads, n_points)
        self.dropout1 = nn.Dropout(dropout)
        self.norm1 = nn.LayerNorm(d_model)
        self.linear1 = nn.Linear(d_model, d_ffn)
        self.dropout2 = nn.Dropout(dropout)
        self.linear2 = nn.Linear(d_ffn, d_model)
        self.dropout3 = nn.Dropout(dropout)
        self.norm2 = nn.LayerNorm(d_model)

    def forward(self, src, pos, reference_points, temporal_shapes, level_start_index, padding_mask=None):
        query = src if pos is None else src + pos
        attended = self.self_attn(query, reference_points, src, temporal_shapes, level_start_index, padding_mask)
        src = self.norm1(src + self.dropout1(attended))
        hidden = _linear(self.dropout2(F.relu(_linear(src, self.linear1))), self.linear2)
        return self.norm2(src + self.dropout3(hidden))


class DeformableTransformerEncoder(nn.Module):
    def __init__(self, encoder_layer, num_layers):
        super().__init__()
        self.layers = _clones(encoder_layer, num_layers)
        self.num_layers = num_layers

    @staticmethod
    def get_reference_points(level_lengths, valid_ratios, device):
        """Centres of every row of every level, (N, S, L, 1) -- deformable_transformer.py:198-208.
        `level_lengths` is a host list (no device read-back)."""
        per_level = []
        for lvl, t in enumerate(level_lengths):
            centres = torch.linspace(0.5, t - 0.5, t, dtype=torch.float32, device=device)
            per_level.append(centres[None] / (valid_ratios[:, None, lvl] * t))
        ref = torch.cat(per_level, 1)
        return (ref[:, :, None] * valid_ratios[:, None])[..., None]

    def forward(self, src, temporal_shapes, level_start_index, valid_ratios, pos=None, padding_mask=None,
                level_lengths=None):
        if level_lengths is None:
            level_lengths = [int(t) for t in temporal_shapes.tolist()]
        reference_points = self.get_reference_points(level_lengths, valid_ratios, src.device)
        out = src
        for layer in self.layers:
            out = layer(out, pos, reference_points, temporal_shapes, level_start_index, padding_mask)
        return out


class DeformableTransformerDecoderLayer(nn.Module):
    def __init__(self, d_model=256, d_ffn=1024, dropout=0.1, n_levels=4, n_heads=8, n_points=4):
        super().__init__()
        self.cross_attn = MSDeformAttn(d_model, n_levels, n_heads, n_points)
        self.dropout1 = nn.Dropout(dropout)
        self.norm1 = nn.LayerNorm(d_model)
        self.self_attn = nn.MultiheadAttention(d_model, n_heads, dropout=dropout)
        self.dropout2 = nn.Dropout(dropout)
        self.norm2 = nn.LayerNorm(d_model)
        self.linear1 = nn.Linear(d_model, d_ffn)
        self.dropout3 = nn.Dropout(dropout)
        self.linear2 = nn.Linear(d_ffn, d_model)
        self.dropout4 = nn.Dropout(dropout)
        self.norm3 = nn.LayerNorm(d_model)

    def forward(self, tgt, query_pos, reference_points, src, src_temporal_shapes, level_start_index,
                src_padding_mask=None, query_mask=None):
        qk = (tgt if query_pos is None else tgt + query_pos).transpose(0, 1)
        key_padding = None if query_mask is None else ~query_mask
        # (need_weights=False -- torch's fused scaled-dot-product attention instead of bmm / softmax / dropout / bmm -- was
        # measured: 4.16 -> 4.20 ms/step at 100 queries, not faster; profiles/r2_notes.md)
        attended = self.self_attn(qk, qk, tgt.transpose(0, 1), key_padding_mask=key_padding)[0].transpose(0, 1)
        tgt = self.norm2(tgt + self.dropout2(attended))
        query = tgt if query_pos is None else tgt + query_pos
        sampled = self.cross_attn(query, reference_points, src, src_temporal_shapes, level_start_index, src_padding_mask)
        tgt = self.norm1(tgt + self.dropout1(sampled))
        hidden = _linear(self.dropout3(F.relu(_linear(tgt, self.linear1))), self.linear2)
        return self.norm3(tgt + self.dropout4(hidden))


class DeformableTransformerDecoder(nn.Module):
    def __init__(self, decoder_layer, num_layers, return_intermediate=False):
        super().__init__()
        self.layers = _clones(decoder_layer, num_layers)
        self.num_layers = num_layers
        self.return_intermediate = return_intermediate
        self.bbox_head = None   # set by the model for iterative refinement (pdvc.py:93-99)

    def forward(self, tgt, reference_points, src, src_temporal_shapes, src_level_start_index, src_valid_ratios,
                query_pos=None, src_padding_mask=None, query_padding_mask=None, disable_iterative_refine=False):
        out = tgt
        states, references = [], []
        for lid, layer in enumerate(self.layers):
            if reference_points.shape[-1] == 2:       # (centre, length) boxes after refinement, :290-292
                ref_in = reference_points[:, :, None] * torch.stack([src_valid_ratios, src_valid_ratios], -1)[:, None]
            else:                                      # centres only, :294-295
                ref_in = reference_points[:, :, None] * src_valid_ratios[:, None, :, None]
            out = layer(out, query_pos, ref_in, src, src_temporal_shapes, src_level_start_index, src_padding_mask,
                        query_padding_mask)
            if not disable_iterative_refine and self.bbox_head is not None:   # :300-314
                delta = self.bbox_head[lid](out)
                if reference_points.shape[-1] == 2:
                    refined = delta + inverse_sigmoid(reference_points)
                else:
                    refined = torch.cat((delta[..., :1] + inverse_sigmoid(reference_points), delta[..., 1:]), -1)
                reference_points = refined.sigmoid().detach()
            if self.return_intermediate:
                states.append(out)
                references.append(reference_points)
        if self.return_intermediate:
            return torch.stack(states), torch.stack(references)
        return out, reference_points


class DeformableTransformer(nn.Module):
    def __init__(self, d_model=256, nhead=8, num_encoder_layers=6, num_decoder_layers=6, dim_feedforward=1024,
                 dropout=0.1, return_intermediate_dec=False, num_feature_levels=4, dec_n_points=4, enc_n_points=4):
        super().__init__()
        self.d_model = d_model
        self.nhead = nhead
        self.no_encoder = num_encoder_layers == 0
        self.num_feature_levels = num_feature_levels
        self.encoder = DeformableTransformerEncoder(
            DeformableTransformerEncoderLayer(d_model, dim_feedforward, dropout, num_feature_levels, nhead, enc_n_points),
            num_encoder_layers)
        self.decoder = DeformableTransformerDecoder(
            DeformableTransformerDecoderLayer(d_model, dim_feedforward, dropout, num_feature_levels, nhead, dec_n_points),
            num_decoder_layers, return_intermediate_dec)
        self.level_embed = nn.Parameter(torch.Tensor(num_feature_levels, d_model))
        self.pos_trans = nn.Linear(d_model, d_model * 2)
        self.pos_trans_norm = nn.LayerNorm(d_model * 2)
        self.reference_points = nn.Linear(d_model, 1)
        self._reset_parameters()

    def _reset_parameters(self):
        # deformable_transformer.py:53-62
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)
        for m in self.modules():
            if isinstance(m, MSDeformAttn):
                m._reset_parameters()
        nn.init.xavier_uniform_(self.reference_points.weight.data, gain=1.0)
        nn.init.constant_(self.reference_points.bias.data, 0.)
        nn.init.normal_(self.level_embed)

    def level_tensors(self, lengths, device):
        """int64 device tensors (T_l) and (start_l) for a pyramid; cached per (lengths, device) so that steady-state
        steps upload nothing (and the step can be captured into a CUDA graph)."""
        cache = self.__dict__.setdefault("_level_cache", {})
        key = (tuple(lengths), str(device))
        if key not in cache:
            starts = [0]
            for t in lengths[:-1]:
                starts.append(starts[-1] + t)
            cache[key] = (torch.as_tensor(lengths, dtype=torch.long, device=device),
                          torch.as_tensor(starts, dtype=torch.long, device=device))
        return cache[key]

    def prepare_encoder_inputs(self, srcs, masks, pos_embeds):
        """Flatten the pyramid (deformable_transformer.py:84-114).  Also returns the level lengths as a host list:
        they are shapes, known without reading the device."""
        lengths = [int(s.shape[-1]) for s in srcs]
        src_flatten = torch.cat([s.transpose(1, 2) for s in srcs], 1)
        mask_flatten = torch.cat(list(masks), 1)
        pos_flatten = torch.cat([p.transpose(1, 2) + self.level_embed[lvl].view(1, 1, -1)
                                 for lvl, p in enumerate(pos_embeds)], 1)
        temporal_shapes, level_start_index = self.level_tensors(lengths, src_flatten.device)
        valid_ratios = torch.stack([(~m).sum(1).float() / m.shape[1] for m in masks], 1)
        return src_flatten, temporal_shapes, level_start_index, valid_ratios, pos_flatten, mask_flatten, lengths

    def forward_encoder(self, src_flatten, temporal_shapes, level_start_index, valid_ratios, pos_flatten, mask_flatten,
                        level_lengths=None):
        if self.no_encoder:
            return src_flatten
        return self.encoder(src_flatten, temporal_shapes, level_start_index, valid_ratios, pos_flatten, mask_flatten,
                            level_lengths)

    def prepare_decoder_input_query(self, memory, query_embed):
        """:127-134 -- learnt queries; the layer-0 reference points carry a gradient."""
        bs = memory.shape[0]
        query_pos, tgt = torch.chunk(query_embed, 2, dim=1)
        query_pos = query_pos.unsqueeze(0).expand(bs, -1, -1)
        tgt = tgt.unsqueeze(0).expand(bs, -1, -1)
        reference_points = self.reference_points(query_pos).sigmoid()
        return reference_points, tgt, reference_points, query_pos

    def get_proposal_pos_embed(self, proposals):
        """:65-78."""
        num_pos_feats, temperature = 256, 10000
        dim_t = torch.arange(num_pos_feats, dtype=torch.float32, device=proposals.device)
        dim_t = temperature ** (2 * (dim_t // 2) / num_pos_feats)
        pos = (proposals.sigmoid() * (2 * math.pi))[:, :, :, None] / dim_t
        return torch.stack((pos[:, :, :, 0::2].sin(), pos[:, :, :, 1::2].cos()), dim=4).flatten(2)

    def prepare_decoder_input_proposal(self, gt_reference_points):
        """:136-142 -- ground-truth proposals as queries."""
        unact = inverse_sigmoid(gt_reference_points)
        query_pos, tgt = torch.chunk(self.pos_trans_norm(self.pos_trans(self.get_proposal_pos_embed(unact))), 2, dim=2)
        return gt_reference_points, tgt, gt_reference_points, query_pos

    def forward_decoder(self, *args):
        return self.decoder(*args)
