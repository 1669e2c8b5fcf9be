"""Set-prediction loss of PDVC: Hungarian matching on (focal class cost, L1, 1-D GIoU), focal classification loss,
event-counter loss, L1 + GIoU segment losses.  Restated from pdvc/matcher.py:71-143, pdvc/criterion.py:44-198 and
:200-249, misc/detr_utils/box_ops.py:8-48.

One deliberate host-side change: the reference moves one cost matrix per decoder layer to the CPU
(matcher.py:115, called once per layer from criterion.py:162 and :183); here all layers' matrices go in a single
transfer, so the step has one device read-back for matching instead of `dec_layers`.  The assignment itself is scipy's
`linear_sum_assignment` on the host, as in the reference (the two-stage training step, pdvc/graphed.py, runs the same
algorithm on the device instead: csrc/msda_match.cuh).  The many-to-one assignment the reference also solves
(matcher.py:118-121) never reaches a loss (criterion.py:49, :103 unpack and drop it) and is not computed.
"""
import torch
import torch.nn.functional as F
from scipy.optimize import linear_sum_assignment
from torch import nn

# empirical distribution of events per video used to weight the counter loss (criterion.py:33-41)
COUNTER_CLASS_RATE = [
    0.00000000e+00, 0.00000000e+00, 1.93425917e-01, 4.12129084e-01, 1.88929963e-01, 7.81296833e-02, 5.09541413e-02,
    3.12718553e-02, 1.84833650e-02, 8.39244680e-03, 6.59406534e-03, 4.49595364e-03, 2.19802178e-03, 1.79838146e-03,
    5.99460486e-04, 4.99550405e-04, 4.99550405e-04, 1.99820162e-04, 2.99730243e-04, 3.99640324e-04, 2.99730243e-04,
    0.00000000e+00, 1.99820162e-04, 0.00000000e+00, 0.00000000e+00, 0.00000000e+00, 9.99100809e-05, 9.99100809e-05]


def segments(box):
    """(centre, length) -> (start, end)."""
    c, l = box.unbind(-1)
    return torch.stack((c - 0.5 * l, c + 0.5 * l), -1)


def pairwise_iou(a, b):
    """a (A,2), b (B,2) as (start, end) -> IoU (A,B), union (A,B)."""
    len_a, len_b = a[:, 1] - a[:, 0], b[:, 1] - b[:, 0]
    inter = (torch.min(a[:, None, 1], b[:, 1]) - torch.max(a[:, None, 0], b[:, 0])).clamp(min=0)
    union = len_a[:, None] + len_b - inter
    return inter / (union + 1e-5), union


def pairwise_giou(a, b):
    iou, union = pairwise_iou(a, b)
    hull = (torch.max(a[:, None, 1], b[:, 1]) - torch.min(a[:, None, 0], b[:, 0])).clamp(min=0)
    return iou - (hull - union) / (hull + 1e-5)


class HungarianMatcher(nn.Module):
    def __init__(self, cost_class=1.0, cost_bbox=1.0, cost_giou=1.0, cost_alpha=0.25, cost_gamma=2.0):
        super().__init__()
        assert cost_class != 0 or cost_bbox != 0 or cost_giou != 0, "all costs cant be 0"
        self.cost_class, self.cost_bbox, self.cost_giou = cost_class, cost_bbox, cost_giou
        self.cost_alpha, self.cost_gamma = cost_alpha, cost_gamma

    @torch.no_grad()
    def cost_matrix(self, pred_logits, pred_boxes, tgt_ids, tgt_boxes):
        """(B,Q,C) logits, (B,Q,2) boxes -> (B*Q, G_total) matching cost on the device."""
        prob = pred_logits.flatten(0, 1).sigmoid()
        box = pred_boxes.flatten(0, 1)
        a, g = self.cost_alpha, self.cost_gamma
        neg = (1 - a) * (prob ** g) * (-(1 - prob + 1e-8).log())
        pos = a * ((1 - prob) ** g) * (-(prob + 1e-8).log())
        cost = self.cost_class * (pos[:, tgt_ids] - neg[:, tgt_ids])
        cost = cost + self.cost_bbox * torch.cdist(box, tgt_boxes, p=1)
        cost = cost - self.cost_giou * pairwise_giou(segments(box), segments(tgt_boxes))
        return cost

    @torch.no_grad()
    def forward(self, layer_outputs, targets):
        """layer_outputs: list of dicts with pred_logits / pred_boxes (one per decoder layer).
        Returns, per layer, a list over the batch of (query_idx, target_idx) int64 host tensors."""
        tgt_ids = torch.cat([t["labels"] for t in targets])
        tgt_boxes = torch.cat([t["boxes"] for t in targets])
        sizes = [int(t["boxes"].shape[0]) for t in targets]
        bs, nq = layer_outputs[0]["pred_logits"].shape[:2]
        costs = torch.stack([self.cost_matrix(o["pred_logits"], o["pred_boxes"], tgt_ids, tgt_boxes)
                             for o in layer_outputs]).view(len(layer_outputs), bs, nq, -1).cpu()
        result = []
        for layer_cost in costs:
            per_video = []
            for b, block in enumerate(layer_cost.split(sizes, -1)):
                rows, cols = linear_sum_assignment(block[b])
                per_video.append((torch.as_tensor(rows, dtype=torch.int64), torch.as_tensor(cols, dtype=torch.int64)))
            result.append(per_video)
        return result


def sigmoid_focal_loss(inputs, targets, num_boxes, alpha=0.25, gamma=2.0):
    prob = inputs.sigmoid()
    ce = F.binary_cross_entropy_with_logits(inputs, targets, reduction="none")
    p_t = prob * targets + (1 - prob) * (1 - targets)
    loss = ce * ((1 - p_t) ** gamma)
    if alpha >= 0:
        loss = (alpha * targets + (1 - alpha) * (1 - targets)) * loss
    return loss.mean(1).sum() / num_boxes


def counter_loss(pred_count, target_onehot, weight, gau_mask=1, beta=1.0):
    """BCE over the event-count classes, down-weighting classes near the target with a Gaussian (sigma 2) --
    criterion.py:200-220."""
    k = target_onehot.shape[1]
    grid = torch.arange(k, device=pred_count.device, dtype=torch.float32)
    near = torch.exp(-(grid[:, None] - grid[None, :]) ** 2 / (2 * 2 ** 2))
    near = near[target_onehot.argmax(dim=1)]
    loss = F.binary_cross_entropy_with_logits(pred_count, target_onehot, reduction="none", weight=1 - weight)
    if gau_mask:
        coef = target_onehot + ((1 - near) ** beta) * (1 - target_onehot)
    else:
        coef = torch.ones_like(target_onehot)
    return (loss * coef).mean(1).mean()


class SetCriterion(nn.Module):
    def __init__(self, num_classes, matcher, weight_dict, focal_alpha=0.25, focal_gamma=2.0, gau_mask=1, beta=1.0):
        super().__init__()
        self.num_classes = num_classes
        self.matcher = matcher
        self.weight_dict = weight_dict
        self.focal_alpha, self.focal_gamma = focal_alpha, focal_gamma
        self.gau_mask, self.beta = gau_mask, beta
        self.register_buffer("counter_class_rate", torch.tensor(COUNTER_CLASS_RATE), persistent=False)

    @staticmethod
    def device_indices(indices, device):
        """Host (query_idx, target_idx) pairs -> device index tensors: (batch, query, [target per video])."""
        batch = torch.cat([torch.full_like(src, i) for i, (src, _) in enumerate(indices)]).to(device, non_blocking=True)
        src = torch.cat([src for src, _ in indices]).to(device, non_blocking=True)
        tgt = [j.to(device, non_blocking=True) for _, j in indices]
        return batch, src, tgt

    def loss_labels(self, outputs, targets, dev_idx, sizes, num_boxes):
        logits = outputs['pred_logits']
        dev = logits.device
        batch, src, tgt = dev_idx
        matched = torch.cat([t["labels"][j] for t, j in zip(targets, tgt)])
        classes = torch.full(logits.shape[:2], self.num_classes, dtype=torch.int64, device=dev)
        classes[batch, src] = matched
        onehot = torch.zeros(logits.shape[0], logits.shape[1], logits.shape[2] + 1, dtype=logits.dtype, device=dev)
        onehot.scatter_(2, classes.unsqueeze(-1), 1)
        loss_ce = sigmoid_focal_loss(logits, onehot[:, :, :-1], num_boxes, self.focal_alpha, self.focal_gamma) * logits.shape[1]
        pred_count = outputs['pred_count']
        max_len = pred_count.shape[1] - 1
        count_onehot = torch.zeros_like(pred_count)
        for b, t in enumerate(targets):                       # number of events is a shape: no device read
            count_onehot[b, min(int(t['boxes'].shape[0]), max_len)].fill_(1)   # a fill kernel, not a host copy
        weight = self.counter_class_rate[:max_len + 1].to(dev)
        return {'loss_ce': loss_ce,
                'loss_counter': counter_loss(pred_count, count_onehot, weight, self.gau_mask, self.beta)}

    @torch.no_grad()
    def loss_cardinality(self, outputs, targets, dev_idx, sizes, num_boxes):
        logits = outputs['pred_logits']
        want = torch.stack([torch.full((), float(t["labels"].shape[0]), device=logits.device) for t in targets])
        got = (logits.argmax(-1) != logits.shape[-1] - 1).sum(1)
        return {'cardinality_error': F.l1_loss(got.float(), want)}

    def loss_boxes(self, outputs, targets, dev_idx, sizes, num_boxes):
        batch, src, tgt = dev_idx
        pred = outputs['pred_boxes'][batch, src]
        want = torch.cat([t['boxes'][j] for t, j in zip(targets, tgt)], dim=0)
        losses = {'loss_bbox': F.l1_loss(pred, want, reduction='none').sum() / num_boxes,
                  'loss_giou': (1 - torch.diag(pairwise_giou(segments(pred), segments(want)))).sum() / num_boxes}
        # overlap among the matched predictions of each video -- logged, never weighted (criterion.py:121-129)
        overlap = torch.triu(pairwise_iou(segments(pred), segments(pred))[0], diagonal=1)
        total = 0
        for k, cols in enumerate(overlap.split(sizes, -1)):
            total = total + cols.split(sizes, -2)[k].sum() / (0.5 * sizes[k] * (sizes[k] - 1))
        losses['loss_self_iou'] = total
        return losses

    def layer_losses(self, outputs, targets, dev_idx, sizes, num_boxes):
        """All loss terms of one decoder layer.  dev_idx = device_indices(...) (device tensors), sizes = matched pairs per
        video (host ints), num_boxes a host float or a 0-d device tensor: nothing here reads the device or copies from
        the host, so the function can be captured into a CUDA graph."""
        losses = {}
        losses.update(self.loss_labels(outputs, targets, dev_idx, sizes, num_boxes))
        losses.update(self.loss_boxes(outputs, targets, dev_idx, sizes, num_boxes))
        losses.update(self.loss_cardinality(outputs, targets, dev_idx, sizes, num_boxes))
        return losses

    def _layer_losses(self, outputs, targets, indices, num_boxes):
        dev = outputs['pred_logits'].device
        return self.layer_losses(outputs, targets, self.device_indices(indices, dev), [len(i) for i, _ in indices], num_boxes)

    def forward(self, outputs, targets):
        """-> (losses, indices of the last layer, [indices of each auxiliary layer])  (criterion.py:154-198)."""
        aux = outputs.get('aux_outputs', [])
        dev = outputs['pred_logits'].device
        num_boxes = float(sum(int(t["labels"].shape[0]) for t in targets))
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            total = torch.tensor([num_boxes], dtype=torch.float, device=dev)
            torch.distributed.all_reduce(total)           # the one collective of the reference's PDVC path (:169-171)
            matched = self.matcher([outputs] + list(aux), targets)
            num_boxes = max(total.item() / torch.distributed.get_world_size(), 1.0)
        else:
            matched = self.matcher([outputs] + list(aux), targets)
            num_boxes = max(num_boxes, 1.0)
        last, aux_indices = matched[0], matched[1:]
        outputs['matched_indices'] = last
        losses = self._layer_losses(outputs, targets, last, num_boxes)
        for i, (layer_out, idx) in enumerate(zip(aux, aux_indices)):
            losses.update({f'{k}_{i}': v for k, v in self._layer_losses(layer_out, targets, idx, num_boxes).items()})
        return losses, last, aux_indices
