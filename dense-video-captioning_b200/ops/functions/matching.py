"""Hungarian matching on the device: the C-ABI entry point msda_lsap_f32 behind the call the reference makes,
scipy.optimize.linear_sum_assignment (pdvc/matcher.py:115), without its `.cpu()` round trip (matcher.py:113)."""
import torch

from ... import _lib


def lsap_supported(costs):
    """(B, R, C) CUDA tensor the kernel takes (larger matrices stay on the host path)."""
    return costs.is_cuda and costs.dim() == 3 and max(costs.shape[1], costs.shape[2]) <= _lib.LSAP_MAX_DIM


def linear_sum_assignment_cuda(costs, status=None):
    """costs (B, R, C) on the device -> (row_idx, col_idx), each (B, min(R, C)) int64 on the device: for every matrix what
    scipy.optimize.linear_sum_assignment(costs[b]) returns (rows ascending), computed in float64 from the float32 image of
    the costs as the reference's host path does (`C.cpu()` is float32, scipy converts to double).  Nothing is synchronised;
    capturable in a CUDA graph.  `status` (B,) int32, optional: 0 ok, 1 NaN / -inf entries, 2 infeasible -- the cases scipy
    raises ValueError for; the pairing written is then the identity."""
    if not costs.is_cuda:
        raise RuntimeError("linear_sum_assignment_cuda: costs must be a CUDA tensor (no CPU fallback; use scipy on the host)")
    if costs.dim() != 3:
        raise ValueError("costs must be (B, R, C)")
    B, R, C = costs.shape
    cost32 = costs.detach().to(torch.float32).contiguous()
    k = min(R, C)
    rows = torch.empty((B, k), dtype=torch.int64, device=costs.device)
    cols = torch.empty((B, k), dtype=torch.int64, device=costs.device)
    if status is not None and (status.dtype != torch.int32 or status.numel() != B or not status.is_contiguous()):
        raise ValueError("status must be a contiguous (B,) int32 tensor")
    with torch.cuda.device(costs.device):
        rc = _lib.lib().msda_lsap_f32(cost32.data_ptr(), rows.data_ptr(), cols.data_ptr(),
                                      None if status is None else status.data_ptr(), B, R, C,
                                      torch.cuda.current_stream(costs.device).cuda_stream)
    _lib.check(rc, "msda_lsap_f32")
    return rows, cols


def pdvc_match_cuda(pred_logits, pred_boxes, tgt_labels, tgt_boxes, cost_class, cost_bbox, cost_giou, alpha, gamma):
    """HungarianMatcher.forward of one video for all decoder layers in one launch (pdvc/matcher.py:71-117).
    pred_logits (layers, queries, classes), pred_boxes (layers, queries, 2), float32 on the device; tgt_labels (targets,)
    int64, tgt_boxes (targets, 2).  Returns (cost (layers, queries, targets) float32, query_idx, target_idx (layers,
    min(queries, targets)) int64): the cost matrix of matcher.py:84-103 and scipy's assignment on it."""
    for t in (pred_logits, pred_boxes, tgt_labels, tgt_boxes):
        if not t.is_cuda:
            raise RuntimeError("pdvc_match_cuda: tensors must be on a CUDA device (no CPU fallback)")
    if pred_logits.dtype != torch.float32 or pred_boxes.dtype != torch.float32:
        raise ValueError("pdvc_match_cuda takes float32 predictions")
    layers, queries, classes = pred_logits.shape
    targets = tgt_boxes.shape[0]
    if tuple(pred_boxes.shape) != (layers, queries, 2) or tgt_labels.numel() != targets or tgt_boxes.shape[-1] != 2:
        raise ValueError("pdvc_match_cuda: inconsistent shapes")
    logits, boxes = pred_logits.detach().contiguous(), pred_boxes.detach().contiguous()
    labels, tboxes = tgt_labels.detach().to(torch.int64).contiguous(), tgt_boxes.detach().to(torch.float32).contiguous()
    k = min(queries, targets)
    cost = torch.empty((layers, queries, targets), dtype=torch.float32, device=logits.device)
    q_idx = torch.empty((layers, k), dtype=torch.int64, device=logits.device)
    t_idx = torch.empty((layers, k), dtype=torch.int64, device=logits.device)
    with torch.cuda.device(logits.device):
        rc = _lib.lib().msda_pdvc_match_f32(logits.data_ptr(), boxes.data_ptr(), labels.data_ptr(), tboxes.data_ptr(),
                                            cost.data_ptr(), q_idx.data_ptr(), t_idx.data_ptr(), None, layers, queries, classes,
                                            targets, float(cost_class), float(cost_bbox), float(cost_giou), float(alpha),
                                            float(gamma), torch.cuda.current_stream(logits.device).cuda_stream)
    _lib.check(rc, "msda_pdvc_match_f32")
    return cost, q_idx, t_idx
