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
