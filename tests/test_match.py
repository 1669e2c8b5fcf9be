"""Hungarian matching: msda_lsap_f32 / linear_sum_assignment_cuda against the call it replaces,
scipy.optimize.linear_sum_assignment (pdvc/matcher.py:115) -- scipy IS the reference's implementation of this step (an
un-vendored dependency, requirement.txt:13, no version pinned; the version used here is printed), so it is the oracle proper.

  * CPU: the line-by-line restatement of the kernel's algorithm (oracle/lsap_port.py: scan order, tie rule, per-lane scan +
    warp comparator) returns scipy's assignment bit for bit, also on matrices full of ties;
  * GPU: the kernel returns scipy's row_ind / col_ind exactly -- PDVC's shapes (queries x events, one matrix per decoder
    layer), fewer rows than columns, squares, tie-heavy integer costs, +inf entries, matrices too large to stage in shared
    memory -- and flags what scipy raises ValueError for; the PDVC step with device matching reproduces the step with host
    matching.
Index work: the bar is exact equality.
"""
import numpy as np
import pytest
import scipy
import torch
from scipy.optimize import linear_sum_assignment

import cases
from oracle import lsap_port


def _matrix(rs, kind, R, C):
    if kind == "normal":
        return rs.standard_normal((R, C))
    if kind == "ties3":
        return rs.randint(0, 3, (R, C)).astype(np.float64)
    if kind == "ties10":
        return rs.randint(0, 10, (R, C)).astype(np.float64)
    if kind == "coarse":
        return np.round(rs.standard_normal((R, C)), 1)
    if kind == "const":
        return np.zeros((R, C))
    raise KeyError(kind)


KINDS = ["normal", "ties3", "ties10", "coarse", "const"]


def test_port_matches_scipy_including_ties():
    print("scipy", scipy.__version__)
    rs = np.random.RandomState(0)
    for trial in range(400):
        R, C = rs.randint(1, 48), rs.randint(1, 48)
        c = _matrix(rs, KINDS[trial % len(KINDS)], R, C)
        a, b = linear_sum_assignment(c)
        a2, b2 = lsap_port.solve(c)
        assert np.array_equal(a, a2) and np.array_equal(b, b2), (KINDS[trial % len(KINDS)], R, C)


def test_port_error_behaviour_follows_scipy():
    c = np.ones((3, 4))
    c[1, 2] = np.nan
    with pytest.raises(ValueError):
        linear_sum_assignment(c)
    with pytest.raises(ValueError):
        lsap_port.solve(c)
    c = np.full((3, 4), np.inf)
    c[0, 0] = c[2, 1] = 1.0                      # row 1 has no finite entry
    with pytest.raises(ValueError):
        linear_sum_assignment(c)
    with pytest.raises(ValueError):
        lsap_port.solve(c)
    c = np.array([[np.inf, 1.0, 2.0], [0.5, np.inf, np.inf]])     # +inf edges, feasible
    a, b = linear_sum_assignment(c)
    a2, b2 = lsap_port.solve(c)
    assert np.array_equal(a, a2) and np.array_equal(b, b2)


# ------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def dvc():
    import dense_video_captioning_b200 as pkg
    assert torch.cuda.is_available()
    return pkg


def _check_batch(dvc, mats):
    """mats: list of equally shaped float32 matrices -> one batched call, each compared with scipy."""
    from dense_video_captioning_b200.ops.functions import linear_sum_assignment_cuda
    costs = torch.from_numpy(np.stack(mats)).cuda()
    status = torch.full((len(mats),), -1, dtype=torch.int32, device="cuda")
    rows, cols = linear_sum_assignment_cuda(costs, status)
    assert dvc._lib.lib().msda_last_launch_path() == dvc._lib.PATH_MATCH
    rows, cols, status = rows.cpu().numpy(), cols.cpu().numpy(), status.cpu().numpy()
    assert rows.dtype == np.int64 and rows.shape == (len(mats), min(mats[0].shape))
    for b, m in enumerate(mats):
        a, c = linear_sum_assignment(m)          # scipy on the float32 image, as after the reference's C.cpu()
        assert status[b] == 0
        assert np.array_equal(rows[b], a) and np.array_equal(cols[b], c), (b, m.shape, rows[b], a, cols[b], c)


@pytest.mark.gpu
@pytest.mark.parametrize("R,C", [(100, 6), (100, 30), (10, 4), (10, 10), (10, 15), (1, 1), (1, 7), (7, 1), (32, 32), (33, 31),
                                 (300, 100), (64, 200)])
@pytest.mark.parametrize("kind", KINDS)
def test_kernel_matches_scipy(dvc, R, C, kind):
    rs = np.random.RandomState(R * 131 + C)
    _check_batch(dvc, [_matrix(rs, kind, R, C).astype(np.float32) for _ in range(5)])


@pytest.mark.gpu
def test_kernel_matches_scipy_on_pdvc_like_costs(dvc):
    """Costs with the structure of matcher.py:88-100: focal class cost + L1 + GIoU of random segments, 4 decoder layers."""
    rs = np.random.RandomState(3)
    for Q, G in ((100, 6), (100, 17), (10, 4)):
        mats = []
        for _ in range(4):
            prob = rs.uniform(0.01, 0.99, (Q, 1))
            pc, pl = rs.uniform(0, 1, (Q, 1)), rs.uniform(0.02, 0.5, (Q, 1))
            gc, gl = rs.uniform(0, 1, (1, G)), rs.uniform(0.02, 0.5, (1, G))
            cls = 0.25 * (1 - prob) ** 2 * -np.log(prob + 1e-8) - 0.75 * prob ** 2 * -np.log(1 - prob + 1e-8)
            l1 = np.abs(pc - gc) + np.abs(pl - gl)
            lo, hi = np.maximum(pc - pl / 2, gc - gl / 2), np.minimum(pc + pl / 2, gc + gl / 2)
            inter = np.clip(hi - lo, 0, None)
            union = pl + gl - inter
            hull = np.maximum(pc + pl / 2, gc + gl / 2) - np.minimum(pc - pl / 2, gc - gl / 2)
            giou = inter / union - (hull - union) / hull
            mats.append((2 * cls + 0 * l1 - 4 * giou + 5 * l1).astype(np.float32))
        _check_batch(dvc, mats)


@pytest.mark.gpu
def test_kernel_large_matrices_and_limits(dvc):
    from dense_video_captioning_b200.ops.functions import linear_sum_assignment_cuda, lsap_supported
    rs = np.random.RandomState(9)
    _check_batch(dvc, [rs.standard_normal((1024, 24)).astype(np.float32)])                 # staged: 96 KB matrix
    _check_batch(dvc, [rs.standard_normal((400, 400)).astype(np.float32) for _ in range(2)])   # 640 KB: read from global
    _check_batch(dvc, [rs.randint(0, 4, (257, 300)).astype(np.float32)])                    # unstaged, ties, R < C
    big = torch.zeros(1, 1025, 4, device="cuda")
    assert not lsap_supported(big)
    with pytest.raises(RuntimeError):
        linear_sum_assignment_cuda(big)
    lib = dvc._lib.lib()
    assert lib.msda_lsap_f32(big.data_ptr(), big.data_ptr(), big.data_ptr(), None, 1, 1025, 4, None) == -7
    empty_r, empty_c = linear_sum_assignment_cuda(torch.zeros(3, 5, 0, device="cuda"))
    assert empty_r.shape == (3, 0) and empty_c.shape == (3, 0)


@pytest.mark.gpu
def test_kernel_flags_what_scipy_rejects(dvc):
    from dense_video_captioning_b200.ops.functions import linear_sum_assignment_cuda
    ok = np.random.RandomState(1).standard_normal((6, 4)).astype(np.float32)
    nan = ok.copy()
    nan[2, 1] = np.nan
    ninf = ok.copy()
    ninf[0, 0] = -np.inf
    infeasible = np.full((6, 4), np.inf, dtype=np.float32)
    infeasible[:, 0] = 1.0                                    # four columns to assign, only one reachable
    feasible_inf = ok.copy()
    feasible_inf[1, :3] = np.inf
    costs = torch.from_numpy(np.stack([ok, nan, ninf, infeasible, feasible_inf])).cuda()
    status = torch.full((5,), -1, dtype=torch.int32, device="cuda")
    rows, cols = linear_sum_assignment_cuda(costs, status)
    assert status.tolist() == [0, 1, 1, 2, 0]
    for b in (1, 2, 3):
        with pytest.raises(ValueError):
            linear_sum_assignment(costs[b].cpu().numpy())
        assert rows[b].tolist() == [0, 1, 2, 3] and cols[b].tolist() == [0, 1, 2, 3]       # defined output: identity
    for b in (0, 4):
        a, c = linear_sum_assignment(costs[b].cpu().numpy())
        assert rows[b].tolist() == a.tolist() and cols[b].tolist() == c.tolist()


@pytest.mark.gpu
@pytest.mark.parametrize("Q,G,classes,gamma", [(100, 6, 1, 2.0), (100, 30, 1, 2.0), (10, 4, 1, 2.0), (10, 14, 3, 2.0), (50, 9, 2, 1.5)])
def test_fused_cost_and_matching_against_the_torch_expression(dvc, Q, G, classes, gamma):
    """msda_pdvc_match_f32 against HungarianMatcher.cost_matrix (the tensor expression of matcher.py:84-103) + scipy: the cost
    matrix to 2 float32 ulps of its largest entry (each operation is performed in float32 as the expression does; expf / logf /
    powf may differ from the tensor kernels' in the last place), the assignment exactly."""
    from dense_video_captioning_b200.ops.functions import pdvc_match_cuda
    from dense_video_captioning_b200.pdvc.criterion import HungarianMatcher
    g = torch.Generator().manual_seed(Q * 7 + G)
    layers = 3
    logits = (torch.randn(layers, Q, classes, generator=g) * 3).cuda()
    boxes = torch.stack([torch.rand(layers, Q, generator=g), torch.rand(layers, Q, generator=g) * 0.5 + 0.01], -1).cuda()
    labels = torch.randint(0, classes, (G,), generator=g).cuda()
    tboxes = torch.stack([torch.rand(G, generator=g), torch.rand(G, generator=g) * 0.4 + 0.02], -1).cuda()
    m = HungarianMatcher(cost_class=2.0, cost_bbox=0.0 if Q == 10 else 5.0, cost_giou=4.0, cost_alpha=0.25, cost_gamma=gamma)
    cost, q_idx, t_idx = pdvc_match_cuda(logits, boxes, labels, tboxes, m.cost_class, m.cost_bbox, m.cost_giou, m.cost_alpha,
                                         m.cost_gamma)
    assert dvc._lib.lib().msda_last_launch_path() == dvc._lib.PATH_MATCH
    for l in range(layers):
        want = m.cost_matrix(logits[l][None], boxes[l][None], labels, tboxes)
        scale = float(want.abs().max())
        err = float((cost[l] - want).abs().max())
        assert err <= 2 * np.finfo(np.float32).eps * scale, (l, err, scale)
        a, c = linear_sum_assignment(want.cpu().numpy())
        assert q_idx[l].tolist() == a.tolist() and t_idx[l].tolist() == c.tolist()


@pytest.mark.gpu
@pytest.mark.parametrize("graphs", [False, True])
def test_step_with_device_matching_reproduces_host_matching(graphs):
    """The two-stage PDVC step, device matching against host (scipy) matching: same assignment, same losses, same gradients."""
    from conftest import load_golden
    from dense_video_captioning_b200.pdvc import build
    from dense_video_captioning_b200.pdvc.graphed import StagedStep
    from test_pdvc_model import _batch, _cfg
    golden = load_golden("pdvc_yc2shape")
    cfg = _cfg(golden)
    res = {}
    for device_match in (False, True):
        torch.manual_seed(0)
        model, criterion = build(cfg)
        model.cuda().train()
        criterion.cuda()
        cases.fill_parameters(model)
        step = StagedStep(model, criterion, graphs=graphs, device_match=device_match)
        batch = _batch(cfg, torch.float32, "cuda", golden)
        for _ in range(2):                                    # second call: graph replay
            model.zero_grad(set_to_none=True)
            total, terms, (q_idx, t_idx) = step(batch)
            total.backward()
        torch.cuda.synchronize()
        assert torch.is_tensor(q_idx) == device_match
        res[device_match] = (float(total), {k: float(v) for k, v in terms.items()}, np.asarray(q_idx.cpu() if device_match else q_idx),
                             np.asarray(t_idx.cpu() if device_match else t_idx),
                             {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None})
    host, dev = res[False], res[True]
    assert np.array_equal(host[2], dev[2]) and np.array_equal(host[3], dev[3])
    assert dev[2][-1].tolist() == golden["matched_query"].tolist()
    assert abs(host[0] - dev[0]) <= 1e-5 * abs(host[0])
    for k in host[1]:
        assert abs(host[1][k] - dev[1][k]) <= 1e-5 * max(abs(host[1][k]), 1.0), k
    assert set(host[4]) == set(dev[4])
    for n in host[4]:
        # two runs of the same step differ in the order of the backward's fp32 reductions (vector atomics), amplified by
        # the layers below them: up to ~4e-4 of max|g| measured on the input projections; bar = the whole-model tests' 2e-3
        scale = float(host[4][n].abs().max())
        assert float((host[4][n] - dev[4][n]).abs().max()) <= 2e-3 * max(scale, 1e-6), n
