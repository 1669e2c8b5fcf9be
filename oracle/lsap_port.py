"""TEST INFRASTRUCTURE -- CPU restatement of the linear-sum-assignment algorithm behind the reference's Hungarian matcher.

The reference calls scipy.optimize.linear_sum_assignment on every layer's cost matrix (pdvc/matcher.py:115-117); scipy is a
third-party dependency that is not vendored in /root/reference (requirement.txt pins no version; this image has the version
recorded by tests/test_match.py).  Its published algorithm is the shortest-augmenting-path method of D. F. Crouse, "On
implementing 2D rectangular assignment algorithms", IEEE Trans. AES 52(4), 2016 (scipy/optimize/rectangular_lsap): one
Dijkstra-like search per row over the reduced costs `minVal + C[i,j] - u[i] - v[j]`, dual update, path flip.

This file restates it in plain Python exactly as the CUDA kernel (csrc/msda_match.cuh) executes it -- including the two
details that decide the result when costs tie: the order in which the remaining columns are scanned (initially nc-1 ... 0,
swap-removal) and the rule "a strictly smaller value wins; an equal value wins only if its column is unassigned", evaluated
here per lane of 32 and merged with the kernel's comparator.  tests/test_match.py pins it against scipy itself (the oracle
proper for this path); the product never imports it.
"""
import math

import numpy as np

LANES = 32


def _better(a, b):
    """Comparator of the kernel's warp reduction: candidates are (value, unassigned, it).  Reproduces a sequential scan
    in ascending `it` with `v < lowest or (v == lowest and unassigned)`."""
    if a is None:
        return b
    if b is None:
        return a
    if a[0] != b[0]:
        return a if a[0] < b[0] else b
    if a[1] != b[1]:
        return a if a[1] else b
    if a[1]:
        return a if a[2] > b[2] else b       # among unassigned ties the LAST scanned wins
    return a if a[2] < b[2] else b           # among assigned ties the FIRST scanned wins


def solve(cost):
    """cost (R, C) array -> (row_ind, col_ind) as scipy.optimize.linear_sum_assignment returns them; raises ValueError
    like scipy on invalid or infeasible input."""
    cost = np.asarray(cost, dtype=np.float64)
    if cost.ndim != 2:
        raise ValueError("expected a matrix")
    if cost.size == 0:
        return np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64)
    if np.isnan(cost).any() or np.isneginf(cost).any():
        raise ValueError("matrix contains invalid numeric entries")
    transposed = cost.shape[1] < cost.shape[0]
    c = cost.T.copy() if transposed else cost
    nr, nc = c.shape
    u, v = np.zeros(nr), np.zeros(nc)
    col4row = -np.ones(nr, dtype=np.int64)
    row4col = -np.ones(nc, dtype=np.int64)
    for cur in range(nr):
        sp = np.full(nc, np.inf)
        path = -np.ones(nc, dtype=np.int64)
        SR, SC = np.zeros(nr, dtype=bool), np.zeros(nc, dtype=bool)
        remaining = [nc - it - 1 for it in range(nc)]
        n_rem = nc
        min_val, sink, i = 0.0, -1, cur
        while sink == -1:
            SR[i] = True
            best = None
            for lane in range(LANES):                       # lane `lane` scans it = lane, lane + 32, ...
                mine = None
                for it in range(lane, n_rem, LANES):
                    j = remaining[it]
                    r = min_val + c[i, j] - u[i] - v[j]
                    if r < sp[j]:
                        path[j] = i
                        sp[j] = r
                    cand = (sp[j], row4col[j] == -1, it)
                    mine = cand if mine is None else _better(mine, cand)
                best = _better(best, mine)
            if best is None or math.isinf(best[0]):
                raise ValueError("cost matrix is infeasible")
            min_val = best[0]
            j = remaining[best[2]]
            if row4col[j] == -1:
                sink = j
            else:
                i = row4col[j]
            SC[j] = True
            n_rem -= 1
            remaining[best[2]] = remaining[n_rem]
        u[cur] += min_val
        for r_ in range(nr):
            if SR[r_] and r_ != cur:
                u[r_] += min_val - sp[col4row[r_]]
        for j in range(nc):
            if SC[j]:
                v[j] -= min_val - sp[j]
        j = sink
        while True:
            i = path[j]
            row4col[j] = i
            col4row[i], j = j, col4row[i]
            if i == cur:
                break
    if transposed:
        order = np.argsort(col4row, kind="stable")
        return col4row[order].astype(np.int64), order.astype(np.int64)
    return np.arange(nr, dtype=np.int64), col4row.astype(np.int64)
