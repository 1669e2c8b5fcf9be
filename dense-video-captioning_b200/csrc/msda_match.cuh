// msda_match.cuh -- the Hungarian matching of the set-prediction loss on the device.
//
// Replaces the host round trip of pdvc/matcher.py:113-117: the reference copies every decoder layer's (queries x events)
// cost matrix to the host (`C.view(...).cpu()`, :113) and calls scipy.optimize.linear_sum_assignment on it (:115) -- the only
// point of a PDVC training step where the host has to wait for the device.  Here one warp per cost matrix runs the same
// algorithm (shortest augmenting paths over reduced costs with dual variables, D. F. Crouse 2016 -- what scipy implements)
// in float64 like scipy, with scipy's scan order and tie rule, so that the assignment is the one scipy returns also when
// costs tie (oracle/lsap_port.py is the line-by-line CPU restatement; tests/test_match.py pins both against scipy).
//
// One CTA per matrix: four warps form / stage the matrix, then one warp runs the search; all state in shared memory; the cost matrix is staged (transposed so that the scan over
// columns is conflict-free) when it fits.  The scan over the remaining columns is split over the lanes (lane k takes the
// k-th, k+32-th, ... remaining column, four at a time) and merged with a comparator that reproduces the sequential scan.
// A search step is a chain of shared-memory round trips, float64 adds and a 5-round warp reduction, ~0.5 us: the matrices of
// a PDVC step (2 x 100 queries x 6 events, costs formed in the kernel) take ~30 us inside stage 1's graph, 4 x (100 x 30) 78 us,
// 1024 x 64 0.64 ms (tools/match_gap_probe.py) -- latency-bound scalar work; the point is that the host no longer waits.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace msda {
namespace match {

constexpr int kMaxDim = 1024;              // max(rows, cols) of a cost matrix
constexpr unsigned kFullMask = 0xffffffffu;

struct Args {
    float *cost;             // (B, R, C) row-major: read (PlainCost) or written first (PdvcCost)
    long long *row_idx;      // (B, min(R, C)) ascending rows
    long long *col_idx;      // (B, min(R, C)) their columns
    int *status;             // (B) 0 ok, 1 invalid entries (NaN / -inf), 2 infeasible; may be NULL
    int R, C;
    int staged;              // cost matrix staged in shared memory
};

// shared-memory bytes: doubles first, then ints, bytes, and the staged matrix
__host__ __device__ inline size_t state_bytes(int nr, int nc)
{
    size_t b = (size_t)(2 * nc + nr) * sizeof(double) + (size_t)(3 * nc + nr) * sizeof(int) + (size_t)(nc + nr);
    return (b + 15) / 16 * 16;
}

// Candidate of the column scan: reduced cost and a tie rank.  A sequential scan in ascending position `it` keeps its
// current minimum unless `val < lowest || (val == lowest && column unassigned)` (scipy): among equal minima the LAST
// unassigned column wins when there is one, else the FIRST.  Encoded as "larger rank wins": unassigned kRankOpen + it,
// assigned kRankOpen - 1 - it; rank < 0: no candidate.
constexpr int kRankOpen = 1 << 20;
struct Cand {
    double val;
    int rank;
};
__device__ __forceinline__ int cand_rank(int it, bool open) { return open ? kRankOpen + it : kRankOpen - 1 - it; }
__device__ __forceinline__ int cand_pos(int rank) { return rank >= kRankOpen ? rank - kRankOpen : kRankOpen - 1 - rank; }
__device__ __forceinline__ bool better(const Cand &a, const Cand &b)   // a strictly preferable to b
{
    if (b.rank < 0) return a.rank >= 0;
    if (a.rank < 0) return false;
    return a.val < b.val || (a.val == b.val && a.rank > b.rank);
}

// ---- where the costs come from ----------------------------------------------------------------------------------------
struct PlainCost {                                   // the caller's matrix
    __device__ __forceinline__ float operator()(const float *cost, int b, int k, int r, int c) const { return cost[k]; }
    static constexpr bool kWrites = false;
};

// The matching cost of PDVC's set-prediction loss, pdvc/matcher.py:84-103, for one (query r, event c) of decoder layer b:
//     cost_class * (focal_pos - focal_neg)(sigmoid(logit[r, label_c])) + cost_bbox * L1((centre, length)) - cost_giou * GIoU
// evaluated operation by operation in float32 as the reference's tensor expression is (no fused multiply-adds, the same
// expf / logf), so that the matrix -- and with it the assignment -- is the one the torch ops produce.
struct PdvcCost {
    const float *logits;      // (B, R, n_classes)
    const float *boxes;       // (B, R, 2) centre, length
    const long long *labels;  // (C)
    const float *tgt_boxes;   // (C, 2)
    int n_classes, R;
    float w_class, w_bbox, w_giou, alpha, one_minus_alpha, gamma;    // 1 - alpha formed in double by the host, like Python does
    static constexpr bool kWrites = true;

    __device__ __forceinline__ float focal_pow(float x) const { return gamma == 2.f ? __fmul_rn(x, x) : powf(x, gamma); }

    __device__ __forceinline__ float operator()(const float *, int b, int, int r, int c) const
    {
        const float logit = logits[((size_t)b * R + r) * n_classes + (int)labels[c]];
        const float prob = 1.f / (1.f + expf(-logit));
        // neg = (1 - alpha) * prob^gamma * -log(1 - prob + 1e-8);  pos = alpha * (1 - prob)^gamma * -log(prob + 1e-8)
        const float neg = __fmul_rn(__fmul_rn(one_minus_alpha, focal_pow(prob)), -logf(__fadd_rn(__fsub_rn(1.f, prob), 1e-8f)));
        const float pos = __fmul_rn(__fmul_rn(alpha, focal_pow(__fsub_rn(1.f, prob))), -logf(__fadd_rn(prob, 1e-8f)));
        float cost = __fmul_rn(w_class, __fsub_rn(pos, neg));
        const float qc = boxes[((size_t)b * R + r) * 2], ql = boxes[((size_t)b * R + r) * 2 + 1];
        const float tc = tgt_boxes[2 * c], tl = tgt_boxes[2 * c + 1];
        const float l1 = __fadd_rn(fabsf(__fsub_rn(qc, tc)), fabsf(__fsub_rn(ql, tl)));          // torch.cdist(p=1)
        cost = __fadd_rn(cost, __fmul_rn(w_bbox, l1));
        // (centre, length) -> (start, end); 1-D generalised IoU (misc/detr_utils/box_ops.py:8-48)
        const float a0 = __fsub_rn(qc, __fmul_rn(0.5f, ql)), a1 = __fadd_rn(qc, __fmul_rn(0.5f, ql));
        const float b0 = __fsub_rn(tc, __fmul_rn(0.5f, tl)), b1 = __fadd_rn(tc, __fmul_rn(0.5f, tl));
        const float len_a = __fsub_rn(a1, a0), len_b = __fsub_rn(b1, b0);
        const float inter = fmaxf(__fsub_rn(fminf(a1, b1), fmaxf(a0, b0)), 0.f);
        const float uni = __fsub_rn(__fadd_rn(len_a, len_b), inter);
        const float iou = __fdiv_rn(inter, __fadd_rn(uni, 1e-5f));
        const float hull = fmaxf(__fsub_rn(fmaxf(a1, b1), fminf(a0, b0)), 0.f);
        const float giou = __fsub_rn(iou, __fdiv_rn(__fsub_rn(hull, uni), __fadd_rn(hull, 1e-5f)));
        return __fsub_rn(cost, __fmul_rn(w_giou, giou));
    }
};

constexpr int kStageThreads = 128;   // the block: all four warps form / stage the cost matrix, warp 0 alone runs the search
constexpr int kScanBatch = 4;   // columns per lane whose loads are issued together (the scan is a chain of shared-memory round trips)

template <class Source>
__global__ void __launch_bounds__(kStageThreads) msda_lsap_kernel(const Args a, const Source source)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x;
    const int b = blockIdx.x;
    const bool transposed = a.C < a.R;               // scipy: solve the transposed problem when there are fewer columns
    const int nr = transposed ? a.C : a.R, nc = transposed ? a.R : a.C;
    float *cost = a.cost + (size_t)b * a.R * a.C;
    long long *row_idx = a.row_idx + (size_t)b * nr, *col_idx = a.col_idx + (size_t)b * nr;

    double *sp = reinterpret_cast<double *>(smem);   // shortest path cost to each column
    double *v = sp + nc;
    double *u = v + nc;
    int *path = reinterpret_cast<int *>(u + nr);
    int *row4col = path + nc;
    int *remaining = row4col + nc;
    int *col4row = remaining + nc;
    unsigned char *SC = reinterpret_cast<unsigned char *>(col4row + nr);
    unsigned char *SR = SC + nc;
    float *cs = reinterpret_cast<float *>(smem + state_bytes(nr, nc));   // [nr][nc], when staged

    // ---- stage + validate ----
    int invalid = 0;
    // four entries per lane and pass: all their (dependent: label -> logit) global loads are in flight together -- the
    // matrices are small and this loop is a chain of memory round trips
    for (int k0 = threadIdx.x; k0 < a.R * a.C; k0 += kStageThreads * kScanBatch) {
        float c[kScanBatch];
#pragma unroll
        for (int t = 0; t < kScanBatch; ++t) {
            const int k = k0 + kStageThreads * t;
            const int kk = k < a.R * a.C ? k : k0;
            const int r = kk / a.C, q = kk - r * a.C;
            c[t] = source(cost, b, kk, r, q);
        }
#pragma unroll
        for (int t = 0; t < kScanBatch; ++t) {
            const int k = k0 + kStageThreads * t;
            if (k >= a.R * a.C) continue;
            const int r = k / a.C, q = k - r * a.C;
            if (Source::kWrites) cost[k] = c[t];
            invalid |= (c[t] != c[t]) || (c[t] == -CUDART_INF_F);
            if (a.staged) cs[transposed ? q * nc + r : k] = c[t];
        }
    }
    if (Source::kWrites) __threadfence_block();       // an unstaged search reads the matrix back from global memory
    invalid = __syncthreads_or(invalid);              // the matrix is complete; the search below is one warp's work
    if (threadIdx.x >= 32) return;
    for (int j = lane; j < nc; j += 32) { v[j] = 0.0; row4col[j] = -1; }
    for (int i = lane; i < nr; i += 32) { u[i] = 0.0; col4row[i] = -1; }
    __syncwarp();
    int status = invalid ? 1 : 0;

    for (int cur = 0; cur < nr && status == 0; ++cur) {
        for (int j = lane; j < nc; j += 32) { sp[j] = CUDART_INF; path[j] = -1; SC[j] = 0; remaining[j] = nc - j - 1; }
        for (int i = lane; i < nr; i += 32) SR[i] = 0;
        __syncwarp();
        int n_rem = nc, sink = -1, i = cur;
        double min_val = 0.0;
        while (sink < 0) {
            if (lane == 0) SR[i] = 1;
            const double ui = u[i];
            Cand mine{CUDART_INF, -1};
            for (int it0 = lane; it0 < n_rem; it0 += 32 * kScanBatch) {
                int jj[kScanBatch], own[kScanBatch];
                float cc[kScanBatch];
                double vv[kScanBatch], ss[kScanBatch];
#pragma unroll
                for (int k = 0; k < kScanBatch; ++k) jj[k] = it0 + 32 * k < n_rem ? remaining[it0 + 32 * k] : -1;
#pragma unroll
                for (int k = 0; k < kScanBatch; ++k) {
                    const int j = jj[k] < 0 ? 0 : jj[k];
                    cc[k] = a.staged ? cs[i * nc + j] : (transposed ? __ldg(cost + (size_t)j * a.C + i) : __ldg(cost + (size_t)i * a.C + j));
                    vv[k] = v[j];
                    ss[k] = sp[j];
                    own[k] = row4col[j];
                }
#pragma unroll
                for (int k = 0; k < kScanBatch; ++k) {
                    if (jj[k] < 0) continue;
                    const double r = min_val + (double)cc[k] - ui - vv[k];      // scipy's expression, evaluated left to right
                    double s = ss[k];
                    if (r < s) { path[jj[k]] = i; sp[jj[k]] = r; s = r; }
                    const Cand c2{s, cand_rank(it0 + 32 * k, own[k] < 0)};
                    if (better(c2, mine)) mine = c2;
                }
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                Cand o;
                o.val = __shfl_xor_sync(kFullMask, mine.val, off);
                o.rank = __shfl_xor_sync(kFullMask, mine.rank, off);
                if (better(o, mine)) mine = o;
            }
            if (mine.rank < 0 || mine.val == CUDART_INF) { status = 2; break; }
            min_val = mine.val;
            const int pos = cand_pos(mine.rank);
            const int j = remaining[pos];
            const int owner = row4col[j];
            __syncwarp();                              // every lane has read remaining[] / row4col[] before lane 0 edits them
            if (owner < 0) sink = j; else i = owner;
            --n_rem;
            if (lane == 0) { SC[j] = 1; remaining[pos] = remaining[n_rem]; }
            __syncwarp();
        }
        if (status) break;
        // ---- dual update (before the path is flipped: col4row is still the previous assignment) ----
        for (int r = lane; r < nr; r += 32) {
            if (r == cur) u[r] += min_val;
            else if (SR[r]) u[r] += min_val - sp[col4row[r]];
        }
        for (int j = lane; j < nc; j += 32)
            if (SC[j]) v[j] -= min_val - sp[j];
        __syncwarp();
        // ---- flip the augmenting path ----
        if (lane == 0) {
            int j = sink;
            while (true) {
                const int r = path[j];
                row4col[j] = r;
                const int prev = col4row[r];
                col4row[r] = j;
                j = prev;
                if (r == cur) break;
            }
        }
        __syncwarp();
    }

    if (a.status && lane == 0) a.status[b] = status;
    if (status) {                                      // defined output: the identity pairing
        for (int k = lane; k < nr; k += 32) { row_idx[k] = k; col_idx[k] = k; }
        return;
    }
    if (!transposed) {
        for (int k = lane; k < nr; k += 32) { row_idx[k] = k; col_idx[k] = col4row[k]; }
    } else {
        // pairs (row = col4row[i], column = i) in ascending row order: rank by counting (rows are distinct)
        for (int k = lane; k < nr; k += 32) {
            const int mine = col4row[k];
            int rank = 0;
            for (int t = 0; t < nr; ++t) rank += col4row[t] < mine;
            row_idx[rank] = mine;
            col_idx[rank] = k;
        }
    }
}

}  // namespace match
}  // namespace msda
