// msda_pipe.cuh -- persistent, software-pipelined forward of the 1-D temporal MSDeformAttn for sm_100a.
//
// Same tiles, decode and gather as msda_lean.cuh, re-scheduled after clock64 phase timing of the lean
// forward showed (profiles/r1_phases.txt) that a CTA spent 35 % of its life waiting for the raw per-sample
// inputs from HBM and 20 % staging its value window, with the LDS gather (the real work) only 36 %:
//   - a CTA is persistent and walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...
//   - the raw inputs (offsets, logits, reference points) of tile i+1 are loaded into registers and the value
//     window of tile i is copied into shared memory (cp.async, double buffered) while tile i-1 is gathered.
// Two barriers per tile.  Replaces ms_deformable_im2col_gpu_kernel (ms_deform_im2col_cuda.cuh:238-300) and,
// FUSED, pdvc/ops/modules/ms_deform_attn.py:100-116.
#pragma once
#include "msda_lean.cuh"

namespace msda {

#ifdef MSDA_PHASE_TIMING
#define MSDA_LOOP_PHASE_DECL unsigned long long lp_acc[8] = {}; unsigned long long lp_last = clock64();
#define MSDA_LOOP_PHASE(K) if (threadIdx.x == 0) { const unsigned long long now_ = clock64(); lp_acc[K] += now_ - lp_last; lp_last = now_; }
#define MSDA_LOOP_PHASE_FLUSH(W) if (threadIdx.x == 0) for (int i_ = 0; i_ < 8; ++i_) atomicAdd(&g_phase_clk[W][i_], lp_acc[i_]);
#else
#define MSDA_LOOP_PHASE_DECL
#define MSDA_LOOP_PHASE(K)
#define MSDA_LOOP_PHASE_FLUSH(W)
#endif

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct TileId {
    int k, m, n;
};

__device__ __forceinline__ TileId tile_id(const TileArgs &a, int t)
{
    TileId id;
    id.k = t % a.chunks;
    const int r = t / a.chunks;
    id.m = r % a.M;
    id.n = r / a.M;
    return id;
}

// dynamic shared memory: 2 x [ records int4 x Q*LPA | value window win_rows x rb ]
template <typename VT, int LPG, int NCH, int LPA, bool FUSED>
__global__ void __launch_bounds__(kLeanThreads, 2) msda_fwd_pipe_kernel(const TileArgs a)
{
    using IO = RowIO<VT, LPG, NCH>;
    constexpr int CH = Vec16<VT>::CH;
    constexpr int NF = IO::NF;
    constexpr int G = 32 / LPG;     // units per warp in the gather phase
    constexpr int UPW = 32 / LPA;   // units per warp in the decode phase
    constexpr int UPP = kLeanWarps * UPW;
    constexpr int QMAX = UPP * kLeanPasses;
    constexpr int rb = LPG * NCH * 16;   // bytes per row of one head (= D * sizeof(VT))
    constexpr int VPR = rb / 16;
    extern __shared__ __align__(128) unsigned char dsm[];
    __shared__ int win_lo[2][kMaxLevels], win_hi[2][kMaxLevels];
    __shared__ TileRun runs[2][kMaxLevels];
    __shared__ int uq[2][QMAX];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int MD = a.M * a.D;
    const int n_tiles = a.chunks * a.M * a.N;
    const size_t rec_bytes = (size_t)a.Q * LPA * sizeof(int4);
    const size_t buf_bytes = rec_bytes + (size_t)a.win_rows * rb;

    const int g1 = lane / LPA, j1 = lane % LPA;
    const LaneLevel lc = lane_level(a, j1);
    int ulv[kLeanPasses];
#pragma unroll
    for (int t = 0; t < kLeanPasses; ++t) ulv[t] = t * UPP + warp * UPW + g1;

    // ---- prologue: queries and raw inputs of the first tile ----
    int t_cur = blockIdx.x;
    if (t_cur >= n_tiles) return;
    TileId id = tile_id(a, t_cur);
    if (tid < a.L) {
        win_lo[0][tid] = win_lo[1][tid] = INT_MAX;
        win_hi[0][tid] = win_hi[1][tid] = -1;
    }
    if (a.gridK) tile_runs_init(a, id.k, runs[0]);
    __syncthreads();
    int qv[kLeanPasses];
    int Qc = tile_queries<kLeanPasses>(a, id.k, runs[0], ulv, qv);
    RawSample raw[kLeanPasses];
#pragma unroll
    for (int t = 0; t < kLeanPasses; ++t) {
        const bool on = lc.live && qv[t] >= 0;
        const int nq = id.n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
        raw[t] = load_raw<FUSED>(a, lc, nq, (nq * a.M + id.m) * a.LP + j1, on);
    }

    // state of the tile whose gather is pending
    bool have_prev = false, staged_prev = false;
    int Qc_prev = 0;
    TileId id_prev = id;

    MSDA_LOOP_PHASE_DECL
    for (int i = 0;; ++i) {
        const int buf = i & 1;
        MSDA_LOOP_PHASE(0)
        int4 *recs = reinterpret_cast<int4 *>(dsm + buf * buf_bytes);
        unsigned char *win = dsm + buf * buf_bytes + rec_bytes;
        const int t_next = t_cur + gridDim.x;
        const bool has_next = t_next < n_tiles;
        const TileId id_next = tile_id(a, has_next ? t_next : t_cur);

        // ---- A: decode tile i; per-level row ranges ----
        const int passes = (Qc + UPP - 1) / UPP;
        int r0[kLeanPasses], r1[kLeanPasses];
        float c0[kLeanPasses], c1[kLeanPasses];
        int lo = INT_MAX, hi = -1;
#pragma unroll
        for (int t = 0; t < kLeanPasses; ++t) {
            r0[t] = r1[t] = -1;
            c0[t] = c1[t] = 0.f;
            if (t < passes) {   // block-uniform
                const bool on = lc.live && qv[t] >= 0;
                const Decoded d = decode<FUSED, LPA>(a, lc, raw[t], on);
                const float k0 = d.A * d.w0, k1 = d.A * d.w1;
                if (on && (k0 != 0.f || k1 != 0.f)) {
                    r0[t] = d.r0; r1[t] = d.r1; c0[t] = k0; c1[t] = k1;
                    lo = min(lo, d.r0);
                    hi = max(hi, d.r1);
                }
            }
        }
        window_update(win_lo[buf], win_hi[buf], lc.l, lo, hi);
        if (has_next && a.gridK) tile_runs_init(a, id_next.k, runs[buf ^ 1]);
        MSDA_LOOP_PHASE(1)
        __syncthreads();   // (1) ranges of tile i and runs of tile i+1 visible; everybody is done with gather(i-2)

        MSDA_LOOP_PHASE(2)
        // ---- B: stage the value window of tile i (async) and write its records ----
        int tot = 0, my_base = 0;
        for (int l = 0; l < a.L; ++l) {
            const int wl = win_lo[buf][l], wh = win_hi[buf][l];
            if (l == lc.l) my_base = tot - wl;   // window row of slice row r at this lane's level: my_base + r
            if (wh >= wl) tot += wh - wl + 1;
        }
        const bool staged = tot > 0 && tot <= a.win_rows;
        const VT *vslice = static_cast<const VT *>(a.value) + (size_t)id.n * a.S * MD + id.m * a.D;
        if (staged) {
            int cum = 0;
            for (int l = 0; l < a.L; ++l) {
                const int wl = win_lo[buf][l], wh = win_hi[buf][l];
                if (wh < wl) continue;
                const int nv = (wh - wl + 1) * VPR;
                for (int v = tid; v < nv; v += kLeanThreads) {
                    const int rr = v / VPR, jj = v % VPR;
                    cp_async16(win + (cum + rr) * rb + jj * 16,
                               reinterpret_cast<const unsigned char *>(vslice + (size_t)(wl + rr) * MD) + jj * 16);
                }
                cum += wh - wl + 1;
            }
        }
        cp_async_commit();
        {
            const int base = staged ? my_base : 0;
            const int scale = staged ? rb : MD;
#pragma unroll
            for (int t = 0; t < kLeanPasses; ++t) {
                if (t < passes && ulv[t] < Qc) {
                    const bool livet = r0[t] >= 0;
                    recs[ulv[t] * LPA + (j1 ^ rec_swizzle(ulv[t]))] = make_int4(livet ? (base + r0[t]) * scale : 0, livet ? (base + r1[t]) * scale : 0,
                                                        __float_as_int(c0[t]), __float_as_int(c1[t]));
                    if (j1 == 0) uq[buf][ulv[t]] = qv[t];
                }
            }
        }

        MSDA_LOOP_PHASE(3)
        // ---- C: queries and raw inputs of tile i+1 (in flight during the gather below) ----
        int Qc_next = 0;
        if (has_next) {
            Qc_next = tile_queries<kLeanPasses>(a, id_next.k, runs[buf ^ 1], ulv, qv);
#pragma unroll
            for (int t = 0; t < kLeanPasses; ++t) {
                const bool on = lc.live && qv[t] >= 0;
                const int nq = id_next.n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
                raw[t] = load_raw<FUSED>(a, lc, nq, (nq * a.M + id_next.m) * a.LP + j1, on);
            }
        }

        MSDA_LOOP_PHASE(4)
        // ---- D: gather tile i-1 (its window landed while tile i was decoded) ----
        if (tid < a.L) { win_lo[buf ^ 1][tid] = INT_MAX; win_hi[buf ^ 1][tid] = -1; }   // for tile i+1 (last read before (1))
        if (!has_next) cp_async_wait<0>(); else cp_async_wait<1>();
        // the last iteration also waits for its own window and gathers both tiles
        MSDA_LOOP_PHASE(5)
        __syncthreads();   // (2) window + records of tile i-1 (and, in the last iteration, tile i) visible
        MSDA_LOOP_PHASE(6)
        const int first = have_prev ? 0 : 1, last = has_next ? 0 : 1;
        for (int which = first; which <= last; ++which) {
            // which == 0: tile i-1 in the other buffer; which == 1: tile i itself (drain)
            const int gb = which == 0 ? (buf ^ 1) : buf;
            const int4 *grecs = reinterpret_cast<const int4 *>(dsm + gb * buf_bytes);
            const unsigned char *gwin = dsm + gb * buf_bytes + rec_bytes;
            const TileId gid = which == 0 ? id_prev : id;
            const int gQc = which == 0 ? Qc_prev : Qc;
            const bool gstaged = which == 0 ? staged_prev : staged;
            const int g = lane / LPG, j = lane % LPG;
            VT *out = static_cast<VT *>(a.out);
            const unsigned char *wlane = gwin + j * 16;
            const VT *vlane = static_cast<const VT *>(a.value) + (size_t)gid.n * a.S * MD + gid.m * a.D + j * CH;
            const int LP4 = (a.LP + 3) & ~3;
            for (int ul0 = warp * G; ul0 < gQc; ul0 += kLeanWarps * G) {
                const int ul = ul0 + g;
                const bool active = ul < gQc;
                const int4 *rp = grecs + (active ? ul : 0) * LPA;
                const int sw = rec_swizzle(active ? ul : 0);
                float acc[NF];
#pragma unroll
                for (int k = 0; k < NF; ++k) acc[k] = 0.f;
                for (int s = 0; s < LP4; s += 4) {
                    int4 r[4];
#pragma unroll
                    for (int b = 0; b < 4; ++b) r[b] = rp[(s + b) ^ sw];
                    if (gstaged) {
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            float v0[NF], v1[NF];
                            IO::load_smem(wlane + r[b].x, v0);
                            IO::load_smem(wlane + r[b].y, v1);
                            axpy<NF>(acc, __int_as_float(r[b].z), v0);
                            axpy<NF>(acc, __int_as_float(r[b].w), v1);
                        }
                    } else {
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            float v0[NF], v1[NF];
                            IO::load_gmem(vlane + r[b].x, v0);
                            IO::load_gmem(vlane + r[b].y, v1);
                            axpy<NF>(acc, __int_as_float(r[b].z), v0);
                            axpy<NF>(acc, __int_as_float(r[b].w), v1);
                        }
                    }
                }
                if (active)
                    IO::store_gmem(out + (size_t)((gid.n * a.Lq + uq[gb][ul]) * a.M + gid.m) * a.D + j * CH, acc);
            }
        }
        MSDA_LOOP_PHASE(7)
        if (!has_next) break;
        have_prev = true;
        staged_prev = staged;
        Qc_prev = Qc;
        id_prev = id;
        id = id_next;
        Qc = Qc_next;
        t_cur = t_next;
    }
    MSDA_LOOP_PHASE_FLUSH(0)
}

}  // namespace msda
