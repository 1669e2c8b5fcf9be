// msda_lean.cuh -- the production kernels of the 1-D temporal MSDeformAttn hot path for sm_100a.
//
// Same algorithm as msda_tile.cuh (chunk of Q queries of one head per CTA; forward = value window staged
// in shared memory by cp.async.bulk + LDS.128 gather; backward = samples counting-sorted by value row, one
// lane group per row), re-written for instruction economy after ncu showed both tile kernels issue-bound
// (profiles/r1_notes.md): every lane decodes ONE sample per query (L*P <= 32; two, SPL = 2, up to L*P = 64), so level, T_l,
// level start are per-lane constants, all indices are 32-bit, decoded samples stay in registers between
// phases, each lane owns NCH 16-byte chunks of a value row (fewer shuffle steps per dot product and
// fewer address computations per FFMA2).
//
// Replaces, for H_l == 1: ms_deformable_im2col_gpu_kernel (ms_deform_im2col_cuda.cuh:238-300), the
// col2im kernels (cuh:302-921) and, FUSED, pdvc/ops/modules/ms_deform_attn.py:100-116.
#pragma once
#include "msda_tile.cuh"

namespace msda {

// Development aid (-DMSDA_PHASE_TIMING): thread 0 of every CTA stamps clock64() at phase boundaries and adds the
// deltas to a device array read back by msda_dev_phase_clocks().  Compiled out of the shipped library.
#ifdef MSDA_PHASE_TIMING
__device__ unsigned long long g_phase_clk[2][16];
#define MSDA_PHASE_DECL unsigned long long ph_t[16]; int ph_n = 0; if (threadIdx.x == 0) ph_t[ph_n++] = clock64();
#define MSDA_PHASE_MARK if (threadIdx.x == 0) ph_t[ph_n++] = clock64();
#define MSDA_PHASE_FLUSH(K) if (threadIdx.x == 0) for (int i = 1; i < ph_n; ++i) atomicAdd(&g_phase_clk[K][i - 1], ph_t[i] - ph_t[i - 1]);
#else
#define MSDA_PHASE_DECL
#define MSDA_PHASE_MARK
#define MSDA_PHASE_FLUSH(K)
#endif

#ifndef MSDA_BWD_MINB
#define MSDA_BWD_MINB 2
#endif
constexpr int kLeanThreads = 256;
constexpr int kLeanWarps = kLeanThreads / 32;
constexpr int kLeanPasses = 4;   // units decoded per lane group and CTA (Q = kLeanPasses * units per pass)

// per-lane constants of the decode phase: lane j of a unit's group always decodes sample j = (l, p)
struct LaneLevel {
    int l, T, start;
    float tf;
    bool live;   // j < L*P
};

__device__ __forceinline__ LaneLevel lane_level(const TileArgs &a, int j)
{
    LaneLevel c;
    c.live = j < a.LP;
    c.l = c.live ? j / a.P : 0;
    c.T = (int)__ldg(a.shapes + (a.shapes_stride == 2 ? 2 * c.l + 1 : c.l));
    c.start = (int)__ldg(a.lsi + c.l);
    c.tf = (float)c.T;
    return c;
}

// ---- tiles ------------------------------------------------------------------------------------------
// gridK == 0: tile k = the Q consecutive queries [k*Q, k*Q + Q).
// gridK  > 0 (self-attention, Lq == S: the queries ARE the rows of the level pyramid,
//            pdvc/deformable_transformer.py:198-208): tile k takes from EVERY level the rows whose position
//            p satisfies k <= p*K/T_l < k+1, i.e. all queries of one time span.  Their samples then fall into
//            one short row window per level (span + offsets) instead of Q/T_l of the whole pyramid, so the
//            window always fits in shared memory.  Pure scheduling: results do not depend on it.
struct TileRun {
    int q_first;   // first query (= slice row) of this level's run
    int count;
};

__device__ __forceinline__ void tile_runs_init(const TileArgs &a, int k, TileRun *runs)
{
    const int l = threadIdx.x;
    if (l < a.L) {
        const unsigned T = (unsigned)__ldg(a.shapes + (a.shapes_stride == 2 ? 2 * l + 1 : l));
        const unsigned K = (unsigned)a.gridK;
        const unsigned p0 = ((unsigned)k * T + K - 1) / K;          // S < 65536: no overflow
        const unsigned p1 = min(T, ((unsigned)(k + 1) * T + K - 1) / K);
        runs[l].q_first = (int)__ldg(a.lsi + l) + (int)p0;
        runs[l].count = p1 > p0 ? (int)(p1 - p0) : 0;
    }
}

// queries of this thread's NU units (-1: none) and the tile's query count; call after the runs are visible
template <int NU>
__device__ __forceinline__ int tile_queries(const TileArgs &a, int k, const TileRun *runs, const int (&ul)[NU], int (&q)[NU])
{
    if (a.gridK == 0) {
        const int q0 = k * a.Q;
        const int Qc = min(a.Q, a.Lq - q0);
#pragma unroll
        for (int i = 0; i < NU; ++i) q[i] = ul[i] < Qc ? q0 + ul[i] : -1;
        return Qc;
    }
#pragma unroll
    for (int i = 0; i < NU; ++i) q[i] = -1;
    int off = 0;
    for (int l = 0; l < a.L; ++l) {
        const TileRun r = runs[l];
        const int cnt = min(r.count, a.Q - off);   // never clips when sum(T_l) == Lq (host picks K for that)
#pragma unroll
        for (int i = 0; i < NU; ++i) {
            const int rel = ul[i] - off;
            if (rel >= 0 && rel < cnt) q[i] = r.q_first + rel;
        }
        off += cnt;
    }
    return off;
}

// Record slot swizzle: the lane groups of a warp read the records of consecutive units at the same sample
// index; unit records are 16 * LPA bytes apart (same banks), so sample s of unit u lives in slot s ^ 2*(u & 3)
// -- four consecutive units then hit four different 16-byte bank groups (one wavefront instead of four).
__device__ __forceinline__ int rec_swizzle(int unit) { return (unit & 3) << 1; }

// per-level [lo, hi] row range: warp-aggregated (REDUX over the lanes of equal level), then one shared atomic
// per level and warp instead of one per lane
__device__ __forceinline__ void window_update(int *win_lo, int *win_hi, int level, int lo, int hi)
{
    const unsigned peers = __match_any_sync(kFull, level);
    const int wlo = __reduce_min_sync(peers, lo);
    const int whi = __reduce_max_sync(peers, hi);
    if ((threadIdx.x & 31) == __ffs(peers) - 1 && whi >= 0) {
        atomicMin(&win_lo[level], wlo);
        atomicMax(&win_hi[level], whi);
    }
}

// raw per-sample inputs (issued for all passes before anything is consumed)
struct RawSample {
    float x;   // fused: offset            unfused: location
    float y;   // fused: attention logit   unfused: attention weight
    float r0;  // fused: reference point
    float r1;  // fused, R == 2: reference length
};

template <bool FUSED>
__device__ __forceinline__ RawSample load_raw(const TileArgs &a, const LaneLevel &c, int nq, int sidx, bool on)
{
    RawSample r;
    r.x = r.y = r.r0 = r.r1 = 0.f;
    if (on) {
        if (FUSED) {
            r.x = __ldg(a.offs + sidx);
            r.y = __ldg(a.logits + sidx);
            const float *rp = a.ref + (nq * a.L + c.l) * a.R;
            r.r0 = __ldg(rp);
            if (a.R == 2) r.r1 = __ldg(rp + 1);
        } else {
            r.x = __ldg(a.loc + 2 * sidx);
            r.y = __ldg(a.attn + sidx);
        }
    }
    return r;
}

struct Decoded {
    int r0, r1;      // slice rows
    float w0, w1;    // interpolation weights (0 for a missing tap)
    float s0, s1;    // -gate / +gate
    float A;         // attention weight
    float dxdoff;    // fused: d(loc)/d(offset)
};

// sampling geometry of one sample given its attention weight
template <bool FUSED>
__device__ __forceinline__ Decoded decode_geom(const TileArgs &a, const LaneLevel &c, const RawSample &raw, float A)
{
    Decoded d;
    float hi, lo;
    d.A = A;
    if (FUSED) {
        const float p = raw.r0 * c.tf;
        const float plo = fmaf(raw.r0, c.tf, -p);
        float add;
        if (a.R == 1) {   // x = ref0 + off / T_l          (ms_deform_attn.py:103-106)
            add = raw.x - 0.5f;
            d.dxdoff = __fdividef(1.f, c.tf);
        } else {          // x = ref0 + off / P * ref1 * 0.5 (ms_deform_attn.py:107-109)
            d.dxdoff = raw.r1 * (0.5f * a.invP);
            add = fmaf(raw.x * d.dxdoff, c.tf, -0.5f);
        }
        float err;
        two_sum(p, add, hi, err);
        lo = plo + err;
    } else {
        d.dxdoff = 0.f;
        pix_from_loc(raw.x, c.T, hi, lo);
    }
    const Tap1D t = locate_1d(hi, lo, c.T, a.pad_mode);
    d.r0 = c.start + t.i0;
    d.r1 = c.start + t.i1;
    d.w0 = t.w0; d.w1 = t.w1; d.s0 = t.s0; d.s1 = t.s1;
    return d;
}

// softmax over the L*P samples of a unit (FUSED): LPA lanes x SPL samples per lane (lane j owns samples j + s*LPA),
// then the geometry of every sample; all lanes of the warp call it
template <bool FUSED, int LPA, int SPL>
__device__ __forceinline__ void decode(const TileArgs &a, const LaneLevel (&c)[SPL], const RawSample (&raw)[SPL],
                                       const bool (&on)[SPL], Decoded (&d)[SPL])
{
    float A[SPL];
    if (FUSED) {
        float m = -INFINITY;
#pragma unroll
        for (int s = 0; s < SPL; ++s) m = fmaxf(m, on[s] ? raw[s].y : -INFINITY);
        const float mx = group_max<LPA>(m);
        float e[SPL], part = 0.f;
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            e[s] = on[s] ? __expf(raw[s].y - mx) : 0.f;
            part += e[s];
        }
        const float sum = group_sum<LPA>(part);
#pragma unroll
        for (int s = 0; s < SPL; ++s) A[s] = sum > 0.f ? __fdividef(e[s], sum) : 0.f;
    } else {
#pragma unroll
        for (int s = 0; s < SPL; ++s) A[s] = raw[s].y;
    }
#pragma unroll
    for (int s = 0; s < SPL; ++s) d[s] = decode_geom<FUSED>(a, c[s], raw[s], A[s]);
}

// value / grad rows: lane j of a row group owns the 16-byte chunks c*LPG + j, c < NCH
template <typename VT, int LPG, int NCH> struct RowIO {
    static constexpr int CH = Vec16<VT>::CH;
    static constexpr int NF = NCH * CH;   // floats per lane
    __device__ __forceinline__ static void load_smem(const unsigned char *p, float (&v)[NF])
    {
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            float t[CH];
            VecPlain<VT>::load(p + c * LPG * 16, t);
#pragma unroll
            for (int k = 0; k < CH; ++k) v[c * CH + k] = t[k];
        }
    }
    __device__ __forceinline__ static void load_gmem(const VT *p, float (&v)[NF])
    {
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            float t[CH];
            Vec16<VT>::load(p + c * LPG * CH, t);
#pragma unroll
            for (int k = 0; k < CH; ++k) v[c * CH + k] = t[k];
        }
    }
    __device__ __forceinline__ static void store_gmem(VT *p, const float (&v)[NF])
    {
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            float t[CH];
#pragma unroll
            for (int k = 0; k < CH; ++k) t[k] = v[c * CH + k];
            Vec16<VT>::store(p + c * LPG * CH, t);
        }
    }
};

// ---------------------------------------------------------------------------------------------
// forward.  grid = (chunks, M, N); dynamic shared memory = [ records int4 x Q*LPA | value window ]
// ---------------------------------------------------------------------------------------------
template <typename VT, int LPG, int NCH, int LPA, bool FUSED, int SPL = 1>
__global__ void __launch_bounds__(kLeanThreads) msda_fwd_lean_kernel(const TileArgs a)
{
    using IO = RowIO<VT, LPG, NCH>;
    constexpr int CH = Vec16<VT>::CH;
    constexpr int NF = IO::NF;
    constexpr int G = 32 / LPG;     // units per warp in the gather phase
    constexpr int UPW = 32 / LPA;   // units per warp in the decode phase
    constexpr int UPP = kLeanWarps * UPW;
    constexpr int PASSES = kLeanPasses / SPL;   // SPL samples per lane (32 < L*P <= 64): half the units per lane
    constexpr int LPT = LPA * SPL;              // sample slots per unit
    constexpr int QMAX = UPP * PASSES;
    extern __shared__ __align__(128) unsigned char dsm[];
    MSDA_PHASE_DECL
    __shared__ int win_lo[kMaxLevels], win_hi[kMaxLevels], win_base[kMaxLevels];
    __shared__ TileRun runs[kMaxLevels];
    __shared__ int uq[QMAX];   // query of every unit of the tile
    __shared__ int win_ok;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n = blockIdx.z, m = blockIdx.y;
    const int MD = a.M * a.D;
    constexpr int rb = LPG * NCH * 16;   // bytes per row of one head (= D * sizeof(VT))
    int4 *recs = reinterpret_cast<int4 *>(dsm);
    unsigned char *win = dsm + (size_t)a.Q * LPT * sizeof(int4);

    if (tid < a.L) { win_lo[tid] = INT_MAX; win_hi[tid] = -1; }
    if (a.gridK) tile_runs_init(a, blockIdx.x, runs);
    __syncthreads();   // win_lo / win_hi / mbar / runs initialised
    MSDA_PHASE_MARK

    // ---- decode: lane (g, j) handles samples j + s*LPA of units warp*UPW + g + t*UPP, t < passes ----
    const int g1 = lane / LPA, j1 = lane % LPA;
    LaneLevel lc[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) lc[s] = lane_level(a, j1 + s * LPA);
    int ulv[PASSES], qv[PASSES];
#pragma unroll
    for (int t = 0; t < PASSES; ++t) ulv[t] = t * UPP + warp * UPW + g1;
    const int Qc = tile_queries<PASSES>(a, blockIdx.x, runs, ulv, qv);
    const int passes = (Qc + UPP - 1) / UPP;
    RawSample raw[PASSES][SPL];
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
        const int nq = n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            const bool on = lc[s].live && qv[t] >= 0;
            raw[t][s] = load_raw<FUSED>(a, lc[s], nq, (nq * a.M + m) * a.LP + j1 + s * LPA, on);
        }
        if (j1 == 0 && ulv[t] < a.Q) uq[ulv[t]] = qv[t];
    }
    int r0[PASSES][SPL], r1[PASSES][SPL];
    float c0[PASSES][SPL], c1[PASSES][SPL];
    int lo[SPL], hi[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) { lo[s] = INT_MAX; hi[s] = -1; }
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
#pragma unroll
        for (int s = 0; s < SPL; ++s) { r0[t][s] = r1[t][s] = -1; c0[t][s] = c1[t][s] = 0.f; }
        if (t < passes) {   // block-uniform
            bool on[SPL];
#pragma unroll
            for (int s = 0; s < SPL; ++s) on[s] = lc[s].live && qv[t] >= 0;
            Decoded d[SPL];
            decode<FUSED, LPA, SPL>(a, lc, raw[t], on, d);
#pragma unroll
            for (int s = 0; s < SPL; ++s) {
                const float k0 = d[s].A * d[s].w0, k1 = d[s].A * d[s].w1;
                if (on[s] && (k0 != 0.f || k1 != 0.f)) {
                    r0[t][s] = d[s].r0; r1[t][s] = d[s].r1; c0[t][s] = k0; c1[t][s] = k1;
                    lo[s] = min(lo[s], d[s].r0);
                    hi[s] = max(hi[s], d[s].r1);
                }
            }
        }
    }
#pragma unroll
    for (int s = 0; s < SPL; ++s) window_update(win_lo, win_hi, lc[s].l, lo[s], hi[s]);
    __syncthreads();
    MSDA_PHASE_MARK
    if (tid == 0) {
        int tot = 0;
        for (int l = 0; l < a.L; ++l) {
            win_base[l] = tot - win_lo[l];   // window row of slice row r at level l: win_base[l] + r
            if (win_hi[l] >= win_lo[l]) tot += win_hi[l] - win_lo[l] + 1;
        }
        const int ok = (tot > 0 && tot <= a.win_rows) ? 1 : 0;
        win_ok = ok;
    }
    __syncthreads();
    MSDA_PHASE_MARK
    const bool staged = win_ok != 0;
    const VT *vslice = static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D;
    if (staged) {
        // 16-byte cp.async per lane: measured 2x cheaper to issue than one 256-byte cp.async.bulk per row
        // (profiles/r1_phases.txt); the TMA engine serialises small bulk copies
        constexpr int VPR = rb / 16;
        for (int l = 0; l < a.L; ++l) {
            if (win_hi[l] < win_lo[l]) continue;   // level without a live sample (ZEROS padding can drop them all)
            const int nv = (win_hi[l] - win_lo[l] + 1) * VPR;
            for (int v = tid; v < nv; v += kLeanThreads) {
                const int r = win_lo[l] + v / VPR, jj = v % VPR;
                cp_async16(win + (win_base[l] + r) * rb + jj * 16,
                           reinterpret_cast<const unsigned char *>(vslice + (size_t)r * MD) + jj * 16);
            }
        }
    }
    {   // final records: byte offsets into the window (staged) or element offsets into the slice
        const int scale = staged ? rb : MD;
#pragma unroll
        for (int t = 0; t < PASSES; ++t) {
            if (t < passes && ulv[t] < Qc) {
#pragma unroll
                for (int s = 0; s < SPL; ++s) {
                    const int base = staged ? win_base[lc[s].l] : 0;
                    const bool livet = r0[t][s] >= 0;
                    recs[ulv[t] * LPT + s * LPA + (j1 ^ rec_swizzle(ulv[t]))] =
                        make_int4(livet ? (base + r0[t][s]) * scale : 0, livet ? (base + r1[t][s]) * scale : 0,
                                  __float_as_int(c0[t][s]), __float_as_int(c1[t][s]));
                }
            }
        }
    }
    cp_async_commit_wait_all();
    __syncthreads();
    MSDA_PHASE_MARK
    MSDA_PHASE_MARK

    // ---- gather ----
    {
        const int g = lane / LPG, j = lane % LPG;
        VT *out = static_cast<VT *>(a.out);
        const unsigned char *wlane = win + j * 16;
        const VT *vlane = vslice + j * CH;
        const int LP4 = (a.LP + 3) & ~3;
        for (int ul0 = warp * G; ul0 < Qc; ul0 += kLeanWarps * G) {
            const int ul = ul0 + g;
            const bool active = ul < Qc;
            const int4 *rp = recs + (active ? ul : 0) * LPT;
            const int sw = rec_swizzle(active ? ul : 0);
            float acc[NF];
#pragma unroll
            for (int k = 0; k < NF; ++k) acc[k] = 0.f;
            for (int s = 0; s < LP4; s += 4) {
                int4 r[4];
#pragma unroll
                for (int b = 0; b < 4; ++b) r[b] = rp[(s + b) ^ sw];
                if (staged) {
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
            if (active) IO::store_gmem(out + (size_t)((n * a.Lq + uq[ul]) * a.M + m) * a.D + j * CH, acc);
        }
    }
    MSDA_PHASE_MARK
    MSDA_PHASE_FLUSH(0)
}


// ---------------------------------------------------------------------------------------------
// backward.  grid = (chunks, M, N).  Dynamic shared memory (QS = Q * LPA sample slots):
//   Gs    Q x D VT          grad_out rows of the chunk (this head)
//   ent   (2*QS + 2) int4   samples in value-row order, every row's run padded to an even length:
//                           {sample slot | second_tap << 15 | window slot << 16, A*wa, A*wb, row};  ent[2*QS] = no-op entry
//   dots  (QS + 1) float2   {G.V[r], G.V[r+1]} per sample slot (+1 sink for the no-op entries)
//   start (S + 1) int       histogram of first rows -> start of each row's run in ent
//   Vw    win_rows x D VT   value rows of the tile's level windows (falls back to global loads if they do not fit)
// ---------------------------------------------------------------------------------------------
constexpr int kRowFlagShift = 30;

__host__ __device__ inline size_t bwd_lean_smem_bytes(int Q, int LPA, int S, int rb, int win_rows)
{
    const size_t QS = (size_t)Q * LPA;
    return (size_t)Q * rb + (2 * QS + 2) * 16 + (QS + 1 + 1) * 8 + ((size_t)S + 1 + 3) / 4 * 16 + (size_t)win_rows * rb;
}

template <typename VT, int LPG, int NCH, int LPA, bool FUSED, int SPL = 1>
__global__ void __launch_bounds__(kLeanThreads, MSDA_BWD_MINB) msda_bwd_lean_kernel(const TileArgs a)
{
    using IO = RowIO<VT, LPG, NCH>;
    constexpr int CH = Vec16<VT>::CH;
    constexpr int NF = IO::NF;
    constexpr int NG = kLeanThreads / LPG;   // row groups per CTA
    constexpr int UPW = 32 / LPA;
    constexpr int UPP = kLeanWarps * UPW;
    constexpr int PASSES = kLeanPasses / SPL;   // SPL samples per lane (32 < L*P <= 64): half the units per lane
    constexpr int LPT = LPA * SPL;              // sample slots per unit
    constexpr int QMAX = UPP * PASSES;
    constexpr int rb = LPG * NCH * 16;
    extern __shared__ __align__(128) unsigned char dsm[];
    MSDA_PHASE_DECL
    __shared__ int wsum[kLeanWarps];
    __shared__ int n_ent_s;
    __shared__ int win_lo[kMaxLevels], win_hi[kMaxLevels];
    __shared__ TileRun runs[kMaxLevels];
    __shared__ int uq[QMAX];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n = blockIdx.z, m = blockIdx.y;
    const int MD = a.M * a.D;
    const int QS = a.Q * LPT;
    unsigned char *Gs = dsm;
    int4 *ent = reinterpret_cast<int4 *>(Gs + (size_t)a.Q * rb);
    float2 *dots = reinterpret_cast<float2 *>(ent + 2 * QS + 2);
    int *start = reinterpret_cast<int *>(dots + QS + 2);
    unsigned char *Vw = reinterpret_cast<unsigned char *>(start + (a.S + 1 + 3) / 4 * 4);   // value window, win_rows x rb

    if (tid < a.L) { win_lo[tid] = INT_MAX; win_hi[tid] = -1; }
    if (a.gridK) tile_runs_init(a, blockIdx.x, runs);
    if (tid == 0) ent[2 * QS] = make_int4(QS, 0, 0, 0);   // no-op entry: unit 0, zero weights, sink slot
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- raw inputs of this lane's samples ----
    const int g1 = lane / LPA, j1 = lane % LPA;
    LaneLevel lc[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) lc[s] = lane_level(a, j1 + s * LPA);
    int ulv[PASSES], qv[PASSES];
#pragma unroll
    for (int t = 0; t < PASSES; ++t) ulv[t] = t * UPP + warp * UPW + g1;
    const int Qc = tile_queries<PASSES>(a, blockIdx.x, runs, ulv, qv);
    const int passes = (Qc + UPP - 1) / UPP;
    RawSample raw[PASSES][SPL];
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
        const int nq = n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            const bool on = lc[s].live && qv[t] >= 0;
            raw[t][s] = load_raw<FUSED>(a, lc[s], nq, (nq * a.M + m) * a.LP + j1 + s * LPA, on);
        }
        if (j1 == 0 && ulv[t] < a.Q) uq[ulv[t]] = qv[t];
    }
    __syncthreads();   // uq complete
    // grad_out rows of the tile: issued while the raw inputs are still in flight, consumed after the sort
    {
        const unsigned char *gsrc = static_cast<const unsigned char *>(a.grad_out) + ((size_t)n * a.Lq * a.M + m) * rb;
        constexpr int VPR = rb / 16;   // 16-byte vectors per row
        for (int v = tid; v < Qc * VPR; v += kLeanThreads) {
            const int ul = v / VPR, jj = v % VPR;
            cp_async16(Gs + ul * rb + jj * 16, gsrc + (size_t)uq[ul] * a.M * rb + jj * 16);
        }
    }

    // ---- decode; per-level range of first rows ----
    int key[PASSES][SPL], rank[PASSES][SPL];
    float wa[PASSES][SPL], wb[PASSES][SPL], ga[PASSES][SPL], gb[PASSES][SPL], At[PASSES][SPL], dxd[PASSES][SPL];
    bool two[PASSES][SPL];
    int lo[SPL], hi[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) { lo[s] = INT_MAX; hi[s] = -1; }
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            key[t][s] = -1; rank[t][s] = 0; two[t][s] = false;
            wa[t][s] = wb[t][s] = ga[t][s] = gb[t][s] = At[t][s] = dxd[t][s] = 0.f;
        }
        if (t < passes) {   // block-uniform
            bool on[SPL];
#pragma unroll
            for (int s = 0; s < SPL; ++s) on[s] = lc[s].live && qv[t] >= 0;
            Decoded dd[SPL];
            decode<FUSED, LPA, SPL>(a, lc, raw[t], on, dd);
#pragma unroll
            for (int s = 0; s < SPL; ++s) {
                const Decoded &d = dd[s];
                if (on[s]) {
                    float xa = d.w0, xb = d.w1, sa = d.s0, sb = d.s1;
                    if (d.r1 == d.r0) { xa += xb; sa += sb; xb = 0.f; sb = 0.f; }   // one row only
                    const float at = d.A * lc[s].tf;
                    wa[t][s] = xa; wb[t][s] = xb; ga[t][s] = sa * at; gb[t][s] = sb * at; At[t][s] = d.A; dxd[t][s] = d.dxdoff;
                    if (xa != 0.f || xb != 0.f || sa != 0.f || sb != 0.f) {
                        key[t][s] = d.r0;
                        two[t][s] = xb != 0.f || sb != 0.f;
                        lo[s] = min(lo[s], d.r0);
                        hi[s] = max(hi[s], d.r0);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int s = 0; s < SPL; ++s) window_update(win_lo, win_hi, lc[s].l, lo[s], hi[s]);
    __syncthreads();   // uq, win_lo / win_hi complete
    MSDA_PHASE_MARK

    // compact row index of slice row r at level l: wbase_l + r; tot = rows in all level windows (<= S)
    int tot = 0, my_base[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) my_base[s] = 0;
    for (int l = 0; l < a.L; ++l) {
        const int wl = win_lo[l], wh = win_hi[l];
#pragma unroll
        for (int s = 0; s < SPL; ++s)
            if (l == lc[s].l) my_base[s] = tot - wl;
        if (wh >= wl) tot += wh - wl + 1;
    }
    for (int i = tid; i <= tot; i += kLeanThreads) start[i] = 0;
    // value window: rows [lo_l, hi_l + 1] of every touched level (first rows and their second taps), cp.async
    int vtot = 0, my_vbase[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) my_vbase[s] = 0;
    for (int l = 0; l < a.L; ++l) {
#pragma unroll
        for (int s = 0; s < SPL; ++s)
            if (l == lc[s].l) my_vbase[s] = vtot - win_lo[l];
        vtot += win_hi[l] >= win_lo[l] ? win_hi[l] - win_lo[l] + 2 : 0;
    }
    const bool vstaged = vtot <= a.win_rows;
    if (vstaged) {
        constexpr int VPR = rb / 16;
        const unsigned char *vsrc = reinterpret_cast<const unsigned char *>(static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D);
        int cum = 0;
        for (int l = 0; l < a.L; ++l) {
            const int wl = win_lo[l], wh = win_hi[l];
            if (wh < wl) continue;
            const int nr = min(wh + 1, a.S - 1) - wl + 1;
            for (int v = tid; v < nr * VPR; v += kLeanThreads) {
                const int rr = v / VPR, jj = v % VPR;
                cp_async16(Vw + (cum + rr) * rb + jj * 16, vsrc + (size_t)(wl + rr) * MD * sizeof(VT) + jj * 16);
            }
            cum += wh - wl + 2;
        }
    }
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- histogram of first rows ----
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
#pragma unroll
        for (int s = 0; s < SPL; ++s)
            if (key[t][s] >= 0) rank[t][s] = atomicAdd(&start[my_base[s] + key[t][s]], 1);
    }
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- exclusive prefix over the (even-padded) histogram ----
    {
        const int B = (tot + kLeanThreads - 1) / kLeanThreads;
        const int i0 = min(tid * B, tot), i1 = min(i0 + B, tot);
        int cnt = 0;   // padded count of this thread's rows
        for (int i = i0; i < i1; ++i) cnt += ((start[i] & 0xffff) + 1) & ~1;
        int incl = cnt;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(kFull, incl, off);
            if (lane >= off) incl += t;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
    MSDA_PHASE_MARK
        int prev = 0;
        for (int w = 0; w < warp; ++w) prev += wsum[w];
        int pos = prev + incl - cnt;
        for (int i = i0; i < i1; ++i) {
            const int c = start[i] & 0xffff;
            start[i] = pos;
            if (c & 1) {   // pad the row's run to an even length with a no-op entry of the same row
                int r = i, cum = 0;   // compact index -> slice row
                for (int l = 0; l < a.L; ++l) {
                    const int wl = win_lo[l], wh = win_hi[l];
                    const int sz = wh >= wl ? wh - wl + 1 : 0;
                    if (i >= cum && i < cum + sz) r = wl + (i - cum);
                    cum += sz;
                }
                ent[pos + c] = make_int4(QS, 0, 0, r);
            }
            pos += (c + 1) & ~1;
        }
        if (tid == kLeanThreads - 1) n_ent_s = prev + incl;
    }
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- scatter samples into row order ----
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            if (key[t][s] >= 0) {
                const int ul = ulv[t];
                ent[start[my_base[s] + key[t][s]] + rank[t][s]] =
                    make_int4((ul * LPT + s * LPA + j1) | (two[t][s] ? 0x8000 : 0) | ((my_vbase[s] + key[t][s]) << 16),
                              __float_as_int(At[t][s] * wa[t][s]), __float_as_int(At[t][s] * wb[t][s]), key[t][s]);
            }
        }
    }
    cp_async_commit_wait_all();
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- row-ordered sweep: every lane group takes an equal share of the (row-sorted) entry pairs ----
    // A group keeps V[r], V[r+1] and the two grad_value accumulators of its current row in registers and
    // flushes them with one vector reduction per row change; a row split between groups is simply reduced
    // twice.  Balanced by construction (coarse levels have ~6x more samples per row than level 0).
    {
        const int pairs = n_ent_s >> 1;
        // pairs per group, rounded up to 1 (mod 4): the four groups of a warp then read entries 32 (mod 128)
        // bytes apart -- conflict-free LDS.128 of the entry pairs
        const int pp = (((pairs + NG - 1) / NG + 2) & ~3) + 1;
        const int gid = tid / LPG, j = lane % LPG;
        const VT *vslice = static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D + j * CH;
        float *gslice = a.gv_acc + (size_t)n * a.S * MD + m * a.D + j * CH;
        const unsigned char *glane = Gs + j * 16;
        float *dots_f = reinterpret_cast<float *>(dots);
        float v0[NF], v1[NF], acc0[NF], acc1[NF];
#pragma unroll
        for (int k = 0; k < NF; ++k) v0[k] = v1[k] = acc0[k] = acc1[k] = 0.f;
        int cur = -1;
        bool any1 = false;
        auto flush = [&]() {
            if (cur >= 0) {
#pragma unroll
                for (int c = 0; c < NCH; ++c)
#pragma unroll
                    for (int k = 0; k < CH; k += 4)
                        red_add_v4(gslice + cur * MD + c * LPG * CH + k, acc0[c * CH + k], acc0[c * CH + k + 1],
                                   acc0[c * CH + k + 2], acc0[c * CH + k + 3]);
                if (any1) {
#pragma unroll
                    for (int c = 0; c < NCH; ++c)
#pragma unroll
                        for (int k = 0; k < CH; k += 4)
                            red_add_v4(gslice + (cur + 1) * MD + c * LPG * CH + k, acc1[c * CH + k], acc1[c * CH + k + 1],
                                       acc1[c * CH + k + 2], acc1[c * CH + k + 3]);
                }
            }
        };
        for (int it = 0; it < pp; ++it) {
            const int pi = gid * pp + it;
            const bool ok = pi < pairs;
            const int4 e0 = ent[ok ? 2 * pi : 2 * QS];
            const int4 e1 = ent[ok ? 2 * pi + 1 : 2 * QS];
            const int row = ok ? e0.w : cur;
            const bool chg = row != cur;
            if (__any_sync(kFull, chg)) {
                if (chg) {
                    flush();
                    cur = row;
                    any1 = false;
#pragma unroll
                    for (int k = 0; k < NF; ++k) v1[k] = acc0[k] = acc1[k] = 0.f;
                    if (vstaged) {
                        const int slot = (int)((unsigned)e0.x >> 16);
                        IO::load_smem(Vw + slot * rb + j * 16, v0);
                        if (row + 1 < a.S) IO::load_smem(Vw + (slot + 1) * rb + j * 16, v1);
                    } else {
                        IO::load_gmem(vslice + row * MD, v0);
                        if (row + 1 < a.S) IO::load_gmem(vslice + (row + 1) * MD, v1);
                    }
                }
            }
            any1 = any1 || ((e0.x | e1.x) & 0x8000) != 0;
            float gq0[NF], gq1[NF];
            int u0 = (e0.x & 0x7fff) / LPT, u1 = (e1.x & 0x7fff) / LPT;   // no-op entries (slot QS) read unit 0
            u0 = u0 < a.Q ? u0 : 0;
            u1 = u1 < a.Q ? u1 : 0;
            IO::load_smem(glane + u0 * rb, gq0);
            IO::load_smem(glane + u1 * rb, gq1);
            axpy<NF>(acc0, __int_as_float(e0.y), gq0);
            axpy<NF>(acc1, __int_as_float(e0.z), gq0);
            axpy<NF>(acc0, __int_as_float(e1.y), gq1);
            axpy<NF>(acc1, __int_as_float(e1.z), gq1);
            float p[4];
            p[0] = dot<NF>(gq0, v0);
            p[1] = dot<NF>(gq0, v1);
            p[2] = dot<NF>(gq1, v0);
            p[3] = dot<NF>(gq1, v1);
            reduce_nv<LPG, 4>(p, j);
            if (LPG >= 4) {
                constexpr int LPV = LPG / 4;   // lanes per value
                const int vi = j / LPV;
                if (j % LPV == 0) dots_f[2 * ((vi < 2 ? e0.x : e1.x) & 0x7fff) + (vi & 1)] = p[0];
            }
        }
        flush();
    }
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- per-sample gradients (everything but the two dot products is still in registers) ----
    {
        const bool p_pow2 = (a.P & (a.P - 1)) == 0 && a.P <= LPA;
#pragma unroll
        for (int t = 0; t < PASSES; ++t) {
            if (t < passes) {   // block-uniform
                const int ul = ulv[t];
                const int nq = n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
                float gat[SPL], gx[SPL];
                float part = 0.f;
#pragma unroll
                for (int s = 0; s < SPL; ++s) {
                    float2 pd = make_float2(0.f, 0.f);
                    if (key[t][s] >= 0) pd = dots[ul * LPT + s * LPA + j1];
                    gat[s] = wa[t][s] * pd.x + wb[t][s] * pd.y;   // d out / d attention weight
                    gx[s] = ga[t][s] * pd.x + gb[t][s] * pd.y;    // d out / d location
                    part += At[t][s] * gat[s];
                }
                // softmax Jacobian: g_logit_s = A_s * (ga_s - sum_t A_t ga_t)
                const float dotsum = FUSED ? group_sum<LPA>(part) : 0.f;
#pragma unroll
                for (int s = 0; s < SPL; ++s) {
                    const bool on = lc[s].live && qv[t] >= 0;
                    const int sidx = (nq * a.M + m) * a.LP + j1 + s * LPA;
                    if (!FUSED) {
                        if (on) {
                            a.grad_attn[sidx] = gat[s];
                            // dummy y (= 0.5 on an H = 1 level): BORDER clamps it, gradient exactly 0; ZEROS keeps the
                            // row-0 weight 1 - |y*H - 0.5| whose slope the reference kernel reports (cuh:159)
                            const float gy = a.pad_mode == MSDA_PAD_ZEROS ? -At[t][s] * gat[s] : 0.f;
                            *reinterpret_cast<float2 *>(a.grad_loc + 2 * sidx) = make_float2(gx[s], gy);
                        }
                    } else {
                        if (on) {
                            a.grad_offs[sidx] = gx[s] * dxd[t][s];
                            a.grad_logits[sidx] = At[t][s] * (gat[s] - dotsum);
                        }
                        if (a.grad_ref != nullptr) {
                            float r0g = on ? gx[s] : 0.f;   // d(loc)/d(ref0) = 1
                            float r1g = 0.f;                // d(loc)/d(ref1) = off * 0.5 / P
                            if (a.R == 2 && on) r1g = gx[s] * raw[t][s].x * (0.5f * a.invP);
                            if (p_pow2) {   // the P lanes of one level are contiguous and P-aligned
                                for (int off = a.P >> 1; off >= 1; off >>= 1) {
                                    r0g += __shfl_xor_sync(kFull, r0g, off);
                                    r1g += __shfl_xor_sync(kFull, r1g, off);
                                }
                                if (on && (j1 & (a.P - 1)) == 0) {
                                    float *gp = a.grad_ref + (nq * a.L + lc[s].l) * a.R;
                                    atomicAdd(gp, r0g);
                                    if (a.R == 2) atomicAdd(gp + 1, r1g);
                                }
                            } else if (on) {
                                float *gp = a.grad_ref + (nq * a.L + lc[s].l) * a.R;
                                atomicAdd(gp, r0g);
                                if (a.R == 2) atomicAdd(gp + 1, r1g);
                            }
                        }
                    }
                }
            }
        }
    }
    MSDA_PHASE_MARK
    MSDA_PHASE_FLUSH(1)
}

}  // namespace msda
