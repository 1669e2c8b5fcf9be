// msda_lean.cuh -- the production kernels of the 1-D temporal MSDeformAttn hot path for sm_100a.
//
// One CTA = one head m of one batch element n and a tile of up to Q queries (forward: the value rows the tile
// samples are staged per level in shared memory and gathered with LDS.128; backward: the tile's samples are
// counting-sorted by value row and swept in row order with V[r], V[r+1] and the two grad_value accumulators in
// registers).  Every lane decodes ONE sample per query (L*P <= 32; two, SPL = 2, up to L*P = 64), so level, T_l and
// level start are per-lane constants, all indices are 32-bit and decoded samples stay in registers between phases.
//
// Round 2 (profiles/r2_notes.md): both kernels were instruction-issue bound, and two thirds of the instructions
// were plumbing -- tile set-up (per-thread level loops, integer divisions), staging address arithmetic, entry
// unpacking and the row-change path of the sweep.  This version
//   * builds the tile's unit -> query table once per CTA in shared memory (64 threads) instead of per thread,
//   * stages rows with one fixed 16-byte column per thread (pointer increments instead of div/mod per copy),
//   * writes the two dot products of a sample back INTO its sorted entry (no separate dots array), keeps 4 instead
//     of 9 registers per sample across the sweep and sizes the histogram by the window (three CTAs would fit an SM's
//     shared memory; the kernel is nevertheless compiled for two, see MSDA_BWD_MINB below),
//   * carries acc1 -> acc0 and V[r+1] -> V[r] when the sweep moves to the next row (half the vector reductions),
//   * sorts nothing when the tile's rows do not fit the histogram (scattered decoder queries on long videos):
//     the same sweep then takes the samples one by one.
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

// CTAs per SM the backward is compiled for: 3 fit (shared memory) but cap the kernel at 80 registers, which spills 200
// bytes per thread and is 5 % slower than 2 CTAs at 121 registers (0.910 vs 0.865 ms at the headline, tools/r2_ab.sh)
#ifndef MSDA_BWD_MINB
#define MSDA_BWD_MINB 2
#endif
#ifndef MSDA_FWD_MINB
#define MSDA_FWD_MINB 4
#endif
#ifndef MSDA_FWD_MINB2      // two samples per lane (32 < L*P <= 64): fits 64 registers without spills; 4 instead of 2 CTAs/SM
#define MSDA_FWD_MINB2 4    // took the P = 16 forward from 2.40 to 1.50 ms (N = 64, T0 = 1000; profiles/r2_notes.md)
#endif
constexpr int kLeanThreads = 256;
constexpr int kLeanWarps = kLeanThreads / 32;
constexpr int kLeanPasses = 4;   // units decoded per lane group and CTA (Q = kLeanPasses * units per pass)
constexpr int kLeanQMax = 64;    // most queries a tile can hold (LPA = 16, SPL = 1)

// ---- per-CTA tables ------------------------------------------------------------------------------------
// Tiles: gridK == 0: tile k = the Q consecutive queries [k*Q, k*Q + Q).
//        gridK  > 0 (self-attention, Lq == S: the queries ARE the rows of the level pyramid,
//            pdvc/deformable_transformer.py:198-208): tile k takes from EVERY level the rows whose position
//            p satisfies k <= p*K/T_l < k+1, i.e. all queries of one time span.  Their samples then fall into
//            one short row window per level (span + offsets) instead of Q/T_l of the whole pyramid, so the
//            window always fits in shared memory.  Pure scheduling: results do not depend on it.
struct LeanTables {
    int T[kMaxLevels];        // level length
    int start[kMaxLevels];    // first slice row of the level
    int run_first[kMaxLevels], run_cnt[kMaxLevels];   // level-grid tile: this tile's rows of every level
    int win_lo[kMaxLevels], win_hi[kMaxLevels];       // rows the tile's samples touch, per level
    int uq[kLeanQMax];        // query of every unit of the tile (-1: none)
    int Qc;                   // units in use
};

// level table, tile runs and the unit -> query table; two barriers, ends synchronised
__device__ __forceinline__ void lean_setup(const TileArgs &a, int k, LeanTables &tb)
{
    const int tid = threadIdx.x;
    if (tid < a.L) {
        const int T = (int)__ldg(a.shapes + (a.shapes_stride == 2 ? 2 * tid + 1 : tid));
        const int st = (int)__ldg(a.lsi + tid);
        tb.T[tid] = T;
        tb.start[tid] = st;
        tb.win_lo[tid] = INT_MAX;
        tb.win_hi[tid] = -1;
        if (a.gridK) {
            const unsigned K = (unsigned)a.gridK;
            const unsigned p0 = ((unsigned)k * (unsigned)T + K - 1) / K;          // S < 65536: no overflow
            const unsigned p1 = min((unsigned)T, ((unsigned)(k + 1) * (unsigned)T + K - 1) / K);
            tb.run_first[tid] = st + (int)p0;
            tb.run_cnt[tid] = p1 > p0 ? (int)(p1 - p0) : 0;
        }
    }
    __syncthreads();
    if (tid < kLeanQMax) {
        int q = -1, used;
        if (a.gridK == 0) {
            const int q0 = k * a.Q;
            used = min(a.Q, a.Lq - q0);
            if (tid < used) q = q0 + tid;
        } else {
            used = 0;
#pragma unroll 1
            for (int l = 0; l < a.L; ++l) {
                const int cnt = min(tb.run_cnt[l], a.Q - used);   // never clips when sum(T_l) == Lq (host picks K for that)
                const int rel = tid - used;
                if (rel >= 0 && rel < cnt) q = tb.run_first[l] + rel;
                used += cnt;
            }
        }
        tb.uq[tid] = q;
        if (tid == 0) tb.Qc = used;
    }
    __syncthreads();
}

// per-lane constants of the decode phase: lane j of a unit's group always decodes sample j = (l, p)
struct LaneLevel {
    int l, T, start;
    float tf;
    bool live;   // j < L*P
};

__device__ __forceinline__ LaneLevel lane_level(const TileArgs &a, const LeanTables &tb, int j)
{
    LaneLevel c;
    c.live = j < a.LP;
    int l = 0;
#pragma unroll 1
    for (int i = 1; i < a.L; ++i) l += (j >= i * a.P) ? 1 : 0;   // j / P without the division (L is small)
    c.l = l;
    c.T = tb.T[l];
    c.start = tb.start[l];
    c.tf = (float)c.T;
    return c;
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

// Copy `rows` rows of `rb` bytes (row stride `src_stride` bytes in global memory) into consecutive shared-memory rows
// with 16-byte cp.async: thread t owns column t % VPR and every (256 / VPR)-th row -- no div/mod per copy.
template <int RB>
__device__ __forceinline__ void stage_rows(unsigned char *dst, const unsigned char *src, size_t src_stride, int rows)
{
    constexpr int VPR = RB / 16;                 // 16-byte vectors per row
    constexpr int RPI = kLeanThreads / VPR;      // rows per iteration
    const int jj = threadIdx.x % VPR, rr0 = threadIdx.x / VPR;
    dst += rr0 * RB + jj * 16;
    src += (size_t)rr0 * src_stride + jj * 16;
    for (int rr = rr0; rr < rows; rr += RPI) {
        cp_async16(dst, src);
        dst += RPI * RB;
        src += (size_t)RPI * src_stride;
    }
}

// ---------------------------------------------------------------------------------------------
// forward.  grid = (chunks, M, N); dynamic shared memory = [ records int4 x Q*LPT | value window ]
// ---------------------------------------------------------------------------------------------
template <typename VT, int LPG, int NCH, int LPA, bool FUSED, int SPL = 1>
__global__ void __launch_bounds__(kLeanThreads, (SPL == 1 ? MSDA_FWD_MINB : MSDA_FWD_MINB2)) msda_fwd_lean_kernel(const TileArgs a)
{
    using IO = RowIO<VT, LPG, NCH>;
    constexpr int CH = Vec16<VT>::CH;
    constexpr int NF = IO::NF;
    constexpr int G = 32 / LPG;     // units per warp in the gather phase
    constexpr int UPW = 32 / LPA;   // units per warp in the decode phase
    constexpr int UPP = kLeanWarps * UPW;
    constexpr int PASSES = kLeanPasses / SPL;   // SPL samples per lane (32 < L*P <= 64): half the units per lane
    constexpr int LPT = LPA * SPL;              // sample slots per unit
    constexpr int rb = LPG * NCH * 16;          // bytes per row of one head (= D * sizeof(VT))
    extern __shared__ __align__(128) unsigned char dsm[];
    MSDA_PHASE_DECL
    __shared__ LeanTables tb;
    __shared__ int win_base[kMaxLevels];
    __shared__ int win_ok;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n = blockIdx.z, m = blockIdx.y;
    const int MD = a.M * a.D;
    int4 *recs = reinterpret_cast<int4 *>(dsm);
    unsigned char *win = dsm + (size_t)a.Q * LPT * sizeof(int4);

    lean_setup(a, blockIdx.x, tb);
    MSDA_PHASE_MARK

    // ---- decode: lane (g, j) handles samples j + s*LPA of units warp*UPW + g + t*UPP, t < passes ----
    const int g1 = lane / LPA, j1 = lane % LPA;
    LaneLevel lc[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) lc[s] = lane_level(a, tb, j1 + s * LPA);
    const int Qc = tb.Qc;
    const int passes = (Qc + UPP - 1) / UPP;
    const int ul0 = warp * UPW + g1;
    int qv[PASSES];
    RawSample raw[PASSES][SPL];
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
        qv[t] = tb.uq[ul0 + t * UPP];   // < UPP * PASSES <= kLeanQMax; -1 beyond the tile's units
        const int nq = n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            const bool on = lc[s].live && qv[t] >= 0;
            raw[t][s] = load_raw<FUSED>(a, lc[s], nq, (nq * a.M + m) * a.LP + j1 + s * LPA, on);
        }
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
    for (int s = 0; s < SPL; ++s) window_update(tb.win_lo, tb.win_hi, lc[s].l, lo[s], hi[s]);
    __syncthreads();
    MSDA_PHASE_MARK
    if (tid == 0) {
        int tot = 0;
        for (int l = 0; l < a.L; ++l) {
            win_base[l] = tot - tb.win_lo[l];   // window row of slice row r at level l: win_base[l] + r
            if (tb.win_hi[l] >= tb.win_lo[l]) tot += tb.win_hi[l] - tb.win_lo[l] + 1;
        }
        win_ok = (tot > 0 && tot <= a.win_rows) ? 1 : 0;
    }
    __syncthreads();
    MSDA_PHASE_MARK
    const bool staged = win_ok != 0;
    const VT *vslice = static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D;
    if (staged) {
        // 16-byte cp.async per lane: measured 2x cheaper to issue than one 256-byte cp.async.bulk per row
        // (profiles/r1_phases.txt); the TMA engine serialises small bulk copies
        for (int l = 0; l < a.L; ++l) {
            const int wl = tb.win_lo[l], wh = tb.win_hi[l];
            if (wh < wl) continue;   // level without a live sample (ZEROS padding can drop them all)
            stage_rows<rb>(win + (size_t)(win_base[l] + wl) * rb,
                           reinterpret_cast<const unsigned char *>(vslice + (size_t)wl * MD), (size_t)MD * sizeof(VT), wh - wl + 1);
        }
    }
    {   // final records: byte offsets into the window (staged) or element offsets into the slice
        const int scale = staged ? rb : MD;
#pragma unroll
        for (int t = 0; t < PASSES; ++t) {
            const int ul = ul0 + t * UPP;
            if (t < passes && ul < Qc) {
#pragma unroll
                for (int s = 0; s < SPL; ++s) {
                    const int base = staged ? win_base[lc[s].l] : 0;
                    const bool livet = r0[t][s] >= 0;
                    recs[ul * LPT + s * LPA + (j1 ^ rec_swizzle(ul))] =
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
        for (int ulb = warp * G; ulb < Qc; ulb += kLeanWarps * G) {
            const int ul = ulb + g;
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
            if (active) IO::store_gmem(out + (size_t)((n * a.Lq + tb.uq[ul]) * a.M + m) * a.D + j * CH, acc);
        }
    }
    MSDA_PHASE_MARK
    MSDA_PHASE_FLUSH(0)
}


// ---------------------------------------------------------------------------------------------
// backward.  grid = (chunks, M, N).  Dynamic shared memory (QS = Q * LPT sample slots):
//   Gs    Q x D VT              grad_out rows of the tile (this head)
//   ent   ent_cap int4          samples in value-row order, every row's run padded to an even length:
//                               { G row offset | second_tap << 16 | window slot << 17, A*wa, A*wb, row };
//                               the sweep overwrites .y/.z with the sample's two dot products {G.V[r], G.V[r+1]};
//                               ent[ent_cap - 1] = the no-op entry
//   start (hist_cap + 1) int    histogram of first rows -> start of each row's run in ent
//   Vw    win_rows x D VT       value rows of the tile's level windows (falls back to global loads if they do not fit)
// Three modes per tile: rows sorted + window staged; sorted, window read through L1 (it does not fit);
// unsorted (the rows do not fit the histogram: every sample is swept on its own).
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline int bwd_lean_ent_cap(int QS, int hist_cap) { return QS + (hist_cap < QS ? hist_cap : QS) + 2; }

__host__ __device__ inline size_t bwd_lean_smem_bytes(int Q, int LPT, int hist_cap, int rb, int win_rows)
{
    const int QS = Q * LPT;
    return (size_t)Q * rb + (size_t)bwd_lean_ent_cap(QS, hist_cap) * 16 + ((size_t)hist_cap + 1 + 3) / 4 * 16 + (size_t)win_rows * rb;
}

constexpr int kEntTwoBit = 1 << 16;
constexpr int kEntSlotShift = 17;
// per-sample state kept across the sweep: position in ent | (sa + 1) << 12 | sb << 14 | live << 15
constexpr int kMetaPosMask = 0xfff;

template <typename VT, int LPG, int NCH, int LPA, bool FUSED, int SPL = 1>
__global__ void __launch_bounds__(kLeanThreads, (SPL == 1 ? MSDA_BWD_MINB : 2)) msda_bwd_lean_kernel(const TileArgs a)
{
    using IO = RowIO<VT, LPG, NCH>;
    constexpr int CH = Vec16<VT>::CH;
    constexpr int NF = IO::NF;
    constexpr int NG = kLeanThreads / LPG;   // row groups per CTA
    constexpr int UPW = 32 / LPA;
    constexpr int UPP = kLeanWarps * UPW;
    constexpr int PASSES = kLeanPasses / SPL;   // SPL samples per lane (32 < L*P <= 64): half the units per lane
    constexpr int LPT = LPA * SPL;              // sample slots per unit
    constexpr int rb = LPG * NCH * 16;
    extern __shared__ __align__(128) unsigned char dsm[];
    MSDA_PHASE_DECL
    __shared__ LeanTables tb;
    __shared__ int wsum[kLeanWarps];
    __shared__ int n_ent_s;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n = blockIdx.z, m = blockIdx.y;
    const int MD = a.M * a.D;
    const int QS = a.Q * LPT;
    const int ent_cap = bwd_lean_ent_cap(QS, a.hist_cap);
    const int NOOP = ent_cap - 1;
    unsigned char *Gs = dsm;
    int4 *ent = reinterpret_cast<int4 *>(Gs + (size_t)a.Q * rb);
    int *start = reinterpret_cast<int *>(ent + ent_cap);
    unsigned char *Vw = reinterpret_cast<unsigned char *>(start + (a.hist_cap + 1 + 3) / 4 * 4);   // value window, win_rows x rb

    lean_setup(a, blockIdx.x, tb);
    MSDA_PHASE_MARK

    // ---- raw inputs of this lane's samples ----
    const int g1 = lane / LPA, j1 = lane % LPA;
    LaneLevel lc[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) lc[s] = lane_level(a, tb, j1 + s * LPA);
    const int Qc = tb.Qc;
    const int passes = (Qc + UPP - 1) / UPP;
    const int ul0 = warp * UPW + g1;
    int qv[PASSES];
    RawSample raw[PASSES][SPL];
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
        qv[t] = tb.uq[ul0 + t * UPP];   // < UPP * PASSES <= kLeanQMax; -1 beyond the tile's units
        const int nq = n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            const bool on = lc[s].live && qv[t] >= 0;
            raw[t][s] = load_raw<FUSED>(a, lc[s], nq, (nq * a.M + m) * a.LP + j1 + s * LPA, on);
        }
    }
    // grad_out rows of the tile: issued while the raw inputs are still in flight, consumed after the sort
    {
        constexpr int VPR = rb / 16, RPI = kLeanThreads / VPR;
        const int jj = tid % VPR;
        const unsigned char *gsrc = static_cast<const unsigned char *>(a.grad_out) + ((size_t)n * a.Lq * a.M + m) * rb + jj * 16;
        for (int ul = tid / VPR; ul < Qc; ul += RPI)
            cp_async16(Gs + ul * rb + jj * 16, gsrc + (size_t)tb.uq[ul] * a.M * rb);
    }
    if (tid == 0) ent[NOOP] = make_int4(0, 0, 0, -2);   // no-op entry: G row of unit 0, zero weights, no row

    // ---- decode; per-level range of first rows ----
    int key[PASSES][SPL], meta[PASSES][SPL];
    float wa[PASSES][SPL], wb[PASSES][SPL], At[PASSES][SPL];
    int lo[SPL], hi[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) { lo[s] = INT_MAX; hi[s] = -1; }
#pragma unroll
    for (int t = 0; t < PASSES; ++t) {
#pragma unroll
        for (int s = 0; s < SPL; ++s) {
            key[t][s] = -1; meta[t][s] = 0;
            wa[t][s] = wb[t][s] = At[t][s] = 0.f;
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
                    wa[t][s] = xa; wb[t][s] = xb; At[t][s] = d.A;
                    if (xa != 0.f || xb != 0.f || sa != 0.f || sb != 0.f) {
                        key[t][s] = d.r0;
                        meta[t][s] = (((int)sa + 1) << 12) | (sb != 0.f ? 1 << 14 : 0) | (1 << 15);
                        lo[s] = min(lo[s], d.r0);
                        hi[s] = max(hi[s], d.r0);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int s = 0; s < SPL; ++s) window_update(tb.win_lo, tb.win_hi, lc[s].l, lo[s], hi[s]);
    __syncthreads();   // win_lo / win_hi complete
    MSDA_PHASE_MARK

    // compact row index of slice row r at level l: my_base + r; tot = first rows in all level windows.
    // value window: rows [lo_l, hi_l + 1] of every touched level (first rows and their second taps): my_vbase + r
    int tot = 0, vtot = 0, my_base[SPL], my_vbase[SPL];
#pragma unroll
    for (int s = 0; s < SPL; ++s) my_base[s] = my_vbase[s] = 0;
    for (int l = 0; l < a.L; ++l) {
        const int wl = tb.win_lo[l], wh = tb.win_hi[l];
#pragma unroll
        for (int s = 0; s < SPL; ++s)
            if (l == lc[s].l) { my_base[s] = tot - wl; my_vbase[s] = vtot - wl; }
        if (wh >= wl) { tot += wh - wl + 1; vtot += wh - wl + 2; }
    }
    const bool sorted = tot <= a.hist_cap;
    const bool vstaged = vtot <= a.win_rows;
    if (sorted)
        for (int i = tid; i <= tot; i += kLeanThreads) start[i] = 0;
    if (vstaged) {
        const unsigned char *vsrc = reinterpret_cast<const unsigned char *>(static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D);
        int cum = 0;
        for (int l = 0; l < a.L; ++l) {
            const int wl = tb.win_lo[l], wh = tb.win_hi[l];
            if (wh < wl) continue;
            stage_rows<rb>(Vw + (size_t)cum * rb, vsrc + (size_t)wl * MD * sizeof(VT), (size_t)MD * sizeof(VT),
                           min(wh + 1, a.S - 1) - wl + 1);
            cum += wh - wl + 2;
        }
    }
    __syncthreads();
    MSDA_PHASE_MARK

    int n_ent;
    if (sorted) {
        // ---- histogram of first rows ----
        int rank[PASSES][SPL];
#pragma unroll
        for (int t = 0; t < PASSES; ++t) {
#pragma unroll
            for (int s = 0; s < SPL; ++s) {
                rank[t][s] = 0;
                if (key[t][s] >= 0) rank[t][s] = atomicAdd(&start[my_base[s] + key[t][s]], 1);
            }
        }
        __syncthreads();
        MSDA_PHASE_MARK

        // ---- exclusive prefix over the (even-padded) histogram ----
        {
            const int B = (tot + kLeanThreads - 1) / kLeanThreads;
            const int i0 = min(tid * B, tot), i1 = min(i0 + B, tot);
            int cnt = 0;   // padded count of this thread's rows
            for (int i = i0; i < i1; ++i) cnt += (start[i] + 1) & ~1;
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
                const int c = start[i];
                start[i] = pos;
                if (c & 1) ent[pos + c] = make_int4(0, 0, 0, -2);   // pad the row's run to an even length (always a pair's second entry)
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
                    const int ul = ul0 + t * UPP;
                    const int pos = start[my_base[s] + key[t][s]] + rank[t][s];
                    const bool two = wb[t][s] != 0.f || (meta[t][s] & (1 << 14)) != 0;
                    const unsigned slot = vstaged ? (unsigned)(my_vbase[s] + key[t][s]) << kEntSlotShift : 0u;
                    ent[pos] = make_int4((int)((unsigned)(ul * rb) | (two ? kEntTwoBit : 0) | slot),
                                         __float_as_int(At[t][s] * wa[t][s]), __float_as_int(At[t][s] * wb[t][s]), key[t][s]);
                    meta[t][s] |= pos;
                }
            }
        }
        cp_async_commit_wait_all();
        __syncthreads();
        n_ent = n_ent_s;
    } else {
        // ---- unsorted: sample slot = entry; dead slots are no-ops ----
#pragma unroll
        for (int t = 0; t < PASSES; ++t) {
            const int ul = ul0 + t * UPP;
            if (ul < a.Q) {
#pragma unroll
                for (int s = 0; s < SPL; ++s) {
                    const int pos = ul * LPT + s * LPA + j1;
                    int4 e = make_int4(0, 0, 0, -2);
                    if (key[t][s] >= 0) {
                        const bool two = wb[t][s] != 0.f || (meta[t][s] & (1 << 14)) != 0;
                        const unsigned slot = vstaged ? (unsigned)(my_vbase[s] + key[t][s]) << kEntSlotShift : 0u;
                        e = make_int4((int)((unsigned)(ul * rb) | (two ? kEntTwoBit : 0) | slot), __float_as_int(At[t][s] * wa[t][s]),
                                      __float_as_int(At[t][s] * wb[t][s]), key[t][s]);
                        meta[t][s] |= pos;
                    }
                    ent[pos] = e;
                }
            }
        }
        cp_async_commit_wait_all();
        __syncthreads();
        MSDA_PHASE_MARK
        MSDA_PHASE_MARK
        MSDA_PHASE_MARK
        n_ent = QS;
    }
    MSDA_PHASE_MARK

    // ---- row-ordered sweep: every lane group takes an equal share of the (row-sorted) entry pairs ----
    // A group keeps V[r], V[r+1] and the two grad_value accumulators of its current row in registers and
    // flushes them with one vector reduction per row change; when the next entry's row is r+1 the second
    // accumulator and V[r+1] simply become the first (one reduction instead of two, one row load instead of two).
    // A row split between groups is reduced twice.  Balanced by construction (coarse levels have ~6x more samples
    // per row than level 0).  Unsorted tiles: a "pair" is one entry and the no-op.
    {
        const int pairs = sorted ? n_ent >> 1 : n_ent;
        // pairs per group, rounded up to 1 (mod 4): the four groups of a warp then read entries 32 (mod 128)
        // bytes apart -- conflict-free LDS.128 of the entry pairs
        const int pp = (((pairs + NG - 1) / NG + 2) & ~3) + 1;
        const int gid = tid / LPG, j = lane % LPG;
        const VT *vslice = static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D + j * CH;
        float *gslice = a.gv_acc + (size_t)n * a.S * MD + m * a.D + j * CH;
        const unsigned char *glane = Gs + j * 16;
        const unsigned char *vwlane = Vw + j * 16;
        constexpr int LPV = LPG >= 4 ? LPG / 4 : 1;   // lanes per reduced value
        // lane (j % LPV == 0) ends the reduction with value vi = j / LPV: dots of entry vi >> 1, tap vi & 1
        const int vi = j / LPV;
        const bool writer = (j % LPV) == 0;
        float v0[NF], v1[NF], acc0[NF], acc1[NF];
#pragma unroll
        for (int k = 0; k < NF; ++k) v0[k] = v1[k] = acc0[k] = acc1[k] = 0.f;
        int cur = -1;
        bool any1 = false;
        int pi = gid * pp;
        for (int it = 0; it < pp; ++it, ++pi) {
            const bool ok = pi < pairs;
            const int i0 = ok ? (sorted ? 2 * pi : pi) : NOOP;
            const int i1 = (ok && sorted) ? 2 * pi + 1 : NOOP;
            const int4 e0 = ent[i0];
            const int4 e1 = ent[i1];
            const int row = e0.w >= 0 ? e0.w : cur;   // no-op first entries (unsorted dead slots, tail) keep the row
            const bool chg = row != cur;
            if (__any_sync(kFull, chg)) {
                if (chg) {
                    const bool carry = row == cur + 1 && cur >= 0;
                    if (cur >= 0) {
                        float *g0 = gslice + (size_t)cur * MD;
#pragma unroll
                        for (int c = 0; c < NCH; ++c)
#pragma unroll
                            for (int k = 0; k < CH; k += 4)
                                red_add_v4(g0 + c * LPG * CH + k, acc0[c * CH + k], acc0[c * CH + k + 1],
                                           acc0[c * CH + k + 2], acc0[c * CH + k + 3]);
                        if (!carry && any1) {
#pragma unroll
                            for (int c = 0; c < NCH; ++c)
#pragma unroll
                                for (int k = 0; k < CH; k += 4)
                                    red_add_v4(g0 + MD + c * LPG * CH + k, acc1[c * CH + k], acc1[c * CH + k + 1],
                                               acc1[c * CH + k + 2], acc1[c * CH + k + 3]);
                        }
                    }
                    const bool has1 = row + 1 < a.S;
                    float nv1[NF];
#pragma unroll
                    for (int k = 0; k < NF; ++k) nv1[k] = 0.f;
                    if (vstaged) {
                        const unsigned char *vp = vwlane + (size_t)((unsigned)e0.x >> kEntSlotShift) * rb;
                        if (!carry) IO::load_smem(vp, v0);
                        if (has1) IO::load_smem(vp + rb, nv1);
                    } else {
                        if (!carry) IO::load_gmem(vslice + (size_t)row * MD, v0);
                        if (has1) IO::load_gmem(vslice + (size_t)(row + 1) * MD, nv1);
                    }
#pragma unroll
                    for (int k = 0; k < NF; ++k) {
                        if (carry) v0[k] = v1[k];
                        acc0[k] = carry ? acc1[k] : 0.f;
                        acc1[k] = 0.f;
                        v1[k] = nv1[k];
                    }
                    cur = row;
                    any1 = false;
                }
            }
            any1 = any1 || ((e0.x | e1.x) & kEntTwoBit) != 0;
            float gq0[NF], gq1[NF];
            IO::load_smem(glane + (e0.x & 0xffff), gq0);
            IO::load_smem(glane + (e1.x & 0xffff), gq1);
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
                if (writer && ok && (vi < 2 || sorted)) reinterpret_cast<float *>(ent + (vi < 2 ? i0 : i1))[1 + (vi & 1)] = p[0];
            }
        }
        if (cur >= 0) {
            float *g0 = gslice + (size_t)cur * MD;
#pragma unroll
            for (int c = 0; c < NCH; ++c)
#pragma unroll
                for (int k = 0; k < CH; k += 4)
                    red_add_v4(g0 + c * LPG * CH + k, acc0[c * CH + k], acc0[c * CH + k + 1], acc0[c * CH + k + 2],
                               acc0[c * CH + k + 3]);
            if (any1) {
#pragma unroll
                for (int c = 0; c < NCH; ++c)
#pragma unroll
                    for (int k = 0; k < CH; k += 4)
                        red_add_v4(g0 + MD + c * LPG * CH + k, acc1[c * CH + k], acc1[c * CH + k + 1],
                                   acc1[c * CH + k + 2], acc1[c * CH + k + 3]);
            }
        }
    }
    __syncthreads();
    MSDA_PHASE_MARK

    // ---- per-sample gradients: the two dot products come back from the sample's entry ----
    {
        const bool p_pow2 = (a.P & (a.P - 1)) == 0 && a.P <= LPA;
#pragma unroll
        for (int t = 0; t < PASSES; ++t) {
            if (t < passes) {   // block-uniform
                const int nq = n * a.Lq + (qv[t] >= 0 ? qv[t] : 0);
                float gat[SPL], gx[SPL];
                float part = 0.f;
#pragma unroll
                for (int s = 0; s < SPL; ++s) {
                    const int mt = meta[t][s];
                    float p0 = 0.f, p1 = 0.f;
                    if (mt & (1 << 15)) {
                        const int4 e = ent[mt & kMetaPosMask];
                        p0 = __int_as_float(e.y);
                        p1 = __int_as_float(e.z);
                    }
                    const float sa = (float)(((mt >> 12) & 3) - 1), sb = (mt & (1 << 14)) ? 1.f : 0.f;
                    const float t0 = wa[t][s] != 0.f ? wa[t][s] * p0 : 0.f, t1 = wb[t][s] != 0.f ? wb[t][s] * p1 : 0.f;
                    const float u0 = sa != 0.f ? sa * p0 : 0.f, u1 = sb != 0.f ? p1 : 0.f;
                    gat[s] = (mt & (1 << 15)) ? t0 + t1 : 0.f;                            // d out / d attention weight
                    gx[s] = (mt & (1 << 15)) ? At[t][s] * lc[s].tf * (u0 + u1) : 0.f;      // d out / d location
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
                        // d(loc)/d(offset): 1 / T_l, or ref1 * 0.5 / P
                        const float dxd = a.R == 1 ? __fdividef(1.f, lc[s].tf) : raw[t][s].r1 * (0.5f * a.invP);
                        if (on) {
                            a.grad_offs[sidx] = gx[s] * dxd;
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
