// msda_tile.cuh -- chunked ("tile") kernels of the 1-D temporal MSDeformAttn hot path for sm_100a.
//
// Same reference functions as msda_fast.cuh (forward ms_deform_im2col_cuda.cuh:238-300, backward
// cuh:302-921, front-end pdvc/ops/modules/ms_deform_attn.py:100-116) but organised around what the
// B200 measurements in tools/ubench.cu show: a 256-byte-row gather runs at ~33 TB/s from shared memory
// but an L1-missing LDG gather at ~11 TB/s, and red.global.add.v4.f32 tops out at ~5 TB/s of row updates.
//
// Work decomposition: one CTA = one head m of one batch element n and a chunk of Q consecutive queries.
//   phase A  every sample (q, l, p) of the chunk is decoded ONCE by one lane (softmax weight, sampling
//            location, clamped rows, interpolation coefficients) into a shared-memory record.
//   forward  the value rows the chunk touches form, per level, a contiguous window; it is staged in
//            shared memory with cp.async.bulk (one bulk copy per row of the head, completion on an
//            mbarrier) and the gather runs on LDS.128.  If the window does not fit (scattered
//            locations) the same loop gathers through L1 with LDG.128.
//   backward the gather is transposed: samples are counting-sorted by their first value row, and a
//            lane group then owns one row r: V[r] and V[r+1] sit in registers, it walks the row's
//            samples reading grad_out[q] from shared memory, accumulates grad_value[r], grad_value[r+1]
//            in registers and emits the two dot products per sample (for grad_attn / grad_loc).  One
//            vector reduction per touched row and chunk replaces one per tap (the reference issues
//            one scalar atomicAdd per tap and channel, cuh:126-153).
#pragma once
#include "msda_common.cuh"
#include "msda_fast.cuh"   // two_sum

namespace msda {

constexpr int kTileThreads = 256;
constexpr int kTileWarps = kTileThreads / 32;
constexpr int kTileMaxSamples = 2048;   // Q * LPpad bound (shared-memory lists)

struct TileArgs {
    const void *value;
    const int64_t *shapes;
    const int64_t *lsi;
    int shapes_stride;
    const float *loc;     // unfused (N,Lq,M,L,P,2)
    const float *attn;    // unfused (N,Lq,M,L,P)
    const float *ref;     // fused (N,Lq,L,R)
    const float *offs;    // fused (N,Lq,M,L,P)
    const float *logits;  // fused (N,Lq,M,L,P)
    int R;
    void *out;             // forward (N,Lq,M*D)
    const void *grad_out;  // backward (N,Lq,M*D)
    float *gv_acc;         // fp32 accumulation image of grad_value, pre-zeroed
    float *grad_loc;       // unfused
    float *grad_attn;      // unfused
    float *grad_ref;       // fused, pre-zeroed, may be NULL
    float *grad_offs;      // fused
    float *grad_logits;    // fused
    int N, S, M, D, L, Lq, P;
    int pad_mode;
    int LP;        // L * P
    int LPpad;     // LP rounded up to a multiple of 4 (record stride)
    float invP;    // 1 / P
    int Q;         // queries per chunk
    int chunks;    // ceil(Lq / Q)
    int win_rows;  // forward: capacity of the shared-memory value window in rows (0: never stage)
    int gridK;     // lean kernels: 0 = a tile is Q consecutive queries; K > 0 = level-grid tiling into K time spans
    int hist_cap;  // lean backward: rows the sort histogram holds (a tile touching more is swept unsorted)
};

// ---- small PTX wrappers -------------------------------------------------------------------

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// global -> shared bulk copy (TMA engine, SASS UBLKCP); bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all()
{
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// 16-byte loads that are not forced through the read-only path (shared memory or global)
template <typename T> struct VecPlain;
template <> struct VecPlain<float> {
    __device__ __forceinline__ static void load(const void *p, float (&v)[4])
    {
        const float4 r = *reinterpret_cast<const float4 *>(p);
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
};
template <> struct VecPlain<__nv_bfloat16> {
    __device__ __forceinline__ static void load(const void *p, float (&v)[8])
    {
        const uint4 r = *reinterpret_cast<const uint4 *>(p);
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
};

// acc[k] += c * g[k], two channels per FFMA2 (packed fp32 FMA, sm_100+)
template <int CH> __device__ __forceinline__ void axpy(float (&acc)[CH], float c, const float (&g)[CH])
{
    const float2 cc = make_float2(c, c);
#pragma unroll
    for (int k = 0; k < CH; k += 2) {
        const float2 r = __ffma2_rn(cc, make_float2(g[k], g[k + 1]), make_float2(acc[k], acc[k + 1]));
        acc[k] = r.x;
        acc[k + 1] = r.y;
    }
}

template <int CH> __device__ __forceinline__ float dot(const float (&a)[CH], const float (&b)[CH])
{
    float2 s = make_float2(0.f, 0.f);
#pragma unroll
    for (int k = 0; k < CH; k += 2) s = __ffma2_rn(make_float2(a[k], a[k + 1]), make_float2(b[k], b[k + 1]), s);
    return s.x + s.y;
}

// Sum NV per-lane values over the W lanes of a lane group.  Transposing butterfly while more than
// one value is left (each step halves the values a lane carries), plain butterfly afterwards.
// On return lane j of the group holds, in a[0 .. max(NV/W,1)), the totals of the values
//   NV >= W: j*(NV/W) + i        NV < W: j / (W/NV)
template <int W, int NV> __device__ __forceinline__ void reduce_nv(float (&a)[NV], int j)
{
    int n = NV;
#pragma unroll
    for (int off = W / 2; off >= 1; off >>= 1) {
        if (n > 1) {
            n >>= 1;
            const bool upper = (j & off) != 0;
#pragma unroll
            for (int i = 0; i < NV / 2; ++i) {
                if (i < n) {
                    const float send = upper ? a[i] : a[i + n];
                    const float keep = upper ? a[i + n] : a[i];
                    a[i] = keep + __shfl_xor_sync(kFull, send, off);
                }
            }
        } else {
            a[0] += __shfl_xor_sync(kFull, a[0], off);
        }
    }
}

// ---- phase A: one sample -> rows, weights, slopes -----------------------------------------------

struct SampleGeom {
    int r0, r1;      // slice rows (level start + clamped tap index)
    float w0, w1;    // interpolation weights (0 for a tap that does not exist)
    float s0, s1;    // d(sample)/d(pix) selectors: -gate, +gate (0 for a tap that does not exist)
    float A;         // attention weight
    float tf;        // T_l as float
    float dxdoff;    // fused: d(loc)/d(offset)
    int level;
};

template <bool FUSED>
__device__ __forceinline__ SampleGeom sample_geom(const TileArgs &a, const LevelTable &lt, long long u, long long nq,
                                                  int s, float sm_max, float sm_inv)
{
    SampleGeom g;
    const int l = __float2int_rz(((float)s + 0.5f) * a.invP);
    const int T = lt.W[l];
    const long long sidx = u * a.LP + s;
    float hi, lo;
    g.tf = (float)T;
    if (FUSED) {
        const float off = __ldg(a.offs + sidx);
        const float logit = __ldg(a.logits + sidx);
        g.A = __expf(logit - sm_max) * sm_inv;
        const float *rp = a.ref + (nq * a.L + l) * a.R;
        const float r0 = __ldg(rp);
        const float p = r0 * g.tf;
        const float plo = fmaf(r0, g.tf, -p);
        float add;
        if (a.R == 1) {   // x = ref0 + off / T_l          (ms_deform_attn.py:103-106)
            add = off - 0.5f;
            g.dxdoff = 1.f / g.tf;
        } else {          // x = ref0 + off / P * ref1 * 0.5 (ms_deform_attn.py:107-109)
            const float r1 = __ldg(rp + 1);
            g.dxdoff = r1 * (0.5f * a.invP);
            add = fmaf(off * g.dxdoff, g.tf, -0.5f);
        }
        float e;
        two_sum(p, add, hi, e);
        lo = plo + e;
    } else {
        const float x = __ldg(a.loc + 2 * sidx);
        g.A = __ldg(a.attn + sidx);
        pix_from_loc(x, T, hi, lo);
        g.dxdoff = 0.f;
    }
    const Tap1D t = locate_1d(hi, lo, T, a.pad_mode);
    g.r0 = lt.start[l] + t.i0;
    g.r1 = lt.start[l] + t.i1;
    g.w0 = t.w0; g.w1 = t.w1; g.s0 = t.s0; g.s1 = t.s1;
    g.level = l;
    return g;
}

// softmax statistics of one unit's LP logits, computed by the unit's group of W lanes
template <int W>
__device__ __forceinline__ void tile_softmax_stats(const float *logits, bool active, int j, int LP, float &mx, float &inv)
{
    float m = -INFINITY;
    if (active)
        for (int s = j; s < LP; s += W) m = fmaxf(m, __ldg(logits + s));
    m = group_max<W>(m);
    float sum = 0.f;
    if (active)
        for (int s = j; s < LP; s += W) sum += __expf(__ldg(logits + s) - m);
    sum = group_sum<W>(sum);
    mx = active ? m : 0.f;
    inv = active ? 1.f / sum : 0.f;
}


// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
constexpr int kRecLevelShift = 26;                 // rec.x = row | level << 26 | dead << 31 until the fix-up pass
constexpr int kRecRowMask = (1 << kRecLevelShift) - 1;

// dynamic shared memory: [ records int4 x Q*LPpad | value window win_rows x D*sizeof(VT) ]
template <typename VT, int LPG, int LPA, bool FUSED>
__global__ void __launch_bounds__(kTileThreads) msda_fwd_tile_kernel(const TileArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    constexpr int G = 32 / LPG;     // units per warp in the gather phase
    constexpr int UPW = 32 / LPA;   // units per warp in the record phase
    extern __shared__ __align__(128) unsigned char dsm[];
    __shared__ LevelTable lt;
    __shared__ int win_lo[kMaxLevels], win_hi[kMaxLevels], win_base[kMaxLevels];
    __shared__ unsigned long long mbar;
    __shared__ int win_ok;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int LP = a.LP, LPpad = a.LPpad;
    const int chunk = blockIdx.x % a.chunks;
    const int nm = blockIdx.x / a.chunks;
    const int n = nm / a.M, m = nm - n * a.M;
    const int q0 = chunk * a.Q;
    const int Qc = min(a.Q, a.Lq - q0);
    const int MD = a.M * a.D;
    int4 *recs = reinterpret_cast<int4 *>(dsm);
    unsigned char *win = dsm + (size_t)a.Q * LPpad * sizeof(int4);
    const int rb = a.D * (int)sizeof(VT);   // bytes per row of one head

    load_level_table(lt, a.shapes, a.lsi, a.L, a.shapes_stride);
    if (tid < a.L) { win_lo[tid] = INT_MAX; win_hi[tid] = -1; }
    if (tid == 0) mbar_init(&mbar, 1);
    __syncthreads();

    // ---- phase A: records ----
    {
        const int g = lane / LPA, j = lane % LPA;
        for (int ul0 = 0; ul0 < Qc; ul0 += kTileWarps * UPW) {
            const int ul = ul0 + warp * UPW + g;
            const bool active = ul < Qc;
            const long long nq = (long long)n * a.Lq + q0 + (active ? ul : 0);
            const long long u = nq * a.M + m;
            float sm_max = 0.f, sm_inv = 0.f;
            if (FUSED) tile_softmax_stats<LPA>(a.logits + u * LP, active, j, LP, sm_max, sm_inv);
            for (int sc = 0; sc < LPpad; sc += LPA) {
                const int s = sc + j;
                int4 rec = make_int4((int)0x80000000u, 0, 0, 0);   // dead record: weight 0, any staged row
                int lvl = -1, r0 = 0, r1 = 0;
                if (active && s < LP) {
                    const SampleGeom sg = sample_geom<FUSED>(a, lt, u, nq, s, sm_max, sm_inv);
                    const float c0 = sg.A * sg.w0, c1 = sg.A * sg.w1;
                    if (c0 != 0.f || c1 != 0.f) {
                        lvl = sg.level; r0 = sg.r0; r1 = sg.r1;
                        rec = make_int4(r0 | (lvl << kRecLevelShift), r1, __float_as_int(c0), __float_as_int(c1));
                    }
                }
                if (active && s < LPpad) recs[ul * LPpad + s] = rec;
                // rows this warp pass needs, per level
                for (int l = 0; l < a.L; ++l) {
                    const int mn = __reduce_min_sync(kFull, lvl == l ? r0 : INT_MAX);
                    const int mx = __reduce_max_sync(kFull, lvl == l ? r1 : -1);
                    if (lane == 0 && mx >= 0) { atomicMin(&win_lo[l], mn); atomicMax(&win_hi[l], mx); }
                }
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        int tot = 0;
        for (int l = 0; l < a.L; ++l) {
            win_base[l] = tot - win_lo[l];   // window row of slice row r at level l: win_base[l] + r
            if (win_hi[l] >= win_lo[l]) tot += win_hi[l] - win_lo[l] + 1;
        }
        const int ok = (tot > 0 && tot <= a.win_rows) ? 1 : 0;
        win_ok = ok;
        if (ok) mbar_expect_tx(&mbar, (unsigned)tot * (unsigned)rb);
    }
    __syncthreads();
    const bool staged = win_ok != 0;
    const VT *vslice = static_cast<const VT *>(a.value) + (long long)n * a.S * MD + m * a.D;
    if (staged) {
        for (int l = 0; l < a.L; ++l)
            for (int r = win_lo[l] + tid; r <= win_hi[l]; r += kTileThreads)
                bulk_g2s(win + (size_t)(win_base[l] + r) * rb, vslice + (long long)r * MD, (unsigned)rb, &mbar);
    }
    // fix-up: rows -> byte offsets into the window (staged) or element offsets into the slice (not staged)
    for (int i = tid; i < Qc * LPpad; i += kTileThreads) {
        int4 rec = recs[i];
        if (rec.x < 0) {
            rec.x = rec.y = 0;
        } else {
            const int l = rec.x >> kRecLevelShift;
            const int r0 = rec.x & kRecRowMask;
            rec.x = staged ? (win_base[l] + r0) * rb : r0 * MD;
            rec.y = staged ? (win_base[l] + rec.y) * rb : rec.y * MD;
        }
        recs[i] = rec;
    }
    __syncthreads();
    if (staged) mbar_wait(&mbar, 0);

    // ---- gather ----
    {
        const int g = lane / LPG, j = lane % LPG;
        VT *out = static_cast<VT *>(a.out);
        const unsigned char *wlane = win + j * 16;
        const VT *vlane = vslice + j * CH;
        for (int ul0 = warp * G; ul0 < Qc; ul0 += kTileWarps * G) {
            const int ul = ul0 + g;
            const bool active = ul < Qc;
            const int4 *rp = recs + (active ? ul : 0) * LPpad;
            float acc[CH];
#pragma unroll
            for (int k = 0; k < CH; ++k) acc[k] = 0.f;
            for (int s = 0; s < LPpad; s += 4) {
                int4 r[4];
#pragma unroll
                for (int b = 0; b < 4; ++b) r[b] = rp[s + b];
                float v0[4][CH], v1[4][CH];
                if (staged) {
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        VecPlain<VT>::load(wlane + r[b].x, v0[b]);
                        VecPlain<VT>::load(wlane + r[b].y, v1[b]);
                    }
                } else {
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        Vec16<VT>::load(vlane + r[b].x, v0[b]);
                        Vec16<VT>::load(vlane + r[b].y, v1[b]);
                    }
                }
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    axpy<CH>(acc, __int_as_float(r[b].z), v0[b]);
                    axpy<CH>(acc, __int_as_float(r[b].w), v1[b]);
                }
            }
            if (active) {
                const long long nq = (long long)n * a.Lq + q0 + ul;
                Vec16<VT>::store(out + (nq * a.M + m) * a.D + j * CH, acc);
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------------
// dynamic shared memory (QL = Q * LPpad):
//   Gs    Q x D VT        grad_out rows of the chunk (this head)
//   sp    QL float4       {wa, wb, sa*A*T, sb*A*T}: weights / location slopes on the dots with rows r, r+1
//   ent   QL int4         samples sorted by first row: {idx | unit << 16, A*wa, A*wb, -}
//   dots  QL float2       {G.V[r], G.V[r+1]} per sample (first used as {row, rank | unit << 16})
//   Aarr  QL float        attention weight
//   start (S+1) int       histogram of first rows -> exclusive prefix
//   rows  QL int          compacted list of rows that own at least one sample
struct BwdSmem {
    unsigned char *Gs;
    float4 *sp;
    int4 *ent;
    float2 *dots;
    float *Aarr;
    int *start;
    int *rows;
};

__host__ __device__ inline size_t bwd_tile_smem_bytes(int Q, int LPpad, int S, int rb)
{
    const size_t QL = (size_t)Q * LPpad;
    return (size_t)Q * rb + QL * (16 + 16 + 8 + 4 + 4) + ((size_t)S + 1 + 3) / 4 * 16;
}

__device__ __forceinline__ BwdSmem bwd_carve(unsigned char *base, int Q, int LPpad, int S, int rb)
{
    const size_t QL = (size_t)Q * LPpad;
    BwdSmem b;
    b.Gs = base; base += (size_t)Q * rb;
    b.sp = reinterpret_cast<float4 *>(base); base += QL * 16;
    b.ent = reinterpret_cast<int4 *>(base); base += QL * 16;
    b.dots = reinterpret_cast<float2 *>(base); base += QL * 8;
    b.Aarr = reinterpret_cast<float *>(base); base += QL * 4;
    b.rows = reinterpret_cast<int *>(base); base += QL * 4;
    b.start = reinterpret_cast<int *>(base);
    return b;
}

template <typename VT, int LPG, int LPA, bool FUSED>
__global__ void __launch_bounds__(kTileThreads) msda_bwd_tile_kernel(const TileArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    constexpr int G = 32 / LPG;            // row groups per warp
    constexpr int NG = kTileThreads / LPG; // row groups per CTA
    constexpr int UPW = 32 / LPA;
    constexpr int BS = LPG == 32 ? 8 : 4;  // samples per reduction batch
    constexpr int NV = 2 * BS;             // dot products per batch
    constexpr int VPL = NV >= LPG ? NV / LPG : 1;
    extern __shared__ __align__(128) unsigned char dsm[];
    __shared__ LevelTable lt;
    __shared__ int wsum[kTileWarps];
    __shared__ int n_active_s;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int LP = a.LP, LPpad = a.LPpad;
    const int chunk = blockIdx.x % a.chunks;
    const int nm = blockIdx.x / a.chunks;
    const int n = nm / a.M, m = nm - n * a.M;
    const int q0 = chunk * a.Q;
    const int Qc = min(a.Q, a.Lq - q0);
    const int MD = a.M * a.D;
    const int rb = a.D * (int)sizeof(VT);
    const BwdSmem sm = bwd_carve(dsm, a.Q, LPpad, a.S, rb);

    load_level_table(lt, a.shapes, a.lsi, a.L, a.shapes_stride);
    for (int i = tid; i <= a.S; i += kTileThreads) sm.start[i] = 0;
    {   // grad_out rows of the chunk -> shared memory (asynchronously; needed in phase B)
        const unsigned char *gsrc = static_cast<const unsigned char *>(a.grad_out) +
                                    (((long long)n * a.Lq + q0) * a.M + m) * (long long)rb;
        for (int v = tid; v < Qc * LPG; v += kTileThreads) {
            const int ul = v / LPG, jj = v % LPG;
            cp_async16(sm.Gs + ul * rb + jj * 16, gsrc + (long long)ul * a.M * rb + jj * 16);
        }
    }
    __syncthreads();

    // ---- phase A: decode samples, histogram of first rows ----
    {
        const int g = lane / LPA, j = lane % LPA;
        for (int ul0 = 0; ul0 < Qc; ul0 += kTileWarps * UPW) {
            const int ul = ul0 + warp * UPW + g;
            const bool active = ul < Qc;
            const long long nq = (long long)n * a.Lq + q0 + (active ? ul : 0);
            const long long u = nq * a.M + m;
            float sm_max = 0.f, sm_inv = 0.f;
            if (FUSED) tile_softmax_stats<LPA>(a.logits + u * LP, active, j, LP, sm_max, sm_inv);
            for (int sc = 0; sc < LPpad; sc += LPA) {
                const int s = sc + j;
                if (active && s < LPpad) {
                    const int idx = ul * LPpad + s;
                    float4 spv = make_float4(0.f, 0.f, 0.f, 0.f);
                    float A = 0.f;
                    int2 tmp = make_int2(-1, 0);
                    if (s < LP) {
                        const SampleGeom sg = sample_geom<FUSED>(a, lt, u, nq, s, sm_max, sm_inv);
                        float wa = sg.w0, wb = sg.w1, sa = sg.s0, sb = sg.s1;
                        if (sg.r1 == sg.r0) { wa += wb; sa += sb; wb = 0.f; sb = 0.f; }   // one row only
                        const float at = sg.A * sg.tf;
                        A = sg.A;
                        spv = make_float4(wa, wb, sa * at, sb * at);
                        if (wa != 0.f || wb != 0.f || sa != 0.f || sb != 0.f) {
                            const int rank = atomicAdd(&sm.start[sg.r0], 1);
                            tmp = make_int2(sg.r0, rank | (ul << 16));
                        }
                    }
                    sm.sp[idx] = spv;
                    sm.Aarr[idx] = A;
                    *reinterpret_cast<int2 *>(&sm.dots[idx]) = tmp;
                }
            }
        }
    }
    __syncthreads();

    // ---- exclusive prefix over the histogram + list of occupied rows ----
    {
        const int B = (a.S + kTileThreads - 1) / kTileThreads;
        const int i0 = min(tid * B, a.S), i1 = min(i0 + B, a.S);
        int packed = 0;   // count in the low 16 bits, occupied rows in the high 16 (both <= 2048)
        for (int i = i0; i < i1; ++i) {
            const int c = sm.start[i];
            packed += c + (c > 0 ? 65536 : 0);
        }
        int incl = packed;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(kFull, incl, off);
            if (lane >= off) incl += t;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int prev = 0;
        for (int w = 0; w < warp; ++w) prev += wsum[w];
        int run = prev + incl - packed;
        int pos = run & 0xffff, rpos = run >> 16;
        for (int i = i0; i < i1; ++i) {
            const int c = sm.start[i];
            sm.start[i] = pos;
            if (c > 0) sm.rows[rpos++] = i;
            pos += c;
        }
        if (tid == kTileThreads - 1) {
            const int tot = prev + incl;
            sm.start[a.S] = tot & 0xffff;
            n_active_s = tot >> 16;
        }
    }
    __syncthreads();

    // ---- scatter samples into row order ----
    for (int idx = tid; idx < Qc * LPpad; idx += kTileThreads) {
        const int2 t = *reinterpret_cast<const int2 *>(&sm.dots[idx]);
        sm.dots[idx] = make_float2(0.f, 0.f);
        if (t.x >= 0) {
            const float4 spv = sm.sp[idx];
            const float A = sm.Aarr[idx];
            sm.ent[sm.start[t.x] + (t.y & 0xffff)] =
                make_int4(idx | (t.y & 0xffff0000), __float_as_int(A * spv.x), __float_as_int(A * spv.y), 0);
        }
    }
    cp_async_commit_wait_all();
    __syncthreads();

    // ---- phase B: one lane group per occupied row ----
    {
        const int n_active = n_active_s;
        const int gid = tid / LPG, j = lane % LPG;
        const VT *vslice = static_cast<const VT *>(a.value) + (long long)n * a.S * MD + m * a.D + j * CH;
        float *gslice = a.gv_acc + (long long)n * a.S * MD + m * a.D + j * CH;
        const unsigned char *glane = sm.Gs + j * 16;
        float *dots_f = reinterpret_cast<float *>(sm.dots);
        for (int ri0 = 0; ri0 < n_active; ri0 += NG) {
            const int ri = ri0 + gid;
            const bool valid = ri < n_active;
            const int r = valid ? sm.rows[ri] : 0;
            const int beg = valid ? sm.start[r] : 0;
            const int end = valid ? sm.start[r + 1] : 0;
            bool has1 = false;
            for (int l = 0; l < a.L; ++l) {
                const int e = lt.start[l] + lt.W[l];
                if (r >= lt.start[l] && r < e) has1 = r + 1 < e;
            }
            has1 = has1 && valid;
            float v0[CH], v1[CH], acc0[CH], acc1[CH];
#pragma unroll
            for (int k = 0; k < CH; ++k) v0[k] = v1[k] = acc0[k] = acc1[k] = 0.f;
            if (valid) Vec16<VT>::load(vslice + (long long)r * MD, v0);
            if (has1) Vec16<VT>::load(vslice + (long long)(r + 1) * MD, v1);
            const int maxcnt = __reduce_max_sync(kFull, end - beg);
            for (int t = 0; t < maxcnt; t += BS) {
                float p[NV];
#pragma unroll
                for (int k = 0; k < BS; ++k) {
                    const int tk = beg + t + k;
                    const bool in = tk < end;
                    int4 e = sm.ent[in ? tk : 0];
                    if (!in) e.y = e.z = 0;
                    float gq[CH];
                    VecPlain<VT>::load(glane + (e.x >> 16) * rb, gq);
                    axpy<CH>(acc0, __int_as_float(e.y), gq);
                    axpy<CH>(acc1, __int_as_float(e.z), gq);
                    p[2 * k] = dot<CH>(gq, v0);
                    p[2 * k + 1] = dot<CH>(gq, v1);
                }
                reduce_nv<LPG, NV>(p, j);
#pragma unroll
                for (int i = 0; i < VPL; ++i) {
                    const int vi = NV >= LPG ? j * VPL + i : j / (LPG / (NV < LPG ? NV : LPG));
                    const bool writer = NV >= LPG ? true : (j % (LPG / (NV < LPG ? NV : LPG))) == 0;
                    const int tk = beg + t + (vi >> 1);
                    if (writer && tk < end) dots_f[2 * (sm.ent[tk].x & 0xffff) + (vi & 1)] = p[i];
                }
            }
            if (valid) {
#pragma unroll
                for (int k = 0; k < CH; k += 4)
                    red_add_v4(gslice + (long long)r * MD + k, acc0[k], acc0[k + 1], acc0[k + 2], acc0[k + 3]);
            }
            if (has1) {
#pragma unroll
                for (int k = 0; k < CH; k += 4)
                    red_add_v4(gslice + (long long)(r + 1) * MD + k, acc1[k], acc1[k + 1], acc1[k + 2], acc1[k + 3]);
            }
        }
    }
    __syncthreads();

    // ---- phase C: per-sample gradients ----
    {
        const int g = lane / LPA, j = lane % LPA;
        const bool p_pow2 = (a.P & (a.P - 1)) == 0 && a.P <= LPA;
        for (int ul0 = 0; ul0 < Qc; ul0 += kTileWarps * UPW) {
            const int ul = ul0 + warp * UPW + g;
            const bool active = ul < Qc;
            const long long nq = (long long)n * a.Lq + q0 + (active ? ul : 0);
            const long long u = nq * a.M + m;
            float dotsum = 0.f;
            for (int sc = 0; sc < LP; sc += LPA) {
                const int s = sc + j;
                const bool on = active && s < LP;
                const int idx = (active ? ul : 0) * LPpad + (s < LP ? s : 0);
                const float2 pd = sm.dots[idx];
                const float4 spv = sm.sp[idx];
                const float A = sm.Aarr[idx];
                const float ga = spv.x * pd.x + spv.y * pd.y;
                const float gx = spv.z * pd.x + spv.w * pd.y;
                const long long sidx = u * LP + s;
                if (!FUSED) {
                    if (on) {
                        a.grad_attn[sidx] = ga;
                        // dummy y (= 0.5 on an H = 1 level): BORDER clamps it, gradient exactly 0; ZEROS keeps the
                        // row-0 weight 1 - |y*H - 0.5| whose slope the reference kernel reports (cuh:159)
                        const float gy = a.pad_mode == MSDA_PAD_ZEROS ? -A * ga : 0.f;
                        *reinterpret_cast<float2 *>(a.grad_loc + 2 * sidx) = make_float2(gx, gy);
                    }
                } else {
                    const int l = __float2int_rz(((float)(s < LP ? s : 0) + 0.5f) * a.invP);
                    float dxdoff = 0.f, r1w = 0.f;
                    if (on) {
                        if (a.R == 1) {
                            dxdoff = 1.f / (float)lt.W[l];
                        } else {
                            dxdoff = __ldg(a.ref + (nq * a.L + l) * a.R + 1) * (0.5f * a.invP);
                            r1w = __ldg(a.offs + sidx) * (0.5f * a.invP);   // d(loc)/d(ref1) = off * 0.5 / P
                        }
                        a.grad_offs[sidx] = gx * dxdoff;
                        dotsum = fmaf(A, ga, dotsum);
                    }
                    if (a.grad_ref != nullptr) {
                        float r0g = on ? gx : 0.f;           // d(loc)/d(ref0) = 1
                        float r1g = on ? gx * r1w : 0.f;
                        if (p_pow2) {   // the P lanes of one level are contiguous and P-aligned
                            for (int off = a.P >> 1; off >= 1; off >>= 1) {
                                r0g += __shfl_xor_sync(kFull, r0g, off);
                                r1g += __shfl_xor_sync(kFull, r1g, off);
                            }
                            if (on && (s & (a.P - 1)) == 0) {
                                float *gp = a.grad_ref + (nq * a.L + l) * a.R;
                                atomicAdd(gp, r0g);
                                if (a.R == 2) atomicAdd(gp + 1, r1g);
                            }
                        } else if (on) {
                            float *gp = a.grad_ref + (nq * a.L + l) * a.R;
                            atomicAdd(gp, r0g);
                            if (a.R == 2) atomicAdd(gp + 1, r1g);
                        }
                    }
                }
            }
            if (FUSED) {
                // softmax Jacobian: g_logit_s = A_s * (ga_s - sum_t A_t ga_t)
                dotsum = group_sum<LPA>(dotsum);
                if (active)
                    for (int s = j; s < LP; s += LPA) {
                        const int idx = ul * LPpad + s;
                        const float2 pd = sm.dots[idx];
                        const float4 spv = sm.sp[idx];
                        const float ga = spv.x * pd.x + spv.y * pd.y;
                        a.grad_logits[u * LP + s] = sm.Aarr[idx] * (ga - dotsum);
                    }
            }
        }
    }
}

}  // namespace msda
