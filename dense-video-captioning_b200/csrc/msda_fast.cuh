// msda_fast.cuh -- the 1-D temporal MSDeformAttn hot path for sm_100a.
//
// Replaces, for H_l == 1 (the only way PDVC calls the op, pdvc/ops/modules/ms_deform_attn.py:114-117):
//   forward : ms_deformable_im2col_gpu_kernel            pdvc/ops/src/cuda/ms_deform_im2col_cuda.cuh:238-300
//   backward: ms_deformable_col2im_gpu_kernel_shm_*      cuh:302-921 (D=64 -> blocksize_aware_reduce_v2, :407-511)
// and, with FUSED = true, the softmax / sampling-location arithmetic in front of them
//   (pdvc/ops/modules/ms_deform_attn.py:100-116).
//
// Work decomposition (differs from the reference's thread-per-output-channel design):
//   unit   = one (n, q, m): all L*P samples of one head of one query.
//   group  = LPG lanes (power of two) that own one unit; every lane owns 16 contiguous bytes of
//            the head's D channels (4 x f32 or 8 x bf16), so a group reads a value row with one
//            fully coalesced 16-byte-per-lane request.  A warp carries 32/LPG units.
//   sample parameters (row offsets, attention * interpolation weights) are computed ONCE per
//   sample by one lane (lane j <-> sample j of the current chunk of LPG samples), parked in a
//   512-byte per-warp shared-memory table and re-read as a broadcast -- the reference recomputes
//   them in every one of the D channel threads (cuh:275-290).
//   backward: the per-sample dot products over D (grad_attn, grad_loc) are reduced with a
//   transposing butterfly (2*LPG-2 shuffles per chunk, common.cuh) instead of the reference's
//   per-sample shared-memory tree with __syncthreads (cuh:484-494); grad_value rows are added with
//   one 16-byte vector reduction per lane (REDG.E.ADD.F32x4) instead of scalar atomicAdd (cuh:126-153).
//   CTAs own contiguous unit ranges, so neighbouring queries (which sample neighbouring rows in the
//   encoder) hit the same L1 lines.
#pragma once
#include "msda_common.cuh"

namespace msda {

struct FastArgs {
    const void *value;
    const int64_t *shapes;
    const int64_t *lsi;
    int shapes_stride;
    // unfused front-end
    const float *loc;    // (N,Lq,M,L,P,2)
    const float *attn;   // (N,Lq,M,L,P)
    // fused front-end
    const float *ref;     // (N,Lq,L,R)
    const float *offs;    // (N,Lq,M,L,P)
    const float *logits;  // (N,Lq,M,L,P)
    int R;
    // forward output
    void *out;  // (N,Lq,M*D)
    // backward
    const void *grad_out;  // (N,Lq,M*D)
    float *gv_acc;         // fp32 accumulation image of grad_value (N,S,M,D), pre-zeroed
    float *grad_loc;       // unfused (N,Lq,M,L,P,2)
    float *grad_attn;      // unfused (N,Lq,M,L,P)
    float *grad_ref;       // fused (N,Lq,L,R), pre-zeroed, may be NULL
    float *grad_offs;      // fused
    float *grad_logits;    // fused
    int N, S, M, D, L, Lq, P;
    int pad_mode;
    int total_units;     // N*Lq*M (< 2^31, checked by the launcher)
    int units_per_cta;
    int head_major;      // iterate units as (n, m, q) instead of storage order (n, q, m)
};

constexpr int kFastThreads = 128;
constexpr int kFastWarps = kFastThreads / 32;

__device__ __forceinline__ void two_sum(float a, float b, float &s, float &e)
{
    s = a + b;
    const float bb = s - a;
    e = (a - (s - bb)) + (b - bb);
}

// Everything one lane needs to know about "its" sample of the current chunk.
struct OwnSample {
    float w0, w1;   // interpolation weights (0 for missing taps)
    float s0, s1;   // d(out)/d(loc) coefficients on dot0 / dot1, incl. attention weight and T_l
    float A;        // attention weight
    float dxdoff;   // fused: d(loc)/d(offset)
    int level;
};

// Computes the parameter record of sample `s` of unit `u` and the owner-lane extras.
template <bool FUSED>
__device__ __forceinline__ int4 make_record(const FastArgs &a, const LevelTable &lt, int u, int nq,
                                            int s, int LP, float sm_max, float sm_inv, OwnSample &own)
{
    const int l = s / a.P;
    const int T = lt.W[l];
    const long long sidx = (long long)u * LP + s;
    float hi, lo, A;
    if (FUSED) {
        const float off = __ldg(a.offs + sidx);
        const float logit = __ldg(a.logits + sidx);
        A = __expf(logit - sm_max) * sm_inv;
        const float *rp = a.ref + ((long long)nq * a.L + l) * a.R;
        const float r0 = __ldg(rp);
        const float tf = (float)T;
        const float p = r0 * tf;
        float plo = fmaf(r0, tf, -p);
        float add;
        if (a.R == 1) {   // x = ref0 + off / T_l          (ms_deform_attn.py:103-106)
            add = off - 0.5f;
            own.dxdoff = 1.f / tf;
        } else {          // x = ref0 + off / P * ref1 * 0.5 (ms_deform_attn.py:107-109)
            const float r1 = __ldg(rp + 1);
            own.dxdoff = r1 * (0.5f / (float)a.P);
            add = fmaf(off * own.dxdoff, tf, -0.5f);
        }
        float e;
        two_sum(p, add, hi, e);
        lo = plo + e;
    } else {
        const float x = __ldg(a.loc + 2 * sidx);
        A = __ldg(a.attn + sidx);
        pix_from_loc(x, T, hi, lo);
        own.dxdoff = 0.f;
    }
    const Tap1D t = locate_1d(hi, lo, T, a.pad_mode);
    const int MD = a.M * a.D;
    own.w0 = t.w0;
    own.w1 = t.w1;
    const float at = A * (float)T;
    own.s0 = t.s0 * at;
    own.s1 = t.s1 * at;
    own.A = A;
    own.level = l;
    int4 rec;
    rec.x = (lt.start[l] + t.i0) * MD;
    rec.y = (lt.start[l] + t.i1) * MD;
    rec.z = __float_as_int(A * t.w0);
    rec.w = __float_as_int(A * t.w1);
    return rec;
}

// softmax statistics of one unit's L*P logits, computed by the unit's group
template <int LPG>
__device__ __forceinline__ void softmax_stats(const float *logits, bool active, int j, int LP, float &mx, float &inv)
{
    float m = -INFINITY;
    if (active)
        for (int s = j; s < LP; s += LPG) m = fmaxf(m, __ldg(logits + s));
    m = group_max<LPG>(m);
    float sum = 0.f;
    if (active)
        for (int s = j; s < LP; s += LPG) sum += __expf(__ldg(logits + s) - m);
    sum = group_sum<LPG>(sum);
    mx = active ? m : 0.f;
    inv = active ? 1.f / sum : 0.f;
}

constexpr int kBatch = 4;   // samples gathered per batch (LPG is a multiple of it)

template <typename VT, int G>
__device__ __forceinline__ void fwd_batch(const int4 *prm_w, int t0, int g, const VT *vbase,
                                          float (&acc)[Vec16<VT>::CH])
{
    constexpr int CH = Vec16<VT>::CH;
    int4 r[kBatch];
#pragma unroll
    for (int b = 0; b < kBatch; ++b) r[b] = prm_w[(t0 + b) * G + g];
    float v0[kBatch][CH], v1[kBatch][CH];
#pragma unroll
    for (int b = 0; b < kBatch; ++b) {
        Vec16<VT>::load(vbase + r[b].x, v0[b]);
        Vec16<VT>::load(vbase + r[b].y, v1[b]);
    }
#pragma unroll
    for (int b = 0; b < kBatch; ++b) {
        const float c0 = __int_as_float(r[b].z), c1 = __int_as_float(r[b].w);
#pragma unroll
        for (int k = 0; k < CH; ++k) acc[k] = fmaf(c1, v1[b][k], fmaf(c0, v0[b][k], acc[k]));
    }
}

// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
template <typename VT, int LPG, bool FUSED>
__global__ void __launch_bounds__(kFastThreads) msda_fwd_fast_kernel(const FastArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    constexpr int G = 32 / LPG;
    __shared__ LevelTable lt;
    __shared__ int4 prm[kFastWarps][32];
    load_level_table(lt, a.shapes, a.lsi, a.L, a.shapes_stride);
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane / LPG, j = lane % LPG;
    const int LP = a.L * a.P;
    const int MD = a.M * a.D;
    const int cta_begin = blockIdx.x * a.units_per_cta;   // grid * units_per_cta < 2^31 (launcher)
    const int cta_end = min(cta_begin + a.units_per_cta, a.total_units);
    const VT *value = static_cast<const VT *>(a.value);
    VT *out = static_cast<VT *>(a.out);
    const int d_step = LPG * CH;

    for (int ubase = cta_begin + warp * G; ubase < cta_end; ubase += kFastWarps * G) {
        const int wi = ubase + g;
        const bool active = wi < cta_end;
        int u, nq, m;
        if (a.head_major) {   // work order (n, m, q): neighbouring queries of ONE head share value rows
            const int nm = active ? wi / a.Lq : 0;
            const int q = active ? wi - nm * a.Lq : 0;
            const int nn = nm / a.M;
            m = nm - nn * a.M;
            nq = nn * a.Lq + q;
            u = nq * a.M + m;
        } else {
            u = wi;
            nq = active ? u / a.M : 0;
            m = active ? u - nq * a.M : 0;
        }
        const long long n = nq / a.Lq;
        float sm_max = 0.f, sm_inv = 0.f;
        if (FUSED) softmax_stats<LPG>(a.logits + (long long)u * LP, active, j, LP, sm_max, sm_inv);

        for (int d0 = 0; d0 < a.D; d0 += d_step) {
            const int ch0 = d0 + j * CH;
            const bool lane_on = active && ch0 < a.D;
            // lanes without work still issue the (branch-free, batched) loads: they read channel 0
            // of valid rows with whatever coefficients the group has and never store.
            const VT *vbase = value + n * (long long)a.S * MD + m * a.D + (lane_on ? ch0 : 0);
            float acc[CH];
#pragma unroll
            for (int k = 0; k < CH; ++k) acc[k] = 0.f;

            for (int sc = 0; sc < LP; sc += LPG) {
                const int s = sc + j;
                int4 rec = make_int4(0, 0, 0, 0);   // offset 0, weight 0: a harmless tap
                if (active && s < LP) {
                    OwnSample own;
                    rec = make_record<FUSED>(a, lt, u, nq, s, LP, sm_max, sm_inv, own);
                }
                __syncwarp();
                prm[warp][j * G + g] = rec;
                __syncwarp();
                // samples are consumed in batches of kBatch so that 2*kBatch 16-byte gathers are in
                // flight per lane; the tail of a short chunk is padded by the zero records above
                const int cnt = min(LPG, LP - sc);
                const int nb = (cnt + kBatch - 1) / kBatch;
                if (nb == LPG / kBatch) {
#pragma unroll
                    for (int b = 0; b < LPG / kBatch; ++b) fwd_batch<VT, G>(prm[warp], b * kBatch, g, vbase, acc);
                } else {
                    for (int b = 0; b < nb; ++b) fwd_batch<VT, G>(prm[warp], b * kBatch, g, vbase, acc);
                }
            }
            if (lane_on) Vec16<VT>::store(out + (long long)u * a.D + ch0, acc);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------------
template <int CH> __device__ __forceinline__ void red_row(float *p, float c, const float (&gvec)[CH])
{
#pragma unroll
    for (int k = 0; k < CH; k += 4) red_add_v4(p + k, c * gvec[k], c * gvec[k + 1], c * gvec[k + 2], c * gvec[k + 3]);
}

template <typename VT, int LPG, bool FUSED>
__global__ void __launch_bounds__(kFastThreads) msda_bwd_fast_kernel(const FastArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    constexpr int G = 32 / LPG;
    __shared__ LevelTable lt;
    __shared__ int4 prm[kFastWarps][32];
    load_level_table(lt, a.shapes, a.lsi, a.L, a.shapes_stride);
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane / LPG, j = lane % LPG;
    const int LP = a.L * a.P;
    const int MD = a.M * a.D;
    const int cta_begin = blockIdx.x * a.units_per_cta;   // grid * units_per_cta < 2^31 (launcher)
    const int cta_end = min(cta_begin + a.units_per_cta, a.total_units);
    const VT *value = static_cast<const VT *>(a.value);
    const VT *grad_out = static_cast<const VT *>(a.grad_out);
    const int d_step = LPG * CH;
    const bool p_pow2 = (a.P & (a.P - 1)) == 0 && a.P <= LPG;

    for (int ubase = cta_begin + warp * G; ubase < cta_end; ubase += kFastWarps * G) {
        const int wi = ubase + g;
        const bool active = wi < cta_end;
        int u, nq, m;
        if (a.head_major) {   // work order (n, m, q): neighbouring queries of ONE head share value rows
            const int nm = active ? wi / a.Lq : 0;
            const int q = active ? wi - nm * a.Lq : 0;
            const int nn = nm / a.M;
            m = nm - nn * a.M;
            nq = nn * a.Lq + q;
            u = nq * a.M + m;
        } else {
            u = wi;
            nq = active ? u / a.M : 0;
            m = active ? u - nq * a.M : 0;
        }
        const long long n = nq / a.Lq;
        float sm_max = 0.f, sm_inv = 0.f;
        if (FUSED) softmax_stats<LPG>(a.logits + (long long)u * LP, active, j, LP, sm_max, sm_inv);
        float dotsum = 0.f;   // fused: sum_s A_s * grad_attn_s

        for (int sc = 0; sc < LP; sc += LPG) {
            const int s = sc + j;
            const bool own_on = active && s < LP;
            int4 rec = make_int4(0, 0, 0, 0);
            OwnSample own;
            own.w0 = own.w1 = own.s0 = own.s1 = own.A = own.dxdoff = 0.f;
            own.level = 0;
            if (own_on) rec = make_record<FUSED>(a, lt, u, nq, s, LP, sm_max, sm_inv, own);
            __syncwarp();
            prm[warp][j * G + g] = rec;
            __syncwarp();
            const int cnt = min(LPG, LP - sc);

            float part[2 * LPG];
#pragma unroll
            for (int i = 0; i < 2 * LPG; ++i) part[i] = 0.f;

            for (int d0 = 0; d0 < a.D; d0 += d_step) {
                const int ch0 = d0 + j * CH;
                const bool lane_on = active && ch0 < a.D;
                const long long voff = n * (long long)a.S * MD + m * a.D + (lane_on ? ch0 : 0);
                const VT *vbase = value + voff;
                float *gvbase = a.gv_acc + voff;
                float gvec[CH];
#pragma unroll
                for (int k = 0; k < CH; ++k) gvec[k] = 0.f;
                if (lane_on) Vec16<VT>::load(grad_out + (long long)u * a.D + ch0, gvec);
#pragma unroll
                for (int t0 = 0; t0 < LPG; t0 += kBatch) {
                    if (t0 < cnt) {   // warp-uniform; short chunks are padded with zero records
                        int4 r[kBatch];
#pragma unroll
                        for (int b = 0; b < kBatch; ++b) r[b] = prm[warp][(t0 + b) * G + g];
                        float v0[kBatch][CH], v1[kBatch][CH];
#pragma unroll
                        for (int b = 0; b < kBatch; ++b) {
                            Vec16<VT>::load(vbase + r[b].x, v0[b]);
                            Vec16<VT>::load(vbase + r[b].y, v1[b]);
                        }
#pragma unroll
                        for (int b = 0; b < kBatch; ++b) {
                            float p0 = 0.f, p1 = 0.f;
#pragma unroll
                            for (int k = 0; k < CH; ++k) {
                                p0 = fmaf(gvec[k], v0[b][k], p0);
                                p1 = fmaf(gvec[k], v1[b][k], p1);
                            }
                            part[2 * (t0 + b)] += lane_on ? p0 : 0.f;
                            part[2 * (t0 + b) + 1] += lane_on ? p1 : 0.f;
                            const float c0 = lane_on ? __int_as_float(r[b].z) : 0.f;
                            const float c1 = lane_on ? __int_as_float(r[b].w) : 0.f;
                            if (c0 != 0.f) red_row<CH>(gvbase + r[b].x, c0, gvec);
                            if (c1 != 0.f) red_row<CH>(gvbase + r[b].y, c1, gvec);
                        }
                    }
                }
            }
            transpose_reduce<LPG>(part, j);   // part[0] = G.V[i0], part[1] = G.V[i1] of the lane's own sample
            const float ga = own.w0 * part[0] + own.w1 * part[1];
            const float gx = own.s0 * part[0] + own.s1 * part[1];
            const long long sidx = (long long)u * LP + s;
            if (!FUSED) {
                if (own_on) {
                    a.grad_attn[sidx] = ga;
                    // dummy y (= 0.5 on an H = 1 level): BORDER clamps it, gradient exactly 0; ZEROS keeps
                    // the row-0 weight 1 - |y*H - 0.5| whose slope the reference kernel reports (cuh:159)
                    const float gy = a.pad_mode == MSDA_PAD_ZEROS ? -own.A * ga : 0.f;
                    *reinterpret_cast<float2 *>(a.grad_loc + 2 * sidx) = make_float2(gx, gy);
                }
            } else {
                if (own_on) {
                    a.grad_offs[sidx] = gx * own.dxdoff;
                    a.grad_logits[sidx] = ga;   // provisional; turned into the softmax gradient below
                    dotsum = fmaf(own.A, ga, dotsum);
                }
                if (a.grad_ref != nullptr) {
                    // d(loc)/d(ref0) = 1 ; d(loc)/d(ref1) = off * 0.5 / P  (R == 2)
                    float r0g = own_on ? gx : 0.f;
                    float r1g = 0.f;
                    if (a.R == 2 && own_on) r1g = gx * __ldg(a.offs + sidx) * (0.5f / (float)a.P);
                    if (p_pow2) {   // lanes j .. j+P-1 of one level are contiguous and P-aligned
                        for (int off = a.P >> 1; off >= 1; off >>= 1) {
                            r0g += __shfl_xor_sync(kFull, r0g, off);
                            r1g += __shfl_xor_sync(kFull, r1g, off);
                        }
                        if (own_on && (s & (a.P - 1)) == 0) {
                            float *gp = a.grad_ref + ((long long)nq * a.L + own.level) * a.R;
                            atomicAdd(gp, r0g);
                            if (a.R == 2) atomicAdd(gp + 1, r1g);
                        }
                    } else if (own_on) {
                        float *gp = a.grad_ref + ((long long)nq * a.L + own.level) * a.R;
                        atomicAdd(gp, r0g);
                        if (a.R == 2) atomicAdd(gp + 1, r1g);
                    }
                }
            }
        }
        if (FUSED) {
            // softmax Jacobian: g_logit_s = A_s * (ga_s - sum_t A_t ga_t)
            dotsum = group_sum<LPG>(dotsum);
            if (active)
                for (int s = j; s < LP; s += LPG) {
                    const long long sidx = (long long)u * LP + s;
                    const float A = __expf(__ldg(a.logits + sidx) - sm_max) * sm_inv;
                    a.grad_logits[sidx] = A * (a.grad_logits[sidx] - dotsum);
                }
        }
    }
}

// fp32 accumulation image -> bf16 grad_value
__global__ void msda_cvt_f32_to_bf16_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst, long long n8)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        const float4 a = __ldcs(reinterpret_cast<const float4 *>(src) + 2 * i);
        const float4 b = __ldcs(reinterpret_cast<const float4 *>(src) + 2 * i + 1);
        const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        Vec16<__nv_bfloat16>::store(dst + 8 * i, v);
    }
}

}  // namespace msda
