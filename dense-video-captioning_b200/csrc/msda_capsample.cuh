// msda_capsample.cuh -- fused sampler for the caption head.
//
// Replaces, for the 1-D temporal case, what MSDeformAttnCap.forward does between its Linear
// projections and its return (pdvc/ops/modules/ms_deform_attn_for_caption.py:100-123):
//   x = ref0 + off / T_l                      (reference points with 1 coordinate, :104-106)
//   x = ref0 + off / P * ref1 * 0.5           (2 coordinates, :107-109)
//   dummy y = 0.5, H = 1                      (:114-118)
//   samples = ms_deform_attn_core_pytorch(..., return_value=True)     (:120-121; func.py:64-65)
// in ONE kernel each way, reading the raw `sampling_offsets` Linear output.  The reference runs ~10
// small torch kernels plus F.grid_sample per level here, once per generated word.
//
// Layout: samples are written as (N, Lq, M, L*P, D) -- D contiguous, so stores (and the loads of
// grad_samples) are coalesced 16-byte vectors; the Python module returns the reference's
// (N*M, D, Lq, L, P) shape as a permuted VIEW of that buffer, and the caption head's own
// permute-back (pdvc/CaptioningHead/LSTM_DSA.py:241-242) lands on contiguous memory again.
//
// One thread per (sample, 16-byte channel vector).  The shapes are tiny (6 events x 16 samples x
// 512 channels): the kernels are launch-latency sized, what matters is that there are two of them
// instead of ~25.
#pragma once
#include "msda_common.cuh"

namespace msda {

struct CapArgs {
    const void *value;        // (N,S,M,D) VT
    const int64_t *shapes;    // (L,) temporal lengths
    const int64_t *lsi;       // (L,)
    const float *ref;         // (N,Lq,L,R)
    const float *off;         // (N,Lq,M,L,P) raw offsets
    void *samples;            // (N,Lq,M,L*P,D) VT                     forward out
    const void *grad_samples; // same layout                            backward in
    float *gv_acc;            // (N,S,M,D) fp32 accumulation image, pre-zeroed
    float *grad_ref;          // (N,Lq,L,R) pre-zeroed, may be NULL
    float *grad_off;          // (N,Lq,M,L,P) pre-zeroed
    int N, S, M, D, L, Lq, P, R;
    int pad_mode;
    long long total;          // N*Lq*M*L*P * (D / CH)
};

struct CapSample {
    Tap1D tap;
    long long row0, row1;   // element offsets of the two value rows (channel 0 of head m)
    float tl;           // T_l
    float dxdoff;       // d(loc)/d(offset)
    float off;
    long long nq;
    int l;
};

__device__ __forceinline__ void two_sum_f(float a, float b, float &s, float &e)
{
    s = a + b;
    const float bb = s - a;
    e = (a - (s - bb)) + (b - bb);
}

// location of sample `sidx` = ((nq*M + m)*L + l)*P + p, same split-precision arithmetic as the fused
// attention kernels (msda_fast.cuh make_record)
__device__ __forceinline__ CapSample cap_locate(const CapArgs &a, const LevelTable &lt, long long sidx)
{
    CapSample s;
    const int LP = a.L * a.P;
    const int lp = (int)(sidx % LP);
    const long long u = sidx / LP;              // (nq, m)
    const int m = (int)(u % a.M);
    s.nq = u / a.M;
    s.l = lp / a.P;
    const long long n = s.nq / a.Lq;
    const int T = lt.W[s.l];
    const float tf = (float)T;
    s.tl = tf;
    s.off = __ldg(a.off + sidx);
    const float *rp = a.ref + (s.nq * a.L + s.l) * a.R;
    const float r0 = __ldg(rp);
    const float p = r0 * tf;
    const float plo = fmaf(r0, tf, -p);
    float add;
    if (a.R == 1) {
        add = s.off - 0.5f;
        s.dxdoff = 1.f / tf;
    } else {
        const float r1 = __ldg(rp + 1);
        s.dxdoff = r1 * (0.5f / (float)a.P);
        add = fmaf(s.off * s.dxdoff, tf, -0.5f);
    }
    float hi, e;
    two_sum_f(p, add, hi, e);
    s.tap = locate_1d(hi, plo + e, T, a.pad_mode);
    const long long MD = (long long)a.M * a.D;
    const long long base = n * a.S * MD + (long long)m * a.D;
    s.row0 = base + (long long)(lt.start[s.l] + s.tap.i0) * MD;
    s.row1 = base + (long long)(lt.start[s.l] + s.tap.i1) * MD;
    return s;
}

template <typename VT>
__global__ void __launch_bounds__(256) msda_cap_fwd_kernel(const CapArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    __shared__ LevelTable lt;
    load_level_table(lt, a.shapes, a.lsi, a.L, 1);
    __syncthreads();
    const VT *value = static_cast<const VT *>(a.value);
    VT *out = static_cast<VT *>(a.samples);
    const int DV = a.D / CH;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < a.total; idx += stride) {
        const int vec = (int)(idx % DV);
        const long long sidx = idx / DV;
        const CapSample s = cap_locate(a, lt, sidx);
        float v0[CH], v1[CH], r[CH];
        Vec16<VT>::load(value + s.row0 + vec * CH, v0);
        Vec16<VT>::load(value + s.row1 + vec * CH, v1);
#pragma unroll
        for (int c = 0; c < CH; ++c) r[c] = s.tap.w0 * v0[c] + s.tap.w1 * v1[c];
        Vec16<VT>::store(out + sidx * a.D + vec * CH, r);
    }
}

template <typename VT>
__global__ void __launch_bounds__(256) msda_cap_bwd_kernel(const CapArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    __shared__ LevelTable lt;
    load_level_table(lt, a.shapes, a.lsi, a.L, 1);
    __syncthreads();
    const VT *value = static_cast<const VT *>(a.value);
    const VT *gs = static_cast<const VT *>(a.grad_samples);
    const int DV = a.D / CH;
    // DV % 32 == 0: every warp lies inside one sample -> shuffle-reduce the location gradient
    const bool warp_in_sample = (DV % 32) == 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long rounds = (a.total + stride - 1) / stride;
    for (long long it = 0; it < rounds; ++it) {
        const long long idx = it * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const bool live = idx < a.total;
        float part = 0.f;
        CapSample s{};
        long long sidx = 0;
        if (live) {
            const int vec = (int)(idx % DV);
            sidx = idx / DV;
            s = cap_locate(a, lt, sidx);
            float g[CH], v0[CH], v1[CH];
            Vec16<VT>::load(gs + sidx * a.D + vec * CH, g);
            Vec16<VT>::load(value + s.row0 + vec * CH, v0);
            Vec16<VT>::load(value + s.row1 + vec * CH, v1);
#pragma unroll
            for (int c = 0; c < CH; ++c) part += g[c] * (s.tap.s0 * v0[c] + s.tap.s1 * v1[c]);
            float *p0 = a.gv_acc + s.row0 + vec * CH;
            float *p1 = a.gv_acc + s.row1 + vec * CH;
#pragma unroll
            for (int c = 0; c < CH; c += 4) {
                if (s.tap.w0 != 0.f)
                    red_add_v4(p0 + c, s.tap.w0 * g[c], s.tap.w0 * g[c + 1], s.tap.w0 * g[c + 2], s.tap.w0 * g[c + 3]);
                if (s.tap.w1 != 0.f)
                    red_add_v4(p1 + c, s.tap.w1 * g[c], s.tap.w1 * g[c + 1], s.tap.w1 * g[c + 2], s.tap.w1 * g[c + 3]);
            }
        }
        bool writer = live;
        if (warp_in_sample) {           // total % 32 == 0 as well, so the whole warp is live or dead together
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) part += __shfl_xor_sync(kFull, part, o);
            writer = live && (threadIdx.x & 31) == 0;
        }
        if (writer && part != 0.f) {
            const float gpix = part * s.tl;                  // d/d(loc): pix = loc * T_l - 0.5
            atomicAdd(a.grad_off + sidx, gpix * s.dxdoff);
            if (a.grad_ref) {
                float *gr = a.grad_ref + (s.nq * a.L + s.l) * a.R;
                atomicAdd(gr, gpix);
                if (a.R == 2) atomicAdd(gr + 1, gpix * s.off * (0.5f / (float)a.P));
            }
        }
    }
}

}  // namespace msda
