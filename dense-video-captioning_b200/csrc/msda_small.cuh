// msda_small.cuh -- single-phase kernels for launch-latency-bound problems (the model's own shapes: BASELINE configs[0],
// N = 1, 100-375 queries, where a call moves 3-5 MB that live in L2 and the lean kernels' six-barrier pipeline -- window
// discovery, staging, records / sort, scan -- costs more than the arithmetic).
//
// One CTA per (n, q), one warp per head: lane j decodes sample j = (l, p) of its (query, head) -- softmax over the warp,
// sampling location, the two value rows and their weights -- and the warp then walks its L*P samples with the taps read
// straight from global memory (L1 / L2 resident): 32 / HL lane groups of HL lanes (16 bytes each) cover both rows of
// 16 / HL samples per step.  Backward: same walk; the two dot products of a sample come back to its decode lane through
// shared memory, grad_value goes out as one vector reduction per lane and tap, and the reference-point gradient is summed
// over heads and points inside the CTA (no atomics, no zero-fill).
//
// Replaces, like the lean kernels, ms_deformable_im2col_gpu_kernel / the col2im kernels (ms_deform_im2col_cuda.cuh:238-300,
// :302-921) for H_l == 1, and FUSED the module front-end (ms_deform_attn.py:100-116).
#pragma once
#include "msda_lean.cuh"

namespace msda {

constexpr int kSmallMaxHeads = 16;   // warps per CTA (512 threads: 128 registers each, room for a batch of 8 row loads in flight)
constexpr int kSmallBatch = 8;       // row loads issued back to back before the first one is consumed

// decode of lane j's sample: attention weight (softmax over the warp when FUSED) and sampling geometry.  These kernels are
// a chain of dependent memory round trips (cold: ~0.7 us each), so everything independent is issued together: the lane's
// level length / start (read-only path, the same few words for every warp) and its raw sample inputs go out first.
template <bool FUSED>
__device__ __forceinline__ Decoded small_decode(const TileArgs &a, int nq, int m, int lane, LaneLevel &c, RawSample &raw)
{
    c.live = lane < a.LP;
    int l = lane / a.P;
    l = l < a.L ? l : a.L - 1;
    c.l = l;
    const long long Tl = __ldg(a.shapes + (a.shapes_stride == 2 ? 2 * l + 1 : l));
    const long long st = __ldg(a.lsi + l);
    raw = load_raw<FUSED>(a, c, nq, (nq * a.M + m) * a.LP + lane, c.live);
    c.T = (int)Tl;
    c.start = (int)st;
    c.tf = (float)c.T;
    float A = raw.y;
    if (FUSED) {
        const float mx = group_max<32>(c.live ? raw.y : -INFINITY);
        const float e = c.live ? __expf(raw.y - mx) : 0.f;
        const float sum = group_sum<32>(e);
        A = sum > 0.f ? __fdividef(e, sum) : 0.f;
    }
    return decode_geom<FUSED>(a, c, raw, A);
}

// grid = N * Lq, block = 32 * M
template <typename VT, int HL, bool FUSED>
__global__ void __launch_bounds__(32 * kSmallMaxHeads) msda_fwd_small_kernel(const TileArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    constexpr int G = 32 / HL;      // lane groups per warp
    constexpr int SPI = G / 2;      // samples per step (two taps each)
    const int m = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = blockIdx.x, n = nq / a.Lq;
    LaneLevel c;
    RawSample raw;
    const Decoded d = small_decode<FUSED>(a, nq, m, lane, c, raw);
    const float k0 = c.live ? d.A * d.w0 : 0.f, k1 = c.live ? d.A * d.w1 : 0.f;

    const int grp = lane / HL, ch = lane % HL, tap = grp & 1, sub = grp >> 1;
    const int MD = a.M * a.D;
    const VT *vslice = static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D + ch * CH;
    float acc[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) acc[k] = 0.f;
    const int steps = (a.LP + SPI - 1) / SPI;
    for (int it0 = 0; it0 < steps; it0 += kSmallBatch) {
        float v[kSmallBatch][CH], kk[kSmallBatch];
#pragma unroll
        for (int u = 0; u < kSmallBatch; ++u) {             // all row loads of the batch first (dead taps read row 0 of the slice)
            const int s = (it0 + u) * SPI + sub;
            const int ra = __shfl_sync(kFull, d.r0, s & 31), rb = __shfl_sync(kFull, d.r1, s & 31);
            const float ka = __shfl_sync(kFull, k0, s & 31), kb = __shfl_sync(kFull, k1, s & 31);
            kk[u] = s < a.LP ? (tap ? kb : ka) : 0.f;
            const int r = kk[u] != 0.f ? (tap ? rb : ra) : 0;
            Vec16<VT>::load(vslice + (size_t)r * MD, v[u]);
        }
#pragma unroll
        for (int u = 0; u < kSmallBatch; ++u)
#pragma unroll
            for (int k = 0; k < CH; ++k) acc[k] = fmaf(kk[u], v[u][k], acc[k]);
    }
    // lanes with equal channel chunk (one per group) hold partial sums
#pragma unroll
    for (int off = HL; off < 32; off <<= 1)
#pragma unroll
        for (int k = 0; k < CH; ++k) acc[k] += __shfl_xor_sync(kFull, acc[k], off);
    if (grp == 0) Vec16<VT>::store(static_cast<VT *>(a.out) + (size_t)(nq * a.M + m) * a.D + ch * CH, acc);
}

// grid = N * Lq, block = 32 * M.  gv_acc (fp32 image of grad_value) is zero-filled by the host wrapper.
template <typename VT, int HL, bool FUSED>
__global__ void __launch_bounds__(32 * kSmallMaxHeads) msda_bwd_small_kernel(const TileArgs a)
{
    constexpr int CH = Vec16<VT>::CH;
    constexpr int G = 32 / HL;
    constexpr int SPI = G / 2;
    __shared__ float sdots[kSmallMaxHeads][64];            // [head][sample][tap]
    __shared__ float sref[2 * kMaxLevels];                 // [level][R]: sums over heads and points
    const int m = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nq = blockIdx.x, n = nq / a.Lq;
    for (int i = threadIdx.x; i < 2 * kMaxLevels; i += blockDim.x) sref[i] = 0.f;
    __syncthreads();
    LaneLevel c;
    RawSample raw;
    const Decoded d = small_decode<FUSED>(a, nq, m, lane, c, raw);
    const float k0 = c.live ? d.A * d.w0 : 0.f, k1 = c.live ? d.A * d.w1 : 0.f;

    const int grp = lane / HL, ch = lane % HL, tap = grp & 1, sub = grp >> 1;
    const int MD = a.M * a.D;
    const VT *vslice = static_cast<const VT *>(a.value) + (size_t)n * a.S * MD + m * a.D + ch * CH;
    float *gslice = a.gv_acc + (size_t)n * a.S * MD + m * a.D + ch * CH;
    float g[CH];
    Vec16<VT>::load(static_cast<const VT *>(a.grad_out) + (size_t)(nq * a.M + m) * a.D + ch * CH, g);
    const int steps = (a.LP + SPI - 1) / SPI;
    for (int it0 = 0; it0 < steps; it0 += kSmallBatch) {
        float v[kSmallBatch][CH], kk[kSmallBatch];
        int rr[kSmallBatch];
#pragma unroll
        for (int u = 0; u < kSmallBatch; ++u) {             // all row loads of the batch first
            const int s = (it0 + u) * SPI + sub;
            const int ra = __shfl_sync(kFull, d.r0, s & 31), rb = __shfl_sync(kFull, d.r1, s & 31);
            const float ka = __shfl_sync(kFull, k0, s & 31), kb = __shfl_sync(kFull, k1, s & 31);
            kk[u] = s < a.LP ? (tap ? kb : ka) : 0.f;
            rr[u] = s < a.LP ? (tap ? rb : ra) : 0;
            Vec16<VT>::load(vslice + (size_t)rr[u] * MD, v[u]);
        }
#pragma unroll
        for (int u = 0; u < kSmallBatch; ++u) {
            const int s = (it0 + u) * SPI + sub;
            if (kk[u] != 0.f) {
#pragma unroll
                for (int k = 0; k < CH; k += 4)
                    red_add_v4(gslice + (size_t)rr[u] * MD + k, kk[u] * g[k], kk[u] * g[k + 1], kk[u] * g[k + 2], kk[u] * g[k + 3]);
            }
            float p = 0.f;
#pragma unroll
            for (int k = 0; k < CH; ++k) p = fmaf(g[k], v[u][k], p);
#pragma unroll
            for (int off = HL / 2; off >= 1; off >>= 1) p += __shfl_xor_sync(kFull, p, off);
            if (ch == 0 && s < a.LP) sdots[m][2 * s + tap] = p;
        }
    }
    __syncwarp();
    // ---- per-sample gradients (lane j = sample j) ----
    const float p0 = c.live ? sdots[m][2 * lane] : 0.f, p1 = c.live ? sdots[m][2 * lane + 1] : 0.f;
    const float gat = c.live ? d.w0 * p0 + d.w1 * p1 : 0.f;                                   // d out / d attention weight
    const float gx = c.live ? d.A * c.tf * (d.s0 * p0 + d.s1 * p1) : 0.f;                     // d out / d location
    const int sidx = (nq * a.M + m) * a.LP + lane;
    if (!FUSED) {
        if (c.live) {
            a.grad_attn[sidx] = gat;
            const float gy = a.pad_mode == MSDA_PAD_ZEROS ? -d.A * gat : 0.f;                 // see msda_lean.cuh (cuh:159)
            *reinterpret_cast<float2 *>(a.grad_loc + 2 * sidx) = make_float2(gx, gy);
        }
    } else {
        const float dotsum = group_sum<32>(d.A * gat);                                        // softmax Jacobian
        const float dxd = a.R == 1 ? __fdividef(1.f, c.tf) : raw.r1 * (0.5f * a.invP);
        if (c.live) {
            a.grad_offs[sidx] = gx * dxd;
            a.grad_logits[sidx] = d.A * (gat - dotsum);
        }
        if (a.grad_ref != nullptr) {
            if (c.live) {
                atomicAdd(&sref[c.l * 2], gx);                                                 // d(loc)/d(ref0) = 1
                if (a.R == 2) atomicAdd(&sref[c.l * 2 + 1], gx * raw.x * (0.5f * a.invP));     // d(loc)/d(ref1) = off * 0.5 / P
            }
            __syncthreads();
            if ((int)threadIdx.x < a.L * a.R) {
                const int l = threadIdx.x / a.R, r = threadIdx.x % a.R;
                a.grad_ref[(nq * a.L + l) * a.R + r] = sref[l * 2 + r];
            }
        }
    }
}

}  // namespace msda
