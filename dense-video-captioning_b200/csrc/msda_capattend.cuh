// msda_capattend.cuh -- soft attention over the deformable samples of the caption head.
//
// After MSDeformAttnCap has sampled the L*P clip features of an event, ShowAttendTellCore weighs them
// (pdvc/CaptioningHead/LSTM_DSA.py:247-257):
//     score_k  = alpha_net(tanh(ctx2att(clip_k) + h2att(h)))          k = 0 .. L*P-1
//     weight   = softmax_k(score)
//     attended = sum_k weight_k * clip_k
// In torch that is add, tanh, gemv, softmax, bmm forward and ~12 kernels backward, per generated word, on 6-30 rows.
// Here: one kernel each way, one CTA per (event, head).  The sampler (msda_capsample.cuh) is run on [value | ctx2att(value)]
// so that one sample row carries clip_k (D channels) followed by ctx2att(clip_k) (A channels); the backward kernel
// writes the gradient of that whole row, ready for the sampler's backward.
#pragma once
#include "msda_common.cuh"

namespace msda {

constexpr int kAttendMaxK = 64;
constexpr int kAttendThreads = 512;
constexpr int kAttendKG = 4;          // the K samples are walked by KG thread groups in parallel, partial sums meet in smem

struct AttendArgs {
    const float *samples;     // (BM, K, D + A)
    const float *hproj;       // (B, A): h2att(h), shared by the M heads of an event
    const float *alpha;       // (A,)  alpha_net.weight
    const float *alpha_bias;  // (1,)
    float *attended;          // (BM, D)                    forward out
    float *weight;            // (BM, K) softmax weights    forward out / backward in
    const float *grad_att;    // (BM, D)                    backward in
    float *grad_samples;      // (BM, K, D + A)             backward out
    float *grad_hproj;        // (BM, A)                    backward out (caller sums the heads)
    float *grad_alpha;        // (BM, A + 1): per-CTA partial of d alpha (A) and d alpha_bias (1)
    int BM, M, K, D, A;
};

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

__device__ __forceinline__ float4 ldg4(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }

// D % 4 == 0, A % 4 == 0; dynamic shared memory: KG * D floats (forward), 2 * KG * A floats (backward)
__global__ void __launch_bounds__(kAttendThreads) msda_cap_attend_fwd_kernel(const AttendArgs a)
{
    extern __shared__ float4 dyn4[];
    __shared__ float score[kAttendMaxK];
    const int bm = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = a.D + a.A;
    const float *s = a.samples + (long long)bm * a.K * row;
    const float *hp = a.hproj + (long long)(bm / a.M) * a.A;
    for (int k = warp; k < a.K; k += kAttendThreads / 32) {
        const float *ctx = s + (long long)k * row + a.D;
        float acc = 0.f;
        for (int h = lane * 4; h < a.A; h += 128) {
            const float4 c = ldg4(ctx + h), hv = ldg4(hp + h), al = ldg4(a.alpha + h);
            acc = fmaf(al.x, tanhf(c.x + hv.x), acc);
            acc = fmaf(al.y, tanhf(c.y + hv.y), acc);
            acc = fmaf(al.z, tanhf(c.z + hv.z), acc);
            acc = fmaf(al.w, tanhf(c.w + hv.w), acc);
        }
        acc = warp_sum(acc);
        if (lane == 0) score[k] = acc + __ldg(a.alpha_bias);
    }
    __syncthreads();
    if (warp == 0) {                       // softmax over K <= 64: two values per lane
        const float v0 = lane < a.K ? score[lane] : -INFINITY;
        const float v1 = lane + 32 < a.K ? score[lane + 32] : -INFINITY;
        float mx = fmaxf(v0, v1);
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(kFull, mx, o));
        const float e0 = lane < a.K ? expf(v0 - mx) : 0.f;
        const float e1 = lane + 32 < a.K ? expf(v1 - mx) : 0.f;
        const float inv = 1.f / warp_sum(e0 + e1);
        if (lane < a.K) score[lane] = e0 * inv;
        if (lane + 32 < a.K) score[lane + 32] = e1 * inv;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < a.K; k += kAttendThreads) a.weight[(long long)bm * a.K + k] = score[k];
    const int nvec = a.D >> 2, per = (a.K + kAttendKG - 1) / kAttendKG;
    for (int idx = threadIdx.x; idx < nvec * kAttendKG; idx += kAttendThreads) {
        const int kg = idx / nvec, d = (idx - kg * nvec) * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = kg * per; k < min(a.K, (kg + 1) * per); ++k) {
            const float4 v = ldg4(s + (long long)k * row + d);
            const float wk = score[k];
            acc.x = fmaf(wk, v.x, acc.x); acc.y = fmaf(wk, v.y, acc.y); acc.z = fmaf(wk, v.z, acc.z); acc.w = fmaf(wk, v.w, acc.w);
        }
        dyn4[idx] = acc;
    }
    __syncthreads();
    for (int v = threadIdx.x; v < nvec; v += kAttendThreads) {
        float4 acc = dyn4[v];
#pragma unroll
        for (int kg = 1; kg < kAttendKG; ++kg) {
            const float4 p = dyn4[kg * nvec + v];
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        reinterpret_cast<float4 *>(a.attended + (long long)bm * a.D)[v] = acc;
    }
}

__global__ void __launch_bounds__(kAttendThreads) msda_cap_attend_bwd_kernel(const AttendArgs a)
{
    extern __shared__ float4 dyn4[];
    __shared__ float w[kAttendMaxK], dscore[kAttendMaxK];
    __shared__ float dbias;
    const int bm = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = a.D + a.A;
    const float *s = a.samples + (long long)bm * a.K * row;
    float *gs = a.grad_samples + (long long)bm * a.K * row;
    const float *hp = a.hproj + (long long)(bm / a.M) * a.A;
    const float *ga = a.grad_att + (long long)bm * a.D;
    for (int k = threadIdx.x; k < a.K; k += kAttendThreads) w[k] = __ldg(a.weight + (long long)bm * a.K + k);
    // d weight_k = <grad_att, clip_k>
    for (int k = warp; k < a.K; k += kAttendThreads / 32) {
        const float *clip = s + (long long)k * row;
        float acc = 0.f;
        for (int d = lane * 4; d < a.D; d += 128) {
            const float4 g = ldg4(ga + d), c = ldg4(clip + d);
            acc = fmaf(g.x, c.x, fmaf(g.y, c.y, fmaf(g.z, c.z, fmaf(g.w, c.w, acc))));
        }
        acc = warp_sum(acc);
        if (lane == 0) dscore[k] = acc;
    }
    __syncthreads();
    if (warp == 0) {                       // softmax backward: dscore_k = w_k (dw_k - sum_j w_j dw_j)
        const float p0 = lane < a.K ? w[lane] * dscore[lane] : 0.f;
        const float p1 = lane + 32 < a.K ? w[lane + 32] * dscore[lane + 32] : 0.f;
        const float dot = warp_sum(p0 + p1);
        const float d0 = lane < a.K ? w[lane] * (dscore[lane] - dot) : 0.f;
        const float d1 = lane + 32 < a.K ? w[lane + 32] * (dscore[lane + 32] - dot) : 0.f;
        __syncwarp();
        if (lane < a.K) dscore[lane] = d0;
        if (lane + 32 < a.K) dscore[lane + 32] = d1;
        const float tot = warp_sum(d0 + d1);
        if (lane == 0) dbias = tot;
    }
    __syncthreads();
    // score_k = sum_h alpha_h tanh(ctx_kh + hproj_h): thread (h4, kg) walks its share of the K samples
    const int nvecA = a.A >> 2, nvecD = a.D >> 2, per = (a.K + kAttendKG - 1) / kAttendKG;
    for (int idx = threadIdx.x; idx < nvecA * kAttendKG; idx += kAttendThreads) {
        const int kg = idx / nvecA, h = (idx - kg * nvecA) * 4;
        const float4 al = ldg4(a.alpha + h), hv = ldg4(hp + h);
        float4 g_h = make_float4(0.f, 0.f, 0.f, 0.f), g_al = g_h;
        for (int k = kg * per; k < min(a.K, (kg + 1) * per); ++k) {
            const float4 c = ldg4(s + (long long)k * row + a.D + h);
            const float ds = dscore[k];
            const float tx = tanhf(c.x + hv.x), ty = tanhf(c.y + hv.y), tz = tanhf(c.z + hv.z), tw = tanhf(c.w + hv.w);
            float4 g;
            g.x = ds * al.x * (1.f - tx * tx); g.y = ds * al.y * (1.f - ty * ty);
            g.z = ds * al.z * (1.f - tz * tz); g.w = ds * al.w * (1.f - tw * tw);
            *reinterpret_cast<float4 *>(gs + (long long)k * row + a.D + h) = g;
            g_h.x += g.x; g_h.y += g.y; g_h.z += g.z; g_h.w += g.w;
            g_al.x = fmaf(ds, tx, g_al.x); g_al.y = fmaf(ds, ty, g_al.y); g_al.z = fmaf(ds, tz, g_al.z); g_al.w = fmaf(ds, tw, g_al.w);
        }
        dyn4[idx] = g_h;
        dyn4[nvecA * kAttendKG + idx] = g_al;
    }
    // attended = sum_k w_k clip_k
    for (int idx = threadIdx.x; idx < nvecD * a.K; idx += kAttendThreads) {
        const int k = idx / nvecD, d = (idx - k * nvecD) * 4;
        const float4 g = ldg4(ga + d);
        const float wk = w[k];
        *reinterpret_cast<float4 *>(gs + (long long)k * row + d) = make_float4(wk * g.x, wk * g.y, wk * g.z, wk * g.w);
    }
    __syncthreads();
    for (int v = threadIdx.x; v < nvecA; v += kAttendThreads) {
        float4 g_h = dyn4[v], g_al = dyn4[nvecA * kAttendKG + v];
#pragma unroll
        for (int kg = 1; kg < kAttendKG; ++kg) {
            const float4 p = dyn4[kg * nvecA + v], q = dyn4[nvecA * kAttendKG + kg * nvecA + v];
            g_h.x += p.x; g_h.y += p.y; g_h.z += p.z; g_h.w += p.w;
            g_al.x += q.x; g_al.y += q.y; g_al.z += q.z; g_al.w += q.w;
        }
        reinterpret_cast<float4 *>(a.grad_hproj + (long long)bm * a.A)[v] = g_h;
        float *gal = a.grad_alpha + (long long)bm * (a.A + 1) + 4 * v;       // row stride A+1: scalar stores
        gal[0] = g_al.x; gal[1] = g_al.y; gal[2] = g_al.z; gal[3] = g_al.w;
    }
    if (threadIdx.x == 0) a.grad_alpha[(long long)bm * (a.A + 1) + a.A] = dbias;
}

}  // namespace msda
