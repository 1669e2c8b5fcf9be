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
constexpr int kAttendThreads = 256;

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

__global__ void __launch_bounds__(kAttendThreads) msda_cap_attend_fwd_kernel(const AttendArgs a)
{
    __shared__ float score[kAttendMaxK];
    const int bm = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row = a.D + a.A;
    const float *s = a.samples + (long long)bm * a.K * row;
    const float *hp = a.hproj + (long long)(bm / a.M) * a.A;
    for (int k = warp; k < a.K; k += kAttendThreads / 32) {
        const float *ctx = s + (long long)k * row + a.D;
        float acc = 0.f;
        for (int h = lane; h < a.A; h += 32) acc = fmaf(__ldg(a.alpha + h), tanhf(__ldg(ctx + h) + __ldg(hp + h)), acc);
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
    for (int d = threadIdx.x; d < a.D; d += kAttendThreads) {
        float acc = 0.f;
        for (int k = 0; k < a.K; ++k) acc = fmaf(score[k], __ldg(s + (long long)k * row + d), acc);
        a.attended[(long long)bm * a.D + d] = acc;
    }
}

__global__ void __launch_bounds__(kAttendThreads) msda_cap_attend_bwd_kernel(const AttendArgs a)
{
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
        for (int d = lane; d < a.D; d += 32) acc = fmaf(__ldg(ga + d), __ldg(clip + d), acc);
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
    // score_k = sum_h alpha_h tanh(ctx_kh + hproj_h): one thread per h walks the K samples
    for (int h = threadIdx.x; h < a.A; h += kAttendThreads) {
        const float al = __ldg(a.alpha + h), hv = __ldg(hp + h);
        float g_h = 0.f, g_alpha = 0.f;
        for (int k = 0; k < a.K; ++k) {
            const float t = tanhf(__ldg(s + (long long)k * row + a.D + h) + hv);
            const float g = dscore[k] * al * (1.f - t * t);
            gs[(long long)k * row + a.D + h] = g;
            g_h += g;
            g_alpha = fmaf(dscore[k], t, g_alpha);
        }
        a.grad_hproj[(long long)bm * a.A + h] = g_h;
        a.grad_alpha[(long long)bm * (a.A + 1) + h] = g_alpha;
    }
    if (threadIdx.x == 0) a.grad_alpha[(long long)bm * (a.A + 1) + a.A] = dbias;
    // attended = sum_k w_k clip_k
    for (int d = threadIdx.x; d < a.D; d += kAttendThreads) {
        const float g = __ldg(ga + d);
        for (int k = 0; k < a.K; ++k) gs[(long long)k * row + d] = w[k] * g;
    }
}

}  // namespace msda
