// msda_optim.cuh -- gradient-norm clipping + Adam of the PDVC training step in two launches (the harness around
// the hot path).
//
// The reference's step ends with `torch.nn.utils.clip_grad_norm_(model.parameters(), opt.grad_clip)` and
// `optimizer.step()` of torch.optim.Adam with weight decay (train.py:185-187, optimizer :123).  On ~130 parameter tensors
// that is ~12 launches (per-tensor norms, stack, norm, clamp, scaling, four multi-tensor Adam chunks) and 0.28 ms of a 5.4 ms
// step.  The trainer keeps every gradient in flat buckets (pdvc/train.py, `sync="overlap"`), so the same arithmetic is:
//   msda_sumsq_kernel   per-block partial sums of g^2 over the buckets (no atomics: deterministic)
//   msda_adam_kernel    every block adds up the partials -> total norm -> clip coefficient (clip_grad_norm_'s
//                       max_norm / (norm + 1e-6), capped at 1), then walks its chunks of (param, grad, exp_avg, exp_avg_sq)
//                       with torch's fused-Adam update rule.  The gradients themselves are left unscaled.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace msda {
namespace optim {

constexpr int kThreads = 256;
constexpr int kChunk = 8192;          // elements per table entry (one block pass)
constexpr int kMaxPartials = 1024;    // blocks of the sum-of-squares kernel

struct Segment {                      // sum-of-squares input: one flat gradient bucket
    const float *g;
    long long n;
};

struct Chunk {                        // <= kChunk consecutive elements of one parameter
    float *param;
    const float *grad;
    float *exp_avg;
    float *exp_avg_sq;
    int n;
    int pad;
};

// derived on the host in double precision and rounded once (1 - 0.999f computed in float is off by 1e-5 relative)
struct Hyper {
    float step_size;                 // lr / (1 - beta1^t)
    float beta2, one_minus_beta1, one_minus_beta2;
    float eps, weight_decay;
    float bias_correction2_sqrt;     // sqrt(1 - beta2^t)
    float max_norm;                  // <= 0: no clipping
};

__device__ __forceinline__ float block_sum(float v, float *scratch)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < kThreads / 32) t = scratch[threadIdx.x];
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = kThreads / 64; o >= 1; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t;   // valid in thread 0
}

// partials[blockIdx.x] = sum of g^2 over this block's grid-stride share of all segments (16-byte loads where the segment
// is aligned, four in flight per thread: the pass is a pure HBM read)
__global__ void __launch_bounds__(kThreads) msda_sumsq_kernel(const Segment *segs, int n_segs, float *partials)
{
    __shared__ float scratch[kThreads / 32];
    float acc = 0.f;
    const long long stride = (long long)gridDim.x * kThreads;
    const long long first = (long long)blockIdx.x * kThreads + threadIdx.x;
    for (int s = 0; s < n_segs; ++s) {
        const float *g = segs[s].g;
        const long long n = segs[s].n;
        long long done = 0;
        if ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) {
            const float4 *g4 = reinterpret_cast<const float4 *>(g);
            const long long n4 = n >> 2;
            long long i = first;
            for (; i + 3 * stride < n4; i += 4 * stride) {
                const float4 a = g4[i], b = g4[i + stride], c = g4[i + 2 * stride], d = g4[i + 3 * stride];
                acc = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, acc))));
                acc = fmaf(b.x, b.x, fmaf(b.y, b.y, fmaf(b.z, b.z, fmaf(b.w, b.w, acc))));
                acc = fmaf(c.x, c.x, fmaf(c.y, c.y, fmaf(c.z, c.z, fmaf(c.w, c.w, acc))));
                acc = fmaf(d.x, d.x, fmaf(d.y, d.y, fmaf(d.z, d.z, fmaf(d.w, d.w, acc))));
            }
            for (; i < n4; i += stride) {
                const float4 a = g4[i];
                acc = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, acc))));
            }
            done = n4 << 2;
        }
        for (long long i = done + first; i < n; i += stride) { const float a = g[i]; acc = fmaf(a, a, acc); }
    }
    const float t = block_sum(acc, scratch);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

// torch's fused-Adam rule for one element (fused_adam_utils.cuh, ADAM_MODE::ORIGINAL, no amsgrad)
__device__ __forceinline__ void adam_update(float &p, float &m, float &v, float grad, float coef, const Hyper &h)
{
    float g = grad * coef;
    if (h.weight_decay != 0.f) g = fmaf(h.weight_decay, p, g);        // Adam's L2 form: grad += wd * param
    m = fmaf(h.one_minus_beta1, g - m, m);                             // lerp(m, g, 1 - beta1)
    v = fmaf(h.beta2, v, h.one_minus_beta2 * g * g);
    const float denom = sqrtf(v) / h.bias_correction2_sqrt + h.eps;
    p -= h.step_size * m / denom;
}

__global__ void __launch_bounds__(kThreads) msda_adam_kernel(const Chunk *chunks, int n_chunks, const float *partials,
                                                             int n_partials, const Hyper h, float *norm_out)
{
    __shared__ float scratch[kThreads / 32];
    __shared__ float coef_s;
    float coef = 1.f;
    if (h.max_norm > 0.f) {
        float acc = 0.f;
        for (int i = threadIdx.x; i < n_partials; i += kThreads) acc += partials[i];
        const float t = block_sum(acc, scratch);
        if (threadIdx.x == 0) {
            const float norm = sqrtf(t);
            coef_s = fminf(h.max_norm / (norm + 1e-6f), 1.f);          // torch.nn.utils.clip_grad_norm_
            if (blockIdx.x == 0 && norm_out) *norm_out = norm;
        }
        __syncthreads();
        coef = coef_s;
    }
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const Chunk ck = chunks[c];
        int done = 0;
        if (((reinterpret_cast<uintptr_t>(ck.param) | reinterpret_cast<uintptr_t>(ck.grad) | reinterpret_cast<uintptr_t>(ck.exp_avg) |
              reinterpret_cast<uintptr_t>(ck.exp_avg_sq)) & 15u) == 0) {
            // 16-byte path: the update moves 28 bytes per parameter and nothing else -- keep the loads wide and many in flight
            float4 *p4 = reinterpret_cast<float4 *>(ck.param), *m4 = reinterpret_cast<float4 *>(ck.exp_avg),
                   *v4 = reinterpret_cast<float4 *>(ck.exp_avg_sq);
            const float4 *g4 = reinterpret_cast<const float4 *>(ck.grad);
            const int n4 = ck.n >> 2;
#pragma unroll 2
            for (int i = threadIdx.x; i < n4; i += kThreads) {
                float4 p = p4[i], m = m4[i], v = v4[i];
                const float4 g = g4[i];
                adam_update(p.x, m.x, v.x, g.x, coef, h);
                adam_update(p.y, m.y, v.y, g.y, coef, h);
                adam_update(p.z, m.z, v.z, g.z, coef, h);
                adam_update(p.w, m.w, v.w, g.w, coef, h);
                p4[i] = p; m4[i] = m; v4[i] = v;
            }
            done = n4 << 2;
        }
        for (int i = done + threadIdx.x; i < ck.n; i += kThreads) {
            float p = ck.param[i], m = ck.exp_avg[i], v = ck.exp_avg_sq[i];
            adam_update(p, m, v, ck.grad[i], coef, h);
            ck.param[i] = p; ck.exp_avg[i] = m; ck.exp_avg_sq[i] = v;
        }
    }
}

}  // namespace optim
}  // namespace msda
