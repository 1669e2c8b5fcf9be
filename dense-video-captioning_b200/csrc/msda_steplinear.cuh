// msda_steplinear.cuh -- Linear layers of the caption head's word step.
//
// Inside the caption LSTM's word loop two Linear layers sit on the recurrence in front of the sampler: the
// `sampling_offsets` projection of MSDeformAttnCap (ms_deform_attn_for_caption.py:100, 2*d_model -> M*L*P) and
// `h2att` (LSTM_DSA.py:249, rnn_size -> att_hid_size).  Their input has one row per matched (query, caption)
// pair -- 4 to ~30 rows -- so the product is a handful of dot products per weight row, bound by reading the
// weight matrix once (L2 resident, 1-2 MB).  cuBLAS runs these as a 32x32x16 SIMT GEMM on 1..16 CTAs with a
// serial K loop: 12 us per call, 2-4 calls per word each way.  Here one warp owns NW weight rows, its lanes
// stride K in float4s, and all 148 SMs read the weight in parallel.
//
//   y[m, n] = sum_k x[m, k] * w[n, k] (+ bias[n])        x (M,K), w (N,K), y (M,N), fp32, K % 4 == 0
//
// The backward data gradient is the same product against the transposed weight (made once per backward pass
// by the caller); the weight gradient is deferred to one GEMM over all word steps (pdvc/step_linear.py).
#pragma once
#include "msda_common.cuh"

namespace msda {

constexpr int kSkinnyRows = 8;   // rows of x per CTA (blockIdx.y tiles M)
constexpr int kSkinnyNW = 2;     // weight rows per warp
constexpr int kSkinnyWarps = 8;

__global__ void __launch_bounds__(kSkinnyWarps * 32)
msda_skinny_linear_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ bias,
                          float *__restrict__ y, int M, int N, int K)
{
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int n0 = (blockIdx.x * kSkinnyWarps + warp) * kSkinnyNW;
    const int m0 = blockIdx.y * kSkinnyRows;
    if (n0 >= N) return;
    const int K4 = K >> 2;
    const float4 *x4 = reinterpret_cast<const float4 *>(x);
    const float4 *w4 = reinterpret_cast<const float4 *>(w);
    float acc[kSkinnyRows][kSkinnyNW];
#pragma unroll
    for (int r = 0; r < kSkinnyRows; ++r)
#pragma unroll
        for (int j = 0; j < kSkinnyNW; ++j) acc[r][j] = 0.f;
    const int rows = min(kSkinnyRows, M - m0);
    for (int k4 = lane; k4 < K4; k4 += 32) {
        float4 wv[kSkinnyNW];
#pragma unroll
        for (int j = 0; j < kSkinnyNW; ++j)
            wv[j] = (n0 + j < N) ? __ldg(w4 + (long long)(n0 + j) * K4 + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < kSkinnyRows; ++r) {
            if (r < rows) {
                const float4 xv = __ldg(x4 + (long long)(m0 + r) * K4 + k4);
#pragma unroll
                for (int j = 0; j < kSkinnyNW; ++j)
                    acc[r][j] = fmaf(xv.x, wv[j].x, fmaf(xv.y, wv[j].y, fmaf(xv.z, wv[j].z, fmaf(xv.w, wv[j].w, acc[r][j]))));
            }
        }
    }
    // butterfly over the 32 lanes; afterwards every lane holds the full sums, lane (r*NW + j) stores its own
    float mine = 0.f;
#pragma unroll
    for (int r = 0; r < kSkinnyRows; ++r)
#pragma unroll
        for (int j = 0; j < kSkinnyNW; ++j) {
            float v = acc[r][j];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
            if (lane == r * kSkinnyNW + j) mine = v;
        }
    const int r = lane / kSkinnyNW, j = lane % kSkinnyNW;
    if (lane < kSkinnyRows * kSkinnyNW && r < rows && n0 + j < N)
        y[(long long)(m0 + r) * N + n0 + j] = mine + (bias ? __ldg(bias + n0 + j) : 0.f);
}

}  // namespace msda
