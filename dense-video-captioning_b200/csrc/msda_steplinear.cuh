// msda_steplinear.cuh -- Linear layers of the caption head's word step.
//
// Inside the caption LSTM's word loop two Linear layers sit on the recurrence in front of the sampler: the
// `sampling_offsets` projection of MSDeformAttnCap (ms_deform_attn_for_caption.py:100, 2*d_model -> M*L*P) and
// `h2att` (LSTM_DSA.py:249, rnn_size -> att_hid_size).  Their input has one row per matched (query, caption)
// pair -- 4 to ~30 rows -- so the product is a handful of dot products per weight row, bound by reading the
// weight matrix once (L2 resident, 1-2 MB).  cuBLAS runs these as a 32x32x16 SIMT GEMM on 1..16 CTAs with a
// serial K loop: 6-12 us per call, 2-4 calls per word each way.  Here a warp (or KS warps splitting K) owns NW weight
// rows, its lanes stride K in float4s, and all 148 SMs read the weight in parallel.
//
//   y[m, n] = sum_k x[m, k] * w[n, k] (+ bias[n])        x (M,K), w (N,K), y (M,N), fp32, K % 4 == 0
//
// The backward data gradient is the same product against the transposed weight (made once per backward pass
// by the caller); the weight gradient is deferred to one GEMM over all word steps (pdvc/step_linear.py).
#pragma once
#include "msda_common.cuh"
#include "msda_tile.cuh"   // reduce_nv

namespace msda {

constexpr int kSkinnyNW = 2;     // weight rows per warp
constexpr int kSkinnyWarps = 8;

// KS = warps that share one group of NW weight rows and split K between them (short N / long K: more CTAs, one or two
// float4 per lane); their partial sums meet in shared memory.
// ROWS = rows of x per CTA (blockIdx.y tiles M): 8, or 16 so that up to 16 rows read the weight once.
template <int KS, int ROWS>
__global__ void __launch_bounds__(kSkinnyWarps * 32)
msda_skinny_linear_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ bias,
                          float *__restrict__ y, int M, int N, int K)
{
    constexpr int GROUPS = kSkinnyWarps / KS;
    constexpr int kSkinnyRows = ROWS;
    constexpr int OUT = kSkinnyRows * kSkinnyNW;
    __shared__ float part[GROUPS][KS][OUT];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int slice = warp % KS, group = warp / KS;
    const int n0 = (blockIdx.x * GROUPS + group) * kSkinnyNW;
    const int m0 = blockIdx.y * kSkinnyRows;
    const int K4 = K >> 2;
    const float4 *x4 = reinterpret_cast<const float4 *>(x);
    const float4 *w4 = reinterpret_cast<const float4 *>(w);
    float acc[kSkinnyRows][kSkinnyNW];
#pragma unroll
    for (int r = 0; r < kSkinnyRows; ++r)
#pragma unroll
        for (int j = 0; j < kSkinnyNW; ++j) acc[r][j] = 0.f;
    const int rows = min(kSkinnyRows, M - m0);
    if (n0 < N) {
        for (int k4 = slice * 32 + lane; k4 < K4; k4 += 32 * KS) {
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
    }
    // sum over the 32 lanes with the transposing butterfly (31 shuffles for 32 sums instead of 5 each: this kernel runs as
    // one wave of ~28 warps per SM and the shuffle pipe, one warp instruction per clock and SM, was a fifth of its time);
    // lane l ends up with the total of sum number l / (32 / OUT), numbered r * NW + j
    float flat[OUT];
#pragma unroll
    for (int r = 0; r < kSkinnyRows; ++r)
#pragma unroll
        for (int j = 0; j < kSkinnyNW; ++j) flat[r * kSkinnyNW + j] = acc[r][j];
    reduce_nv<32, OUT>(flat, lane);
    constexpr int LPV = 32 / OUT;            // lanes that hold the same total
    const int vidx = lane / LPV;
    const bool writer = lane % LPV == 0;
    float mine = flat[0];
    if (KS > 1) {
        if (writer) part[group][slice][vidx] = mine;
        __syncthreads();
        if (slice != 0) return;
        mine = 0.f;
#pragma unroll
        for (int s = 0; s < KS; ++s) mine += part[group][s][vidx];
    }
    const int r = vidx / kSkinnyNW, j = vidx % kSkinnyNW;
    if (writer && r < rows && n0 + j < N)
        y[(long long)(m0 + r) * N + n0 + j] = mine + (bias ? __ldg(bias + n0 + j) : 0.f);
}

// ---- LSTM cell of the word step (nn.LSTM with one layer and one time step, LSTM_DSA.py:206, :261) -------------------
// gates = pre + rec, (B, 4H) in torch's order i, f, g, o.  `pre` = the part of W_ih x that does not depend on the
// recurrence (word embedding, event query: computed for all word steps in one GEMM by the caller), row stride
// pre_stride; `rec` = [attended, h_prev] against [W_ih(attended columns) | W_hh] (skinny linear above).
__device__ __forceinline__ float sigmoid_f(float v) { return 1.f / (1.f + expf(-v)); }

__global__ void __launch_bounds__(256)
msda_lstm_cell_fwd_kernel(const float *__restrict__ pre, long long pre_stride, const float *__restrict__ rec,
                          const float *__restrict__ c_prev, float *__restrict__ c, float *__restrict__ h,
                          float *__restrict__ act, int B, int H)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * H) return;
    const int b = idx / H, j = idx - b * H;
    const float *p = pre + (long long)b * pre_stride;
    const float *r = rec + (long long)b * 4 * H;
    const float gi = sigmoid_f(__ldg(p + j) + __ldg(r + j));
    const float gf = sigmoid_f(__ldg(p + H + j) + __ldg(r + H + j));
    const float gg = tanhf(__ldg(p + 2 * H + j) + __ldg(r + 2 * H + j));
    const float go = sigmoid_f(__ldg(p + 3 * H + j) + __ldg(r + 3 * H + j));
    const float cn = fmaf(gf, __ldg(c_prev + idx), gi * gg);
    c[idx] = cn;
    h[idx] = go * tanhf(cn);
    float *a = act + (long long)b * 4 * H;
    a[j] = gi; a[H + j] = gf; a[2 * H + j] = gg; a[3 * H + j] = go;
}

// grad_c may be NULL (no gradient reaches the cell state of the last word step)
__global__ void __launch_bounds__(256)
msda_lstm_cell_bwd_kernel(const float *__restrict__ grad_h, const float *__restrict__ grad_c,
                          const float *__restrict__ act, const float *__restrict__ c,
                          const float *__restrict__ c_prev, float *__restrict__ grad_gates,
                          float *__restrict__ grad_c_prev, int B, int H)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * H) return;
    const int b = idx / H, j = idx - b * H;
    const float *a = act + (long long)b * 4 * H;
    const float gi = __ldg(a + j), gf = __ldg(a + H + j), gg = __ldg(a + 2 * H + j), go = __ldg(a + 3 * H + j);
    const float tc = tanhf(__ldg(c + idx));
    const float dh = grad_h ? __ldg(grad_h + idx) : 0.f;
    const float dc = (grad_c ? __ldg(grad_c + idx) : 0.f) + dh * go * (1.f - tc * tc);
    float *g = grad_gates + (long long)b * 4 * H;
    g[j] = dc * gg * gi * (1.f - gi);
    g[H + j] = dc * __ldg(c_prev + idx) * gf * (1.f - gf);
    g[2 * H + j] = dc * gi * (1.f - gg * gg);
    g[3 * H + j] = dh * tc * go * (1.f - go);
    grad_c_prev[idx] = dc * gf;
}

}  // namespace msda
