// msda_common.cuh -- device helpers shared by the sm_100a MSDeformAttn kernels.
//
// Math reproduced here (1-D temporal case; 2-D in msda_generic.cuh):
//   pix = loc * T_l - 0.5            grid = 2*loc-1, align_corners=False
//                                    (pdvc/ops/functions/ms_deform_attn_func.py:49, :58-59;
//                                     pdvc/ops/src/cuda/ms_deform_im2col_cuda.cuh:286-287)
//   BORDER: gate = 0 < pix < T-1 ; pix clamped to [0, T-1]        (func.py:58-59)
//   ZEROS : taps outside [0, T-1] contribute 0                    (cuh:289, :56-79)
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/msda.h"

namespace msda {

constexpr int kMaxLevels = MSDA_MAX_LEVELS;
constexpr unsigned kFull = 0xffffffffu;

// Level table staged in shared memory once per CTA from the int64 device arrays.
struct LevelTable {
    int H[kMaxLevels];
    int W[kMaxLevels];
    int start[kMaxLevels];
};

// shapes_stride = 2 for (L,2) [H,W] spatial shapes, 1 for (L,) temporal shapes.
__device__ __forceinline__ void load_level_table(LevelTable &t, const int64_t *__restrict__ shapes,
                                                 const int64_t *__restrict__ lsi, int L,
                                                 int shapes_stride)
{
    for (int l = threadIdx.x; l < L; l += blockDim.x) {
        if (shapes_stride == 2) {
            t.H[l] = (int)shapes[2 * l];
            t.W[l] = (int)shapes[2 * l + 1];
        } else {
            t.H[l] = 1;
            t.W[l] = (int)shapes[l];
        }
        t.start[l] = (int)lsi[l];
    }
}

// One 1-D tap pair: rows i0 and i1 = i0+1 (already clamped into [0, T-1]) with their weights
// (0 for a tap that does not exist) and the d(pix)/d(loc) gate.
struct Tap1D {
    int i0, i1;
    float w0, w1;   // interpolation weights, zeroed for out-of-range taps
    float s0, s1;   // d(sample)/d(pix) coefficients of dot(G, V[i0]) and dot(G, V[i1]): -+gate
};

// `hi` + `lo` is the (unevaluated) pixel coordinate: hi carries x*T - 0.5 and lo the rounding
// residual of the product, so the interpolation fraction keeps ~fp32 precision even at T ~ 2000
// where a plain fp32 x*T-0.5 has lost 11 bits.
__device__ __forceinline__ Tap1D locate_1d(float hi, float lo, int T, int pad_mode)
{
    Tap1D t;
    const float pix = hi + lo;
    const float tmax = (float)(T - 1);
    if (pad_mode == MSDA_PAD_BORDER) {
        // borders themselves count as out of range for the gradient (ATen clip_coordinates_set_grad)
        const bool inside = (pix > 0.f) && (pix < tmax);
        float fl = floorf(pix);
        float w1 = (hi - fl) + lo;
        if (!(pix > 0.f)) { fl = 0.f; w1 = 0.f; }            // also catches NaN
        if (pix >= tmax) { fl = tmax; w1 = 0.f; }
        w1 = fminf(fmaxf(w1, 0.f), 1.f);
        const int i0 = (int)fl;
        t.i0 = i0;
        t.i1 = min(i0 + 1, T - 1);
        t.w1 = w1;
        t.w0 = 1.f - w1;
        t.s0 = inside ? -1.f : 0.f;
        t.s1 = inside ? 1.f : 0.f;
    } else {
        const bool valid = (pix > -1.f) && (pix < (float)T);
        const float fl = floorf(pix);
        float w1 = (hi - fl) + lo;
        w1 = fminf(fmaxf(w1, 0.f), 1.f);
        int i0 = valid ? (int)fl : 0;
        const bool ok0 = valid && (i0 >= 0);
        const bool ok1 = valid && (i0 + 1 <= T - 1);
        t.w0 = ok0 ? 1.f - w1 : 0.f;
        t.w1 = ok1 ? w1 : 0.f;
        t.s0 = ok0 ? -1.f : 0.f;
        t.s1 = ok1 ? 1.f : 0.f;
        t.i0 = max(i0, 0);
        t.i1 = min(i0 + 1, T - 1);
        t.i1 = max(t.i1, 0);
    }
    return t;
}

// pix = x*T - 0.5 split into (hi, lo)
__device__ __forceinline__ void pix_from_loc(float x, int T, float &hi, float &lo)
{
    const float tf = (float)T;
    const float p = x * tf;
    lo = fmaf(x, tf, -p);
    hi = p - 0.5f;
}

// ---- 16-byte vector access ---------------------------------------------------------------

template <typename T> struct Vec16;   // CH channels per 16-byte lane slot

template <> struct Vec16<float> {
    static constexpr int CH = 4;
    __device__ __forceinline__ static void load(const float *p, float (&v)[4])
    {
        const float4 r = __ldg(reinterpret_cast<const float4 *>(p));
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
    __device__ __forceinline__ static void store(float *p, const float (&v)[4])
    {
        *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};

template <> struct Vec16<__nv_bfloat16> {
    static constexpr int CH = 8;
    __device__ __forceinline__ static void load(const __nv_bfloat16 *p, float (&v)[8])
    {
        const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p));
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ __forceinline__ static void store(__nv_bfloat16 *p, const float (&v)[8])
    {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            w[i] = *reinterpret_cast<const uint32_t *>(&h);
        }
        *reinterpret_cast<uint4 *>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// fp32 vector reduction into global memory (native REDG.E.ADD.F32x4 on sm_90+).
__device__ __forceinline__ void red_add_v4(float *p, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
                 "f"(d)
                 : "memory");
}

// Sum `a[0..2W)` across the W lanes of a (power-of-two, lane-aligned) group so that lane j of the
// group ends with the totals of a[2j] and a[2j+1] in a[0], a[1].  2W-2 shuffles instead of
// 2W*log2(W).
template <int W> __device__ __forceinline__ void transpose_reduce(float (&a)[2 * W], int lane_in_group)
{
#pragma unroll
    for (int off = W / 2, n = W; off >= 1; off >>= 1, n >>= 1) {
        const bool upper = (lane_in_group & off) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const float send = upper ? a[i] : a[i + n];
            const float keep = upper ? a[i + n] : a[i];
            a[i] = keep + __shfl_xor_sync(kFull, send, off);
        }
    }
}

template <int W> __device__ __forceinline__ float group_max(float v)
{
#pragma unroll
    for (int off = W / 2; off >= 1; off >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, off));
    return v;
}

template <int W> __device__ __forceinline__ float group_sum(float v)
{
#pragma unroll
    for (int off = W / 2; off >= 1; off >>= 1) v += __shfl_xor_sync(kFull, v, off);
    return v;
}

}  // namespace msda
