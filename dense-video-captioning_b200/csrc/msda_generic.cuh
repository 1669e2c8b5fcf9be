// msda_generic.cuh -- shape-generic kernels: any head dim D (the reference's gradcheck list
// 30, 71, 1025, ... pdvc/ops/test.py:85-86), true 2-D levels (H_l > 1), f32 / f64 / bf16.
// One thread per (n, q, m, d) output channel.  These are the correctness fall-backs ON THE GPU for
// shapes the vectorised 1-D kernels (msda_fast.cuh) do not take; they are not a CPU path.
//
// Covers the same reference functions as msda_fast.cuh (ms_deform_im2col_cuda.cuh:238-300 forward,
// :302-921 backward, helpers :34-235) plus the return_value=True sampler branch of
// ms_deform_attn_core_pytorch (pdvc/ops/functions/ms_deform_attn_func.py:64-65).
#pragma once
#include "msda_common.cuh"

namespace msda {

template <typename T> struct Compute { using type = float; };
template <> struct Compute<double> { using type = double; };

template <typename T> __device__ __forceinline__ typename Compute<T>::type to_ct(T v) { return v; }
template <> __device__ __forceinline__ float to_ct<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T, typename CT> __device__ __forceinline__ T from_ct(CT v) { return (T)v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_ct<__nv_bfloat16, float>(float v) { return __float2bfloat16_rn(v); }

// Index-side types: locations / weights are CT (float for f32 and bf16 values, double for f64).
template <typename CT> struct AxisTap {
    int i0;          // first tap (clamped into range)
    int i1;          // second tap (clamped)
    CT w0, w1;       // weights, 0 for non-existent taps
    CT s0, s1;       // d(weight)/d(pix) * gate: -1/+1 or 0
};

template <typename CT> __device__ __forceinline__ AxisTap<CT> locate_axis(CT loc, int size, int pad_mode)
{
    AxisTap<CT> t;
    // ((2*loc-1 + 1) * size - 1) / 2, written as loc*size - 0.5
    const CT pix = loc * (CT)size - (CT)0.5;
    const CT smax = (CT)(size - 1);
    if (pad_mode == MSDA_PAD_BORDER) {
        const bool inside = (pix > (CT)0) && (pix < smax);
        CT fl = floor(pix);
        CT w1 = pix - fl;
        if (!(pix > (CT)0)) { fl = (CT)0; w1 = (CT)0; }
        if (pix >= smax) { fl = smax; w1 = (CT)0; }
        t.i0 = (int)fl;
        t.i1 = min(t.i0 + 1, size - 1);
        t.w1 = w1;
        t.w0 = (CT)1 - w1;
        t.s0 = inside ? (CT)-1 : (CT)0;
        t.s1 = inside ? (CT)1 : (CT)0;
    } else {
        const bool valid = (pix > (CT)-1) && (pix < (CT)size);
        const CT fl = floor(pix);
        const CT w1 = pix - fl;
        const int i0 = valid ? (int)fl : 0;
        const bool ok0 = valid && i0 >= 0;
        const bool ok1 = valid && (i0 + 1 <= size - 1);
        t.w0 = ok0 ? (CT)1 - w1 : (CT)0;
        t.w1 = ok1 ? w1 : (CT)0;
        t.s0 = ok0 ? (CT)-1 : (CT)0;
        t.s1 = ok1 ? (CT)1 : (CT)0;
        t.i0 = max(i0, 0);
        t.i1 = max(min(i0 + 1, size - 1), 0);
    }
    return t;
}

struct GenericArgs {
    const void *value;
    const int64_t *shapes;   // (L,2) or (L,) per shapes_stride
    const int64_t *lsi;
    int shapes_stride;
    const void *loc;         // (N,Lq,M,L,P,2)  CT
    const void *attn;        // (N,Lq,M,L,P)    CT   (NULL for the sampler)
    void *out;               // (N,Lq,M*D) VT   or samples (N*M,D,Lq,L,P) VT
    const void *grad_out;    // VT, same layout as out
    void *gv_acc;            // accumulation image of grad_value: CT-typed (float for bf16), pre-zeroed
    void *grad_loc;          // CT, pre-zeroed
    void *grad_attn;         // CT, pre-zeroed (NULL for the sampler)
    int N, S, M, D, L, Lq, P;
    int pad_mode;
    int assume_1d;           // ignore loc[...,1]; treat every level as H = 1
    long long total;         // N*Lq*M*D
};

template <typename VT, bool SAMPLER>
__global__ void __launch_bounds__(256) msda_fwd_generic_kernel(const GenericArgs a)
{
    using CT = typename Compute<VT>::type;
    __shared__ LevelTable lt;
    load_level_table(lt, a.shapes, a.lsi, a.L, a.shapes_stride);
    __syncthreads();
    const VT *value = static_cast<const VT *>(a.value);
    const CT *loc = static_cast<const CT *>(a.loc);
    const CT *attn = static_cast<const CT *>(a.attn);
    VT *out = static_cast<VT *>(a.out);
    const int LP = a.L * a.P;
    const long long MD = (long long)a.M * a.D;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < a.total; idx += stride) {
        const int d = (int)(idx % a.D);
        const long long u = idx / a.D;
        const int m = (int)(u % a.M);
        const long long nq = u / a.M;
        const int q = (int)(nq % a.Lq);
        const long long n = nq / a.Lq;
        const VT *vb = value + n * a.S * MD + (long long)m * a.D + d;
        CT acc = (CT)0;
        for (int l = 0; l < a.L; ++l) {
            const int H = a.assume_1d ? 1 : lt.H[l], W = lt.W[l];
            const VT *vl = vb + (long long)lt.start[l] * MD;
            for (int p = 0; p < a.P; ++p) {
                const long long sidx = u * LP + l * a.P + p;
                const AxisTap<CT> tx = locate_axis<CT>(loc[2 * sidx], W, a.pad_mode);
                AxisTap<CT> ty;
                if (a.assume_1d) { ty.i0 = ty.i1 = 0; ty.w0 = (CT)1; ty.w1 = (CT)0; ty.s0 = a.pad_mode == MSDA_PAD_ZEROS ? (CT)-1 : (CT)0; ty.s1 = (CT)0; }
                else ty = locate_axis<CT>(loc[2 * sidx + 1], H, a.pad_mode);
                CT sv = (CT)0;
                if (ty.w0 != (CT)0) {
                    const VT *row = vl + (long long)ty.i0 * W * MD;
                    if (tx.w0 != (CT)0) sv += ty.w0 * tx.w0 * (CT)to_ct(row[(long long)tx.i0 * MD]);
                    if (tx.w1 != (CT)0) sv += ty.w0 * tx.w1 * (CT)to_ct(row[(long long)tx.i1 * MD]);
                }
                if (ty.w1 != (CT)0) {
                    const VT *row = vl + (long long)ty.i1 * W * MD;
                    if (tx.w0 != (CT)0) sv += ty.w1 * tx.w0 * (CT)to_ct(row[(long long)tx.i0 * MD]);
                    if (tx.w1 != (CT)0) sv += ty.w1 * tx.w1 * (CT)to_ct(row[(long long)tx.i1 * MD]);
                }
                if (SAMPLER) {
                    // (N*M, D, Lq, L, P)
                    out[(((n * a.M + m) * a.D + d) * a.Lq + q) * LP + l * a.P + p] = from_ct<VT, CT>(sv);
                } else {
                    acc += attn[sidx] * sv;
                }
            }
        }
        if (!SAMPLER) out[idx] = from_ct<VT, CT>(acc);
    }
}

template <typename CT> __device__ __forceinline__ CT warp_sum(CT v)
{
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(kFull, v, off);
    return v;
}

// Backward: grad_value taps by atomicAdd into the (pre-zeroed) accumulation image; the per-sample
// reductions over D by warp shuffle when a warp lies inside one unit (D % 32 == 0), else by
// per-thread atomicAdd into the (pre-zeroed) outputs.
template <typename VT, bool SAMPLER>
__global__ void __launch_bounds__(256) msda_bwd_generic_kernel(const GenericArgs a)
{
    using CT = typename Compute<VT>::type;
    __shared__ LevelTable lt;
    load_level_table(lt, a.shapes, a.lsi, a.L, a.shapes_stride);
    __syncthreads();
    const VT *value = static_cast<const VT *>(a.value);
    const CT *loc = static_cast<const CT *>(a.loc);
    const CT *attn = static_cast<const CT *>(a.attn);
    const VT *grad_out = static_cast<const VT *>(a.grad_out);
    CT *gv = static_cast<CT *>(a.gv_acc);
    CT *grad_loc = static_cast<CT *>(a.grad_loc);
    CT *grad_attn = static_cast<CT *>(a.grad_attn);
    const int LP = a.L * a.P;
    const long long MD = (long long)a.M * a.D;
    // D % 32 == 0  =>  every warp lies inside one unit (and total % 32 == 0): shuffle-reduce over d
    const bool warp_uniform_unit = (a.D % 32) == 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < a.total; idx += stride) {
        const int d = (int)(idx % a.D);
        const long long u = idx / a.D;
        const int m = (int)(u % a.M);
        const long long nq = u / a.M;
        const int q = (int)(nq % a.Lq);
        const long long n = nq / a.Lq;
        const long long vofs = n * a.S * MD + (long long)m * a.D + d;
        const CT g_out = SAMPLER ? (CT)0 : (CT)to_ct(grad_out[idx]);
        for (int l = 0; l < a.L; ++l) {
            const int H = a.assume_1d ? 1 : lt.H[l], W = lt.W[l];
            const long long lofs = vofs + (long long)lt.start[l] * MD;
            for (int p = 0; p < a.P; ++p) {
                const long long sidx = u * LP + l * a.P + p;
                const AxisTap<CT> tx = locate_axis<CT>(loc[2 * sidx], W, a.pad_mode);
                AxisTap<CT> ty;
                if (a.assume_1d) { ty.i0 = ty.i1 = 0; ty.w0 = (CT)1; ty.w1 = (CT)0; ty.s0 = a.pad_mode == MSDA_PAD_ZEROS ? (CT)-1 : (CT)0; ty.s1 = (CT)0; }
                else ty = locate_axis<CT>(loc[2 * sidx + 1], H, a.pad_mode);
                CT gs;   // gradient reaching this sample's interpolated value, for channel d
                if (SAMPLER) gs = (CT)to_ct(grad_out[(((n * a.M + m) * a.D + d) * a.Lq + q) * LP + l * a.P + p]);
                else gs = attn[sidx] * g_out;
                CT sv = (CT)0, gix = (CT)0, giy = (CT)0;
                const int ys[2] = {ty.i0, ty.i1};
                const int xs[2] = {tx.i0, tx.i1};
                const CT wy[2] = {ty.w0, ty.w1}, wx[2] = {tx.w0, tx.w1};
                const CT sy[2] = {ty.s0, ty.s1}, sx[2] = {tx.s0, tx.s1};
#pragma unroll
                for (int cy = 0; cy < 2; ++cy)
#pragma unroll
                    for (int cx = 0; cx < 2; ++cx) {
                        // tap indices are clamped into range, so a tap that does not exist simply
                        // carries zero weight and zero slope
                        const CT w = wy[cy] * wx[cx];
                        const CT dx = sx[cx] * wy[cy];
                        const CT dy = sy[cy] * wx[cx];
                        if (w == (CT)0 && dx == (CT)0 && dy == (CT)0) continue;
                        const long long vi = lofs + ((long long)ys[cy] * W + xs[cx]) * MD;
                        const CT v = (CT)to_ct(value[vi]);
                        sv += w * v;
                        gix += dx * v;
                        giy += dy * v;
                        if (w != (CT)0) atomicAdd(gv + vi, w * gs);
                    }
                CT ga = g_out * sv;
                CT glx = (CT)W * gix * gs;
                CT gly = (CT)H * giy * gs;
                if (warp_uniform_unit) {
                    glx = warp_sum(glx);
                    gly = warp_sum(gly);
                    if (!SAMPLER) ga = warp_sum(ga);
                    if ((threadIdx.x & 31) == 0) {
                        atomicAdd(grad_loc + 2 * sidx, glx);
                        atomicAdd(grad_loc + 2 * sidx + 1, gly);
                        if (!SAMPLER) atomicAdd(grad_attn + sidx, ga);
                    }
                } else {
                    atomicAdd(grad_loc + 2 * sidx, glx);
                    atomicAdd(grad_loc + 2 * sidx + 1, gly);
                    if (!SAMPLER) atomicAdd(grad_attn + sidx, ga);
                }
            }
        }
    }
}

// accumulation image -> VT (identity copy is skipped by the launcher when CT == VT)
__global__ void msda_cvt_f32_to_bf16_scalar_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst,
                                                   long long n)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        dst[i] = __float2bfloat16_rn(src[i]);
}

}  // namespace msda
