// tools/ubench.cu -- on-chip bandwidth micro-benchmarks that size the MSDeformAttn kernel design on B200.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench tools/ubench.cu && tools/ubench
// Every test moves "rows" of 256 B (one head of D=64 fp32) per half-warp, the access shape of the
// 1-D deformable-attention gather / scatter.  Prints aggregate TB/s.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ unsigned hash32(unsigned x)
{
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

constexpr int kRows = 512;           // rows of 256 B resident per CTA (128 KB)
constexpr int kIters = 4096;

// mode 0: LDS.128 gather from shared memory, mode 1: LDG.128 gather (L1-resident window of global memory)
template <int MODE, int UNROLL>
__global__ void __launch_bounds__(1024) gather_kernel(const float4 *__restrict__ g, float4 *out, int spread)
{
    extern __shared__ float4 sm[];
    const float4 *base = g + (size_t)blockIdx.x * kRows * 16;
    if (MODE == 0) {
        for (int i = threadIdx.x; i < kRows * 16; i += blockDim.x) sm[i] = base[i];
        __syncthreads();
    }
    const int lane = threadIdx.x & 31, half = lane >> 4, col = lane & 15;
    const int warp = threadIdx.x >> 5;
    float4 acc = make_float4(0, 0, 0, 0);
    unsigned seed = blockIdx.x * 977 + warp * 131 + half * 7;
    for (int it = 0; it < kIters; it += UNROLL) {
        float4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            // rows wander inside a window of `spread` rows around a slowly moving centre (encoder-like)
            seed = seed * 1664525u + 1013904223u;   // one IMAD per index: keep the loop load-bound
            const int row = ((it >> 4) + ((seed >> 16) & (spread - 1))) & (kRows - 1);
            if (MODE == 0) v[u] = sm[row * 16 + col];
            else v[u] = __ldg(base + row * 16 + col);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
    }
    if (acc.x == 123.456f) out[threadIdx.x] = acc;
}

// mode 0: red.global.add.v4.f32 into an L2-resident region; mode 1: non-atomic LDS.128/FADD/STS.128 RMW in smem
template <int MODE>
__global__ void __launch_bounds__(1024) scatter_kernel(float *g, int region_rows, int spread)
{
    extern __shared__ float4 sm[];
    if (MODE == 1) {
        for (int i = threadIdx.x; i < kRows * 16; i += blockDim.x) sm[i] = make_float4(0, 0, 0, 0);
        __syncthreads();
    }
    const int lane = threadIdx.x & 31, half = lane >> 4, col = lane & 15;
    const int warp = threadIdx.x >> 5;
    unsigned seed = blockIdx.x * 977 + warp * 131 + half * 7;
    const int cta_base = (int)(((long long)blockIdx.x * region_rows) / gridDim.x);
    for (int it = 0; it < kIters; ++it) {
        seed = seed * 1664525u + 1013904223u;
        const unsigned h = seed >> 16;
        if (MODE == 0) {
            const int row = (cta_base + (it >> 4) + (h & (spread - 1))) & (region_rows - 1);
            float *p = g + ((size_t)row * 16 + col) * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
        } else {
            const int row = ((it >> 4) + (h & (spread - 1)) + warp * 64) & (kRows - 1);   // warps own disjoint-ish ranges
            float4 v = sm[row * 16 + col];
            v.x += 1.f; v.y += 2.f; v.z += 3.f; v.w += 4.f;
            sm[row * 16 + col] = v;
        }
    }
    if (MODE == 1) {
        __syncthreads();
        if (sm[threadIdx.x].x == 123.456f) g[threadIdx.x] = sm[threadIdx.x].y;
    }
}

template <typename F> float time_ms(F f, int reps = 5)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f(); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; ++i) f();
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    return ms / reps;
}

int main()
{
    int dev = 0, sms = 0, clk = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev));
    printf("SMs %d, max clock %d MHz\n", sms, clk / 1000);
    const int smem = kRows * 256;
    float4 *g; float4 *out; float *red;
    const int region_rows = 1 << 16;   // 16 MB: L2 resident
    for (int ctas_per_sm = 1; ctas_per_sm <= 1; ++ctas_per_sm) for (int threads : {256, 512, 1024}) {
        const int grid = sms * ctas_per_sm;
        CK(cudaMalloc(&g, (size_t)grid * smem));
        CK(cudaMemset(g, 0, (size_t)grid * smem));
        CK(cudaMalloc(&out, 4096 * 16));
        CK(cudaMalloc(&red, (size_t)region_rows * 256));
        CK(cudaMemset(red, 0, (size_t)region_rows * 256));
        CK(cudaFuncSetAttribute(gather_kernel<0, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CK(cudaFuncSetAttribute(gather_kernel<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CK(cudaFuncSetAttribute(scatter_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        const double bytes = (double)grid * (threads / 32) * kIters * 512.0;
        printf("-- %d threads per CTA, 1 CTA per SM\n", threads);
        for (int spread : {16, 64, 512}) {
            float ms;
            ms = time_ms([&] { gather_kernel<0, 8><<<grid, threads, smem>>>(g, out, spread); });
            printf("LDS.128 gather  unroll8 spread %3d : %7.2f TB/s  (%.1f B/clk/SM at max clock)\n", spread, bytes / ms / 1e9, bytes / ms / 1e3 / sms / (clk / 1e3) );
            ms = time_ms([&] { gather_kernel<0, 1><<<grid, threads, smem>>>(g, out, spread); });
            printf("LDS.128 gather  unroll1 spread %3d : %7.2f TB/s\n", spread, bytes / ms / 1e9);
            ms = time_ms([&] { gather_kernel<1, 8><<<grid, threads, 0>>>(g, out, spread); });
            printf("LDG.128 gather  unroll8 spread %3d : %7.2f TB/s  (%.1f B/clk/SM at max clock)\n", spread, bytes / ms / 1e9, bytes / ms / 1e3 / sms / (clk / 1e3));
            ms = time_ms([&] { gather_kernel<1, 1><<<grid, threads, 0>>>(g, out, spread); });
            printf("LDG.128 gather  unroll1 spread %3d : %7.2f TB/s\n", spread, bytes / ms / 1e9);
            ms = time_ms([&] { scatter_kernel<1><<<grid, threads, smem>>>(red, region_rows, spread); });
            printf("smem RMW (LDS+STS.128)  spread %3d : %7.2f TB/s of row updates\n", spread, bytes / ms / 1e9);
            ms = time_ms([&] { scatter_kernel<0><<<grid, threads, 0>>>(red, region_rows, spread); });
            printf("red.global.add.v4.f32   spread %3d : %7.2f TB/s of row updates\n", spread, bytes / ms / 1e9);
            ms = time_ms([&] { scatter_kernel<0><<<grid * 4, threads, 0>>>(red, region_rows, spread); });
            printf("red.global.add.v4.f32 4 CTA/SM  %3d : %7.2f TB/s of row updates\n", spread, 4 * bytes / ms / 1e9);
        }
        CK(cudaFree(g)); CK(cudaFree(out)); CK(cudaFree(red));
    }
    return 0;
}
