// tools/ubench_tmem.cu -- can Tensor Memory serve as the gather cache of the deformable-attention forward?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_tmem tools/ubench_tmem.cu && tools/ubench_tmem
// Layout under test: lane = channel d, column = value row r, so that ONE tcgen05.ld.32x32b.x2 at column r
// hands lane d the two interpolation taps V[r][d], V[r+1][d] of a sample (32 channels per warp).
// Measures tcgen05.ld / tcgen05.st throughput per SM with warp-uniform, data-dependent column addresses.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int kIters = 8192;

__device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tmem_ld_x2(uint32_t taddr, float &a, float &b)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=f"(a), "=f"(b) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_x1(uint32_t taddr, float &a)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=f"(a) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_x4(uint32_t taddr, float &a, float &b, float &c, float &d)
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_st_x1(uint32_t taddr, float a)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr), "f"(a) : "memory");
}
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const float (&v)[8])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "f"(v[0]), "f"(v[1]),
                 "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// MODE 0: ld.x2 gather; 1: ld.x1 gather; 2: ld.x4 gather; 3: st.x8 stream; 4: correctness of the (lane, column) layout
template <int MODE, int UNROLL>
__global__ void __launch_bounds__(1024) tmem_kernel(float *out, int spread)
{
    __shared__ uint32_t base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&base_s, 512);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = base_s;
    const uint32_t quad = (uint32_t)(warp & 3) * 32u;           // a warp may only touch lanes 32*(warp%4) .. +31
    const uint32_t wbase = base + (quad << 16);
    // fill: column c of lane l holds 1000*l + c  (each quadrant written by warps 0..3)
    if (warp < 4) {
        for (int c = 0; c < 512; c += 8) {
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = 1000.f * (float)(quad + lane) + (float)(c + k);
            tmem_st_x8(wbase + c, v);
        }
        tmem_wait_st();
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    float acc = 0.f;
    if (MODE == 4) {
        float a, b;
        const int col = 17 + warp;
        tmem_ld_x2(wbase + col, a, b);
        tmem_wait_ld();
        const float ea = 1000.f * (float)(quad + lane) + (float)col;
        if (a != ea || b != ea + 1.f) atomicAdd(out, 1.f);   // counts mismatching lanes
    } else if (MODE == 3) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = (float)(lane + k);
        for (int it = 0; it < kIters; ++it) {
            tmem_st_x8(wbase + ((it * 8) & 511), v);
        }
        tmem_wait_st();
    } else {
        unsigned seed = blockIdx.x * 977 + warp * 131;
        for (int it = 0; it < kIters; it += UNROLL) {
            float a[UNROLL], b[UNROLL], c[UNROLL], d[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                seed = seed * 1664525u + 1013904223u;   // warp-uniform, data-dependent column
                const uint32_t col = ((it >> 4) + ((seed >> 16) & (spread - 1))) & 507u;
                if (MODE == 0) tmem_ld_x2(wbase + col, a[u], b[u]);
                if (MODE == 1) tmem_ld_x1(wbase + col, a[u]);
                if (MODE == 2) tmem_ld_x4(wbase + col, a[u], b[u], c[u], d[u]);
            }
            tmem_wait_ld();
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                acc += a[u];
                if (MODE != 1) acc += b[u];
                if (MODE == 2) acc += c[u] + d[u];
            }
        }
    }
    if (acc == 123.456f) out[threadIdx.x + 1] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) tmem_dealloc(base, 512);
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
    int sms = 0, clk = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0));
    printf("SMs %d, max clock %d MHz\n", sms, clk / 1000);
    float *out;
    CK(cudaMalloc(&out, 8192));
    CK(cudaMemset(out, 0, 8192));
    tmem_kernel<4, 1><<<sms, 512>>>(out, 16);
    CK(cudaDeviceSynchronize());
    float bad = -1;
    CK(cudaMemcpy(&bad, out, 4, cudaMemcpyDeviceToHost));
    printf("layout check (lane = channel, column = row; ld.x2 returns columns c, c+1): %g mismatching lanes\n", bad);
    for (int threads : {128, 256, 512, 1024}) {
        const double insts = (double)sms * (threads / 32) * kIters;
        printf("-- %d threads per CTA, 1 CTA per SM\n", threads);
        for (int spread : {16, 256}) {
            float ms;
            ms = time_ms([&] { tmem_kernel<1, 8><<<sms, threads>>>(out, spread); });
            printf("tcgen05.ld.32x32b.x1 unroll8 spread %3d: %6.2f TB/s  %.3f inst/clk/SM  %.1f B/clk/SM\n", spread, insts * 128 / ms / 1e9, insts / sms / (ms * 1e-3 * clk * 1e3), insts * 128 / sms / (ms * 1e-3 * clk * 1e3));
            ms = time_ms([&] { tmem_kernel<0, 8><<<sms, threads>>>(out, spread); });
            printf("tcgen05.ld.32x32b.x2 unroll8 spread %3d: %6.2f TB/s  %.3f inst/clk/SM  %.1f B/clk/SM\n", spread, insts * 256 / ms / 1e9, insts / sms / (ms * 1e-3 * clk * 1e3), insts * 256 / sms / (ms * 1e-3 * clk * 1e3));
            ms = time_ms([&] { tmem_kernel<0, 2><<<sms, threads>>>(out, spread); });
            printf("tcgen05.ld.32x32b.x2 unroll2 spread %3d: %6.2f TB/s  %.3f inst/clk/SM\n", spread, insts * 256 / ms / 1e9, insts / sms / (ms * 1e-3 * clk * 1e3));
            ms = time_ms([&] { tmem_kernel<2, 8><<<sms, threads>>>(out, spread); });
            printf("tcgen05.ld.32x32b.x4 unroll8 spread %3d: %6.2f TB/s  %.3f inst/clk/SM  %.1f B/clk/SM\n", spread, insts * 512 / ms / 1e9, insts / sms / (ms * 1e-3 * clk * 1e3), insts * 512 / sms / (ms * 1e-3 * clk * 1e3));
        }
        float ms = time_ms([&] { tmem_kernel<3, 1><<<sms, threads>>>(out, 16); });
        printf("tcgen05.st.32x32b.x8 stream            : %6.2f TB/s  %.3f inst/clk/SM  %.1f B/clk/SM\n", insts * 1024 / ms / 1e9, insts / sms / (ms * 1e-3 * clk * 1e3), insts * 1024 / sms / (ms * 1e-3 * clk * 1e3));
    }
    return 0;
}
