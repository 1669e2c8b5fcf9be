// msda_proj.cuh -- the Linear projections of MSDeformAttn on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// The four nn.Linear layers around the deformable-attention core (pdvc/ops/modules/ms_deform_attn.py:95-100, :125):
//     value_proj (+ bias, rows of padded frames zeroed: :96-98), sampling_offsets || attention_weights (ONE GEMM from the
//     query, written as the two raw tensors the fused MSDA kernel reads: :99-100), output_proj (+ bias: :125),
// with their data gradients, weight gradients and bias gradients.  All three GEMMs of a Linear layer are one kernel,
//     D[m, n] = sum_k A(m, k) * B(n, k)            (accumulator in tensor memory, fp32)
// instantiated for the three operand arrangements (K = reduction axis of the GEMM, "K-major" = reduction axis contiguous):
//     FWD    y  = x  W^T + b      A = x  (rows x Cin,  K-major)   B = W  (Cout x Cin, K-major)
//     DGRAD  dx = dy W            A = dy (rows x Cout, K-major)   B = W  (Cout x Cin, MN-major: reduction over Cout)
//     WGRAD  dW = dy^T x, db      A = dy (rows x Cout, MN-major)  B = x  (rows x Cin, MN-major: reduction over rows)
// so no operand is ever transposed in memory: the UMMA shared-memory descriptors take both majors.
//
// Arithmetic: tcgen05.mma.kind::f16 (bf16 x bf16 -> fp32 in TMEM) for both data types.  bf16 data is multiplied as it
// is.  fp32 data is split while it is staged into THREE bf16 terms, x = hi + mid + lo (each the bf16 rounding of what the
// previous ones left: 3 x 8 mantissa bits cover fp32's 24), and the tile product is the six largest cross terms
//     lo*hi + hi*lo + mid*mid + mid*hi + hi*mid + hi*hi            (dropped: mid*lo, lo*mid, lo*lo <= 2^-24 relative)
// accumulated in fp32 -- fp32-level results (plain TF32, 2^-11 per product, would be below the reference's fp32 matmul;
// tests/test_proj_gpu.py states the achieved error against float64).  kind::tf32 was tried first (hi/lo split, three
// products): it multiplies K-major operands correctly but returns zeros for MN-major ones in the no-swizzle layout, and
// the gradients need MN-major operands; bf16 takes both majors and runs at twice the TF32 rate, so the six bf16 products
// cost what three TF32 products would.
//
// Structure of one CTA (256 threads, one 128 x BN output tile, one slice of the reduction axis):
//   fp32 data, and the weight gradient (msda_proj_kernel):
//     all threads global -> registers (16-byte loads, a ring of 2-4 reduction blocks in flight) -> [bf16 x 3 split] -> shared
//                 memory in the 128-byte-swizzled UMMA layout (K-major or MN-major tiles, 64 reduction elements per block)
//     thread 0    fence.proxy.async + tcgen05.mma for the staged block (four TMEM accumulators used round-robin),
//                 tcgen05.commit -> the stage's mbarrier
//   bf16 forward / data gradient (msda_proj_tma_kernel):
//     warp 0 lane 0  producer: cp.async.bulk.tensor.2d boxes of the row-major matrices straight into the swizzled tiles (the
//                 TMA swizzle IS the UMMA layout), full / empty mbarriers, four stages
//     warp 1 lane 0  issues the MMAs and commits
//   all threads   epilogue: tcgen05.ld the accumulator rows (warp w reads TMEM lanes 32*(w%4)...), sum of the four
//                 accumulators, cluster split-K reduce-scatter through distributed shared memory (the CTAs of a cluster own
//                 consecutive slices of the reduction axis and each reduces its share of the tile), bias / padding mask /
//                 bias gradient, coalesced 16-byte stores.
// The GEMMs here are small (100-375 rows, 256-512 columns, reduction 100-512): they are latency-bound, which is why
// the reduction axis is split over a cluster -- more SMs stream the operands -- rather than tiles being made larger.
#pragma once
#include <cuda.h>          // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/msda.h"

namespace msda {
namespace proj {

constexpr int kThreads = 256;
constexpr int BM = 128;    // UMMA M (TMEM lanes)
constexpr int BN = 64;     // UMMA N (TMEM columns of the accumulator)
constexpr int kStages = 2;
constexpr int kMaxSplit = 4;   // CTAs per cluster along the reduction axis
constexpr int kBK = 64;        // reduction elements per staged block: one 128-byte swizzle row of bf16
constexpr int kAcc = 4;        // independent accumulators (TMEM column blocks of BN): consecutive MMAs go round-robin, so that
                               // an MMA never waits for the accumulate of the previous one (a dependent chain runs at the
                               // MMA latency, ~110 clk per 128x64x16 instruction measured, not at its throughput)

// Development aid (-DMSDA_PROJ_TIMING): thread 0 of CTA (0,0,0) stamps clock64() into Args::debug
#ifdef MSDA_PROJ_TIMING
#define MSDA_PROJ_STAMP(I) if (a.debug && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) a.debug[I] = clock64();
#else
#define MSDA_PROJ_STAMP(I)
#endif
enum Mode { FWD = 0, DGRAD = 1, WGRAD = 2 };

struct Args {
    // FWD  : x (rows,K)  w[s] (Ns,K)  bias[s] (Ns)  mask (rows)  -> y[s] (rows,Ns)
    // DGRAD: dy[s] (rows,Ns)  w[s] (Ns,K)  mask  -> dx (rows,K)
    // WGRAD: dy[s] (rows,Ns)  x (rows,K)  mask  -> dw[s] (Ns,K) fp32  db[s] (Ns) fp32
    const void *x;
    const void *w[2];
    const void *bias[2];      // fp32 (FWD)
    const void *dy[2];
    const unsigned char *mask;   // one byte per row, non-zero = padded frame: y[0] row := 0 (FWD), dy[0] row := 0 (DGRAD, WGRAD)
    void *y[2];
    void *dx;
    void *dw[2];
    float *db[2];
    int rows, K, N0, N1;      // N1 == 0: one segment
    int ksplit;               // cluster size along the reduction axis
    unsigned long long *debug;   // development builds only
};

// ---- PTX wrappers -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t done = 0;
    const uint32_t addr = smem_u32(bar);
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(addr), "r"(parity)
                     : "memory");
    }
}
__device__ __forceinline__ void tmem_alloc(uint32_t *slot, uint32_t cols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
                     "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
                     : "memory");
}
// all tcgen05.mma issued so far by this thread -> one arrival on `bar` when they have completed
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 lanes x 16 consecutive 32-bit columns: thread t of the warp receives row (lane base + t)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16])
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `p` (own shared memory) in the shared memory of CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(const void *p, uint32_t rank)
{
    uint32_t out;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(smem_u32(p)), "r"(rank));
    return out;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d)
{
    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ---- shared-memory operand tiles in the canonical 128-byte-swizzled UMMA layouts ------------------------------------
// (the no-swizzle "interleaved" layouts multiply correctly but the tensor core fetches them at a fraction of the rate:
//  158-285 clk per 128x64x16 MMA measured, against ~48 clk of shared-memory bandwidth; profiles/r2_notes.md)
// A swizzle atom is 8 rows of 128 bytes; the 16-byte chunk c of row r is stored at chunk position c ^ (r % 8).
//   K-major  tile (R rows x 64 reduction elements = 128 bytes per row): chunk (row r, K-chunk c) at
//            (r/8)*1024 + (r%8)*128 + ((c ^ r%8) * 16); descriptor SBO = 1024 (next 8 rows); one MMA consumes 32 bytes of a
//            row: the start address advances by 32 per K step.
//   MN-major tile (R MN-elements x 64 reduction elements): the 128-byte rows run along MN (64 elements), 8 consecutive
//            reduction indices form an atom; chunk (reduction index kk, MN-chunk c) at
//            (c/8)*LBO + (kk/8)*1024 + (kk%8)*128 + ((c%8 ^ kk%8) * 16) with SBO = 1024 (next 8 reduction indices) and
//            LBO = 8192 (next 64 MN-elements); one MMA consumes 16 reduction indices = 2048 bytes.
// Tiles always hold bf16 (fp32 sources are split into three bf16 tiles): a chunk is 8 elements; BK = 64.
template <bool MN_MAJOR, int R, int BK> struct Tile {
    static_assert(BK == 64, "one 128-byte swizzle row of bf16 per K-major tile row");
    static constexpr int EPC = 8;                                     // bf16 elements per 16-byte chunk
    static constexpr int NCK = BK / EPC;                              // K-chunks per row (K-major)
    static constexpr int NCM = R / EPC;                               // MN-chunks (MN-major)
    static constexpr int LBO = MN_MAJOR ? 8192 : 16;
    static constexpr int SBO = 1024;
    static constexpr int BYTES = R * BK * 2;
    static constexpr int CHUNKS = R * BK / EPC;
    static constexpr int PER_THREAD = CHUNKS / kThreads;
    static constexpr int MMA_K = 16;                                  // reduction elements per bf16 MMA
    static constexpr int KSTEP_BYTES = MN_MAJOR ? 2048 : 32;
    static_assert(CHUNKS % kThreads == 0, "tile chunks must divide among the threads");

    // chunk i of the tile: MN index / reduction index of its first element and its byte offset in shared memory
    __device__ __forceinline__ static void locate(int i, int &mn, int &kk, int &off)
    {
        if (MN_MAJOR) {
            const int c = i % NCM;
            kk = i / NCM;
            mn = c * EPC;
            off = (c >> 3) * LBO + (kk >> 3) * 1024 + (kk & 7) * 128 + (((c & 7) ^ (kk & 7)) << 4);
        } else {
            const int c = i % NCK, r = i / NCK;
            mn = r;
            kk = c * EPC;
            off = (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4);
        }
    }
    __device__ __forceinline__ static uint64_t descriptor(uint32_t smem_addr)
    {
        // start address [0,14), LBO [16,30), SBO [32,46) (all >> 4); version 1 at [46,48); layout type 2 = SWIZZLE_128B at [61,64)
        return (uint64_t)((smem_addr & 0x3ffff) >> 4) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(SBO >> 4) << 32) | (1ull << 46) |
               (2ull << 61);
    }
};

template <typename T> struct Elem;
template <> struct Elem<float> {
    static constexpr bool SPLIT = true;
    static constexpr int COPIES = 3;   // hi, mid and lo tiles
    static constexpr int LOADS = 2;    // 16-byte global loads per 8-element chunk
    static constexpr int PREFETCH = 2; // reduction blocks whose global loads are in flight (register ring: 12 x 16 bytes each)
};
template <> struct Elem<__nv_bfloat16> {
    static constexpr bool SPLIT = false;
    static constexpr int COPIES = 1;
    static constexpr int LOADS = 1;
    static constexpr int PREFETCH = 4;
};
constexpr uint32_t kFormatBF16 = 1;

template <typename T, int MODE> struct OutType { using type = T; };
template <typename T> struct OutType<T, WGRAD> { using type = float; };

template <typename T, int MODE, int BK> struct Config {
    static constexpr bool A_MN = MODE == WGRAD;
    static constexpr bool B_MN = MODE != FWD;
    using TA = Tile<A_MN, BM, BK>;
    using TB = Tile<B_MN, BN, BK>;
    static constexpr int STAGE_BYTES = (TA::BYTES + TB::BYTES) * Elem<T>::COPIES;
    // behind the stages: the epilogue / split-K receive buffer [BM][BN + 4] floats + bias-gradient partials [kMaxSplit][BM]
    static constexpr int EP_OFFSET = kStages * STAGE_BYTES;
    static constexpr int SMEM_BYTES = EP_OFFSET + (BM * (BN + 4) + kMaxSplit * BM) * 4;
    // instruction descriptor: D fp32, A/B format, majors, N >> 3 at [17,23), M >> 4 at [24,29)
    static constexpr uint32_t IDESC = (1u << 4) | (kFormatBF16 << 7) | (kFormatBF16 << 10) | ((A_MN ? 1u : 0u) << 15) |
                                      ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
};

__device__ __forceinline__ uint4 zero4() { return make_uint4(0, 0, 0, 0); }

// x = hi + mid + lo, each term the bf16 rounding of what the previous ones left; two fp32 values -> three packed pairs
__device__ __forceinline__ void split_bf16x3(float x0, float x1, uint32_t &hi, uint32_t &mid, uint32_t &lo)
{
    const __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
    const float2 hf = __bfloat1622float2(h);
    const float r0 = x0 - hf.x, r1 = x1 - hf.y;
    const __nv_bfloat162 m = __floats2bfloat162_rn(r0, r1);
    const float2 mf = __bfloat1622float2(m);
    const __nv_bfloat162 l = __floats2bfloat162_rn(r0 - mf.x, r1 - mf.y);
    hi = *reinterpret_cast<const uint32_t *>(&h);
    mid = *reinterpret_cast<const uint32_t *>(&m);
    lo = *reinterpret_cast<const uint32_t *>(&l);
}

// One operand view: element (mn, kk) of the GEMM operand lives at base[mn * s_mn + kk * s_k] (one of the strides is 1);
// mn < lim_mn and kk < lim_k are in range, everything else reads as zero.  `mask` zeroes whole MN rows (mask_axis 1) or
// whole reduction indices (mask_axis 2).
template <typename T> struct View {
    const T *base;
    long s_mn, s_k;
    int lim_mn, lim_k;
    const unsigned char *mask;
    int mask_axis;
};

// registers of one thread's chunks of a tile: 8 source elements per chunk (fp32: two 16-byte loads, bf16: one)
template <typename T, typename TILE> struct Regs {
    uint4 v[TILE::PER_THREAD][Elem<T>::LOADS];
};

template <typename T, typename TILE>
__device__ __forceinline__ void load_tile(const View<T> &v, int mn0, int k0, Regs<T, TILE> &regs)
{
#pragma unroll
    for (int j = 0; j < TILE::PER_THREAD; ++j) {
        int mn, kk, off;
        TILE::locate(threadIdx.x + j * kThreads, mn, kk, off);
        mn += mn0;
        kk += k0;
        bool ok = mn < v.lim_mn && kk < v.lim_k;      // chunks never straddle a limit (limits are multiples of 8)
        if (ok && v.mask_axis == 1) ok = v.mask[mn] == 0;
        if (ok && v.mask_axis == 2) ok = v.mask[kk] == 0;
        const uint4 *src = reinterpret_cast<const uint4 *>(v.base + (long)mn * v.s_mn + (long)kk * v.s_k);
#pragma unroll
        for (int h = 0; h < Elem<T>::LOADS; ++h) regs.v[j][h] = ok ? __ldg(src + h) : zero4();
    }
}

template <typename T, typename TILE>
__device__ __forceinline__ void store_tile(unsigned char *tile, const Regs<T, TILE> &regs)
{
#pragma unroll
    for (int j = 0; j < TILE::PER_THREAD; ++j) {
        int mn, kk, off;
        TILE::locate(threadIdx.x + j * kThreads, mn, kk, off);
        if (Elem<T>::SPLIT) {
            const uint4 &lo4 = regs.v[j][0], &hi4 = regs.v[j][Elem<T>::LOADS - 1];
            const uint32_t w[8] = {lo4.x, lo4.y, lo4.z, lo4.w, hi4.x, hi4.y, hi4.z, hi4.w};
            uint32_t h[4], m[4], l[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) split_bf16x3(__uint_as_float(w[2 * p]), __uint_as_float(w[2 * p + 1]), h[p], m[p], l[p]);
            *reinterpret_cast<uint4 *>(tile + off) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4 *>(tile + TILE::BYTES + off) = make_uint4(m[0], m[1], m[2], m[3]);
            *reinterpret_cast<uint4 *>(tile + 2 * TILE::BYTES + off) = make_uint4(l[0], l[1], l[2], l[3]);
        } else {
            *reinterpret_cast<uint4 *>(tile + off) = regs.v[j][0];
        }
    }
}

// ---- epilogue shared by the register-staged and the TMA-fed kernels ----------------------------------------------------
// TMEM accumulators -> registers, split-K reduce-scatter over the cluster, bias / padding mask, coalesced stores, bias gradient.
template <typename T, int MODE, int BK>
__device__ __forceinline__ void proj_epilogue(const Args &a, unsigned char *smem, int ep_offset, uint64_t &done_bar, float *colsum,
                                              uint32_t tmem_d, int n_mma, int nkb, uint32_t krank, int m0, int n0, int seg_n, int seg_m)
{
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int Ntot = a.N0 + a.N1;
    // ---- epilogue --------------------------------------------------------------------------------------------------
    // warp w reads TMEM lanes 32*(w%4) .. +31 (the only ones it may access), columns (w/4)*BN/2 .. +BN/2 of every
    // accumulator; the kAcc accumulators (independent MMA chains) are summed in registers
    const int q = warp & 3, half = warp >> 2;
    const int row_in_tile = 32 * q + lane;
    constexpr int CPT = BN / 2;                               // columns per thread
    float acc[CPT];
#pragma unroll
    for (int c = 0; c < CPT; ++c) acc[c] = 0.f;
    if (nkb > 0) {
        mbar_wait(&done_bar, 0);
        fence_after_sync();
#pragma unroll
        for (int j = 0; j < kAcc; ++j) {
            if (j < n_mma) {                                  // accumulators that received at least one product
#pragma unroll
                for (int c = 0; c < CPT; c += 16) {
                    float v[16];
                    tmem_ld16(tmem_d + ((uint32_t)(32 * q) << 16) + (uint32_t)(j * BN + half * CPT + c), v);
#pragma unroll
                    for (int k = 0; k < 16; ++k) acc[c + k] += v[k];
                }
            }
        }
    }
    // split-K as a reduce-scatter over the cluster: CTA r finishes rows [r * RPC, (r + 1) * RPC) of the tile, RPC = BM / ksplit.
    // Every CTA sends each accumulator row to the CTA that owns it (distributed shared memory; its own rows stay local):
    // receive buffer [source rank][RPC rows][BN + 4] floats = BM * (BN + 4) floats whatever the split.
    constexpr int EPS = BN + 4;                               // padded row: conflict-free float4 accesses
    float *ep = reinterpret_cast<float *>(smem + ep_offset);
    const int rpc = BM / a.ksplit;
    {
        const uint32_t owner = (uint32_t)(row_in_tile / rpc);
        const int r_local = row_in_tile - (int)owner * rpc;
        const uint32_t dst = (a.ksplit > 1 ? map_to_cta(ep, owner) : smem_u32(ep)) +
                             (uint32_t)((((int)krank * rpc + r_local) * EPS + half * CPT) * 4);
#pragma unroll
        for (int c = 0; c < CPT; c += 4) {
            if (a.ksplit > 1) st_cluster_v4(dst + c * 4, acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
            else *reinterpret_cast<float4 *>(ep + row_in_tile * EPS + half * CPT + c) = make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
        }
    }
    float *redc = ep + BM * EPS;                              // bias-gradient partials [ksplit][BM], gathered in rank 0
    if (MODE == WGRAD && blockIdx.y == 0 && tid < BM) {
        if (a.ksplit > 1) {
            const uint32_t dstc = map_to_cta(redc, 0) + (uint32_t)((krank * BM + tid) * 4);
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dstc), "f"(colsum[tid]) : "memory");
        } else {
            redc[tid] = colsum[tid];
        }
    }
    if (a.ksplit > 1) cluster_sync_all();
    else __syncthreads();
    {
        using OutT = typename OutType<T, MODE>::type;         // dx / y: the layer's data type; dw: always fp32
        const int seg = MODE == FWD ? seg_n : seg_m;
        const int base = seg ? a.N0 : 0;
        const int row_lim = MODE == WGRAD ? (seg_m ? Ntot : a.N0) : a.rows;
        const int col_lim = MODE == FWD ? (seg_n ? Ntot : a.N0) : a.K;
        const long ld = MODE == FWD ? (seg_n ? a.N1 : a.N0) : a.K;
        OutT *out = MODE == FWD ? static_cast<OutT *>(a.y[seg_n]) - base
                  : MODE == DGRAD ? static_cast<OutT *>(a.dx)
                                  : static_cast<OutT *>(a.dw[seg_m]) - (long)base * a.K;
        const float *bias = MODE == FWD ? static_cast<const float *>(a.bias[seg_n]) : nullptr;
        for (int idx = tid; idx < rpc * (BN / 4); idx += kThreads) {
            const int r = idx / (BN / 4), c = (idx % (BN / 4)) * 4;
            const int m = m0 + (int)krank * rpc + r, n = n0 + c;
            if (m >= row_lim || n >= col_lim) continue;
            float4 v = *reinterpret_cast<const float4 *>(ep + r * EPS + c);
            for (int src = 1; src < a.ksplit; ++src) {
                const float4 p = *reinterpret_cast<const float4 *>(ep + (src * rpc + r) * EPS + c);
                v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
            }
            if (MODE == FWD) {
                if (bias != nullptr) {
                    const float4 bv = __ldg(reinterpret_cast<const float4 *>(bias + (n - base)));
                    v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
                }
                if (a.mask != nullptr && seg_n == 0 && a.mask[m] != 0) v = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            OutT *dst = out + (long)m * ld + n;
            if (sizeof(OutT) == 4) {
                *reinterpret_cast<float4 *>(dst) = v;
            } else {
                const __nv_bfloat162 p0 = __floats2bfloat162_rn(v.x, v.y), p1 = __floats2bfloat162_rn(v.z, v.w);
                *reinterpret_cast<uint2 *>(dst) = make_uint2(*reinterpret_cast<const uint32_t *>(&p0), *reinterpret_cast<const uint32_t *>(&p1));
            }
        }
    }
    if (MODE == WGRAD && blockIdx.y == 0 && krank == 0 && tid < BM) {
        // bias gradient: rank 0 holds every k-slice's column sums
        float sum = 0.f;
        for (int src = 0; src < a.ksplit; ++src) sum += redc[src * BM + tid];
        const int n = m0 + tid;
        const int lim = seg_m ? Ntot : a.N0;
        if (n < lim && a.db[seg_m] != nullptr) a.db[seg_m][n - (seg_m ? a.N0 : 0)] = sum;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// grid = (M tiles, N tiles, ksplit), cluster = (1, 1, ksplit); dynamic shared memory = Config::SMEM_BYTES
// ---------------------------------------------------------------------------------------------------------------------
template <typename T, int MODE, int BK>
__global__ void __launch_bounds__(kThreads, 1) msda_proj_kernel(const Args a)
{
    using CFG = Config<T, MODE, BK>;
    using TA = typename CFG::TA;
    using TB = typename CFG::TB;
    constexpr bool SPLIT = Elem<T>::SPLIT;
    constexpr int kPrefetch = Elem<T>::PREFETCH;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t stage_bar[kStages];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t tmem_slot;
    __shared__ float colsum[BM];                 // WGRAD: per-output-row sums of dy (bias gradient)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    MSDA_PROJ_STAMP(0)
    const uint32_t krank = a.ksplit > 1 ? cluster_ctarank() : 0;

    // ---- GEMM extents and operand views for this tile -----------------------------------------------------------
    // FWD: M = rows, N = Cout (N0 + N1), Kred = Cin.   DGRAD: M = rows, N = Cin, Kred = Cout.   WGRAD: M = Cout, N = Cin, Kred = rows
    const int Ntot = a.N0 + a.N1;
    const int Kred = MODE == FWD ? a.K : MODE == DGRAD ? Ntot : a.rows;
    // reduction blocks of this CTA: consecutive slices over the cluster
    const int nkb_all = (Kred + BK - 1) / BK;
    const int per = (nkb_all + a.ksplit - 1) / a.ksplit;
    const int kb_lo = min((int)krank * per, nkb_all), kb_hi = min(kb_lo + per, nkb_all);
    const int nkb = kb_hi - kb_lo;

    View<T> va{}, vb{};
    int seg_m = 0, seg_n = 0;     // segment of the output tile (FWD: columns; WGRAD: rows)
    if (MODE == FWD) {
        seg_n = n0 >= a.N0 ? 1 : 0;
        va = {static_cast<const T *>(a.x), (long)a.K, 1, a.rows, a.K, nullptr, 0};
        vb = {static_cast<const T *>(a.w[seg_n]) - (long)(seg_n ? a.N0 : 0) * a.K, (long)a.K, 1, seg_n ? Ntot : a.N0, a.K, nullptr, 0};
    } else if (MODE == WGRAD) {
        seg_m = m0 >= a.N0 ? 1 : 0;
        const int Ns = seg_m ? a.N1 : a.N0;
        // A(mn = output channel n, kk = row m) = dy[m][n]
        va = {static_cast<const T *>(a.dy[seg_m]) - (seg_m ? a.N0 : 0), 1, (long)Ns, seg_m ? Ntot : a.N0, a.rows, a.mask, (a.mask && seg_m == 0) ? 2 : 0};
        // B(mn = input channel k, kk = row m) = x[m][k]
        vb = {static_cast<const T *>(a.x), 1, (long)a.K, a.K, a.rows, nullptr, 0};
    }
    // DGRAD views depend on the reduction block (the reduction axis is the segmented one): built per block below

    // ---- one-time setup ---------------------------------------------------------------------------------------------
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) mbar_init(&stage_bar[s], 1);
        mbar_init(&done_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (MODE == WGRAD && tid < BM) colsum[tid] = 0.f;
    if (warp == 0) tmem_alloc(&tmem_slot, kAcc * BN);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_d = tmem_slot;
    int n_mma = 0;                                            // MMAs issued so far (every thread counts: block-uniform)
    float bsum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // WGRAD: this thread's share of the bias gradient (8 channels)

    // ---- main loop over the reduction blocks of this CTA ----------------------------------------------------------
    // register ring: the global loads of kPrefetch reduction blocks are in flight at any time (these GEMMs are short and
    // latency-bound: with one block in flight an iteration costs a full L2 round trip, ~1 us)
    Regs<T, TA> ra[kPrefetch];
    Regs<T, TB> rb[kPrefetch];
    auto load_block = [&](int kb, Regs<T, TA> &qa, Regs<T, TB> &qb) {
        int k0 = kb * BK;
        if (MODE == DGRAD) {
            const int s = k0 >= a.N0 ? 1 : 0;             // blocks never straddle the segment boundary (N0 % BK == 0)
            const int Ns = s ? a.N1 : a.N0, base = s ? a.N0 : 0;
            // A(mn = row m, kk = output channel n) = dy[m][n];  B(mn = input channel k, kk = n) = w[n][k]
            va = {static_cast<const T *>(a.dy[s]) - base, (long)Ns, 1, a.rows, base + Ns, a.mask, (a.mask && s == 0) ? 1 : 0};
            vb = {static_cast<const T *>(a.w[s]) - (long)base * a.K, 1, (long)a.K, a.K, base + Ns, nullptr, 0};
        }
        load_tile<T, TA>(va, m0, k0, qa);
        load_tile<T, TB>(vb, n0, k0, qb);
    };
    MSDA_PROJ_STAMP(1)
#pragma unroll
    for (int u = 0; u < kPrefetch; ++u)
        if (u < nkb) load_block(kb_lo + u, ra[u], rb[u]);
    if (a.ksplit > 1) cluster_sync_all();   // every CTA of the cluster is running before any remote shared-memory store
    for (int i0 = 0; i0 < nkb; i0 += kPrefetch) {
#pragma unroll
    for (int u = 0; u < kPrefetch; ++u) {
        const int i = i0 + u;
        if (i >= nkb) break;                                // block-uniform
        const int s = i % kStages;
        unsigned char *sa = smem + s * CFG::STAGE_BYTES;
        unsigned char *sb = sa + TA::BYTES * Elem<T>::COPIES;
        if (i >= kStages) mbar_wait(&stage_bar[s], ((i / kStages) - 1) & 1);   // the MMAs that read this stage are done
        store_tile<T, TA>(sa, ra[u]);
        store_tile<T, TB>(sb, rb[u]);
        if (MODE == WGRAD && blockIdx.y == 0) {
            // bias gradient: db[n] = sum over rows of dy[m][n] -- column sums of the A tile.  The MN-chunk of a thread's tile
            // chunks is fixed (chunk i -> MN-chunk i % NCM, and kThreads % NCM == 0), so the sums of its 8 channels stay in
            // registers over the whole reduction loop; they meet in shared memory once, after the loop.
            static_assert(kThreads % TA::NCM == 0, "a thread keeps one MN-chunk");
#pragma unroll
            for (int j = 0; j < TA::PER_THREAD; ++j) {
                if (SPLIT) {
#pragma unroll
                    for (int h = 0; h < Elem<T>::LOADS; ++h) {
                        bsum[4 * h + 0] += __uint_as_float(ra[u].v[j][h].x);
                        bsum[4 * h + 1] += __uint_as_float(ra[u].v[j][h].y);
                        bsum[4 * h + 2] += __uint_as_float(ra[u].v[j][h].z);
                        bsum[4 * h + 3] += __uint_as_float(ra[u].v[j][h].w);
                    }
                } else {
                    const uint32_t wv[4] = {ra[u].v[j][0].x, ra[u].v[j][0].y, ra[u].v[j][0].z, ra[u].v[j][0].w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        bsum[2 * q] += __uint_as_float(wv[q] << 16);
                        bsum[2 * q + 1] += __uint_as_float(wv[q] & 0xffff0000u);
                    }
                }
            }
        }
        if (i + kPrefetch < nkb) load_block(kb_lo + i + kPrefetch, ra[u], rb[u]);   // refill this slot of the ring
        fence_async_smem();                                 // generic-proxy stores -> visible to the tensor core's reads
        __syncthreads();
        if (tid == 0) {
            fence_after_sync();
            const uint32_t a_addr = smem_u32(sa), b_addr = smem_u32(sb);
            int j = n_mma;
            auto issue = [&](uint64_t da, uint64_t db) {
                umma(tmem_d + (uint32_t)((j % kAcc) * BN), da, db, CFG::IDESC, j >= kAcc ? 1u : 0u);
                ++j;
            };
#pragma unroll
            for (int ks = 0; ks < BK / TA::MMA_K; ++ks) {
                const uint64_t ah = TA::descriptor(a_addr + ks * TA::KSTEP_BYTES);
                const uint64_t bh = TB::descriptor(b_addr + ks * TB::KSTEP_BYTES);
                if (SPLIT) {
                    const uint64_t am = TA::descriptor(a_addr + TA::BYTES + ks * TA::KSTEP_BYTES);
                    const uint64_t al = TA::descriptor(a_addr + 2 * TA::BYTES + ks * TA::KSTEP_BYTES);
                    const uint64_t bm = TB::descriptor(b_addr + TB::BYTES + ks * TB::KSTEP_BYTES);
                    const uint64_t bl = TB::descriptor(b_addr + 2 * TB::BYTES + ks * TB::KSTEP_BYTES);
                    issue(al, bh);
                    issue(ah, bl);
                    issue(am, bm);
                    issue(am, bh);
                    issue(ah, bm);
                    issue(ah, bh);
                } else {
                    issue(ah, bh);
                }
            }
            umma_commit(&stage_bar[s]);                     // frees the stage when these MMAs have read it
            if (i == nkb - 1) umma_commit(&done_bar);       // accumulator complete
        }
        n_mma += (BK / TA::MMA_K) * (SPLIT ? 6 : 1);
    }
    }
    if (MODE == WGRAD && blockIdx.y == 0) {
        // threads tid, tid + NCM, ... hold the same 8 channels (MN-chunk tid % NCM): lanes l and l + 16 of a warp first
        // (NCM = 16), then one shared-memory atomic per warp and channel
        static_assert(TA::NCM == 16, "lane pairing below assumes 16 MN-chunks (BM = 128)");
#pragma unroll
        for (int k = 0; k < 8; ++k) bsum[k] += __shfl_xor_sync(0xffffffffu, bsum[k], 16);
        if (lane < 16) {
#pragma unroll
            for (int k = 0; k < 8; ++k) atomicAdd(&colsum[lane * 8 + k], bsum[k]);
        }
        __syncthreads();
    }
    MSDA_PROJ_STAMP(2)

    // ---- epilogue --------------------------------------------------------------------------------------------------
    proj_epilogue<T, MODE, BK>(a, smem, CFG::EP_OFFSET, done_bar, colsum, tmem_d, n_mma, nkb, krank, m0, n0, seg_n, seg_m);
    MSDA_PROJ_STAMP(6)
    // ---- teardown ---------------------------------------------------------------------------------------------------
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_d, kAcc * BN);
}


// ---------------------------------------------------------------------------------------------------------------------
// bf16 operands fed by the Tensor Memory Accelerator (forward and data gradient; the weight gradient keeps the register
// path because it also needs the operand values for the bias gradient, and so does a data gradient with a padding mask).
// One elected thread issues cp.async.bulk.tensor boxes straight into the 128-byte-swizzled UMMA tiles (the TMA swizzle
// IS the UMMA layout: a {64 x rows} box of a row-major matrix lands as rows of 128 bytes with chunk c of row r at c ^ r%8),
// a second one issues the MMAs; full / empty mbarriers per stage, four stages; out-of-range rows and reduction tails are
// zero-filled by the TMA.  Epilogue as above.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int kTmaStages = 4;
struct TmaMaps {
    CUtensorMap a[2];   // FWD: a[0] = x (rows, K).            DGRAD: a[s] = dy_s (rows, N_s)
    CUtensorMap b[2];   // both: b[s] = w_s (N_s, K) -- K-major operand of the forward, MN-major operand of the data gradient
};
constexpr int kTmaStageBytes = (BM + BN) * kBK * 2;
constexpr int kTmaEpOffset = kTmaStages * kTmaStageBytes;
constexpr int kTmaSmemBytes = kTmaEpOffset + (BM * (BN + 4) + kMaxSplit * BM) * 4;

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst)),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(kThreads, 1) msda_proj_tma_kernel(const Args a, const __grid_constant__ TmaMaps maps)
{
    using T = __nv_bfloat16;
    constexpr int BK = kBK;
    using CFG = Config<T, MODE, BK>;
    using TA = typename CFG::TA;
    using TB = typename CFG::TB;
    static_assert(MODE == FWD || MODE == DGRAD, "TMA-fed kernel: forward and data gradient");
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t full_bar[kTmaStages];
    __shared__ __align__(8) uint64_t empty_bar[kTmaStages];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t tmem_slot;
    __shared__ float colsum[BM];   // unused here (epilogue signature)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const uint32_t krank = a.ksplit > 1 ? cluster_ctarank() : 0;
    const int Ntot = a.N0 + a.N1;
    const int Kred = MODE == FWD ? a.K : Ntot;
    const int nkb_all = (Kred + BK - 1) / BK;
    const int per = (nkb_all + a.ksplit - 1) / a.ksplit;
    const int kb_lo = min((int)krank * per, nkb_all), kb_hi = min(kb_lo + per, nkb_all);
    const int nkb = kb_hi - kb_lo;
    const int seg_n = (MODE == FWD && n0 >= a.N0) ? 1 : 0;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kTmaStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&done_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&maps.a[0])) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&maps.b[0])) : "memory");
    }
    if (warp == 0) tmem_alloc(&tmem_slot, kAcc * BN);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_d = tmem_slot;
    if (a.ksplit > 1) cluster_sync_all();   // every CTA of the cluster is running before any remote shared-memory store

    if (warp == 0 && lane == 0) {
        // ===== TMA producer =====
        for (int i = 0; i < nkb; ++i) {
            const int s = i % kTmaStages;
            if (i >= kTmaStages) mbar_wait(&empty_bar[s], ((i / kTmaStages) - 1) & 1);
            unsigned char *sa = smem + s * kTmaStageBytes, *sb = sa + TA::BYTES;
            const int k0 = (kb_lo + i) * BK;
            mbar_expect_tx(&full_bar[s], kTmaStageBytes);
            if (MODE == FWD) {
                tma_load_2d(sa, &maps.a[0], k0, m0, &full_bar[s]);
                tma_load_2d(sb, &maps.b[seg_n], k0, n0 - (seg_n ? a.N0 : 0), &full_bar[s]);
            } else {
                const int sg = k0 >= a.N0 ? 1 : 0;          // reduction blocks never straddle the segments (N0 % 128 == 0)
                const int kl = k0 - (sg ? a.N0 : 0);
                tma_load_2d(sa, &maps.a[sg], kl, m0, &full_bar[s]);      // dy_s: box {64 channels n, 128 rows}
                tma_load_2d(sb, &maps.b[sg], n0, kl, &full_bar[s]);      // w_s : box {64 columns k, 64 rows n} = MN-major tile
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ===== MMA issuer =====
        int j = 0;
        for (int i = 0; i < nkb; ++i) {
            const int s = i % kTmaStages;
            mbar_wait(&full_bar[s], (i / kTmaStages) & 1);
            fence_after_sync();
            const uint32_t a_addr = smem_u32(smem + s * kTmaStageBytes), b_addr = a_addr + TA::BYTES;
#pragma unroll
            for (int ks = 0; ks < BK / TA::MMA_K; ++ks) {
                umma(tmem_d + (uint32_t)((j % kAcc) * BN), TA::descriptor(a_addr + ks * TA::KSTEP_BYTES),
                     TB::descriptor(b_addr + ks * TB::KSTEP_BYTES), CFG::IDESC, j >= kAcc ? 1u : 0u);
                ++j;
            }
            umma_commit(&empty_bar[s]);
            if (i == nkb - 1) umma_commit(&done_bar);
        }
    }
    __syncwarp();
    const int n_mma = nkb * (BK / TA::MMA_K);
    proj_epilogue<T, MODE, BK>(a, smem, kTmaEpOffset, done_bar, colsum, tmem_d, n_mma, nkb, krank, m0, n0, seg_n, 0);
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_d, kAcc * BN);
}

}  // namespace proj
}  // namespace msda
