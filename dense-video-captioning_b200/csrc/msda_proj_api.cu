// msda_proj_api.cu -- extern "C" entry points of the tcgen05 projection GEMMs (csrc/msda_proj.cuh; include/msda.h).
// Replaces, for MSDeformAttn, torch's nn.Linear calls of pdvc/ops/modules/ms_deform_attn.py:95-100 and :125.
#include "msda_host.cuh"
#include "msda_proj.cuh"

using namespace msda;
using namespace msda::host;

namespace {

// ---- projection GEMMs (csrc/msda_proj.cuh) -----------------------------------------------------------------------
int proj_check(int rows, int K, int N0, int N1, int dtype)
{
    if (dtype != MSDA_F32 && dtype != MSDA_BF16) return MSDA_ERR_DTYPE;
    if (rows <= 0 || K <= 0 || N0 <= 0 || N1 < 0) return MSDA_ERR_SHAPE;
    if ((K & 7) || (N0 & 7) || (N1 & 7)) return MSDA_ERR_UNSUPPORTED;      // 16-byte chunks of bf16 / fp32 rows
    if (N1 > 0 && (N0 % 128)) return MSDA_ERR_UNSUPPORTED;                  // tiles and reduction blocks never straddle the segments
    return MSDA_OK;
}

template <typename T, int MODE, int BK> int launch_proj_inst(proj::Args a, int gx, int gy, int kred, cudaStream_t st)
{
    using CFG = proj::Config<T, MODE, BK>;
    static unsigned long long done = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 63) dev = 63, done &= ~(1ull << 63);
    if (!(done & (1ull << dev))) {
        cudaError_t e = cudaFuncSetAttribute(proj::msda_proj_kernel<T, MODE, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, CFG::SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        done |= 1ull << dev;
    }
    // the GEMMs are latency-bound: split the reduction axis over a cluster until the SMs are covered
    const int nkb = (kred + BK - 1) / BK;
    // ... but keep all clusters in one wave with room to place them (clusters are scheduled per GPC: 36 tiles x 4 = 144 CTAs
    // on 148 SMs took 12.0 us where 36 x 2 takes 7.5)
    int ks = (num_sms() * 7 / 10) / (gx * gy > 0 ? gx * gy : 1);
    ks = ks >= 4 ? 4 : ks >= 2 ? 2 : 1;                 // BM / ks rows per CTA in the reduce-scatter: powers of two
    while (ks > nkb) ks >>= 1;
    static const int ks_env = env_int("MSDA_PROJ_KSPLIT", 0);
    if ((ks_env == 1 || ks_env == 2 || ks_env == 4) && ks_env <= nkb) ks = ks_env;
    a.ksplit = ks;
#ifdef MSDA_PROJ_TIMING
    static const char *dbg = getenv("MSDA_PROJ_DEBUG_PTR");
    a.debug = dbg ? reinterpret_cast<unsigned long long *>(strtoull(dbg, nullptr, 0)) : nullptr;
#endif
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(gx, gy, ks);
    cfg.blockDim = dim3(proj::kThreads);
    cfg.dynamicSmemBytes = CFG::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = ks;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, proj::msda_proj_kernel<T, MODE, BK>, a);
    ++g_launches;
    g_path = MSDA_PATH_PROJ;
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

// ---- TMA-fed bf16 kernels: tensor maps are built per call with the driver's encoder (fetched once through the runtime;
// libcuda is not linked) and travel as a __grid_constant__ kernel parameter
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn tensor_map_encoder()
{
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// 2-D bf16 row-major matrix (rows, cols), box {64 columns, box_rows rows}, 128-byte swizzle, zero fill outside
bool make_map(CUtensorMap *map, const void *base, int rows, int cols, int box_rows)
{
    const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
    const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1u, 1u};
    return tensor_map_encoder()(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), gdim, gstride, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int MODE> int launch_proj_tma(proj::Args a, int gx, int gy, int kred, cudaStream_t st)
{
    static unsigned long long done = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 63) dev = 63, done &= ~(1ull << 63);
    if (!(done & (1ull << dev))) {
        cudaError_t e = cudaFuncSetAttribute(proj::msda_proj_tma_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, proj::kTmaSmemBytes);
        if (e != cudaSuccess) return (int)e;
        done |= 1ull << dev;
    }
    proj::TmaMaps maps;
    bool ok = true;
    if (MODE == proj::FWD) {
        ok = make_map(&maps.a[0], a.x, a.rows, a.K, proj::BM);
        maps.a[1] = maps.a[0];
    } else {
        ok = make_map(&maps.a[0], a.dy[0], a.rows, a.N0, proj::BM);
        if (a.N1 > 0) ok = ok && make_map(&maps.a[1], a.dy[1], a.rows, a.N1, proj::BM);
        else maps.a[1] = maps.a[0];
    }
    ok = ok && make_map(&maps.b[0], a.w[0], a.N0, a.K, 64);
    if (a.N1 > 0) ok = ok && make_map(&maps.b[1], a.w[1], a.N1, a.K, 64);
    else maps.b[1] = maps.b[0];
    if (!ok) return MSDA_ERR_UNSUPPORTED;
    const int nkb = (kred + proj::kBK - 1) / proj::kBK;
    // ... but keep all clusters in one wave with room to place them (clusters are scheduled per GPC: 36 tiles x 4 = 144 CTAs
    // on 148 SMs took 12.0 us where 36 x 2 takes 7.5)
    int ks = (num_sms() * 7 / 10) / (gx * gy > 0 ? gx * gy : 1);
    ks = ks >= 4 ? 4 : ks >= 2 ? 2 : 1;
    while (ks > nkb) ks >>= 1;
    static const int ks_env = env_int("MSDA_PROJ_KSPLIT", 0);
    if ((ks_env == 1 || ks_env == 2 || ks_env == 4) && ks_env <= nkb) ks = ks_env;
    a.ksplit = ks;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(gx, gy, ks);
    cfg.blockDim = dim3(proj::kThreads);
    cfg.dynamicSmemBytes = proj::kTmaSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = ks;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, proj::msda_proj_tma_kernel<MODE>, a, maps);
    ++g_launches;
    g_path = MSDA_PATH_PROJ;
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

template <int MODE> int launch_proj(const proj::Args &a, int gx, int gy, int kred, int dtype, cudaStream_t st)
{
    static const int use_tma = env_int("MSDA_PROJ_TMA", 1);
    if (use_tma && dtype == MSDA_BF16 && tensor_map_encoder() != nullptr) {
        if (MODE == proj::FWD) return launch_proj_tma<proj::FWD>(a, gx, gy, kred, st);
        if (MODE == proj::DGRAD && a.mask == nullptr) return launch_proj_tma<proj::DGRAD>(a, gx, gy, kred, st);
    }
    return dtype == MSDA_F32 ? launch_proj_inst<float, MODE, proj::kBK>(a, gx, gy, kred, st)
                             : launch_proj_inst<__nv_bfloat16, MODE, proj::kBK>(a, gx, gy, kred, st);
}

}  // namespace

extern "C" {

int msda_proj_forward(const void *x, const void *w0, const float *b0, const void *w1, const float *b1,
                      const unsigned char *row_mask, void *y0, void *y1, int rows, int K, int N0, int N1, int dtype, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (int e = proj_check(rows, K, N0, N1, dtype)) return e;
    if (!x || !w0 || !y0 || (N1 > 0 && (!w1 || !y1))) return MSDA_ERR_NULL;
    if (!aligned16(x) || !aligned16(w0) || !aligned16(y0) || !aligned16(w1) || !aligned16(y1)) return MSDA_ERR_ALIGN;
    proj::Args a{};
    a.x = x; a.w[0] = w0; a.w[1] = w1; a.bias[0] = b0; a.bias[1] = b1; a.mask = row_mask; a.y[0] = y0; a.y[1] = y1;
    a.rows = rows; a.K = K; a.N0 = N0; a.N1 = N1;
    return launch_proj<proj::FWD>(a, (rows + proj::BM - 1) / proj::BM, (N0 + N1 + proj::BN - 1) / proj::BN, K, dtype,
                                  static_cast<cudaStream_t>(stream));
}

int msda_proj_dgrad(const void *dy0, const void *dy1, const void *w0, const void *w1, const unsigned char *row_mask,
                    void *dx, int rows, int K, int N0, int N1, int dtype, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (int e = proj_check(rows, K, N0, N1, dtype)) return e;
    if (!dy0 || !w0 || !dx || (N1 > 0 && (!w1 || !dy1))) return MSDA_ERR_NULL;
    if (!aligned16(dy0) || !aligned16(w0) || !aligned16(dx) || !aligned16(w1) || !aligned16(dy1)) return MSDA_ERR_ALIGN;
    proj::Args a{};
    a.dy[0] = dy0; a.dy[1] = dy1; a.w[0] = w0; a.w[1] = w1; a.mask = row_mask; a.dx = dx;
    a.rows = rows; a.K = K; a.N0 = N0; a.N1 = N1;
    return launch_proj<proj::DGRAD>(a, (rows + proj::BM - 1) / proj::BM, (K + proj::BN - 1) / proj::BN, N0 + N1, dtype,
                                    static_cast<cudaStream_t>(stream));
}

int msda_proj_wgrad(const void *dy0, const void *dy1, const void *x, const unsigned char *row_mask, float *dw0,
                    float *db0, float *dw1, float *db1, int rows, int K, int N0, int N1, int dtype, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (int e = proj_check(rows, K, N0, N1, dtype)) return e;
    if (!dy0 || !x || !dw0 || (N1 > 0 && (!dy1 || !dw1))) return MSDA_ERR_NULL;
    if (!aligned16(dy0) || !aligned16(x) || !aligned16(dw0) || !aligned16(dy1) || !aligned16(dw1)) return MSDA_ERR_ALIGN;
    proj::Args a{};
    a.dy[0] = dy0; a.dy[1] = dy1; a.x = x; a.mask = row_mask; a.dw[0] = dw0; a.dw[1] = dw1; a.db[0] = db0; a.db[1] = db1;
    a.rows = rows; a.K = K; a.N0 = N0; a.N1 = N1;
    return launch_proj<proj::WGRAD>(a, (N0 + N1 + proj::BM - 1) / proj::BM, (K + proj::BN - 1) / proj::BN, rows, dtype,
                                    static_cast<cudaStream_t>(stream));
}

}  // extern "C"
