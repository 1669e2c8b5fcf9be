// msda_api.cu -- extern "C" entry points of libmsda_sm100.so (declared in include/msda.h).
//
// Host-side role of the reference's ms_deform_attn_cuda_forward / _backward
// (pdvc/ops/src/cuda/ms_deform_attn_cuda.cu:20-80, :83-153): validate, pick a kernel, launch on the
// caller's stream.  Unlike the reference there is no ATen, no allocation, no per-64-sample host loop
// (cu:61-75) and no silent printf on launch failure (ms_deform_im2col_cuda.cuh:949-953): errors come
// back as return codes.
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "msda_fast.cuh"
#include "msda_generic.cuh"
#include "msda_tile.cuh"
#include "msda_lean.cuh"
#include "msda_small.cuh"
#include "msda_capsample.cuh"
#include "msda_steplinear.cuh"
#include "msda_capattend.cuh"

#include "msda_host.cuh"

using namespace msda;
using namespace msda::host;

namespace msda {
namespace host {
thread_local int g_launches = 0;
thread_local int g_path = MSDA_PATH_NONE;
}  // namespace host
}  // namespace msda

namespace {

inline int head_major_env()
{
    static const int v = env_int("MSDA_HEAD_MAJOR", 1);
    return v;
}

inline int check_launch(int path = 0)
{
    ++g_launches;
    if (path) g_path = path;
    const cudaError_t e = cudaGetLastError();
    return (int)e;
}

struct Dims {
    int N, S, M, D, L, Lq, P;
};

int check_dims(const Dims &d, int dtype, int pad_mode)
{
    if (dtype != MSDA_F32 && dtype != MSDA_BF16 && dtype != MSDA_F64) return MSDA_ERR_DTYPE;
    if (pad_mode != MSDA_PAD_BORDER && pad_mode != MSDA_PAD_ZEROS) return MSDA_ERR_PAD;
    if (d.N <= 0 || d.S <= 0 || d.M <= 0 || d.D <= 0 || d.L <= 0 || d.Lq <= 0 || d.P <= 0) return MSDA_ERR_SHAPE;
    if (d.L > MSDA_MAX_LEVELS) return MSDA_ERR_SHAPE;
    // per-sample offsets are 32-bit inside the kernels
    if ((long long)d.S * d.M * d.D >= (1ll << 31)) return MSDA_ERR_SHAPE;
    if ((long long)d.L * d.P >= (1 << 20)) return MSDA_ERR_SHAPE;
    // unit indices (n,q,m) are 32-bit inside the vectorised kernels
    if ((long long)d.N * d.Lq * d.M >= (1ll << 30)) return MSDA_ERR_SHAPE;
    return MSDA_OK;
}


// lanes per unit for the vectorised path, or 0 if (D, dtype) is not eligible
int fast_lpg(int D, int dtype)
{
    const int ch = dtype == MSDA_BF16 ? 8 : 4;
    if (dtype == MSDA_F64 || D % ch != 0) return 0;
    const int lanes = D / ch;
    if (lanes < 4) return 0;
    if (lanes >= 32) return (D % (32 * ch) == 0) ? 32 : 0;   // D chunks of 32 lanes each
    int p = 4;
    while (p < lanes) p <<= 1;
    return p;   // 4, 8, 16 or 32 (with masked lanes if lanes is not a power of two)
}

void fast_grid(int total_units, int lpg, int &grid, int &units_per_cta)
{
    const int g = 32 / lpg;
    const long long per_iter = (long long)kFastWarps * g;   // units one CTA covers per loop trip
    long long ctas = (total_units + per_iter - 1) / per_iter;
    const long long cap = (long long)num_sms() * 16;          // <= 16 resident 128-thread CTAs per SM
    if (ctas > cap) ctas = cap;
    long long upc = (total_units + ctas - 1) / ctas;
    upc = (upc + per_iter - 1) / per_iter * per_iter;
    grid = (int)((total_units + upc - 1) / upc);
    units_per_cta = (int)upc;
}

template <typename VT, bool FUSED> int launch_fast_fwd(const FastArgs &a, int lpg, int grid, cudaStream_t st)
{
#ifdef MSDA_DEV_FAST
    return MSDA_ERR_UNSUPPORTED;
#else
    switch (lpg) {
        case 4: msda_fwd_fast_kernel<VT, 4, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
        case 8: msda_fwd_fast_kernel<VT, 8, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
        case 16: msda_fwd_fast_kernel<VT, 16, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
        default: msda_fwd_fast_kernel<VT, 32, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
    }
    return check_launch(MSDA_PATH_FAST);
#endif
}

template <typename VT, bool FUSED> int launch_fast_bwd(const FastArgs &a, int lpg, int grid, cudaStream_t st)
{
#ifdef MSDA_DEV_FAST
    return MSDA_ERR_UNSUPPORTED;
#else
    switch (lpg) {
        case 4: msda_bwd_fast_kernel<VT, 4, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
        case 8: msda_bwd_fast_kernel<VT, 8, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
        case 16: msda_bwd_fast_kernel<VT, 16, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
        default: msda_bwd_fast_kernel<VT, 32, FUSED><<<grid, kFastThreads, 0, st>>>(a); break;
    }
    return check_launch(MSDA_PATH_FAST);
#endif
}

// ---- chunked ("tile") kernels: planning ----------------------------------------------------------

struct TilePlan {
    bool ok;
    int lpg, lpa, Q, chunks, LPpad, win_rows;
    size_t smem;
};

constexpr size_t kMaxDynSmem = 227 * 1024 - 2048;   // opt-in limit minus the kernels' static shared memory

// lanes per value row (16 bytes each): the tile kernels need a power of two between 4 and 32
int tile_lpg(int D, int dtype)
{
    const int ch = dtype == MSDA_BF16 ? 8 : 4;
    if (dtype == MSDA_F64 || D % ch != 0) return 0;
    const int lanes = D / ch;
    return (lanes == 4 || lanes == 8 || lanes == 16 || lanes == 32) ? lanes : 0;
}

TilePlan tile_plan(const Dims &d, int dtype, bool backward)
{
    static const int use_tile = env_int("MSDA_TILE", 1);
    static const int q_env = env_int("MSDA_Q", 0);
    static const int win_kb = env_int("MSDA_WIN_KB", 64);
    TilePlan p{};
    p.lpg = tile_lpg(d.D, dtype);
    if (!use_tile || !p.lpg) return p;
    const int LP = d.L * d.P;
    p.LPpad = (LP + 3) / 4 * 4;
    if (p.LPpad > 256) return p;
    p.lpa = LP <= 16 ? 16 : 32;
    int Q = 1024 / p.LPpad;
    Q = Q > 64 ? 64 : (Q < 4 ? 4 : Q);
    // small problems: more, smaller chunks so that every SM gets a CTA
    const int sms = num_sms();
    while (Q > 8 && (long long)d.N * d.M * ((d.Lq + Q - 1) / Q) < 2 * sms) Q >>= 1;
    if (q_env > 0) Q = q_env;
    if (Q > d.Lq) Q = d.Lq;
    if (Q * p.LPpad > kTileMaxSamples) return p;
    p.Q = Q;
    p.chunks = (d.Lq + Q - 1) / Q;
    if ((long long)d.N * d.M * p.chunks >= (1ll << 31)) return p;
    const int rb = d.D * (dtype == MSDA_BF16 ? 2 : 4);
    if (backward) {
        p.smem = bwd_tile_smem_bytes(Q, p.LPpad, d.S, rb);
        if (p.smem > kMaxDynSmem || d.S >= (1 << kRecLevelShift)) return p;
    } else {
        const size_t recs = (size_t)Q * p.LPpad * sizeof(int4);
        size_t win = (size_t)win_kb * 1024;
        if (recs + win > kMaxDynSmem) win = kMaxDynSmem > recs ? kMaxDynSmem - recs : 0;
        p.win_rows = (int)(win / rb);
        if (p.win_rows > d.S) p.win_rows = d.S;   // the whole head slice fits
        p.smem = recs + (size_t)p.win_rows * rb;
        if (p.smem > kMaxDynSmem || d.S >= (1 << kRecLevelShift)) return p;
    }
    p.ok = true;
    return p;
}

void fill_tile_args(TileArgs &a, const Dims &d, const TilePlan &p, int pad_mode)
{
    a.N = d.N; a.S = d.S; a.M = d.M; a.D = d.D; a.L = d.L; a.Lq = d.Lq; a.P = d.P;
    a.pad_mode = pad_mode;
    a.LP = d.L * d.P;
    a.LPpad = p.LPpad;
    a.invP = 1.f / (float)d.P;
    a.Q = p.Q;
    a.chunks = p.chunks;
    a.win_rows = p.win_rows;
}

// opt in to > 48 KB of dynamic shared memory once per kernel instantiation and device
template <typename K> int allow_big_smem(K kernel, unsigned long long &done_mask)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev > 63) dev = 63, done_mask &= ~(1ull << 63);
    if (done_mask & (1ull << dev)) return 0;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
    if (e != cudaSuccess) return (int)e;
    done_mask |= 1ull << dev;
    return 0;
}

template <typename VT, int LPG, int LPA, bool FUSED, bool BWD>
int launch_tile_inst(const TileArgs &a, int grid, size_t smem, cudaStream_t st)
{
    static unsigned long long done = 0;
    if (BWD) {
        if (int e = allow_big_smem(msda_bwd_tile_kernel<VT, LPG, LPA, FUSED>, done)) return e;
        msda_bwd_tile_kernel<VT, LPG, LPA, FUSED><<<grid, kTileThreads, smem, st>>>(a);
    } else {
        if (int e = allow_big_smem(msda_fwd_tile_kernel<VT, LPG, LPA, FUSED>, done)) return e;
        msda_fwd_tile_kernel<VT, LPG, LPA, FUSED><<<grid, kTileThreads, smem, st>>>(a);
    }
    return check_launch(MSDA_PATH_TILE);
}

template <typename VT, bool FUSED, bool BWD>
int launch_tile_lanes(const TileArgs &a, const TilePlan &p, int grid, cudaStream_t st)
{
#ifdef MSDA_DEV_FAST
    return MSDA_ERR_UNSUPPORTED;
#else
#define MSDA_TILE_CASE(LPG_)                                                                                  \
    case LPG_:                                                                                                \
        return p.lpa == 16 ? launch_tile_inst<VT, LPG_, 16, FUSED, BWD>(a, grid, p.smem, st)                  \
                           : launch_tile_inst<VT, LPG_, 32, FUSED, BWD>(a, grid, p.smem, st);
    switch (p.lpg) {
        MSDA_TILE_CASE(4)
        MSDA_TILE_CASE(8)
        MSDA_TILE_CASE(16)
        default: MSDA_TILE_CASE(32)
    }
#undef MSDA_TILE_CASE
#endif
}

template <bool FUSED, bool BWD> int launch_tile(const TileArgs &a, const TilePlan &p, int dtype, cudaStream_t st)
{
    const int grid = a.N * a.M * a.chunks;
    return dtype == MSDA_F32 ? launch_tile_lanes<float, FUSED, BWD>(a, p, grid, st)
                             : launch_tile_lanes<__nv_bfloat16, FUSED, BWD>(a, p, grid, st);
}

// ---- lean kernels (L*P <= 64): planning and launch ---------------------------------------------------

struct LeanPlan {
    bool ok;
    int lpg, nch, lpa, spl, Q, chunks, win_rows, gridK, hist_cap;
    size_t smem;
};

LeanPlan lean_plan(const Dims &d, int dtype, bool backward)
{
    static const int use_lean = env_int("MSDA_LEAN", 1);
    static const int win_kb = env_int("MSDA_WIN_KB", 32);
    static const int passes_env = env_int("MSDA_PASSES", 0);
    LeanPlan p{};
    if (!use_lean || dtype == MSDA_F64) return p;
    const int LP = d.L * d.P;
    const int rb = d.D * (dtype == MSDA_BF16 ? 2 : 4);
    if (LP > 64 || rb % 16 != 0) return p;   // lane j of a unit's group owns samples j, j + 32 (spl = 2) when L*P > 32
    const int chunks16 = rb / 16;
    if (chunks16 == 4) { p.lpg = 4; p.nch = 1; }
    else if (chunks16 == 8 || chunks16 == 16 || chunks16 == 32) { p.lpg = 8; p.nch = chunks16 / 8; }
    else return p;
    if (d.N > 65535 || d.M > 65535 || d.S >= (1 << 26)) return p;
    if ((long long)d.N * d.Lq * d.M * LP >= (1ll << 30)) return p;   // 32-bit sample indices
    p.lpa = LP <= 16 ? 16 : 32;
    p.spl = LP > 32 ? 2 : 1;
    const int upp = kLeanWarps * (32 / p.lpa);
    int passes = kLeanPasses / p.spl;
    // small problems: fewer queries per CTA so that every SM gets work
    const int sms = num_sms();
    while (passes > 1 && (long long)d.N * d.M * ((d.Lq + upp * passes - 1) / (upp * passes)) < 2 * sms) passes >>= 1;
    if (passes_env > 0 && passes_env <= kLeanPasses / p.spl) passes = passes_env;
    p.Q = upp * passes;
    p.chunks = (d.Lq + p.Q - 1) / p.Q;
    // self-attention (the queries are the rows of the level pyramid): tile by time span, see tile_queries()
    static const int use_grid = env_int("MSDA_GRID", 1);
    if (use_grid && d.Lq == d.S && d.L > 1 && d.S < 65536 && p.Q >= 4 * d.L) {
        p.gridK = (d.S + (p.Q - d.L) - 1) / (p.Q - d.L);   // sum_l ceil(T_l / K) <= S / K + L <= Q
        p.chunks = p.gridK;
    }
    if (backward) {
        static const int bwin_kb = env_int("MSDA_BWIN_KB", 34);
        p.win_rows = bwin_kb * 1024 / rb;
        if (p.win_rows > d.S + d.L) p.win_rows = d.S + d.L;
        // sort histogram: level-grid tiles touch one short window per level (no more rows than the value window
        // holds); tiles of consecutive queries may touch the whole slice.  Tiles that exceed it are swept unsorted.
        p.hist_cap = p.gridK ? (p.win_rows < d.S ? p.win_rows : d.S) : (d.S < 8192 ? d.S : 8192);
        p.smem = bwd_lean_smem_bytes(p.Q, p.lpa * p.spl, p.hist_cap, rb, p.win_rows);
        if (p.smem > kMaxDynSmem) {   // no room for a window: gather value rows through L1
            p.win_rows = 0;
            p.smem = bwd_lean_smem_bytes(p.Q, p.lpa * p.spl, p.hist_cap, rb, 0);
        }
        if (p.smem > kMaxDynSmem) return p;
    } else {
        const size_t recs = (size_t)p.Q * p.lpa * p.spl * sizeof(int4);
        const size_t win = (size_t)win_kb * 1024;
        p.win_rows = (int)(win / rb);
        if (p.win_rows > d.S) p.win_rows = d.S;   // the whole head slice fits
        p.smem = recs + (size_t)p.win_rows * rb;
        if (p.smem > kMaxDynSmem) return p;
    }
    p.ok = true;
    return p;
}

void fill_lean_args(TileArgs &a, const Dims &d, const LeanPlan &p, int pad_mode)
{
    a.N = d.N; a.S = d.S; a.M = d.M; a.D = d.D; a.L = d.L; a.Lq = d.Lq; a.P = d.P;
    a.pad_mode = pad_mode;
    a.LP = d.L * d.P;
    a.LPpad = p.lpa;
    a.invP = 1.f / (float)d.P;
    a.Q = p.Q;
    a.chunks = p.chunks;
    a.win_rows = p.win_rows;
    a.gridK = p.gridK;
    a.hist_cap = p.hist_cap;
}

template <typename VT, int LPG, int NCH, int LPA, bool FUSED, bool BWD, int SPL = 1>
int launch_lean_inst(const TileArgs &a, const LeanPlan &p, cudaStream_t st)
{
    static unsigned long long done = 0;
    const dim3 grid(p.chunks, a.M, a.N);
    if (BWD) {
        if (int e = allow_big_smem(msda_bwd_lean_kernel<VT, LPG, NCH, LPA, FUSED, SPL>, done)) return e;
        msda_bwd_lean_kernel<VT, LPG, NCH, LPA, FUSED, SPL><<<grid, kLeanThreads, p.smem, st>>>(a);
    } else {
        if (int e = allow_big_smem(msda_fwd_lean_kernel<VT, LPG, NCH, LPA, FUSED, SPL>, done)) return e;
        msda_fwd_lean_kernel<VT, LPG, NCH, LPA, FUSED, SPL><<<grid, kLeanThreads, p.smem, st>>>(a);
    }
    return check_launch(MSDA_PATH_LEAN);
}

template <typename VT, bool FUSED, bool BWD> int launch_lean_shape(const TileArgs &a, const LeanPlan &p, cudaStream_t st)
{
#define MSDA_LEAN_CASE(LPG_, NCH_)                                                                       \
    if (p.lpg == LPG_ && p.nch == NCH_)                                                                  \
        return p.lpa == 16 ? launch_lean_inst<VT, LPG_, NCH_, 16, FUSED, BWD>(a, p, st)                  \
             : p.spl == 1  ? launch_lean_inst<VT, LPG_, NCH_, 32, FUSED, BWD>(a, p, st)                  \
                           : launch_lean_inst<VT, LPG_, NCH_, 32, FUSED, BWD, 2>(a, p, st);
#ifndef MSDA_DEV_FAST   // development builds instantiate the headline shapes only (fp32 and bf16, D = 64)
    MSDA_LEAN_CASE(4, 1)
    MSDA_LEAN_CASE(8, 4)
#endif
    MSDA_LEAN_CASE(8, 1)
    MSDA_LEAN_CASE(8, 2)
#undef MSDA_LEAN_CASE
    return MSDA_ERR_UNSUPPORTED;
}

template <bool FUSED, bool BWD> int launch_lean(const TileArgs &a, const LeanPlan &p, int dtype, cudaStream_t st)
{
    return dtype == MSDA_F32 ? launch_lean_shape<float, FUSED, BWD>(a, p, st)
                             : launch_lean_shape<__nv_bfloat16, FUSED, BWD>(a, p, st);
}

// ---- small problems: one CTA per (n, q), one warp per head (csrc/msda_small.cuh) -------------------------------------
// lanes per value row, or 0 when the single-phase kernels do not take the problem
int small_lanes(const Dims &d, int dtype, unsigned flags, bool backward = false)
{
    if (flags & MSDA_FLAG_NO_SMALL) return 0;
    // (n, q) pairs up to which the single-phase kernels win (graph-replayed, YC2 shapes: forward 3.6 / 4.5 us at 100 / 375
    // queries against 5.5 / 5.3 us staged; backward 5.6 us at 100 queries against 11.7, but 11.4 against 10.3 at 375: one
    // vector reduction per lane and tap is 25 MB of L2 atomics there, the staged kernel's per-row accumulation wins)
    static const int max_fwd = env_int("MSDA_SMALL_MAX", 1024), max_bwd = env_int("MSDA_SMALL_MAX_BWD", 256);
    const int max_units = backward ? max_bwd : max_fwd;
    if (dtype != MSDA_F32 && dtype != MSDA_BF16) return 0;
    const int rb = d.D * (dtype == MSDA_BF16 ? 2 : 4);
    if (rb != 64 && rb != 128 && rb != 256) return 0;
    if (d.L * d.P > 32 || d.M > kSmallMaxHeads || (long long)d.N * d.Lq > max_units) return 0;
    return rb / 16;
}

template <typename VT, bool FUSED, bool BWD> int launch_small(TileArgs &a, const Dims &d, int hl, int pad_mode, cudaStream_t st)
{
    a.N = d.N; a.S = d.S; a.M = d.M; a.D = d.D; a.L = d.L; a.Lq = d.Lq; a.P = d.P;
    a.pad_mode = pad_mode;
    a.LP = d.L * d.P;
    a.invP = 1.f / (float)d.P;
    const int grid = d.N * d.Lq, block = 32 * d.M;
#define MSDA_SMALL_CASE(HL_)                                                                     \
    case HL_:                                                                                    \
        if (BWD) msda_bwd_small_kernel<VT, HL_, FUSED><<<grid, block, 0, st>>>(a);               \
        else msda_fwd_small_kernel<VT, HL_, FUSED><<<grid, block, 0, st>>>(a);                   \
        break;
    switch (hl) {
        MSDA_SMALL_CASE(4)
        MSDA_SMALL_CASE(8)
        default: MSDA_SMALL_CASE(16)
    }
#undef MSDA_SMALL_CASE
    return check_launch(MSDA_PATH_SMALL);
}

int generic_grid(long long total)
{
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)num_sms() * 32;
    return (int)(blocks > cap ? cap : blocks);
}

template <bool SAMPLER> int launch_generic_fwd(const GenericArgs &a, int dtype, cudaStream_t st)
{
    const int grid = generic_grid(a.total);
    if (dtype == MSDA_F32) msda_fwd_generic_kernel<float, SAMPLER><<<grid, 256, 0, st>>>(a);
    else if (dtype == MSDA_F64) msda_fwd_generic_kernel<double, SAMPLER><<<grid, 256, 0, st>>>(a);
    else msda_fwd_generic_kernel<__nv_bfloat16, SAMPLER><<<grid, 256, 0, st>>>(a);
    return check_launch(MSDA_PATH_GENERIC);
}

template <bool SAMPLER> int launch_generic_bwd(const GenericArgs &a, int dtype, cudaStream_t st)
{
    const int grid = generic_grid(a.total);
    if (dtype == MSDA_F32) msda_bwd_generic_kernel<float, SAMPLER><<<grid, 256, 0, st>>>(a);
    else if (dtype == MSDA_F64) msda_bwd_generic_kernel<double, SAMPLER><<<grid, 256, 0, st>>>(a);
    else msda_bwd_generic_kernel<__nv_bfloat16, SAMPLER><<<grid, 256, 0, st>>>(a);
    return check_launch(MSDA_PATH_GENERIC);
}

// fp32 accumulation image the bf16 backward passes need (the part of the workspace that is mandatory)
size_t required_workspace(int N, int S, int M, int D, int dtype)
{
    return dtype == MSDA_BF16 ? (size_t)N * S * M * D * sizeof(float) : 0;
}
int memset_async(void *p, size_t bytes, cudaStream_t st)
{
    ++g_launches;
    return (int)cudaMemsetAsync(p, 0, bytes, st);
}

// bf16: turn the fp32 accumulation image into grad_value
int finish_bf16(const float *acc, void *grad_value, long long n, cudaStream_t st)
{
    if (n % 8 == 0) {
        const long long n8 = n / 8;
        long long blocks = (n8 + 255) / 256;
        const int cap = num_sms() * 16;
        if (blocks > cap) blocks = cap;
        msda_cvt_f32_to_bf16_kernel<<<(int)blocks, 256, 0, st>>>(acc, static_cast<__nv_bfloat16 *>(grad_value), n8);
    } else {
        msda_cvt_f32_to_bf16_scalar_kernel<<<generic_grid(n), 256, 0, st>>>(acc, static_cast<__nv_bfloat16 *>(grad_value), n);
    }
    return check_launch();
}

}  // namespace

extern "C" {

int msda_version(void) { return MSDA_VERSION; }
const char *msda_arch(void) { return "sm_100a"; }
int msda_last_launch_count(void) { return g_launches; }
int msda_last_launch_path(void) { return g_path; }
#ifdef MSDA_PHASE_TIMING
int msda_dev_phase_clocks(unsigned long long *out32, int reset)
{
    cudaDeviceSynchronize();
    cudaError_t e = cudaMemcpyFromSymbol(out32, g_phase_clk, sizeof(unsigned long long) * 32);
    if (reset) { unsigned long long z[32] = {}; cudaMemcpyToSymbol(g_phase_clk, z, sizeof(z)); }
    return (int)e;
}
#endif

const char *msda_error_string(int code)
{
    switch (code) {
        case MSDA_OK: return "success";
        case MSDA_ERR_DTYPE: return "msda: unsupported dtype for this entry point";
        case MSDA_ERR_SHAPE: return "msda: invalid or out-of-range dimension";
        case MSDA_ERR_NULL: return "msda: required pointer is NULL";
        case MSDA_ERR_ALIGN: return "msda: tensor base pointer must be 16-byte aligned";
        case MSDA_ERR_WORKSPACE: return "msda: workspace too small (see msda_backward_workspace_bytes)";
        case MSDA_ERR_PAD: return "msda: unknown pad_mode";
        case MSDA_ERR_UNSUPPORTED: return "msda: combination not supported";
        case MSDA_ERR_STATIC_SCRATCH:
            return "msda: N*S*M*D*4 exceeds MSDA_STATIC_SCRATCH_BYTES (use msda_fused_backward with a workspace)";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "msda: unknown error";
    }
}

size_t msda_backward_workspace_bytes(int N, int S, int M, int D, int dtype)
{
    return required_workspace(N, S, M, D, dtype);
}

int msda_ms_deform_attn_forward(const void *value, const int64_t *spatial_shapes, const int64_t *level_start_index,
                                const void *sampling_loc, const void *attn_weight, void *output, int N, int S, int M,
                                int D, int L, int Lq, int P, int dtype, int pad_mode, unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (!value || !spatial_shapes || !level_start_index || !sampling_loc || !attn_weight || !output) return MSDA_ERR_NULL;
    if (!aligned16(value) || !aligned16(sampling_loc) || !aligned16(attn_weight) || !aligned16(output)) return MSDA_ERR_ALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool one_d = (flags & MSDA_FLAG_ASSUME_1D) != 0;
    const int lpg = (flags & MSDA_FLAG_FORCE_GENERIC) ? 0 : fast_lpg(D, dtype);
    if (one_d && !(flags & MSDA_FLAG_FORCE_GENERIC)) {
        if (const int hl = small_lanes(d, dtype, flags)) {
            TileArgs a{};
            a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
            a.loc = static_cast<const float *>(sampling_loc); a.attn = static_cast<const float *>(attn_weight);
            a.out = output;
            return dtype == MSDA_F32 ? launch_small<float, false, false>(a, d, hl, pad_mode, st)
                                     : launch_small<__nv_bfloat16, false, false>(a, d, hl, pad_mode, st);
        }
        const LeanPlan lp = lean_plan(d, dtype, false);
        const TilePlan tp = lp.ok ? TilePlan{} : tile_plan(d, dtype, false);
        if (lp.ok || tp.ok) {
            TileArgs a{};
            a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
            a.loc = static_cast<const float *>(sampling_loc); a.attn = static_cast<const float *>(attn_weight);
            a.out = output;
            if (lp.ok) {
                fill_lean_args(a, d, lp, pad_mode);
                return launch_lean<false, false>(a, lp, dtype, st);
            }
            fill_tile_args(a, d, tp, pad_mode);
            return launch_tile<false, false>(a, tp, dtype, st);
        }
    }
    if (one_d && lpg) {
        FastArgs a{};
        a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
        a.loc = static_cast<const float *>(sampling_loc); a.attn = static_cast<const float *>(attn_weight);
        a.out = output;
        a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
        a.total_units = N * Lq * M;
        int grid;
        fast_grid(a.total_units, lpg, grid, a.units_per_cta);
        a.head_major = head_major_env();
        return dtype == MSDA_F32 ? launch_fast_fwd<float, false>(a, lpg, grid, st)
                                 : launch_fast_fwd<__nv_bfloat16, false>(a, lpg, grid, st);
    }
    GenericArgs a{};
    a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
    a.loc = sampling_loc; a.attn = attn_weight; a.out = output;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
    a.assume_1d = one_d ? 1 : 0;
    a.total = (long long)N * Lq * M * D;
    return launch_generic_fwd<false>(a, dtype, st);
}

int msda_ms_deform_attn_backward(const void *value, const int64_t *spatial_shapes, const int64_t *level_start_index,
                                 const void *sampling_loc, const void *attn_weight, const void *grad_output,
                                 void *grad_value, void *grad_sampling_loc, void *grad_attn_weight, void *workspace,
                                 size_t workspace_bytes, int N, int S, int M, int D, int L, int Lq, int P, int dtype,
                                 int pad_mode, unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (!value || !spatial_shapes || !level_start_index || !sampling_loc || !attn_weight || !grad_output || !grad_value ||
        !grad_sampling_loc || !grad_attn_weight)
        return MSDA_ERR_NULL;
    if (!aligned16(value) || !aligned16(sampling_loc) || !aligned16(attn_weight) || !aligned16(grad_output) ||
        !aligned16(grad_value) || !aligned16(grad_sampling_loc) || !aligned16(grad_attn_weight))
        return MSDA_ERR_ALIGN;
    const size_t need = required_workspace(N, S, M, D, dtype);
    if (need && (!workspace || workspace_bytes < need || !aligned16(workspace))) return MSDA_ERR_WORKSPACE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long nv = (long long)N * S * M * D;
    const long long ns = (long long)N * Lq * M * L * P;
    const size_t ct = dtype == MSDA_F64 ? 8 : 4;   // element size of loc / attn and their grads
    void *acc = dtype == MSDA_BF16 ? workspace : grad_value;
    const bool one_d = (flags & MSDA_FLAG_ASSUME_1D) != 0;
    const int lpg = (flags & MSDA_FLAG_FORCE_GENERIC) ? 0 : fast_lpg(D, dtype);
    int rc;
    const bool try_chunked = one_d && !(flags & MSDA_FLAG_FORCE_GENERIC);
    if (int e = memset_async(acc, (size_t)nv * (dtype == MSDA_F64 ? 8 : 4), st)) return e;
    if (const int hl = try_chunked ? small_lanes(d, dtype, flags, true) : 0) {
        TileArgs a{};
        a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
        a.loc = static_cast<const float *>(sampling_loc); a.attn = static_cast<const float *>(attn_weight);
        a.grad_out = grad_output; a.gv_acc = static_cast<float *>(acc);
        a.grad_loc = static_cast<float *>(grad_sampling_loc); a.grad_attn = static_cast<float *>(grad_attn_weight);
        rc = dtype == MSDA_F32 ? launch_small<float, false, true>(a, d, hl, pad_mode, st)
                               : launch_small<__nv_bfloat16, false, true>(a, d, hl, pad_mode, st);
        if (rc) return rc;
        if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
        return MSDA_OK;
    }
    const LeanPlan lp = try_chunked ? lean_plan(d, dtype, true) : LeanPlan{};
    const TilePlan tp = (try_chunked && !lp.ok) ? tile_plan(d, dtype, true) : TilePlan{};
    if (lp.ok || tp.ok) {
        TileArgs a{};
        a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
        a.loc = static_cast<const float *>(sampling_loc); a.attn = static_cast<const float *>(attn_weight);
        a.grad_out = grad_output; a.gv_acc = static_cast<float *>(acc);
        a.grad_loc = static_cast<float *>(grad_sampling_loc); a.grad_attn = static_cast<float *>(grad_attn_weight);
        if (lp.ok) {
            fill_lean_args(a, d, lp, pad_mode);
            rc = launch_lean<false, true>(a, lp, dtype, st);
        } else {
            fill_tile_args(a, d, tp, pad_mode);
            rc = launch_tile<false, true>(a, tp, dtype, st);
        }
    } else if (one_d && lpg) {
        FastArgs a{};
        a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
        a.loc = static_cast<const float *>(sampling_loc); a.attn = static_cast<const float *>(attn_weight);
        a.grad_out = grad_output; a.gv_acc = static_cast<float *>(acc);
        a.grad_loc = static_cast<float *>(grad_sampling_loc); a.grad_attn = static_cast<float *>(grad_attn_weight);
        a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
        a.total_units = N * Lq * M;
        int grid;
        fast_grid(a.total_units, lpg, grid, a.units_per_cta);
        a.head_major = head_major_env();
        rc = dtype == MSDA_F32 ? launch_fast_bwd<float, false>(a, lpg, grid, st)
                               : launch_fast_bwd<__nv_bfloat16, false>(a, lpg, grid, st);
    } else {
        if (int e = memset_async(grad_sampling_loc, (size_t)ns * 2 * ct, st)) return e;
        if (int e = memset_async(grad_attn_weight, (size_t)ns * ct, st)) return e;
        GenericArgs a{};
        a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
        a.loc = sampling_loc; a.attn = attn_weight; a.grad_out = grad_output;
        a.gv_acc = acc; a.grad_loc = grad_sampling_loc; a.grad_attn = grad_attn_weight;
        a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
        a.assume_1d = one_d ? 1 : 0;
        a.total = (long long)N * Lq * M * D;
        rc = launch_generic_bwd<false>(a, dtype, st);
    }
    if (rc) return rc;
    if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
    return MSDA_OK;
}

int msda_fused_forward(const void *value, const int64_t *temporal_shapes, const int64_t *level_start_index,
                       const float *reference_points, const float *sampling_offsets, const float *attn_logits,
                       void *output, int N, int S, int M, int D, int L, int Lq, int P, int R, int dtype, int pad_mode,
                       unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (R != 1 && R != 2) return MSDA_ERR_SHAPE;
    if (!value || !temporal_shapes || !level_start_index || !reference_points || !sampling_offsets || !attn_logits || !output)
        return MSDA_ERR_NULL;
    if (!aligned16(value) || !aligned16(output)) return MSDA_ERR_ALIGN;
    const int lpg = fast_lpg(D, dtype);
    if (!lpg) return MSDA_ERR_UNSUPPORTED;   // callers fall back to the unfused entry points (still CUDA)
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (const int hl = small_lanes(d, dtype, flags)) {
        TileArgs t{};
        t.value = value; t.shapes = temporal_shapes; t.lsi = level_start_index; t.shapes_stride = 1;
        t.ref = reference_points; t.offs = sampling_offsets; t.logits = attn_logits; t.R = R;
        t.out = output;
        return dtype == MSDA_F32 ? launch_small<float, true, false>(t, d, hl, pad_mode, st)
                                 : launch_small<__nv_bfloat16, true, false>(t, d, hl, pad_mode, st);
    }
    {
        const LeanPlan lp = lean_plan(d, dtype, false);
        const TilePlan tp = lp.ok ? TilePlan{} : tile_plan(d, dtype, false);
        if (lp.ok || tp.ok) {
            TileArgs t{};
            t.value = value; t.shapes = temporal_shapes; t.lsi = level_start_index; t.shapes_stride = 1;
            t.ref = reference_points; t.offs = sampling_offsets; t.logits = attn_logits; t.R = R;
            t.out = output;
            if (lp.ok) {
                fill_lean_args(t, d, lp, pad_mode);
                return launch_lean<true, false>(t, lp, dtype, st);
            }
            fill_tile_args(t, d, tp, pad_mode);
            return launch_tile<true, false>(t, tp, dtype, st);
        }
    }
    FastArgs a{};
    a.value = value; a.shapes = temporal_shapes; a.lsi = level_start_index; a.shapes_stride = 1;
    a.ref = reference_points; a.offs = sampling_offsets; a.logits = attn_logits; a.R = R;
    a.out = output;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
    a.total_units = N * Lq * M;
    int grid;
    fast_grid(a.total_units, lpg, grid, a.units_per_cta);
    a.head_major = head_major_env();
    return dtype == MSDA_F32 ? launch_fast_fwd<float, true>(a, lpg, grid, st)
                             : launch_fast_fwd<__nv_bfloat16, true>(a, lpg, grid, st);
}

int msda_fused_backward(const void *value, const int64_t *temporal_shapes, const int64_t *level_start_index,
                        const float *reference_points, const float *sampling_offsets, const float *attn_logits,
                        const void *grad_output, void *grad_value, float *grad_reference_points,
                        float *grad_sampling_offsets, float *grad_attn_logits, void *workspace, size_t workspace_bytes,
                        int N, int S, int M, int D, int L, int Lq, int P, int R, int dtype, int pad_mode, unsigned flags,
                        void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (R != 1 && R != 2) return MSDA_ERR_SHAPE;
    if (!value || !temporal_shapes || !level_start_index || !reference_points || !sampling_offsets || !attn_logits ||
        !grad_output || !grad_value || !grad_sampling_offsets || !grad_attn_logits)
        return MSDA_ERR_NULL;   // grad_reference_points may be NULL (reference points without grad)
    if (!aligned16(value) || !aligned16(grad_output) || !aligned16(grad_value)) return MSDA_ERR_ALIGN;
    const int lpg = fast_lpg(D, dtype);
    if (!lpg) return MSDA_ERR_UNSUPPORTED;
    const size_t need = required_workspace(N, S, M, D, dtype);
    if (need && (!workspace || workspace_bytes < need || !aligned16(workspace))) return MSDA_ERR_WORKSPACE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long nv = (long long)N * S * M * D;
    void *acc = dtype == MSDA_BF16 ? workspace : grad_value;
    if (int e = memset_async(acc, (size_t)nv * 4, st)) return e;
    if (const int hl = small_lanes(d, dtype, flags, true)) {
        // the reference-point gradient is summed inside the CTA of its query: no zero-fill, no atomics
        TileArgs t{};
        t.value = value; t.shapes = temporal_shapes; t.lsi = level_start_index; t.shapes_stride = 1;
        t.ref = reference_points; t.offs = sampling_offsets; t.logits = attn_logits; t.R = R;
        t.grad_out = grad_output; t.gv_acc = static_cast<float *>(acc);
        t.grad_ref = grad_reference_points; t.grad_offs = grad_sampling_offsets; t.grad_logits = grad_attn_logits;
        const int src = dtype == MSDA_F32 ? launch_small<float, true, true>(t, d, hl, pad_mode, st)
                                          : launch_small<__nv_bfloat16, true, true>(t, d, hl, pad_mode, st);
        if (src) return src;
        if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
        return MSDA_OK;
    }
    if (grad_reference_points)
        if (int e = memset_async(grad_reference_points, (size_t)N * Lq * L * R * 4, st)) return e;
    {
        const LeanPlan lp = lean_plan(d, dtype, true);
        const TilePlan tp = lp.ok ? TilePlan{} : tile_plan(d, dtype, true);
        if (lp.ok || tp.ok) {
            TileArgs t{};
            t.value = value; t.shapes = temporal_shapes; t.lsi = level_start_index; t.shapes_stride = 1;
            t.ref = reference_points; t.offs = sampling_offsets; t.logits = attn_logits; t.R = R;
            t.grad_out = grad_output; t.gv_acc = static_cast<float *>(acc);
            t.grad_ref = grad_reference_points; t.grad_offs = grad_sampling_offsets; t.grad_logits = grad_attn_logits;
            int trc;
            if (lp.ok) {
                fill_lean_args(t, d, lp, pad_mode);
                trc = launch_lean<true, true>(t, lp, dtype, st);
            } else {
                fill_tile_args(t, d, tp, pad_mode);
                trc = launch_tile<true, true>(t, tp, dtype, st);
            }
            if (trc) return trc;
            if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
            return MSDA_OK;
        }
    }
    FastArgs a{};
    a.value = value; a.shapes = temporal_shapes; a.lsi = level_start_index; a.shapes_stride = 1;
    a.ref = reference_points; a.offs = sampling_offsets; a.logits = attn_logits; a.R = R;
    a.grad_out = grad_output; a.gv_acc = static_cast<float *>(acc);
    a.grad_ref = grad_reference_points; a.grad_offs = grad_sampling_offsets; a.grad_logits = grad_attn_logits;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
    a.total_units = N * Lq * M;
    int grid;
    fast_grid(a.total_units, lpg, grid, a.units_per_cta);
    a.head_major = head_major_env();
    const int rc = dtype == MSDA_F32 ? launch_fast_bwd<float, true>(a, lpg, grid, st)
                                     : launch_fast_bwd<__nv_bfloat16, true>(a, lpg, grid, st);
    if (rc) return rc;
    if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
    return MSDA_OK;
}

int msda_sample_forward(const void *value, const int64_t *spatial_shapes, const int64_t *level_start_index,
                        const void *sampling_loc, void *samples, int N, int S, int M, int D, int L, int Lq, int P,
                        int dtype, int pad_mode, unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (!value || !spatial_shapes || !level_start_index || !sampling_loc || !samples) return MSDA_ERR_NULL;
    GenericArgs a{};
    a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
    a.loc = sampling_loc; a.out = samples;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
    a.assume_1d = (flags & MSDA_FLAG_ASSUME_1D) ? 1 : 0;
    a.total = (long long)N * Lq * M * D;
    return launch_generic_fwd<true>(a, dtype, static_cast<cudaStream_t>(stream));
}

int msda_cap_sample_forward(const void *value, const int64_t *temporal_shapes, const int64_t *level_start_index,
                            const float *reference_points, const float *sampling_offsets, void *samples, int N, int S,
                            int M, int D, int L, int Lq, int P, int R, int dtype, int pad_mode, unsigned flags, void *stream)
{
    (void)flags;
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (R != 1 && R != 2) return MSDA_ERR_SHAPE;
    if (!value || !temporal_shapes || !level_start_index || !reference_points || !sampling_offsets || !samples) return MSDA_ERR_NULL;
    if (!aligned16(value) || !aligned16(samples)) return MSDA_ERR_ALIGN;
    const int ch = dtype == MSDA_BF16 ? 8 : 4;
    if (dtype == MSDA_F64 || D % ch) return MSDA_ERR_UNSUPPORTED;
    CapArgs a{};
    a.value = value; a.shapes = temporal_shapes; a.lsi = level_start_index; a.ref = reference_points;
    a.off = sampling_offsets; a.samples = samples;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.R = R; a.pad_mode = pad_mode;
    a.total = (long long)N * Lq * M * L * P * (D / ch);
    long long blocks = (a.total + 255) / 256;
    const int cap = num_sms() * 8;
    if (blocks > cap) blocks = cap;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == MSDA_F32) msda_cap_fwd_kernel<float><<<(int)blocks, 256, 0, st>>>(a);
    else msda_cap_fwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>(a);
    return check_launch();
}

static int attend_check(int BM, int M, int K, int D, int A)
{
    if (BM <= 0 || M <= 0 || BM % M || K <= 0 || K > kAttendMaxK || D <= 0 || A <= 0) return MSDA_ERR_SHAPE;
    if ((D & 3) || (A & 3) || D > 3072 || A > 1536) return MSDA_ERR_UNSUPPORTED;      // float4 rows, <= 48 KB of partial sums
    return 0;
}

int msda_cap_attend_forward_f32(const float *samples, const float *hproj, const float *alpha, const float *alpha_bias,
                                float *attended, float *weight, int BM, int M, int K, int D, int A, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (int e = attend_check(BM, M, K, D, A)) return e;
    if (!samples || !hproj || !alpha || !alpha_bias || !attended || !weight) return MSDA_ERR_NULL;
    AttendArgs a{};
    a.samples = samples; a.hproj = hproj; a.alpha = alpha; a.alpha_bias = alpha_bias; a.attended = attended; a.weight = weight;
    a.BM = BM; a.M = M; a.K = K; a.D = D; a.A = A;
    if (!aligned16(samples) || !aligned16(hproj) || !aligned16(alpha) || !aligned16(attended)) return MSDA_ERR_ALIGN;
    msda_cap_attend_fwd_kernel<<<BM, kAttendThreads, sizeof(float) * kAttendKG * D, static_cast<cudaStream_t>(stream)>>>(a);
    return check_launch();
}

int msda_cap_attend_backward_f32(const float *samples, const float *hproj, const float *alpha, const float *weight,
                                 const float *grad_attended, float *grad_samples, float *grad_hproj, float *grad_alpha,
                                 int BM, int M, int K, int D, int A, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (int e = attend_check(BM, M, K, D, A)) return e;
    if (!samples || !hproj || !alpha || !weight || !grad_attended || !grad_samples || !grad_hproj || !grad_alpha) return MSDA_ERR_NULL;
    AttendArgs a{};
    a.samples = samples; a.hproj = hproj; a.alpha = alpha; a.weight = const_cast<float *>(weight); a.grad_att = grad_attended;
    a.grad_samples = grad_samples; a.grad_hproj = grad_hproj; a.grad_alpha = grad_alpha;
    a.BM = BM; a.M = M; a.K = K; a.D = D; a.A = A;
    if (!aligned16(samples) || !aligned16(hproj) || !aligned16(alpha) || !aligned16(grad_attended) || !aligned16(grad_samples) ||
        !aligned16(grad_hproj))
        return MSDA_ERR_ALIGN;
    msda_cap_attend_bwd_kernel<<<BM, kAttendThreads, sizeof(float) * 2 * kAttendKG * A, static_cast<cudaStream_t>(stream)>>>(a);
    return check_launch();
}

int msda_skinny_linear_f32(const float *x, const float *weight, const float *bias, float *y, int M, int N, int K,
                           void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (M <= 0 || N <= 0 || K <= 0 || (K & 3)) return MSDA_ERR_SHAPE;
    if (!x || !weight || !y) return MSDA_ERR_NULL;
    if (!aligned16(x) || !aligned16(weight)) return MSDA_ERR_ALIGN;
    const int K4 = K >> 2;
    const int rows = M > 8 ? 16 : 8;
    const int mtiles = (M + rows - 1) / rows;
    // K split: every lane of every slice gets a float4, and no more than ~4 CTAs per SM (each CTA should still have work)
    int ks = K4 >= 256 ? 8 : K4 >= 128 ? 4 : K4 >= 64 ? 2 : 1;
    const int sms = num_sms();
    while (ks > 1 && (long long)((N + (kSkinnyWarps / ks) * kSkinnyNW - 1) / ((kSkinnyWarps / ks) * kSkinnyNW)) * mtiles > 4 * sms)
        ks >>= 1;
    const int per_cta = (kSkinnyWarps / ks) * kSkinnyNW;
    const dim3 grid((N + per_cta - 1) / per_cta, mtiles);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
#define MSDA_SKINNY(KS_, ROWS_) msda_skinny_linear_kernel<KS_, ROWS_><<<grid, kSkinnyWarps * 32, 0, st>>>(x, weight, bias, y, M, N, K)
    if (rows == 16) {
        switch (ks) {
        case 8: MSDA_SKINNY(8, 16); break;
        case 4: MSDA_SKINNY(4, 16); break;
        case 2: MSDA_SKINNY(2, 16); break;
        default: MSDA_SKINNY(1, 16); break;
        }
    } else {
        switch (ks) {
        case 8: MSDA_SKINNY(8, 8); break;
        case 4: MSDA_SKINNY(4, 8); break;
        case 2: MSDA_SKINNY(2, 8); break;
        default: MSDA_SKINNY(1, 8); break;
        }
    }
#undef MSDA_SKINNY
    return check_launch();
}

int msda_lstm_cell_forward_f32(const float *pre, long long pre_row_stride, const float *rec, const float *c_prev,
                               float *c, float *h, float *act, int B, int H, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (B <= 0 || H <= 0 || pre_row_stride < 4LL * H) return MSDA_ERR_SHAPE;
    if (!pre || !rec || !c_prev || !c || !h || !act) return MSDA_ERR_NULL;
    msda_lstm_cell_fwd_kernel<<<(B * H + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        pre, pre_row_stride, rec, c_prev, c, h, act, B, H);
    return check_launch();
}

int msda_lstm_cell_backward_f32(const float *grad_h, const float *grad_c, const float *act, const float *c,
                                const float *c_prev, float *grad_gates, float *grad_c_prev, int B, int H, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (B <= 0 || H <= 0) return MSDA_ERR_SHAPE;
    if (!act || !c || !c_prev || !grad_gates || !grad_c_prev) return MSDA_ERR_NULL;
    msda_lstm_cell_bwd_kernel<<<(B * H + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        grad_h, grad_c, act, c, c_prev, grad_gates, grad_c_prev, B, H);
    return check_launch();
}

int msda_cap_sample_backward(const void *value, const int64_t *temporal_shapes, const int64_t *level_start_index,
                             const float *reference_points, const float *sampling_offsets, const void *grad_samples,
                             void *grad_value, float *grad_reference_points, float *grad_sampling_offsets,
                             void *workspace, size_t workspace_bytes, int N, int S, int M, int D, int L, int Lq, int P,
                             int R, int dtype, int pad_mode, unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (R != 1 && R != 2) return MSDA_ERR_SHAPE;
    if (!value || !temporal_shapes || !level_start_index || !reference_points || !sampling_offsets || !grad_samples ||
        !grad_value || !grad_sampling_offsets)
        return MSDA_ERR_NULL;
    // MSDA_FLAG_ACCUMULATE: grad_value / grad_reference_points already hold the sum of earlier calls (the word steps of one
    // decoding run sample the same value tensor around the same reference points): add to them instead of zero-filling
    const bool accumulate = (flags & MSDA_FLAG_ACCUMULATE) != 0;
    if (accumulate && dtype != MSDA_F32) return MSDA_ERR_UNSUPPORTED;
    if (!aligned16(value) || !aligned16(grad_samples) || !aligned16(grad_value)) return MSDA_ERR_ALIGN;
    const int ch = dtype == MSDA_BF16 ? 8 : 4;
    if (dtype == MSDA_F64 || D % ch) return MSDA_ERR_UNSUPPORTED;
    const size_t need = required_workspace(N, S, M, D, dtype);
    if (need && (!workspace || workspace_bytes < need || !aligned16(workspace))) return MSDA_ERR_WORKSPACE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long nv = (long long)N * S * M * D;
    const long long ns = (long long)N * Lq * M * L * P;
    void *acc = dtype == MSDA_BF16 ? workspace : grad_value;
    if (!accumulate)
        if (int e = memset_async(acc, (size_t)nv * 4, st)) return e;
    if (int e = memset_async(grad_sampling_offsets, (size_t)ns * 4, st)) return e;
    if (grad_reference_points && !accumulate)
        if (int e = memset_async(grad_reference_points, (size_t)N * Lq * L * R * 4, st)) return e;
    CapArgs a{};
    a.value = value; a.shapes = temporal_shapes; a.lsi = level_start_index; a.ref = reference_points;
    a.off = sampling_offsets; a.grad_samples = grad_samples;
    a.gv_acc = static_cast<float *>(acc); a.grad_ref = grad_reference_points; a.grad_off = grad_sampling_offsets;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.R = R; a.pad_mode = pad_mode;
    a.total = ns * (D / ch);
    // grid-stride with whole-warp liveness: the grid covers `total` in full rounds of the same stride
    long long blocks = (a.total + 255) / 256;
    const int cap = num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (dtype == MSDA_F32) msda_cap_bwd_kernel<float><<<(int)blocks, 256, 0, st>>>(a);
    else msda_cap_bwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>(a);
    if (int rc = check_launch()) return rc;
    if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
    return MSDA_OK;
}

int msda_sample_backward(const void *value, const int64_t *spatial_shapes, const int64_t *level_start_index,
                         const void *sampling_loc, const void *grad_samples, void *grad_value, void *grad_sampling_loc,
                         void *workspace, size_t workspace_bytes, int N, int S, int M, int D, int L, int Lq, int P,
                         int dtype, int pad_mode, unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    const Dims d{N, S, M, D, L, Lq, P};
    if (int e = check_dims(d, dtype, pad_mode)) return e;
    if (!value || !spatial_shapes || !level_start_index || !sampling_loc || !grad_samples || !grad_value || !grad_sampling_loc)
        return MSDA_ERR_NULL;
    const size_t need = required_workspace(N, S, M, D, dtype);
    if (need && (!workspace || workspace_bytes < need || !aligned16(workspace))) return MSDA_ERR_WORKSPACE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long long nv = (long long)N * S * M * D;
    const long long ns = (long long)N * Lq * M * L * P;
    const size_t ct = dtype == MSDA_F64 ? 8 : 4;
    void *acc = dtype == MSDA_BF16 ? workspace : grad_value;
    if (int e = memset_async(acc, (size_t)nv * ct, st)) return e;
    if (int e = memset_async(grad_sampling_loc, (size_t)ns * 2 * ct, st)) return e;
    GenericArgs a{};
    a.value = value; a.shapes = spatial_shapes; a.lsi = level_start_index; a.shapes_stride = 2;
    a.loc = sampling_loc; a.grad_out = grad_samples;
    a.gv_acc = acc; a.grad_loc = grad_sampling_loc;
    a.N = N; a.S = S; a.M = M; a.D = D; a.L = L; a.Lq = Lq; a.P = P; a.pad_mode = pad_mode;
    a.assume_1d = (flags & MSDA_FLAG_ASSUME_1D) ? 1 : 0;
    a.total = (long long)N * Lq * M * D;
    const int rc = launch_generic_bwd<true>(a, dtype, st);
    if (rc) return rc;
    if (dtype == MSDA_BF16) return finish_bf16(static_cast<const float *>(acc), grad_value, nv, st);
    return MSDA_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// msda_fused_backward_bf16_static: the bf16 fused backward with its fp32 accumulation image in a static device array.
// No counterpart in the reference (see include/msda.h).  In host code `g_static_scratch` is only a handle for the runtime's
// symbol table, not a device address: the address each device maps it to comes from cudaGetSymbolAddress, asked once per
// device with that device current, and that pointer is what the kernels receive.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_static_scratch[MSDA_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
float *g_scratch_ptr[kMaxDevices];   // [device ordinal] -> device address of g_static_scratch there (NULL: not resolved yet)

int static_scratch(float **out)
{
    int dev = -1;
    if (cudaError_t e = cudaGetDevice(&dev)) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return MSDA_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[dev]) {
        void *p = nullptr;
        if (cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch)) return (int)e;
        if (!p || !aligned16(p)) return MSDA_ERR_ALIGN;
        g_scratch_ptr[dev] = static_cast<float *>(p);
    }
    *out = g_scratch_ptr[dev];
    return MSDA_OK;
}

}  // namespace

extern "C" int msda_fused_backward_bf16_static(const void *value, const int64_t *temporal_shapes,
                                               const int64_t *level_start_index, const float *reference_points,
                                               const float *sampling_offsets, const float *attn_logits,
                                               const void *grad_output, void *grad_value, float *grad_reference_points,
                                               float *grad_sampling_offsets, float *grad_attn_logits, int N, int S, int M,
                                               int D, int L, int Lq, int P, int R, int pad_mode, unsigned flags, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (int e = check_dims(Dims{N, S, M, D, L, Lq, P}, MSDA_BF16, pad_mode)) return e;
    if ((unsigned long long)N * S * M * D * sizeof(float) > MSDA_STATIC_SCRATCH_BYTES) return MSDA_ERR_STATIC_SCRATCH;
    float *scratch = nullptr;
    if (int e = static_scratch(&scratch)) return e;
    // the shared path zero-fills the N*S*M*D floats it accumulates into on `stream` first: what an earlier call left is gone
    return msda_fused_backward(value, temporal_shapes, level_start_index, reference_points, sampling_offsets, attn_logits,
                               grad_output, grad_value, grad_reference_points, grad_sampling_offsets, grad_attn_logits,
                               scratch, MSDA_STATIC_SCRATCH_BYTES, N, S, M, D, L, Lq, P, R, MSDA_BF16, pad_mode, flags, stream);
}
