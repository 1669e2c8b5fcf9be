// msda_match_api.cu -- extern "C" entry point of the device-side Hungarian matching (include/msda.h: msda_lsap_f32).
#include "msda_match.cuh"

#include "msda_host.cuh"

using namespace msda::host;
using namespace msda::match;

extern "C" int msda_lsap_f32(const float *cost, long long *row_idx, long long *col_idx, int *status, int B, int R, int C,
                             void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (B < 0 || R < 0 || C < 0) return MSDA_ERR_SHAPE;
    if (B == 0 || R == 0 || C == 0) return MSDA_OK;              // nothing to pair
    if (!cost || !row_idx || !col_idx) return MSDA_ERR_NULL;
    if (R > kMaxDim || C > kMaxDim) return MSDA_ERR_UNSUPPORTED;
    const int nr = R < C ? R : C, nc = R < C ? C : R;
    Args a{cost, row_idx, col_idx, status, R, C, 0};
    size_t smem = state_bytes(nr, nc);
    const size_t with_matrix = smem + (size_t)R * C * sizeof(float);
    if (with_matrix <= 200 * 1024) { a.staged = 1; smem = with_matrix; }
    if (smem > 48 * 1024)
        if (cudaError_t e = cudaFuncSetAttribute(msda_lsap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))
            return (int)e;
    msda_lsap_kernel<<<B, 32, smem, static_cast<cudaStream_t>(stream)>>>(a);
    ++g_launches;
    g_path = MSDA_PATH_MATCH;
    return (int)cudaGetLastError();
}
