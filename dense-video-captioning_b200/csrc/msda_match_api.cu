// msda_match_api.cu -- extern "C" entry points of the device-side Hungarian matching (include/msda.h: msda_lsap_f32,
// msda_pdvc_match_f32).
#include "msda_match.cuh"

#include "msda_host.cuh"

using namespace msda::host;
using namespace msda::match;

namespace {

template <class Source> int launch_lsap(Args a, const Source &src, int B, int nr, int nc, void *stream)
{
    size_t smem = state_bytes(nr, nc);
    const size_t with_matrix = smem + (size_t)a.R * a.C * sizeof(float);
    if (with_matrix <= 200 * 1024) { a.staged = 1; smem = with_matrix; }
    if (smem > 48 * 1024)
        if (cudaError_t e = cudaFuncSetAttribute(msda_lsap_kernel<Source>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))
            return (int)e;
    msda_lsap_kernel<Source><<<B, kStageThreads, smem, static_cast<cudaStream_t>(stream)>>>(a, src);
    ++g_launches;
    g_path = MSDA_PATH_MATCH;
    return (int)cudaGetLastError();
}

}  // namespace

extern "C" int msda_lsap_f32(const float *cost, long long *row_idx, long long *col_idx, int *status, int B, int R, int C,
                             void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (B < 0 || R < 0 || C < 0) return MSDA_ERR_SHAPE;
    if (B == 0 || R == 0 || C == 0) return MSDA_OK;              // nothing to pair
    if (!cost || !row_idx || !col_idx) return MSDA_ERR_NULL;
    if (R > kMaxDim || C > kMaxDim) return MSDA_ERR_UNSUPPORTED;
    const int nr = R < C ? R : C, nc = R < C ? C : R;
    Args a{const_cast<float *>(cost), row_idx, col_idx, status, R, C, 0};
    return launch_lsap(a, PlainCost{}, B, nr, nc, stream);
}

extern "C" int msda_pdvc_match_f32(const float *pred_logits, const float *pred_boxes, const long long *tgt_labels,
                                   const float *tgt_boxes, float *cost, long long *query_idx, long long *target_idx,
                                   int *status, int layers, int queries, int classes, int targets, double cost_class,
                                   double cost_bbox, double cost_giou, double alpha, double gamma, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (layers < 0 || queries < 0 || targets < 0 || classes < 1) return MSDA_ERR_SHAPE;
    if (layers == 0 || queries == 0 || targets == 0) return MSDA_OK;
    if (!pred_logits || !pred_boxes || !tgt_labels || !tgt_boxes || !cost || !query_idx || !target_idx) return MSDA_ERR_NULL;
    if (queries > kMaxDim || targets > kMaxDim) return MSDA_ERR_UNSUPPORTED;
    const int nr = queries < targets ? queries : targets, nc = queries < targets ? targets : queries;
    Args a{cost, query_idx, target_idx, status, queries, targets, 0};
    PdvcCost src{pred_logits, pred_boxes, tgt_labels, tgt_boxes, classes, queries, (float)cost_class, (float)cost_bbox, (float)cost_giou,
                 (float)alpha, (float)(1.0 - alpha), (float)gamma};
    return launch_lsap(a, src, layers, nr, nc, stream);
}
