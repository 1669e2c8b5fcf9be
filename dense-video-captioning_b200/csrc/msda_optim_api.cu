// msda_optim_api.cu -- extern "C" entry point of the fused clip + Adam step (include/msda.h: msda_adam_clip_step).
#include <cmath>

#include "msda_optim.cuh"

#include "msda_host.cuh"

using namespace msda::host;
using namespace msda::optim;

static_assert(sizeof(Segment) == 16 && sizeof(Chunk) == 40, "table layouts are part of the C ABI (include/msda.h)");

extern "C" int msda_adam_clip_step(const void *segments, int n_segments, const void *chunks, int n_chunks, float *partials,
                                   int partials_capacity, float *norm_out, double lr, double beta1, double beta2, double eps,
                                   double weight_decay, int step, double max_norm, void *stream)
{
    g_launches = 0; g_path = MSDA_PATH_NONE;
    if (n_segments < 0 || n_chunks < 0) return MSDA_ERR_SHAPE;
    if (n_chunks == 0) return MSDA_OK;
    if (!chunks) return MSDA_ERR_NULL;
    if (step < 1 || !(beta1 >= 0.0 && beta1 < 1.0) || !(beta2 >= 0.0 && beta2 < 1.0)) return MSDA_ERR_SHAPE;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const double bias_correction1 = 1.0 - pow(beta1, (double)step), bias_correction2 = 1.0 - pow(beta2, (double)step);
    Hyper h{(float)(lr / bias_correction1), (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps,
            (float)weight_decay, (float)sqrt(bias_correction2), (float)max_norm};
    int n_partials = 0;
    if (max_norm > 0.f) {
        if (!segments || !partials || n_segments == 0) return MSDA_ERR_NULL;
        n_partials = num_sms() * 4;
        if (n_partials > kMaxPartials) n_partials = kMaxPartials;
        if (n_partials > partials_capacity) n_partials = partials_capacity;
        if (n_partials <= 0) return MSDA_ERR_WORKSPACE;
        msda_sumsq_kernel<<<n_partials, kThreads, 0, st>>>(static_cast<const Segment *>(segments), n_segments, partials);
        ++g_launches;
        if (cudaError_t e = cudaGetLastError()) return (int)e;
    }
    int grid = num_sms() * 8;
    if (grid > n_chunks) grid = n_chunks;
    msda_adam_kernel<<<grid, kThreads, 0, st>>>(static_cast<const Chunk *>(chunks), n_chunks, partials, n_partials, h, norm_out);
    ++g_launches;
    return (int)cudaGetLastError();
}
