/*
 * include/msda.h -- C ABI of libmsda_sm100.so, the B200 (sm_100a) replacement for the reference's
 * `MultiScaleDeformableAttention` pybind11/ATen extension (pdvc/ops/setup.py:52-61).
 *
 * Conventions (all entry points):
 *   - plain device pointers and sizes; no torch / ATen types.  Every tensor is dense,
 *     row-major ("contiguous" in the reference's asserts, ms_deform_attn_cuda.cu:28-32) and
 *     16-byte aligned at its base.
 *   - the CALLER owns and allocates every buffer (outputs need no pre-zeroing; the library
 *     zero-fills what it accumulates into).  The library allocates nothing, keeps no pointer
 *     after return, reads no environment variable and holds no mutable global state beyond two
 *     thread-local diagnostic integers (msda_last_launch_count / _path) and a cache of immutable
 *     device properties (SM count, shared-memory opt-in done).
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*; the reference uses
 *     at::cuda::getCurrentCUDAStream(), ms_deform_attn_cuda.cu:65,:135).  No entry point
 *     synchronises the host; all are CUDA-graph capturable.
 *   - return value: 0 on success; > 0 a cudaError_t from the launch (the reference only
 *     printf()s these, ms_deform_im2col_cuda.cuh:949-953); < 0 an MSDA_ERR_* argument error
 *     (the reference raises through AT_ASSERTM, ms_deform_attn_cuda.cu:28-52).
 *     msda_error_string() maps either to text.
 *   - `spatial_shapes` (L,2) int64 [H_l, W_l] and `level_start_index` (L,) int64 live in DEVICE
 *     memory exactly as the reference passes them (ms_deform_attn_cuda.cu:66-67) and are read by
 *     the kernels; no host copy is made.
 *
 * dtype: element type of value / output / grad_output / grad_value.
 *   MSDA_F32 : everything float32.
 *   MSDA_BF16: value/output/grad_output/grad_value bfloat16; sampling locations, attention
 *              weights, offsets, logits, reference points and their gradients stay float32;
 *              accumulation is float32.
 *   MSDA_F64 : everything float64 (generic kernels only; lets the reference's gradcheck,
 *              pdvc/ops/test.py:63-78, run against this library).
 */
#ifndef MSDA_H_
#define MSDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSDA_VERSION 100 /* 0.1.0 */

enum { MSDA_F32 = 0, MSDA_BF16 = 1, MSDA_F64 = 2 };

/* Out-of-range handling of sampling locations.
 * BORDER = the reference's PyTorch path (F.grid_sample padding_mode='border',
 *          pdvc/ops/functions/ms_deform_attn_func.py:58-59) -- the parity oracle.
 * ZEROS  = the reference's CUDA kernel (ms_deform_im2col_cuda.cuh:286-289, :39-84). */
enum { MSDA_PAD_BORDER = 0, MSDA_PAD_ZEROS = 1 };

/* flags */
#define MSDA_FLAG_ASSUME_1D 1u   /* caller guarantees H_l == 1 and sampling_loc[...,1] == 0.5 for every
                                    sample (PDVC always: pdvc/ops/modules/ms_deform_attn.py:114-117).
                                    sampling_loc[...,1] is not read; grad_sampling_loc[...,1] is what the
                                    reference computes for y = 0.5: exactly 0 under BORDER, and
                                    -attn * grad_attn under ZEROS (ms_deform_im2col_cuda.cuh:159) */
#define MSDA_FLAG_FORCE_GENERIC 2u /* testing: take the generic (any-D, 2-D capable) kernels */
#define MSDA_FLAG_ACCUMULATE 8u    /* msda_cap_sample_backward, f32: add into grad_value / grad_reference_points instead of
                                      overwriting them (the caller zero-fills once per decoding run; csrc/msda_api.cu) */
#define MSDA_FLAG_NO_SMALL 4u      /* testing: never take the single-phase kernels for small problems
                                      (csrc/msda_small.cuh), so that the staged kernels can be checked on small inputs */

enum {
    MSDA_OK = 0,
    MSDA_ERR_DTYPE = -1,      /* unknown dtype code / dtype not supported by this entry point */
    MSDA_ERR_SHAPE = -2,      /* a dimension is <= 0 or exceeds a documented limit */
    MSDA_ERR_NULL = -3,       /* a required pointer is NULL */
    MSDA_ERR_ALIGN = -4,      /* a base pointer is not 16-byte aligned */
    MSDA_ERR_WORKSPACE = -5,  /* workspace smaller than msda_backward_workspace_bytes() */
    MSDA_ERR_PAD = -6,        /* unknown pad_mode */
    MSDA_ERR_UNSUPPORTED = -7,/* combination not implemented (e.g. 2-D shapes with bf16 fast path) */
    MSDA_ERR_STATIC_SCRATCH = -8 /* msda_fused_backward_bf16_static: N*S*M*D*4 > MSDA_STATIC_SCRATCH_BYTES */
};

#define MSDA_MAX_LEVELS 32

int msda_version(void);
/* "sm_100a" -- the only architecture this library contains code for. */
const char *msda_arch(void);
const char *msda_error_string(int code);

/* ---------------------------------------------------------------------------------------------
 * Reference-equivalent entry points.
 * Replace  ms_deform_attn_forward / ms_deform_attn_backward  exported by the reference module
 * (pdvc/ops/src/vision.cpp:14-15; signatures pdvc/ops/src/ms_deform_attn.h:20-27, :41-49;
 * CUDA side pdvc/ops/src/cuda/ms_deform_attn_cuda.cu:20-80, :83-153).
 *
 *   value              (N, S, M, D)
 *   spatial_shapes     (L, 2) int64, device
 *   level_start_index  (L,)   int64, device
 *   sampling_loc       (N, Lq, M, L, P, 2)   [...,0] = x along W, [...,1] = y along H
 *   attn_weight        (N, Lq, M, L, P)
 *   output             (N, Lq, M*D)
 * `im2col_step` of the reference has no counterpart: one launch covers any N (a superset of the
 * reference's `batch % min(batch, im2col_step) == 0` rule, ms_deform_attn_cuda.cu:50-52).
 * ------------------------------------------------------------------------------------------- */
int msda_ms_deform_attn_forward(const void *value, const int64_t *spatial_shapes,
                                const int64_t *level_start_index, const void *sampling_loc,
                                const void *attn_weight, void *output, int N, int S, int M, int D,
                                int L, int Lq, int P, int dtype, int pad_mode, unsigned flags,
                                void *stream);

/* Bytes of scratch msda_ms_deform_attn_backward / msda_fused_backward / msda_sample_backward
 * need for this problem (0 for f32/f64; an fp32 accumulation image of grad_value for bf16). */
size_t msda_backward_workspace_bytes(int N, int S, int M, int D, int dtype);

/*   grad_output        (N, Lq, M*D)
 *   grad_value         (N, S, M, D)           fully overwritten
 *   grad_sampling_loc  (N, Lq, M, L, P, 2)    fully overwritten
 *   grad_attn_weight   (N, Lq, M, L, P)       fully overwritten
 * (the reference allocates these with at::zeros_like, ms_deform_attn_cuda.cu:121-123). */
int msda_ms_deform_attn_backward(const void *value, const int64_t *spatial_shapes,
                                 const int64_t *level_start_index, const void *sampling_loc,
                                 const void *attn_weight, const void *grad_output, void *grad_value,
                                 void *grad_sampling_loc, void *grad_attn_weight, void *workspace,
                                 size_t workspace_bytes, int N, int S, int M, int D, int L, int Lq,
                                 int P, int dtype, int pad_mode, unsigned flags, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Fused 1-D temporal path: replaces, in one kernel each way, the softmax over levels x points,
 * the sampling-location arithmetic, the dummy-y stacking and the core op of
 * MSDeformAttn.forward (pdvc/ops/modules/ms_deform_attn.py:100-122).
 *
 *   value              (N, S, M, D)          output of value_proj, padding already zeroed
 *   temporal_shapes    (L,) int64, device    T_l  (the module's input_spatial_shapes, :84)
 *   level_start_index  (L,) int64, device
 *   reference_points   (N, Lq, L, R) f32, R in {1, 2}          (:81-82, branches :103-109)
 *   sampling_offsets   (N, Lq, M, L, P) f32  raw output of the sampling_offsets Linear (:99)
 *   attn_logits        (N, Lq, M, L, P) f32  raw output of the attention_weights Linear (:100)
 *   output             (N, Lq, M*D)
 * Locations: R=1: x = ref0 + off / T_l ; R=2: x = ref0 + off / P * ref1 * 0.5.
 * Weights:   softmax over the flattened (L*P) axis.
 * ------------------------------------------------------------------------------------------- */
int msda_fused_forward(const void *value, const int64_t *temporal_shapes,
                       const int64_t *level_start_index, const float *reference_points,
                       const float *sampling_offsets, const float *attn_logits, void *output, int N,
                       int S, int M, int D, int L, int Lq, int P, int R, int dtype, int pad_mode,
                       unsigned flags, void *stream);

/*   grad_reference_points (N, Lq, L, R) f32, grad_sampling_offsets / grad_attn_logits
 *   (N, Lq, M, L, P) f32, grad_value (N, S, M, D); all fully overwritten.  The softmax and the
 *   locations are recomputed from the raw inputs; nothing is saved by the forward. */
int msda_fused_backward(const void *value, const int64_t *temporal_shapes,
                        const int64_t *level_start_index, const float *reference_points,
                        const float *sampling_offsets, const float *attn_logits,
                        const void *grad_output, void *grad_value, float *grad_reference_points,
                        float *grad_sampling_offsets, float *grad_attn_logits, void *workspace,
                        size_t workspace_bytes, int N, int S, int M, int D, int L, int Lq, int P,
                        int R, int dtype, int pad_mode, unsigned flags, void *stream);

/* msda_fused_backward for bf16 values WITHOUT a caller-provided workspace and without any allocation.
 * No counterpart in the reference: added for callers
 * that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The fp32 accumulation image of grad_value that the bf16 backward needs lives in a
 * file-scope __device__ array of MSDA_STATIC_SCRATCH_BYTES (16 MiB) inside the library; every GPU that
 * loads the library holds that much until the process exits.
 *   - arguments and results as msda_fused_backward with dtype = MSDA_BF16 (value, grad_output,
 *     grad_value bf16; everything else f32); inputs and outputs are still the caller's buffers;
 *   - N*S*M*D*4 > MSDA_STATIC_SCRATCH_BYTES returns MSDA_ERR_STATIC_SCRATCH and launches nothing
 *     (N*S*M*D <= 4 194 304: e.g. 16 videos of the YouCook2 encoder memory, S = 375, M*D = 512);
 *   - calls no cudaMalloc / cudaFree / allocating library routine.  The first call on a device
 *     resolves the array's device address (cudaGetSymbolAddress, that device current) and caches it;
 *   - the array is shared by every call on the device and keeps the previous call's contents: each
 *     call zero-fills the part it uses on `stream` before its kernels.  Calls that overlap on the
 *     GPU (different streams, no ordering between them) race on it -- serialise them. */
#define MSDA_STATIC_SCRATCH_BYTES (16u << 20)
int msda_fused_backward_bf16_static(const void *value, const int64_t *temporal_shapes,
                                    const int64_t *level_start_index, const float *reference_points,
                                    const float *sampling_offsets, const float *attn_logits,
                                    const void *grad_output, void *grad_value,
                                    float *grad_reference_points, float *grad_sampling_offsets,
                                    float *grad_attn_logits, int N, int S, int M, int D, int L, int Lq,
                                    int P, int R, int pad_mode, unsigned flags, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Sampler: the return_value=True branch of ms_deform_attn_core_pytorch
 * (pdvc/ops/functions/ms_deform_attn_func.py:64-65) used by MSDeformAttnCap
 * (pdvc/ops/modules/ms_deform_attn_for_caption.py:120-121): un-weighted sampled values.
 *   samples       (N*M, D, Lq, L, P)
 *   grad_samples  same shape;  grad_value (N,S,M,D), grad_sampling_loc (N,Lq,M,L,P,2) overwritten.
 * ------------------------------------------------------------------------------------------- */
int msda_sample_forward(const void *value, const int64_t *spatial_shapes,
                        const int64_t *level_start_index, const void *sampling_loc, void *samples,
                        int N, int S, int M, int D, int L, int Lq, int P, int dtype, int pad_mode,
                        unsigned flags, void *stream);

int msda_sample_backward(const void *value, const int64_t *spatial_shapes,
                         const int64_t *level_start_index, const void *sampling_loc,
                         const void *grad_samples, void *grad_value, void *grad_sampling_loc,
                         void *workspace, size_t workspace_bytes, int N, int S, int M, int D, int L,
                         int Lq, int P, int dtype, int pad_mode, unsigned flags, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Fused caption sampler: everything MSDeformAttnCap.forward does between its Linear projections and
 * its return for the 1-D temporal case (pdvc/ops/modules/ms_deform_attn_for_caption.py:100-123:
 * sampling locations from reference points + raw offsets, dummy y, ms_deform_attn_core_pytorch with
 * return_value=True) in one kernel each way.  f32 / bf16, D a multiple of 4 (f32) or 8 (bf16).
 *   reference_points (N,Lq,L,R) f32, R in {1,2};  sampling_offsets (N,Lq,M,L,P) f32 (raw Linear output)
 *   samples / grad_samples  (N, Lq, M, L*P, D)  -- D contiguous; the reference's (N*M, D, Lq, L, P)
 *                           is the permutation (0,2,4,1,3) of this buffer
 *   grad_value (N,S,M,D), grad_sampling_offsets, grad_reference_points (NULL to skip) overwritten.
 *   workspace: msda_backward_workspace_bytes(N,S,M,D,dtype) bytes (bf16 only).
 * ------------------------------------------------------------------------------------------- */
int msda_cap_sample_forward(const void *value, const int64_t *temporal_shapes,
                            const int64_t *level_start_index, const float *reference_points,
                            const float *sampling_offsets, void *samples, int N, int S, int M, int D,
                            int L, int Lq, int P, int R, int dtype, int pad_mode, unsigned flags,
                            void *stream);

int msda_cap_sample_backward(const void *value, const int64_t *temporal_shapes,
                             const int64_t *level_start_index, const float *reference_points,
                             const float *sampling_offsets, const void *grad_samples,
                             void *grad_value, float *grad_reference_points,
                             float *grad_sampling_offsets, void *workspace, size_t workspace_bytes,
                             int N, int S, int M, int D, int L, int Lq, int P, int R, int dtype,
                             int pad_mode, unsigned flags, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Soft attention over the L*P deformable samples of an event in the caption head
 * (pdvc/CaptioningHead/LSTM_DSA.py:247-257): score_k = alpha . tanh(ctx_k + hproj) + alpha_bias,
 * weight = softmax_k(score), attended = sum_k weight_k clip_k.  All f32.
 *   samples (BM, K, D+A): per sample the D clip channels followed by the A channels of ctx2att(clip)
 *                         (the sampler run on [value | ctx2att(value)]); BM = events x heads, K <= 64,
 *                         D % 4 == 0 <= 3072, A % 4 == 0 <= 1536, 16-byte aligned buffers
 *   hproj (BM/M, A) = h2att(h);  alpha (A,), alpha_bias (1,) = alpha_net
 *   forward : attended (BM, D), weight (BM, K)
 *   backward: grad_samples (BM, K, D+A) overwritten; grad_hproj (BM, A) (sum over the M heads is the
 *             caller's); grad_alpha (BM, A+1): per-(event, head) partials of d alpha and d alpha_bias.
 * ------------------------------------------------------------------------------------------- */
int msda_cap_attend_forward_f32(const float *samples, const float *hproj, const float *alpha,
                                const float *alpha_bias, float *attended, float *weight, int BM, int M,
                                int K, int D, int A, void *stream);

int msda_cap_attend_backward_f32(const float *samples, const float *hproj, const float *alpha,
                                 const float *weight, const float *grad_attended, float *grad_samples,
                                 float *grad_hproj, float *grad_alpha, int BM, int M, int K, int D,
                                 int A, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Linear layer with a few input rows (the caption head's per-word projections):
 *   y[m,n] = sum_k x[m,k] * weight[n,k] (+ bias[n]);  x (M,K), weight (N,K), y (M,N), all f32,
 *   K % 4 == 0, x and weight 16-byte aligned, bias NULL to skip.  Replaces, for the word loop,
 *   nn.Linear `sampling_offsets` (pdvc/ops/modules/ms_deform_attn_for_caption.py:100) and `h2att`
 *   (pdvc/CaptioningHead/LSTM_DSA.py:249); the data gradient is the same call on the transposed weight.
 * ------------------------------------------------------------------------------------------- */
int msda_skinny_linear_f32(const float *x, const float *weight, const float *bias, float *y, int M,
                           int N, int K, void *stream);

/* ---------------------------------------------------------------------------------------------
 * LSTM cell of the caption head's word step: nn.LSTM with one layer run for one time step
 * (pdvc/CaptioningHead/LSTM_DSA.py:206, :261), gate order i,f,g,o as in torch's weight layout.
 *   gates = pre + rec, both (B,4H) f32; pre has row stride pre_row_stride (>= 4H) elements.
 *   forward : c = f*c_prev + i*g, h = o*tanh(c); act (B,4H) receives the activated gates.
 *   backward: grad_h / grad_c (B,H), either may be NULL (= 0) -> grad_gates (B,4H) (the gradient of
 *             pre and of rec alike) and grad_c_prev (B,H).
 * ------------------------------------------------------------------------------------------- */
int msda_lstm_cell_forward_f32(const float *pre, long long pre_row_stride, const float *rec,
                               const float *c_prev, float *c, float *h, float *act, int B, int H,
                               void *stream);

int msda_lstm_cell_backward_f32(const float *grad_h, const float *grad_c, const float *act,
                                const float *c, const float *c_prev, float *grad_gates,
                                float *grad_c_prev, int B, int H, void *stream);

/* ---------------------------------------------------------------------------------------------
 * The Linear projections of MSDeformAttn (pdvc/ops/modules/ms_deform_attn.py:95-100, :125) on the
 * tcgen05 tensor cores, fp32 accumulation in tensor memory.  dtype MSDA_F32: 3xTF32 split products
 * (fp32-level accuracy); MSDA_BF16: bf16 x bf16 -> fp32.
 * One call covers ONE Linear layer or TWO that share their input (the sampling_offsets and
 * attention_weights layers, :99-100: "segments" 0 and 1 with N0 / N1 output channels; N1 = 0 and
 * NULL pointers for a single layer).
 *   x (rows, K), w_s (N_s, K), b_s (N_s) fp32 or NULL, y_s / dy_s (rows, N_s), dx (rows, K);
 *   dw_s (N_s, K) and db_s (N_s) are ALWAYS fp32 (gradients of the fp32 master parameters);
 *   row_mask (rows) bytes or NULL, non-zero = padded frame: forward writes a zero row of y_0 (the
 *   masked_fill of :97-98), the gradient entry points treat that row of dy_0 as zero (segment 1 is
 *   never masked).
 * Limits: K, N0, N1 multiples of 8; N0 a multiple of 128 when N1 > 0 (else MSDA_ERR_UNSUPPORTED and
 * the caller uses its library GEMM).  All outputs fully overwritten; no workspace.
 * ------------------------------------------------------------------------------------------- */
int msda_proj_forward(const void *x, const void *w0, const float *b0, const void *w1, const float *b1,
                      const unsigned char *row_mask, void *y0, void *y1, int rows, int K, int N0,
                      int N1, int dtype, void *stream);
int msda_proj_dgrad(const void *dy0, const void *dy1, const void *w0, const void *w1,
                    const unsigned char *row_mask, void *dx, int rows, int K, int N0, int N1,
                    int dtype, void *stream);
int msda_proj_wgrad(const void *dy0, const void *dy1, const void *x, const unsigned char *row_mask,
                    float *dw0, float *db0, float *dw1, float *db1, int rows, int K, int N0, int N1,
                    int dtype, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Hungarian matching of the set-prediction loss on the device: replaces the host round
 * trip of pdvc/matcher.py:113-117 -- `C.view(bs, num_queries, -1).cpu()` followed by
 * scipy.optimize.linear_sum_assignment per cost matrix -- the one point of a PDVC training step where
 * the host waits for the device.
 *   cost      (B, R, C) f32, device   B independent matrices (decoder layers); R queries, C events
 *   row_idx   (B, min(R,C)) int64     matched rows in ascending order   (scipy's row_ind)
 *   col_idx   (B, min(R,C)) int64     their columns                      (scipy's col_ind)
 *   status    (B) int32 or NULL       0 ok; 1 a NaN or -inf entry; 2 infeasible (a row with only +inf) --
 *                                     scipy raises ValueError for both; here the pairing written is the
 *                                     identity and the caller reads the flag when it next synchronises.
 * Same algorithm, arithmetic (float64 duals), scan order and tie rule as scipy's rectangular_lsap, so
 * the assignment is scipy's also when costs tie.  max(R, C) <= 1024, else MSDA_ERR_UNSUPPORTED (the
 * Python wrapper then takes the host path).  One warp per matrix; enqueued on `stream`, no
 * synchronisation, graph-capturable.
 * ------------------------------------------------------------------------------------------- */
int msda_lsap_f32(const float *cost, long long *row_idx, long long *col_idx, int *status, int B, int R,
                  int C, void *stream);

/* The whole HungarianMatcher.forward of one video (pdvc/matcher.py:71-117) for all decoder layers in one
 * launch: the cost matrix  cost_class * (focal_pos - focal_neg)(sigmoid(logit)) + cost_bbox * L1 -
 * cost_giou * GIoU  (matcher.py:84-103; ~45 tensor kernels per layer in the reference) is formed inside
 * the matching kernel -- each float32 operation as the reference's expression performs it, no
 * contraction -- written to `cost`, and solved as msda_lsap_f32 does.
 *   pred_logits (layers, queries, classes) f32     pred_boxes (layers, queries, 2) f32 (centre, length)
 *   tgt_labels  (targets) int64                     tgt_boxes  (targets, 2) f32
 *   cost        (layers, queries, targets) f32 out  query_idx / target_idx (layers, min(queries,targets)) int64
 *   status as msda_lsap_f32.  gamma == 2 is evaluated as x*x (torch's pow fast path). */
int msda_pdvc_match_f32(const float *pred_logits, const float *pred_boxes, const long long *tgt_labels,
                        const float *tgt_boxes, float *cost, long long *query_idx, long long *target_idx,
                        int *status, int layers, int queries, int classes, int targets, double cost_class,
                        double cost_bbox, double cost_giou, double alpha, double gamma, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Gradient-norm clipping + Adam in two launches: replaces
 * `torch.nn.utils.clip_grad_norm_(model.parameters(), opt.grad_clip)` + `optimizer.step()` of
 * torch.optim.Adam(lr, weight_decay) at the end of the reference's step (train.py:185-187, :123) for a
 * trainer that keeps its gradients in flat buckets.
 *   segments  device array of n_segments { const float *grad; int64 count }   (16 bytes each): the
 *             buckets whose global L2 norm is clipped to max_norm (max_norm <= 0: no clipping,
 *             segments / partials unused)
 *   chunks    device array of n_chunks { float *param; const float *grad; float *exp_avg;
 *             float *exp_avg_sq; int32 count; int32 pad }                      (40 bytes each): runs of
 *             one parameter tensor, any length (8192 balances the blocks)
 *   partials  device scratch of partials_capacity floats (>= 1; 1024 is always enough)
 *   norm_out  device float or NULL: receives the total gradient norm (what clip_grad_norm_ returns)
 *   step      t >= 1, the number of this update (bias corrections 1 - beta^t)
 * Update rule (torch's fused Adam, fp32): g = coef*grad + weight_decay*p; m = lerp(m, g, 1-beta1);
 * v = beta2*v + (1-beta2)*g*g; p -= lr/(1-beta1^t) * m / (sqrt(v)/sqrt(1-beta2^t) + eps), with
 * coef = min(1, max_norm / (norm + 1e-6)).  The hyper-parameters are doubles: 1-beta and the bias
 * corrections are formed in double precision and rounded once, as torch does.  The gradients are read,
 * not rescaled in place.
 * ------------------------------------------------------------------------------------------- */
int msda_adam_clip_step(const void *segments, int n_segments, const void *chunks, int n_chunks,
                        float *partials, int partials_capacity, float *norm_out, double lr,
                        double beta1, double beta2, double eps, double weight_decay, int step,
                        double max_norm, void *stream);

/* Diagnostics (the library's only state: two thread-local integers, reset by every entry point).
 * msda_last_launch_count: number of kernels / memsets the last call on this thread enqueued (bench.py's
 * gpu_launches).  msda_last_launch_path: which kernel family served the last MSDeformAttn call -- the tests use it
 * to prove that every family in csrc/ is reached. */
enum {
    MSDA_PATH_NONE = 0,
    MSDA_PATH_LEAN = 1,    /* csrc/msda_lean.cuh: L*P <= 64, D*sizeof in {64,128,256,512} B */
    MSDA_PATH_TILE = 2,    /* csrc/msda_tile.cuh: L*P <= 256 */
    MSDA_PATH_FAST = 3,    /* csrc/msda_fast.cuh: any 16-byte divisible D */
    MSDA_PATH_GENERIC = 4, /* csrc/msda_generic.cuh: any D, 2-D levels, f64, sampler */
    MSDA_PATH_SMALL = 5,   /* csrc/msda_small.cuh: single-phase kernels for launch-latency-bound problems */
    MSDA_PATH_PROJ = 6,    /* csrc/msda_proj.cuh: tcgen05 projection GEMMs */
    MSDA_PATH_MATCH = 7    /* csrc/msda_match.cuh: linear sum assignment */
};
int msda_last_launch_count(void);
int msda_last_launch_path(void);

#ifdef __cplusplus
}
#endif
#endif /* MSDA_H_ */
