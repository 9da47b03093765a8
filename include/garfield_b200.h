/* garfield_b200 — C ABI of the native libraries.
 *
 *   libgarfield_b200_cpu.so   host runtime + CPU aggregation rules   (csrc/cpu, g++)
 *   libgarfield_b200_cuda.so  sm_100a kernels + device runtime       (csrc/cuda, nvcc)
 *
 * Every function returns 0 on success unless stated otherwise; on failure a message is
 * available from gb_cpu_last_error() / gb_cuda_last_error() (thread-local).
 * Pointers named `in` are arrays of n row pointers (one per gradient), never a dense matrix,
 * so inputs can live anywhere (including peer-mapped memory of another GPU).
 * "ref:" comments cite the reference component each entry point is the counterpart of
 * (paths relative to /root/reference/pytorch_impl/libs/native unless absolute).
 */
#ifndef GARFIELD_B200_H
#define GARFIELD_B200_H

#include <stdint.h>

#if defined(__GNUC__)
#define GB_API __attribute__((visibility("default")))
#else
#define GB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ host library --- */

GB_API char const* gb_cpu_last_error(void);
GB_API int gb_cpu_num_threads(void);                         /* ref: so_threadpool/threadpool.cpp:101-106 */

/* Multi-Krum, native semantics. ref: py_krum/krum.cpp:50-118 */
GB_API int gb_cpu_krum_f32(float const* const* in, int n, int64_t d, int f, int m, float* out);
GB_API int gb_cpu_krum_f64(double const* const* in, int n, int64_t d, int f, int m, double* out);
/* Bulyan over Multi-Krum, native semantics. ref: py_bulyan/bulyan.cpp:53-193 */
GB_API int gb_cpu_bulyan_f32(float const* const* in, int n, int64_t d, int f, int m, float* out);
GB_API int gb_cpu_bulyan_f64(double const* const* in, int n, int64_t d, int f, int m, double* out);
/* Non-finite-resilient coordinate-wise median. ref: py_median/median.cpp:42-77 */
GB_API int gb_cpu_median_f32(float const* const* in, int n, int64_t d, float* out);
GB_API int gb_cpu_median_f64(double const* const* in, int n, int64_t d, double* out);
/* Python-semantics coordinate-wise rules for n <= 32 (vectorised sorting network): rule 0 = lower median, NaN-propagating
 * (ref: aggregators/median.py:31-39); 1 = arg-trimmed mean; 2 = mean of the arg values closest to the lower median
 * (ref: tensorflow_impl/rsrcs/aggregators/averaged-median.py:30-67). Returns 0; 1 error; 2 = a column held NaN (rule 1) or a
 * non-finite value (rule 2): recompute with the reference formulation; 3 = n too large for the network. */
GB_API int gb_cpu_coordwise_f32(float const* const* in, int n, int64_t d, int rule, int arg, float* out);
GB_API int gb_cpu_coordwise_f64(double const* const* in, int n, int64_t d, int rule, int arg, double* out);
/* Brute / MDA. ref: py_brute/brute.cpp:47-113 */
GB_API int gb_cpu_brute_f32(float const* const* in, int n, int64_t d, int f, float* out);
GB_API int gb_cpu_brute_f64(double const* const* in, int n, int64_t d, int f, double* out);
/* n×n squared distances (row-major doubles, +inf diagonal). ref: include/operations.hpp:41-55 */
GB_API int gb_cpu_pairwise_sqdist_f32(float const* const* in, int n, int64_t d, double* out);
GB_API int gb_cpu_pairwise_sqdist_f64(double const* const* in, int n, int64_t d, double* out);
/* out = mean of in[sel[0..count)]. ref: include/operations.hpp:64-79 */
GB_API int gb_cpu_selection_average_f32(float const* const* in, int64_t d, int const* sel, int count, float* out);
GB_API int gb_cpu_selection_average_f64(double const* const* in, int64_t d, int const* sel, int count, double* out);
GB_API int gb_cpu_weighted_sum_f32(float const* const* in, int64_t d, int const* sel, float const* w, int count, float* out);

/* k-combinations in lexicographic order. ref: include/combinations.hpp:57-178 */
GB_API uint64_t gb_combinations_count(unsigned n, unsigned k, int* overflow);
GB_API void gb_combinations_seek(unsigned n, unsigned k, uint64_t pos, int* idx);
GB_API int gb_combinations_next(unsigned n, unsigned k, int* idx);   /* 1 = advanced, 0 = was last */

/* Thread-pool coverage probe for tests. ref: include/threadpool.hpp:202-222 */
GB_API int gb_cpu_parallel_for_probe(int64_t total, int64_t grain, int32_t* hits);
/* k-th smallest along `axis` of a row-major [shape] array, in place (std::nth_element over csrc/cpu/strided_view.hpp's
 * strided iterator), one value per line -> out. ref: include/array.hpp (N-D strided view + iterator, bulyan.cpp:167-171). */
GB_API int gb_cpu_axis_nth_f32(float* data, int rank, int64_t const* shape, int axis, int64_t k, float* out);

/* Multi-replica register. ref: include/multiregister.hpp:232-552.
 * write/read return 0 ok, 1 closed, 2 timed out; timeout_ns < 0 waits forever. */
GB_API void* gb_mreg_create(int replicas, int64_t nbytes, int consume);
GB_API void gb_mreg_close(void* handle);
GB_API void gb_mreg_destroy(void* handle);
GB_API int gb_mreg_write(void* handle, void const* src, int64_t timeout_ns);
GB_API int gb_mreg_read(void* handle, void* dst, uint64_t* seq, int64_t timeout_ns);


/* ---------------------------------------------------------------- sm_100a library --- */
/* All device entry points take raw device pointers and a cudaStream_t (as void*); they launch
 * asynchronously on that stream and never synchronise or allocate unless stated. */

GB_API char const* gb_cuda_last_error(void);
GB_API int gb_cuda_device_count(void);
GB_API int gb_cuda_sm_count(void);

/* Peer-memory arena over CUDA IPC (replaces the reference's RPC transport of CPU tensors,
 * ref: /root/reference/pytorch_impl/libs/garfieldpp/server.py:112-184; allocation replaces
 * include/cudarray.cu.hpp:43-190 per-call cudaMalloc). */
GB_API int gb_ipc_handle_size(void);
GB_API int gb_arena_alloc(int64_t bytes, void** ptr);
GB_API int gb_arena_free(void* ptr);
GB_API int gb_ipc_export(void* ptr, void* handle_out);
GB_API int gb_ipc_open(void const* handle, void** ptr);
GB_API int gb_ipc_close(void* ptr);
GB_API int gb_enable_peer(int peer);

/* Optional optimizer step fused into an aggregation epilogue (torch.optim.SGD semantics,
 * ref: garfieldpp/server.py:277-287 update_model + optimizer.step) and publication of the
 * updated parameters into up to 8 peer arenas (ref: server.py:127 model shipping). */
typedef struct gb_fused_update {
    int mode;        /* 0 none, 1 SGD */
    int nesterov;
    int first_step;  /* 1: momentum buffer := gradient */
    int n_mirrors;
    float lr, momentum, dampening, weight_decay;
    float* param;
    float* mom;
    float* mirror[8];
    float* multicast; /* optional NVLS multicast address of `param` (symmetric memory): the updated parameters are published to every
                         GPU of the group with ONE multimem.st per 16 bytes instead of a store per mirror; n_mirrors must then be 0 */
} gb_fused_update;

/* Per-coordinate rules. rule: 0 linear (needs W 1×n), 1 lower median (NaN-propagating),
 * 2 upper median of the finite values, 3 f-trimmed mean (rank = f), 4 mean of the `rank`
 * values closest to the lower/upper median, 5 condense (Bernoulli(p) ? median : row 0).
 * theta > 0: the rule is applied to the theta mixes y = W·x (W theta×n row-major, device).
 * ref: include/operations.cu.hpp:57-139, py_median/median.cu:60-83, py_bulyan/bulyan.cu:215-243 */
GB_API int gb_coordwise(float const* const* rows, int n, int64_t d, int rule, int rank, int upper, float p, uint64_t seed,
                        float const* W, int theta, float* out, gb_fused_update const* upd, void* stream);

/* One pass over the rows → n×n squared distances (+inf diagonal, non-finite → +inf) and, in the
 * same kernel, the selection logic. mode: 0 none, 1 Krum(f, m, neighbors, squared), 2 Bulyan
 * python semantics, 3 Bulyan native semantics, 4 MDA (squared=1: native non-finite policy),
 * 5 the m rows closest to the LAST row (Aksel). Outputs (device): dist[n*n], W[theta*n]
 * (theta = n-2f-2 for Bulyan else 1), info[1+n] (info[0] = theta, then round-0 selected ids).
 * `scratch` is caller-provided device memory of at least gb_pairwise_scratch_bytes(n) bytes
 * whose first word must be zero before the first use (the kernel leaves it zero).
 * ref: py_krum/krum.cu:68-158, py_bulyan/bulyan.cu:67-207,256-322, py_brute/brute.cu:64-154 */
GB_API int64_t gb_pairwise_scratch_bytes(int n);
GB_API int gb_pairwise_select(float const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors, int squared,
                              void* scratch, int64_t scratch_bytes, float* dist, float* W, int* info, void* stream);
/* Selection logic alone on a (summed) n×n squared-distance matrix already in device memory — used
 * when the distance pass is sharded over coordinates/GPUs and the partial matrices were all-reduced.
 * `dist` is normalised in place (+inf diagonal, non-finite → +inf). Same modes as gb_pairwise_select.
 * ref: py_krum/krum.cu:104-150 (host sort/nth_element), py_bulyan/bulyan.cu:294-316 */
GB_API int gb_select_only(float* dist, int n, int mode, int f, int m, int neighbors, int squared, float* W, int* info, void* stream);
/* Cap the grid of this thread's next gb_pairwise_select launches at `max_ctas` CTAs (0 = fill the GPU); returns the previous cap.
 * For a distance pass that runs in the background of other kernels (distances of finished layers during the backward pass). */
GB_API int gb_pairwise_cta_limit(int max_ctas);
/* Host-only: which kernel, tile and block-pair assignment gb_pairwise_select uses for n rows (see gar_pairwise.cu). */
GB_API int gb_pairwise_plan(int n, int* kind, int* tile, int* slots, unsigned char* owner);
/* Tensor-core variant of gb_pairwise_select for 2 <= n <= 32 sixteen-byte-aligned rows: the Gram
 * matrix X·Xᵀ is accumulated with tcgen05.mma (kind::tf32) in TMEM from TMA-bulk-copied tiles, then
 * D_ij = G_ii + G_jj − 2·G_ij and the same in-kernel selection. TF32 operand precision (see
 * csrc/cuda/gar_gram_tc.cu); returns an error without launching when the inputs do not qualify.
 * ref: py_krum/krum.cu:87-98, py_bulyan/bulyan.cu:286-293 (K1+K2 launched n(n-1)/2 times) */
GB_API int64_t gb_gram_scratch_bytes(int n);
GB_API int gb_gram_select_tc(float const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors, int squared,
                             void* scratch, int64_t scratch_bytes, float* dist, float* W, int* info, void* stream);
/* gb_pairwise_select for callers that may not allocate device memory (a process whose HBM was
 * handed to its own pool at start-up, a loop under a fixed memory plan, a path that cannot take
 * cudaMalloc's device-wide synchronisation): same outputs, no scratch argument, no allocation of
 * any kind — the scratch is a fixed 303 360-byte device array inside the library, which limits
 * n to 16 rows (larger n returns an error). The first call on each device looks up that array's
 * device address: issue it with the device current and outside any stream capture. Calls share
 * the array, so calls that may overlap on the GPU race — serialise them.
 * This entry point has no counterpart in the reference. */
GB_API int gb_pairwise_select_noalloc(float const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors,
                                      int squared, float* dist, float* W, int* info, void* stream);

/* Double-precision variants for CUDA float64 rows (SIMT; same selection code and semantics; `dist` is n*n doubles, W / info as
 * above; coordwise rules 0-4). ref: aggregate_gpu_double dispatch, native/include/aggregator.hpp:123-133. */
GB_API int64_t gb_pairwise_scratch_bytes_f64(int n);
GB_API int gb_pairwise_select_f64(double const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors, int squared,
                                  void* scratch, int64_t scratch_bytes, double* dist, float* W, int* info, void* stream);
GB_API int gb_coordwise_f64(double const* const* rows, int n, int64_t d, int rule, int rank, int upper, float const* W, int theta,
                            double* out, void* stream);

/* C[M,N] (fp32, row stride ldc) = A[M,K] · B[N,K]^T (+ bias[N]) (ReLU) on tcgen05 tensor cores with TMA
 * tiled loads and a TMEM accumulator. dtype 0: fp32 operands read as TF32; 1: bf16 operands.
 * Operands K-major, 16-byte aligned, row strides (in elements) multiples of 16 bytes.
 * ref: nn.Linear forward/backward through cuBLAS (garfieldpp/models/nets.py:47-57). */
GB_API int gb_gemm_nt(void const* A, void const* B, float* C, float const* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                      int64_t ldc, int dtype, int relu, void* stream);

/* Implicit-GEMM convolutions on tcgen05 tensor cores (TF32 operands, fp32 accumulation in TMEM) over channels-last
 * activations; operand tiles are tiled-TMA boxes of the activation tensor shifted by the filter tap (zero fill = padding,
 * traversal stride = convolution stride), weights are read in OHWI order [cout][kh][kw][cin]. A 1 x 1 geometry makes the
 * same entry points plain GEMMs (Linear layers, explicit patch matrices). Pixel pitches are in floats.
 * ref: cuDNN / cuBLAS behind nn.Conv2d / nn.Linear (garfieldpp/models/nets.py:40-78, garfieldpp/tools.py:69-75,
 * garfieldpp/worker.py:85-91). */
GB_API int gb_conv_fprop_tc(float const* x, float const* w, float const* bias, float* y, int n, int h, int wd, int cin, int x_pitch, int cout,
                            int y_pitch, int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int relu, void* stream);
/* dx of a stride-1 convolution (strided layers: zero-upsample dy first, gb_upsample_zero_nhwc); `addend` (optional, dx's shape
 * and pitch) is summed in the epilogue - the gradient of a residual skip connection joins without a separate add kernel. */
GB_API int gb_conv_dgrad_tc(float const* dy, float const* w, float* dx, int n, int h, int wd, int cin, int dx_pitch, int cout, int dy_pitch,
                            int kh, int kw, int pad_h, int pad_w, float const* addend, void* stream);
/* Per-worker weight gradients: images g*b ... g*b+b-1 belong to worker g; matrix g is written at dw + g*dw_group_stride
 * ([cout][kh*kw*dw_cols], row stride dw_ld) — straight into the workers' gradient rows. */
GB_API int gb_conv_wgrad_tc(float const* x, float const* dy, float* dw, int groups, int b, int h, int wd, int cin, int x_pitch, int cout,
                            int dy_pitch, int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dw_cols, int64_t dw_ld,
                            int64_t dw_group_stride, void* stream);
/* Channels-last layers around them (csrc/cuda/nn_nhwc.cu). ref: ATen/cuDNN behind nn.MaxPool2d, nn.AdaptiveAvgPool2d,
 * nn.CrossEntropyLoss and the strided-convolution backward (garfieldpp/tools.py:47-57,69-75). */
GB_API int gb_im2col_pad_nhwc(float const* x, float* cols, int N, int H, int W, int C, int pitch, int KH, int KW, int stride_h, int stride_w,
                              int pad_h, int pad_w, int kpad, void* stream);
GB_API int gb_pad_rows(float const* src, float* dst, int64_t rows, int k, int64_t ld, int kpad, void* stream);
GB_API int gb_col2im_nhwc(float const* dcols, float* dx, int N, int H, int W, int C, int pitch, int KH, int KW, int stride_h, int stride_w,
                          int pad_h, int pad_w, int kpad, void* stream);
GB_API int gb_upsample_zero_nhwc(float const* in, float* out, int N, int Ho, int Wo, int C, int stride_h, int stride_w, int H, int W, void* stream);
GB_API int gb_maxpool_nhwc_forward(float const* x, float* y, unsigned char* arg, int N, int H, int W, int C, int k, int stride, int pad, void* stream);
GB_API int gb_maxpool_nhwc_backward(float const* dy, unsigned char const* arg, float* dx, int N, int H, int W, int C, int k, int stride, int pad,
                                    void* stream);
/* Grouped cross-entropy: row_loss[row] and dlogits = d(sum over workers of their mean loss)/d(logits), padded columns zeroed. */
GB_API int gb_softmax_ce_grouped(float const* logits, int ld, int64_t const* target, float* row_loss, float* dlogits, int ldd, int rows,
                                 int classes, int rows_per_group, void* stream);
GB_API int gb_colsum_grouped(float const* in, int64_t ld, float* out, int groups, int rows_per_group, int cols, float scale, void* stream);
GB_API int gb_relu_forward(float const* x, float* y, int64_t total, void* stream);
GB_API int gb_relu_backward(float const* dy, float const* y, float* dx, int64_t total, void* stream);
/* out = a + b (dense float arrays of `total` elements; out may alias a or b). */
GB_API int gb_add(float const* a, float const* b, float* out, int64_t total, void* stream);
GB_API int gb_avgpool_nhwc_forward(float const* x, float* y, int64_t n, int hw, int C, void* stream);
GB_API int gb_avgpool_nhwc_backward(float const* dy, float* dx, int64_t n, int hw, int C, void* stream);

/* Layer kernels of the small reference networks (fp32, NCHW). ref: cuDNN/ATen behind nn.Conv2d, MaxPool2d,
 * BatchNorm2d, NLLLoss/CrossEntropyLoss (garfieldpp/models/nets.py:40-144, garfieldpp/tools.py:47-57). */
GB_API int gb_conv2d_forward(float const* x, float const* w, float const* bias, float* y, int n, int cin, int h, int wd, int cout, int kh, int kw,
                             int stride, int pad, int relu, void* stream);
GB_API int gb_conv2d_backward(float const* x, float const* w, float const* dy, float* dx, float* dw, float* db, int n, int cin, int h, int wd,
                              int cout, int kh, int kw, int stride, int pad, void* stream);
GB_API int gb_maxpool2d_forward(float const* x, float* y, int* arg, int planes, int h, int w, int k, int stride, void* stream);
GB_API int gb_maxpool2d_backward(float const* dy, int const* arg, float* dx, int planes, int hw_in, int hw_out, void* stream);
/* mode 0: inputs are log-probabilities (NLLLoss); 1: raw logits (CrossEntropyLoss). dlogits = d(mean loss)/d(input). */
GB_API int gb_softmax_loss(float const* in, int64_t const* target, float* row_loss, float* dlogits, int rows, int classes, int mode, void* stream);
GB_API int gb_batchnorm_forward(float const* x, float const* gamma, float const* beta, float* y, float* save_mean, float* save_invstd,
                                float* running_mean, float* running_var, int n, int c, int hw, float eps, float momentum, int relu, void* stream);
GB_API int gb_batchnorm_backward(float const* x, float const* dy, float const* gamma, float const* save_mean, float const* save_invstd, float* dx,
                                 float* dgamma, float* dbeta, int n, int c, int hw, void* stream);

/* Fused training-mode BatchNorm for channels-last activations [rows = N*H*W, C]: y = [ReLU](BN(x) [+ residual]) in one
 * launch, and its backward (dx, dgamma, dbeta and the masked gradient of the skip path) in one launch.
 * ref: cuDNN/ATen BatchNorm2d + ReLU + residual add in the torchvision ResNet blocks (garfieldpp/tools.py:69-75). */
GB_API int gb_bn_nhwc_forward(float const* x, float const* residual, float const* gamma, float const* beta, float* y, float* save_mean,
                              float* save_invstd, float* running_mean, float* running_var, int rows, int C, float eps, float momentum, int relu,
                              void* stream);
GB_API int gb_bn_nhwc_backward(float const* x, float const* y, float const* dy, float const* gamma, float const* save_mean,
                               float const* save_invstd, float* dx, float* d_residual, float* dgamma, float* dbeta, int rows, int C, int relu,
                               void* stream);
/* The same kernels over `groups` consecutive blocks of `rows` rows, each normalised with its own statistics (save_* and
 * running_*: [groups, C]; gamma/beta shared; dgamma/dbeta: [groups, C]) — the co-located workers of one GPU in one launch. */
GB_API int gb_bn_nhwc_forward_grouped(float const* x, float const* residual, float const* gamma, float const* beta, float* y, float* save_mean,
                                      float* save_invstd, float* running_mean, float* running_var, int groups, int rows, int C, float eps,
                                      float momentum, int relu, void* stream);
GB_API int gb_bn_nhwc_backward_grouped(float const* x, float const* y, float const* dy, float const* gamma, float const* save_mean,
                                       float const* save_invstd, float* dx, float* d_residual, float* dgamma, float* dbeta, int groups, int rows,
                                       int C, int relu, void* stream);
/* cols[(n,ho,wo), (kh,kw,c)] = x[n, ho*s-p+kh, wo*s-p+kw, c] for channels-last x [N,H,W,C] (C % 4 == 0): the patch matrix whose
 * batched GEMM with dY gives per-worker k x k weight gradients directly in OHWI order (worker fusion). No reference counterpart. */
GB_API int gb_im2col_nhwc(float const* x, float* cols, int N, int H, int W, int C, int KH, int KW, int stride_h, int stride_w, int pad_h,
                          int pad_w, void* stream);
GB_API int gb_global_avgpool_forward(float const* x, float* y, int64_t planes, int hw, void* stream);
GB_API int gb_global_avgpool_backward(float const* dy, float* dx, int64_t planes, int hw, void* stream);
/* Inverted dropout with a regenerable Philox mask (apply to dy with the same seed for the backward pass). */
GB_API int gb_dropout(float const* x, float* y, int64_t total, float p, uint64_t seed, void* stream);
/* nn.BCELoss (mean): per-element losses and d(mean loss)/d(prob). ref: garfieldpp/tools.py:52 */
GB_API int gb_bce_loss(float const* prob, float const* target, float* loss, float* dprob, int64_t total, void* stream);

/* Flat-arena optimizer steps (torch.optim semantics). ref: garfieldpp/server.py:277-287,
 * garfieldpp/tools.py:107-123 (sgd, adam, adamw, rmsprop, adagrad — run on the HOST there). */
GB_API int gb_sgd_step(float* param, float const* grad, float* mom, int64_t d, float lr, float momentum, float dampening,
                       float weight_decay, int nesterov, int first_step, float* const* mirrors, int n_mirrors, void* stream);
GB_API int gb_adam_step(float* param, float const* grad, float* exp_avg, float* exp_avg_sq, int64_t d, float lr, float beta1, float beta2,
                        float eps, float weight_decay, int step, int decoupled, void* stream);
GB_API int gb_rmsprop_step(float* param, float const* grad, float* square_avg, float* mom, int64_t d, float lr, float alpha, float eps,
                           float weight_decay, float momentum, void* stream);
GB_API int gb_adagrad_step(float* param, float const* grad, float* state_sum, int64_t d, float lr, float lr_decay, float eps,
                           float weight_decay, int step, void* stream);

/* Byzantine attack generation on device. ref: garfieldpp/byzWorker.py:78-143, byzServer.py:86-107 */
GB_API int gb_attack_random(float* out, int64_t d, uint64_t seed, void* stream);
GB_API int gb_scale(float const* in, float* out, int64_t d, float factor, void* stream);
GB_API int gb_attack_drop(float const* in, float* out, int64_t d, float p, uint64_t seed, void* stream);
GB_API int gb_attack_collude(float const* const* rows, int fw, int64_t d, int kind, float coeff, int nan_single, float* out, void* stream);

/* Cross-GPU readiness flags (system-scope release store / acquire-load spin with a deadline): the
 * device-side counterpart of the reference's RPC completion + "first n-f replies" quorum wait
 * (garfieldpp/server.py:134-155). Only for kernels on different GPUs. */
GB_API int gb_flag_signal(uint32_t* flag, uint32_t value, void* stream);
GB_API int gb_flag_wait(uint32_t* const* flags, int count, uint32_t value, uint64_t timeout_ns, uint32_t* late, void* stream);
/* The same with the epoch in DEVICE memory (a whole-iteration CUDA graph replays unchanged every step): gb_epoch_bump
 * increments it, signal stores *epoch, wait targets *epoch - lag and reports per-flag arrival in arrived[count] (optional).
 * gb_reduce_slots sums `nslots` small peer-visible buffers in slot order (the n x n partial distances of the coordinate
 * shards) - the device-side replacement of the reference's gather over RPC / an NCCL all-reduce. */
GB_API int gb_epoch_bump(uint32_t* epoch, void* stream);
GB_API int gb_flag_signal_ep(uint32_t* flag, uint32_t const* epoch, void* stream);
GB_API int gb_flag_wait_ep(uint32_t* const* flags, int count, uint32_t const* epoch, uint32_t lag, uint64_t timeout_ns, uint32_t* late,
                           uint32_t* arrived, void* stream);
GB_API int gb_reduce_slots(float const* const* slots, int nslots, int count, float* out, void* stream);
/* Quorum decision of the lead server: wait for all `count` flags, or for the deadline with >= `quorum` of them; store the 64-bit
 * arrival mask into words[i][0..1] of every rank and release decision_flags[i] with the epoch.
 * ref: the first-q wait of garfieldpp/server.py:134-155. */
GB_API int gb_quorum_decide(uint32_t* const* flags, int count, uint32_t const* epoch, int quorum, uint64_t timeout_ns, uint32_t* const* words,
                            uint32_t* const* decision_flags, int n_out, uint32_t* late, void* stream);

/* North-star path 4: `rounds` agreement rounds of a per-coordinate rule over the n nodes' double-buffered, peer-mapped vectors
 * inside ONE persistent kernel per GPU (grid barrier + peer flags between rounds; see csrc/cuda/gar_gossip.cu).
 * ref: avg_agree / contract loops of applications/LEARN/trainer.py:208-222,244-258. */
GB_API int gb_gossip_rounds(float const* const* rows0, float const* const* rows1, int n, int64_t d, int rule, int rank, int upper,
                            float* const* out0, float* const* out1, int n_out, int rounds, int first_parity, uint32_t* seq, uint32_t* my_flag,
                            uint32_t* const* peer_flags, int n_peers, uint32_t* counter, uint64_t timeout_ns, uint32_t* late, void* stream);
/* out (multicast address, every rank's copy) = scale * sum over ranks and over `slots` buffers of `pitch` floats of in (multicast
 * address): NVLS in-switch reduction (multimem.ld_reduce) + multicast store (multimem.st) of this rank's coordinate slice
 * [lo, hi) — the `average` contract step of the decentralized topology without any per-peer load. Call between two phase barriers.
 * ref: gar=average in applications/LEARN/trainer.py:244-258. */
GB_API int gb_multimem_average(float const* mc_in, float* mc_out, int slots, int64_t pitch, int64_t lo, int64_t hi, float scale, void* stream);
/* North-star path 3 head: lower median of `replicas` servers' copies of the first layer's weight ([rows_out][k], rows may be
 * peer-mapped) written into the flat model AND as the zero-padded K-major operand [rows_out][kpad] of the first forward GEMM.
 * ref: applications/ByzSGD/trainer.py:237-244. */
GB_API int gb_pull_median_operand(float const* const* rows, int replicas, int rows_out, int k, int kpad, float* flat, float* operand, void* stream);

/* Gather `count` separate device tensors into their slots (element offsets) of one flat vector in one
 * launch per 112 tensors; table passed in kernel parameters (CUDA-graph capturable).
 * ref: torch.cat of the per-parameter gradients, garfieldpp/worker.py:93-94. */
GB_API int gb_multi_copy(float const* const* srcs, uint64_t const* dst_offsets, uint32_t const* sizes, int count, float* dst, void* stream);
GB_API int gb_axpby(float const* x, float* y, int64_t d, float a, float b, void* stream);
GB_API int gb_broadcast_copy(float const* src, float* const* dsts, int n_dst, int64_t d, void* stream);

#ifdef __cplusplus
}
#endif

#endif /* GARFIELD_B200_H */
