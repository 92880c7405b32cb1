/* C ABI of libdcnn_b200.so -- hand-written sm_100a kernels behind plain pointers + a CUDA stream.
 *
 * Conventions: activations are NHWC fp32; every function returns 0 on success, otherwise a non-zero code whose
 * text is available from dcnn_last_error().  All launches are asynchronous on `stream`; nothing here synchronises.
 * Each declaration cites the reference code it replaces (paths relative to the reference repository). */
#ifndef DCNN_B200_H_
#define DCNN_B200_H_
#include <cuda_runtime.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ---- */
const char* dcnn_last_error(void);
const char* dcnn_version(void);
unsigned long long dcnn_launch_count(void);
void dcnn_reset_launch_count(void);
int dcnn_device_sms(void);

/* ---- tensor-core implicit GEMM (tcgen05 + TMEM + TMA); replaces cudnnConvolutionForward/BackwardData/
 *      BackwardFilter and cublasSgemm: src/nn/layers_impl/cuda/cudnn_conv2d_ops.cu:171-232, dense_ops.cu:24-137 ---- */
int dcnn_conv_tc_m_tiles(int N, int OH, int OW, int mul_h, int mul_w);
/* statistics tiles dcnn_conv_tc emits for this problem (== m_tiles unless the 256-pixel kernel is selected; DCNN_CONV_V3=0 disables it) */
/* autotuner hook: force conv_tc2 (2), conv_tc3 (3) or conv_tc5 (5) for this thread's next calls; 0 = built-in rules */
int dcnn_conv_tc_set_variant(int variant);
int dcnn_conv_tc_stat_tiles(int N, int OH, int OW, int mul_h, int mul_w, int Cout, int num_kb);
int dcnn_conv_tc(const float* x, int N, int H, int W, int Cin, const float* w, int Cout, int Ktot, float* out,
                 long long out_sn, long long out_sh, long long out_sw, int OH, int OW, int in_mul_h, int in_mul_w,
                 int num_taps, const int* tap_dh, const int* tap_dw, const int* tap_w, const float* bias,
                 int accumulate, int relu, float* partials, cudaStream_t stream);
/* scratch == NULL: split-K sums by red.global.add (order varies run to run); scratch = dcnn_wgrad_tc_splits * Cout * Ktot
 * floats: per-split slabs summed in index order by a second kernel (bit-reproducible). */
int dcnn_wgrad_tc_splits(int N, int OH, int OW, int Cout, int Cin, int in_mul_h, int in_mul_w, int num_taps);
int dcnn_wgrad_tc(const float* dy, int N, int OH, int OW, int Cout, const float* x, int H, int W, int Cin, float* dw,
                  int Ktot, int in_mul_h, int in_mul_w, int num_taps, const int* tap_dh, const int* tap_dw,
                  const int* tap_w, float* scratch, float* db, cudaStream_t stream);   /* db: fused bias gradient or NULL */
/* measurement tool: persistent CTAs stream TMA operand tiles with no MMA (per-SM / chip-wide ingest rate, effect of a
 * shared B tile); see scripts/tma_ingest_probe.py */
int dcnn_tma_stream_probe(const float* A, long long a_total_rows, long long a_cols, int a_rows, const float* B,
                          long long b_total_rows, long long b_cols, int b_rows, int iters, int stages, int b_shared,
                          int ctas, long long* bytes_per_iter, cudaStream_t stream);
int dcnn_tma_probe(const float* x, int N, int H, int W, int C, int bn, int bh, int bw, int mul_h, int mul_w, int c0,
                   int c1, int c2, int c3, float* out, cudaStream_t stream);

/* ---- exact-FP32 CUDA-core fallbacks (src/nn/layers_impl/cuda/conv2d_ops.cu:19-100, tensor_kernels.cu:18-119) ---- */
int dcnn_conv_fprop_simt(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int C,
                         int K, int R, int S, int sh, int sw, int ph, int pw, cudaStream_t stream);
int dcnn_conv_dgrad_simt(const float* dy, const float* w, float* dx, int N, int H, int W, int C, int K, int R, int S,
                         int sh, int sw, int ph, int pw, int accumulate, cudaStream_t stream);
int dcnn_conv_wgrad_simt(const float* dy, const float* x, float* dw, int N, int H, int W, int C, int K, int R, int S,
                         int sh, int sw, int ph, int pw, cudaStream_t stream);
int dcnn_weight_dgrad_layout(const float* w, float* wt, int K, int R, int S, int C, cudaStream_t stream);
/* the same for `count` weight tensors, 64 per launch: table_host[e] = {w, wt, K, R, S, C} as six 64-bit words (host) */
int dcnn_weight_dgrad_layout_batched(const long long* table_host, int count, cudaStream_t stream);

/* ---- BatchNorm / GroupNorm (batchnorm_ops.cu:108-295, groupnorm_ops.cu:45-252, cudnnConvolutionBackwardBias) ---- */
int dcnn_bn_stat_tiles(long long M);
int dcnn_bn_colstats(const float* x, long long M, int C, float* partials, cudaStream_t stream);
int dcnn_bn_finalize(const float* partials, int tiles, int C, float eps, float momentum, float* running_mean,
                     float* running_var, const float* gamma, const float* beta, float* mean, float* invstd,
                     float* scale, float* shift, cudaStream_t stream);
/* Sync-BatchNorm in ONE launch (no counterpart in the reference, which has no data parallelism): the finalize kernel
 * pushes its per-channel values into every replica's exchange area over NVLink peer mappings, raises per-channel-group
 * epoch flags and merges all replicas' values from local memory.  areas[r] = replica r's area for this layer and
 * direction as mapped in this process (`world` device pointers, dcnn_bn_sync_area_bytes each, zeroed once). */
long long dcnn_bn_sync_area_bytes(int world, int C);
int dcnn_bn_finalize_sync1(const float* partials, int tiles, int C, float eps, float momentum, float* running_mean,
                           float* running_var, const float* gamma, const float* beta, float* mean, float* invstd,
                           float* scale, float* shift, unsigned char* const* areas, int rank, int world,
                           cudaStream_t stream);
int dcnn_bn_eval_coeffs(const float* rm, const float* rv, const float* gamma, const float* beta, float eps, int C,
                        float* scale, float* shift, cudaStream_t stream);
int dcnn_bn_apply(const float* x, const float* scale, const float* shift, const float* residual, float* y,
                  long long M, int C, int relu, cudaStream_t stream);
/* BN apply (+residual) + ReLU that also writes a 4-bit-per-float4 ReLU mask (M*C/4 bytes); pass the mask as `out`
 * with relu = 3 to dcnn_bn_bwd_reduce / dcnn_bn_bwd_apply.  relu modes there: 0 none, 1 mask = out > 0,
 * 2 mask recomputed from x (out -> scale[C] | shift[C]), 3 mask bytes. */
int dcnn_bn_apply_mask(const float* x, const float* scale, const float* shift, const float* residual, float* y,
                       unsigned char* mask, long long M, int C, cudaStream_t stream);
int dcnn_bn_bwd_tiles(long long M, int C);   /* row tiles of dcnn_bn_bwd_reduce: partials = tiles*2*C floats */
int dcnn_bn_bwd_reduce(const float* dout, const float* x, const float* out, const float* mean, const float* invstd,
                       long long M, int C, int relu, float* partials, cudaStream_t stream);
int dcnn_bn_bwd_finalize(const float* partials, int tiles, int C, double count, const float* mean,
                         const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* k1, float* k2,
                         float* k3, cudaStream_t stream);
int dcnn_bn_bwd_finalize_sync1(const float* partials, int tiles, int C, double count, const float* mean,
                               const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* k1,
                               float* k2, float* k3, unsigned char* const* areas, int rank, int world,
                               cudaStream_t stream);
int dcnn_bn_bwd_apply(const float* dout, const float* x, const float* out, const float* k1, const float* k2,
                      const float* k3, float* dx, float* masked, long long M, int C, int relu, cudaStream_t stream);
int dcnn_colsum_accum(const float* dy, long long M, int C, float* db, int deterministic, cudaStream_t stream);
/* Allocation-free variant of dcnn_bn_colstats + dcnn_bn_finalize (no counterpart in the reference).  For callers
 * that may not allocate device memory when they need the statistics (no access to an allocator, or not allowed to
 * take its lock): the per-tile scratch lives in a 4 MiB `static __device__` array inside the library instead of a
 * caller buffer.  x is [M][C] fp32; outputs are [C] each; running_mean/running_var/gamma/beta may be NULL.
 * Problems needing more than dcnn_bn_noalloc_capacity_floats() = tiles*(2*C+1) floats are rejected.  The first call
 * on a device looks the array's address up (do it before any stream capture); calls share the scratch, so calls
 * that overlap on the GPU race -- keep them on one stream or order them with events. */
long long dcnn_bn_noalloc_capacity_floats(void);
int dcnn_bn_batch_stats_noalloc(const float* x, long long M, int C, float eps, float momentum, float* running_mean,
                                float* running_var, const float* gamma, const float* beta, float* mean,
                                float* invstd, float* scale, float* shift, cudaStream_t stream);
int dcnn_gn_forward(const float* x, float* y, const float* gamma, const float* beta, float* mean, float* invstd,
                    int N, int HW, int C, int G, float eps, cudaStream_t stream);
int dcnn_gn_backward(const float* dy, const float* x, const float* mean, const float* invstd, const float* gamma,
                     float* dgamma, float* dbeta, float* dx, int N, int HW, int C, int G, cudaStream_t stream);

/* ---- pooling / activations / dropout / layout (maxpool_ops.cu, avgpool_ops.cu, activations_impl/cuda/*.cu,
 *      dropout.cu, ops/cuda/kernels.cu) ---- */
int dcnn_maxpool_fwd(const float* x, float* y, uint8_t* idx, int N, int H, int W, int C, int OH, int OW, int kh,
                     int kw, int sh, int sw, int ph, int pw, cudaStream_t stream);
int dcnn_maxpool_bwd(const float* dy, const uint8_t* idx, float* dx, int N, int H, int W, int C, int OH, int OW,
                     int kh, int kw, int sh, int sw, int ph, int pw, cudaStream_t stream);
int dcnn_avgpool_fwd(const float* x, float* y, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh,
                     int sw, int ph, int pw, cudaStream_t stream);
int dcnn_avgpool_bwd(const float* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh,
                     int sw, int ph, int pw, cudaStream_t stream);
int dcnn_act_fwd(int kind, const float* x, float* y, long long n, float alpha, cudaStream_t stream);
int dcnn_act_bwd(int kind, const float* x, const float* g, float* dx, long long n, float alpha, cudaStream_t stream);
int dcnn_softmax_fwd(const float* x, float* y, long long rows, int C, cudaStream_t stream);
int dcnn_softmax_bwd(const float* x, const float* g, float* dx, long long rows, int C, cudaStream_t stream);
int dcnn_add(const float* a, const float* b, float* y, long long n, cudaStream_t stream);
int dcnn_dropout(const float* x, float* y, long long n, float rate, unsigned long long seed,
                 unsigned long long offset, cudaStream_t stream);
/* packed-tap im2col for small-Cin convolutions: P[n,oh,ow,Kp], k = (r*S+s)*Cin + c, zero padded to Kp (multiple of 32) */
int dcnn_pack_taps(const float* x, float* P, int N, int H, int W, int Cin, int OH, int OW, int R, int S, int sh, int sw,
                   int ph, int pw, int Kp, cudaStream_t stream);
int dcnn_transpose_batched(const float* in, float* out, int batch, int rows, int cols, cudaStream_t stream);
int dcnn_fill(float* p, long long n, float v, cudaStream_t stream);

/* ---- loss + accuracy, optimizers (loss_ops.cu:18-435, accuracy.cu:15-111, adam_kernels.cu, sgd_kernels.cu) ---- */
int dcnn_softmax_ce(const float* pred, const float* tgt, float* grad, float* out2, int B, int C, int mode, float eps,
                    cudaStream_t stream);
int dcnn_regression_loss(const float* pred, const float* tgt, float* grad, float* out1, long long n, int kind,
                         float delta, cudaStream_t stream);
int dcnn_multi_tensor_update(int mode, float* const* param, float* const* grad, float* const* s1, float* const* s2,
                             const long long* numel, const int* block_tensor, const long long* block_off,
                             int num_blocks, float lr, float a, float b, float eps, float bc1, float bc2, float wd,
                             int zero_grads, const float* hyper /* device {lr,bc1,bc2} or NULL */,
                             cudaStream_t stream);
int dcnn_mt_chunk(void);

/* ---- data-parallel peer-memory path (no counterpart in the reference: it has no data parallelism) ---- */
/* ---- NVLS (NVSwitch multicast) variant of the data-parallel update: csrc/nvls.cu.  Opt-in, see parallel/dp.py ---- */
int dcnn_nvls_supported(int device);
long long dcnn_nvls_padded_size(int world, long long bytes, int device);
int dcnn_nvls_create(int world, long long bytes, unsigned long long* handle_out, int* fd_out);   /* rank 0 */
int dcnn_nvls_import(int fd, unsigned long long* handle_out);                                    /* other ranks */
int dcnn_nvls_add_device(unsigned long long handle, int device);   /* every rank, before any rank binds */
int dcnn_nvls_bind(unsigned long long handle, int device, long long bytes, void** uc_ptr, void** mc_ptr,
                   unsigned long long* mem_handle_out);
int dcnn_nvls_release(unsigned long long handle, unsigned long long mem_handle, void* uc_ptr, void* mc_ptr,
                      long long bytes);   /* every rank, after the last use */
int dcnn_fused_allreduce_update_nvls(const float* mc_grads, float* mc_params, const float* params, float* grads,
                                     float* s1, float* s2, long long n, long long shard_begin, long long shard_end,
                                     unsigned long long* const* start_flag_ptrs,
                                     unsigned long long* const* done_flag_ptrs, unsigned int* block_counter, int rank,
                                     int world, unsigned long long* epoch_counter, int mode, float lr, float a, float b,
                                     float eps, float bc1, float bc2, float wd, const float* hyper, int num_blocks,
                                     cudaStream_t stream);
int dcnn_dev_malloc(void** ptr, long long bytes);
int dcnn_dev_free(void* ptr);
int dcnn_ipc_get_handle(void* ptr, unsigned char* handle64);
int dcnn_ipc_open(const unsigned char* handle64, void** ptr);
int dcnn_ipc_close(void* ptr);
int dcnn_enable_peer_access(int peer_device);
int dcnn_peer_barrier(unsigned long long* const* flag_ptrs, int rank, int world, unsigned long long* epoch_counter,
                      cudaStream_t stream);
/* Push variant: gradient slices are copied into the owners' staging buffers, reduced from local memory, parameters
 * are stored into every replica -- one launch, no load ever crosses NVLink (no counterpart in the reference). */
int dcnn_fused_allreduce_update_push(float* const* grad_ptrs, float* const* param_ptrs, float* const* stage_ptrs,
                                     const long long* sb, const long long* se, const long long* soff, float* s1,
                                     float* s2, long long n, unsigned long long* const* start_flag_ptrs,
                                     unsigned long long* const* done_flag_ptrs, unsigned int* block_counters, int rank,
                                     int world, unsigned long long* epoch_counter, int mode, float lr, float a, float b,
                                     float eps, float bc1, float bc2, float wd, const float* hyper, int num_blocks,
                                     cudaStream_t stream);
int dcnn_fused_allreduce_update(float* const* grad_ptrs, float* const* param_ptrs, float* s1, float* s2, long long n,
                                long long shard_begin, long long shard_end,
                                unsigned long long* const* start_flag_ptrs, unsigned long long* const* done_flag_ptrs,
                                unsigned int* block_counter, int rank, int world, unsigned long long* epoch_counter,
                                int mode, float lr, float a, float b, float eps, float bc1, float bc2, float wd,
                                const float* hyper, int num_blocks, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DCNN_B200_H_ */
