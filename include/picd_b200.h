/* picd_b200 -- C ABI of libpicd_b200.so (sm_100a kernels for the B200-native image-classification trainer).
 *
 * Conventions: every function returns 0 on success; on failure picd_last_error() (thread local) describes it.
 * Pointers are device pointers unless stated.  Activations are NHWC bf16 ("void*"), statistics / master weights /
 * gradients are fp32.  All launches are asynchronous on `stream`.  Citations name the reference behaviour each
 * entry point replaces (file:line in /root/reference or SURVEY.md row).
 */
#ifndef PICD_B200_H
#define PICD_B200_H

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* activation codes */
enum { PICD_ACT_NONE = 0, PICD_ACT_RELU = 1, PICD_ACT_GELU = 2, PICD_ACT_SWISH = 3, PICD_ACT_SIGMOID = 4 };
/* optimizer kinds */
enum { PICD_OPT_SGD = 0, PICD_OPT_ADAM = 1, PICD_OPT_ADAMW = 2 };
/* all-reduce variants */
enum { PICD_AR_LOCAL = 0, PICD_AR_ONE_SHOT = 1, PICD_AR_TWO_SHOT = 2,
       PICD_AR_NVLS = 3,      /* two-shot through the NVSwitch multicast mapping: multimem.ld_reduce + multimem.st */
       PICD_AR_NVLS_LD = 4 }; /* in-switch reduction (multimem.ld_reduce), unicast pushes */

const char* picd_last_error(void);
int picd_version(void);
int picd_num_sms(void);
unsigned long long picd_launch_count(void);   /* kernel launches issued by the library so far */

/* ---- tcgen05 implicit-GEMM convolution / GEMM (replaces cuDNN conv + cuBLAS Linear: SURVEY.md K1-K4,
 *      nn/classifier.py:26-37 via torchvision) ------------------------------------------------------------- */
/* y = conv(x, w) [+ bias][act]; stats (nullable) accumulates per-channel [sum(Cout), sumsq(Cout)] of y for BatchNorm */
int picd_conv_fprop(const void* x, const void* w, void* y, float* stats, const float* bias, int act, int N, int H,
                    int W, int Cin, int Cout, int R, int S, int stride, int pad_h, int pad_w, int OH, int OW,
                    cudaStream_t stream);
int picd_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int R, int S,
                    int stride, int pad_h, int pad_w, int OH, int OW, cudaStream_t stream);
/* dx += ...: same as picd_conv_dgrad but ADDS into the existing bf16 dx with TMA reduce-add (residual-gradient fusion) */
int picd_conv_dgrad_accumulate(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int R,
                               int S, int stride, int pad_h, int pad_w, int OH, int OW, cudaStream_t stream);
/* dgrad whose output is the gradient dz flowing into an upstream conv + BatchNorm(+ReLU) layer: the epilogue also reduces
 * that layer's BatchNorm-backward sums sum(g), sum(g * xhat) into sums[STAT_REPLICAS][2][Cin] (zeroed by the caller), g = dz
 * masked by the upstream ReLU (recomputed from y_up * scale + shift).  Replaces the separate batch_norm_backward_reduce pass
 * (torch/nn/modules/_functions.py:144-153 via train.py:72). */
int picd_conv_dgrad_bnreduce(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int R, int S,
                             int stride, int pad_h, int pad_w, int OH, int OW, const void* y_up, const float* scale,
                             const float* shift, const float* mean, const float* invstd, int relu, float* sums,
                             cudaStream_t stream);
/* Same as picd_conv_wgrad / picd_gemm_tn_f32 plus the bias gradient dbias[Cout] += column sums of dy, computed by the same kernel
 * (one extra N = 16 MMA per k-step against a tile of ones): no separate pass over dy. */
int picd_conv_wgrad_bias(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W, int Cin, int Cout, int R,
                         int S, int stride, int pad_h, int pad_w, int OH, int OW, cudaStream_t stream);
int picd_gemm_tn_f32_bias(const void* a_dy, const void* b_x, float* d, float* dbias, int M, int N, int K, cudaStream_t stream);
/* dw (fp32, [Cout,R,S,Cin]) += ...; accumulates with red.add, so it can target the gradient arena directly */
int picd_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R, int S,
                    int stride, int pad_h, int pad_w, int OH, int OW, cudaStream_t stream);
int picd_gemm_nt(const void* a, const void* b, void* d, const float* bias, int act, int M, int N, int K,
                 cudaStream_t stream);                       /* D[M,N] = act(A[M,K] B[N,K]^T + bias) */
int picd_gemm_nn(const void* a, const void* b, void* d, int M, int N, int K, cudaStream_t stream);
                                                             /* D[M,K] = A[M,N] B[N,K] */
int picd_gemm_tn_f32(const void* a, const void* b, float* d, int M, int N, int K, cudaStream_t stream);
                                                             /* D[N,K] += A[M,N]^T B[M,K] (fp32) */

/* ---- multi-head self-attention, head dim 64, S <= 256 (ViT; the reference has no attention -- BASELINE.json config 4).
 *      qkv is the packed projection output [B*S, 3*H*64]; o is [B*S, H*64]; lse/delta are fp32 [B,H,S]. */
int picd_attention_fwd(const void* qkv, void* o, float* lse, int B, int S, int H, cudaStream_t stream);
int picd_attention_bwd(const void* qkv, const void* o, const void* dout, const float* lse, float* delta, void* dqkv, int B,
                       int S, int H, cudaStream_t stream);

/* ---- depthwise convolution (EfficientNet MBConv; reference backbone 'efficientnet-b3', nn/classifier.py:17-18).
 *      wt / dwt are fp32 tap-major [K*K][C]; asymmetric (TF-"same") padding via pad_t/pad_l + explicit OH/OW. */
int picd_depthwise_fprop(const void* x, const float* wt, void* y, int N, int H, int W, int C, int OH, int OW, int K, int stride,
                         int pad_t, int pad_l, cudaStream_t stream);
int picd_depthwise_dgrad(const void* dy, const float* wt, void* dx, int N, int H, int W, int C, int OH, int OW, int K, int stride,
                         int pad_t, int pad_l, cudaStream_t stream);
int picd_depthwise_wgrad(const void* x, const void* dy, float* dwt, int N, int H, int W, int C, int OH, int OW, int K, int stride,
                         int pad_t, int pad_l, cudaStream_t stream);

/* ---- BatchNorm family (torch/nn/modules/_functions.py:39-179 = SURVEY.md K5-K10,K12) ------------------- */
/* Statistics buffers produced by kernels (conv_fprop `stats`, bn_stats, bn_bwd_reduce `sums`) hold
 * picd_stat_replicas() (= 16) copies of [2C] floats, zeroed by the caller; consumers take the number of copies. */
int picd_stat_replicas(void);
int picd_bn_stats(const void* x, float* stats, long long P, int C, cudaStream_t stream);
int picd_bn_finalize(const float* sums, int replicas, float count, const float* gamma, const float* beta, float* running_mean,
                     float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                     float* invstd, int C, cudaStream_t stream);
int picd_bn_eval_coeffs(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                        float eps, float* scale, float* shift, int C, cudaStream_t stream);
/* ReLU bit mask for layers that add a residual before the activation: picd_bn_apply_mask additionally writes one byte per
 * 16-byte output vector (bit k = output 8*v+k is positive) into relu_mask, which must hold P*C/8 bytes plus 4 KB of slack;
 * the *_mask backward entry points read it instead of streaming z again (z is still used by the small-tensor kernels). */
int picd_bn_apply_mask(const void* y, const float* scale, const float* shift, const void* residual, int act, void* out,
                       void* relu_mask, long long P, int C, cudaStream_t stream);
int picd_bn_bwd_reduce_mask(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale,
                            const float* shift, const float* mean, const float* invstd, int act, float* sums, long long P,
                            int C, cudaStream_t stream);
int picd_bn_bwd_apply_mask(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale,
                           const float* shift, const float* mean, const float* invstd, int act, const float* sums,
                           int sums_replicas, float inv_count, const float* param_sums, int param_replicas, float* dgamma,
                           float* dbeta, void* dy, void* dres, long long P, int C, cudaStream_t stream);
/* As picd_bn_bwd_apply_mask; count_dev (nullable) points at the GLOBAL per-channel element count in device memory and then
 * replaces inv_count (SyncBatchNorm with unequal per-rank batches, torch/nn/modules/_functions.py:59-62,170-179). */
int picd_bn_bwd_apply_mask_cnt(const void* dz, const void* y, const void* z, const void* relu_mask, const float* scale,
                               const float* shift, const float* mean, const float* invstd, int act, const float* sums,
                               int sums_replicas, float inv_count, const float* count_dev, const float* param_sums,
                               int param_replicas, float* dgamma, float* dbeta, void* dy, void* dres, long long P, int C,
                               cudaStream_t stream);
/* picd_bn_finalize + picd_bn_apply_mask in one launch (per-replica BatchNorm in training): scale / shift / mean / invstd
 * are OUTPUTS (the backward entry points take them), running statistics are updated in place. */
int picd_bn_finalize_apply(const void* y, const float* sums, int replicas, float count, const float* gamma, const float* beta,
                           float* running_mean, float* running_var, float momentum, float eps, float* scale, float* shift,
                           float* mean, float* invstd, const void* residual, int act, void* out, void* relu_mask, long long P,
                           int C, cudaStream_t stream);
int picd_bn_apply(const void* y, const float* scale, const float* shift, const void* residual, int act, void* out,
                  long long P, int C, cudaStream_t stream);
int picd_bn_bwd_reduce(const void* dz, const void* y, const void* z, const float* scale, const float* shift,
                       const float* mean, const float* invstd, int act, float* sums, long long P, int C,
                       cudaStream_t stream);
int picd_bn_bwd_apply(const void* dz, const void* y, const void* z, const float* scale, const float* shift,
                      const float* mean, const float* invstd, int act, const float* sums, int sums_replicas,
                      float inv_count, const float* param_sums, int param_replicas, float* dgamma, float* dbeta, void* dy,
                      void* dres, long long P, int C, cudaStream_t stream);
int picd_bn_bwd_eval(const void* dz, const void* y, const void* z, const float* scale, const float* shift, int act,
                     void* dy, void* dres, long long P, int C, cudaStream_t stream);

/* ---- pooling, loss, activations, LayerNorm, small linear, packing, data (SURVEY.md K10-K15,K18-K21) ---- */
int picd_pool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C, int OH, int OW, int kh, int kw,
                  int stride, int pad, int mode /*0 max, 1 avg*/, cudaStream_t stream);
int picd_pool_bwd(const void* dy, const void* argmax, void* dx, int N, int H, int W, int C, int OH, int OW, int kh,
                  int kw, int stride, int pad, int mode, cudaStream_t stream);
int picd_gap_fwd(const void* x, void* out, int N, int HW, int C, cudaStream_t stream);
/* ResNet / DenseNet stem: max_pool3x3/s2/p1(relu(y * scale + shift)) in one pass (argmax: one byte per output element, window tap
 * 0..8), and its backward as two kernels that regenerate the pooled gradient per input pixel (the normalised activation and its
 * gradient are never written).  sums = [STAT_REPLICAS][2C] fp32, zeroed by the caller; bwd_apply folds sums / param_sums in place. */
int picd_bn_relu_maxpool_fwd(const void* y, const float* scale, const float* shift, void* out, void* argmax, int N, int H, int W,
                             int C, int OH, int OW, cudaStream_t stream);
int picd_bn_relu_maxpool_bwd_reduce(const void* dpool, const void* argmax, const void* y, const float* scale, const float* shift,
                                    const float* mean, const float* invstd, float* sums, int N, int H, int W, int C, int OH, int OW,
                                    cudaStream_t stream);
int picd_bn_relu_maxpool_bwd_apply(const void* dpool, const void* argmax, const void* y, const float* scale, const float* shift,
                                   const float* mean, const float* invstd, float* sums, int sums_replicas, float inv_count,
                                   const float* count_dev, float* param_sums, int param_replicas, float* dgamma, float* dbeta,
                                   void* dy, int N, int H, int W, int C, int OH, int OW, cudaStream_t stream);
/* picd_gap_fwd with the pixels of a sample split over blocks (few samples x channels, many pixels); scratch fp32 [N, C], zeroed */
int picd_gap_fwd_split(const void* x, float* scratch, void* out, int N, int HW, int C, cudaStream_t stream);
int picd_gap_bwd(const void* dy, void* dx, int N, int HW, int C, cudaStream_t stream);
/* out = x * s[n, c] * m[n]  (squeeze-excite gate + stochastic-depth mask; m nullable); backward in one pass, ds fp32 zeroed by caller */
int picd_channel_scale_fwd(const void* x, const void* s, const float* m, void* out, int N, int HW, int C, cudaStream_t stream);
int picd_channel_scale_bwd(const void* dy, const void* x, const void* s, const float* m, void* dx, float* ds, int N, int HW, int C,
                           cudaStream_t stream);
/* class-weighted softmax cross-entropy (train.py:157-158): loss[0] and d loss/d logits in one launch */
int picd_cross_entropy(const float* logits, const long long* labels, const float* weight, float* loss,
                       float* dlogits, int B, int nc, cudaStream_t stream);
int picd_act_fwd(const void* x, const float* bias, void* out, long long rows, int C, int act, cudaStream_t stream);
int picd_act_bwd(const void* dy, const void* x, const float* bias, void* dx, long long rows, int C, int act,
                 cudaStream_t stream);
/* picd_act_bwd that also accumulates the bias gradient dbias[c] += sum_r dx[r, c] in the same pass (dbias may be NULL). */
int picd_act_bwd_bias(const void* dy, const void* x, const float* bias, void* dx, float* dbias, long long rows, int C,
                      int act, cudaStream_t stream);
int picd_colsum(const void* x, float* out, long long P, int C, cudaStream_t stream);
/* NHWC channel concat (split = 0) / its backward split (split = 1) of n <= 8 bf16 tensors [P, C_i] <-> [P, sum C_i];
 * `parts` and `channels` are host arrays. */
int picd_concat_nhwc(void* const* parts, const int* channels, int n, void* wide, long long P, int split,
                     cudaStream_t stream);
/* out[b] = (argmax_c logits[b, c] == labels[b]) as int64; logits bf16 or fp32 row-major [B, NC]. */
int picd_top1_hits(const void* logits, int is_bf16, const long long* labels, long long* out, int B, int NC,
                   cudaStream_t stream);
int picd_dropout(const void* x, void* out, long long n, float p, unsigned seed, cudaStream_t stream);
/* graph-replay variant: the seed is mixed with the 32-bit word at step_word (device memory) when the kernel runs */
int picd_dropout_dyn(const void* x, void* out, long long n, float p, unsigned seed, const void* step_word,
                     cudaStream_t stream);
/* Transformer residual add fused into the LayerNorm that follows it: s = x + res (bf16) -> sum_out, y = LayerNorm(s).
 * Backward: dx = LayerNorm-backward(dy, s) + add_in (gradient arriving over the residual stream, nullable). */
int picd_add_layernorm_fwd(const void* x, const void* res, const float* gamma, const float* beta, void* sum_out, void* y,
                           float* mean, float* rstd, long long M, int D, float eps, cudaStream_t stream);
int picd_layernorm_bwd_add(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                           const void* add_in, void* dx, float* dgamma, float* dbeta, long long M, int D, cudaStream_t stream);
int picd_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                       long long M, int D, float eps, cudaStream_t stream);
int picd_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                       void* dx, float* dgamma, float* dbeta, long long M, int D, cudaStream_t stream);
int picd_small_linear_fwd(const void* x, const float* w, const float* bias, void* y, int M, int N, int K, int act,
                          int out_fp32, cudaStream_t stream);
int picd_small_linear_bwd(const void* dy, const void* y, const void* x, const float* w, float* dpre_scratch, void* dx,
                          float* dw, float* db, int M, int N, int K, int act, int dy_fp32, int y_fp32,
                          cudaStream_t stream);
/* No counterpart in the reference (it has no native code and lets PyTorch allocate everything).  Same computation
 * and results as picd_small_linear_bwd, for hosts that may not allocate device memory when they call it (a fixed
 * memory budget handed out at start-up, a plug-in under a foreign allocator, a loop that cannot take cudaMalloc's
 * device-wide synchronisation): the M*N-float `dpre` scratch lives in a 1 MiB array built into the library, so
 * M*N must be <= picd_small_linear_noalloc_max_floats() (262144) or the call fails without launching.  The array
 * is shared by all calls on a device and keeps the previous call's contents (none of it needs to be zero: each
 * call overwrites what it uses first); calls whose kernels overlap on the GPU race on it -- keep them on one stream
 * or order them.  The first call on each device resolves the array's device address and should come before any
 * stream capture.  The call itself never allocates or frees. */
int picd_small_linear_bwd_noalloc(const void* dy, const void* y, const void* x, const float* w, void* dx, float* dw,
                                  float* db, int M, int N, int K, int act, int dy_fp32, int y_fp32,
                                  cudaStream_t stream);
int picd_small_linear_noalloc_max_floats(void);
int picd_pack_s2d(const void* x, void* out, int N, int H, int W, int C, int Hp, int Wp, int Cp, int s, int pad_h,
                  int pad_w, int src_u8, cudaStream_t stream);
/* on-device synthetic batches + ImageNet normalisation (replaces dp/loader.py:39-91 and the H2D copy) */
int picd_synthetic_batch(void* img, long long* labels, int batch, int H, int W, int C, int num_classes, unsigned base,
                         int out_fp32, cudaStream_t stream);
int picd_normalize_u8(const void* in, void* out, long long n_el, int C, int out_fp32, cudaStream_t stream);
int picd_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t stream);

/* ---- NVLink peer-memory communicator (replaces DDP's NCCL all-reduce + optimizer.step + SyncBN collectives:
 *      train.py:72-73,124,128 = SURVEY.md N5-N7, K16) ------------------------------------------------------ */
int picd_slab_alloc(size_t bytes, void** out);
int picd_slab_free(void* p);
int picd_ipc_handle_bytes(void);
int picd_ipc_get_handle(void* p, void* handle_out);
int picd_ipc_open(const void* handle, void** out);
int picd_ipc_close(void* p);
int picd_comm_create(int rank, int world, void* const* grad, void* const* master, void* const* shadow,
                     void* const* xchg, void* const* flags, int num_buckets, void** ctx_out);
/* NVSwitch multicast (NVLS) views of the three regions, as mapped by the host (e.g. torch symmetric memory); enables
 * mode 3 of picd_fused_allreduce_update: multimem.ld_reduce / multimem.st instead of W peer loads / stores. */
int picd_comm_set_multicast(void* ctx, void* mc_grad, void* mc_master, void* mc_shadow);
/* Per-rank module-buffer regions (BatchNorm running statistics; n floats each, n % 4 == 0) inside the peer-mapped slabs.
 * A bucket launch with bcast_buffers != 0 copies rank 0's region into every peer's between its two flag barriers: DDP's
 * per-forward buffer broadcast (torch/nn/parallel/distributed.py:2214-2257 via train.py:49,128) without a collective call. */
int picd_comm_set_buffers(void* ctx, void* const* bufs, long long n);
/* Wire format: gbf = per-rank bf16 staging regions of the gradient arenas (mc_gbf: their multicast view or NULL).  wire_bf16: peers
 * read 2 B per gradient element (packed by the owner inside the bucket kernel) instead of 4; shard_master: slice owners keep
 * their fp32 master slices local and push only the bf16 compute copy (2 B instead of 6) -- the host gathers masters on demand. */
int picd_comm_set_wire(void* ctx, void* const* gbf, void* mc_gbf, int wire_bf16, int shard_master);
/* Seconds a kernel may spin on a peer's flag before it traps (default 600); applies to the current device. */
int picd_comm_set_flag_timeout(double seconds);
int picd_comm_destroy(void* ctx);
int picd_fused_allreduce_update(void* ctx, int bucket, int mode, long long start, long long count, float* state0,
                                float* state1, int kind, float lr, float momentum, float beta1, float beta2, float eps,
                                float weight_decay, float bc1, float bc2, int nesterov, unsigned epoch, int blocks,
                                int bcast_buffers, cudaStream_t stream);
/* Same bucket update for CUDA-graph replay: the step-varying scalars (lr, 1-beta1^t, 1-beta2^t, flag epoch as raw bits)
 * are read from the device array dyn[4] when the kernel runs, so one captured launch serves every later step. */
int picd_fused_allreduce_update_dyn(void* ctx, int bucket, int mode, long long start, long long count, float* state0,
                                    float* state1, int kind, float momentum, float beta1, float beta2, float eps,
                                    float weight_decay, int nesterov, int blocks, const float* dyn, int bcast_buffers,
                                    cudaStream_t stream);
/* First kernel of a captured step: increments dyn[3] (step / flag epoch, raw bits) and dyn[4] (replay index, raw bits) and
 * recomputes dyn[1], dyn[2] = 1 - beta^t on the device, so that a graph replay needs no host-to-device copy. */
int picd_dyn_advance(float* dyn, float beta1, float beta2, cudaStream_t stream);
/* Small collectives (SyncBatchNorm): PUSH exchange -- every rank stores its payload into sub-slot `rank` of the slot half at
 * float offset slot_off in every peer's exchange region (the half spans world * sub floats, sub >= payload), signals, and sums
 * the `world` sub-slots of its own slab. */
int picd_p2p_allreduce_small(void* ctx, const float* in, int in_replicas, float* out, int n, long long slot_off,
                             int flag_slot, unsigned epoch, int sub, cudaStream_t stream);
int picd_bn_finalize_sync(void* ctx, const float* sums, int replicas, float count, const float* gamma, const float* beta,
                          float* running_mean, float* running_var, float momentum, float eps, float* scale,
                          float* shift, float* mean, float* invstd, float* total_count, int C, long long slot_off,
                          int flag_slot, unsigned epoch, int sub, cudaStream_t stream);
/* Graph-replay variants of the two small collectives: slot_off = EVEN half of the double-buffered slot, epoch = the epoch
 * at capture time; the kernels add the device word *dyn_delta (steps since capture) and pick the half from the parity. */
int picd_p2p_allreduce_small_dyn(void* ctx, const float* in, int in_replicas, float* out, int n, long long slot_off,
                                 int flag_slot, unsigned epoch, const void* dyn_delta, int slot_floats, int sub,
                                 cudaStream_t stream);
int picd_bn_finalize_sync_dyn(void* ctx, const float* sums, int replicas, float count, const float* gamma, const float* beta,
                              float* running_mean, float* running_var, float momentum, float eps, float* scale,
                              float* shift, float* mean, float* invstd, float* total_count, int C, long long slot_off,
                              int flag_slot, unsigned epoch, const void* dyn_delta, int slot_floats, int sub,
                              cudaStream_t stream);

/* The library's own symmetric slab on the CUDA VMM API (csrc/comm/vmm_slab.cu): a cuMemCreate allocation per rank exported as a
 * POSIX fd (the host moves the fds between processes, e.g. SCM_RIGHTS), every peer's allocation mapped with cuMemMap, and an
 * NVSwitch multicast object (cuMulticastCreate / AddDevice / BindMem) whose mapping the NVLS modes of the bucket kernel use.
 * Protocol: every rank picd_vmm_create -> exchange fds -> picd_vmm_import_peer per peer; rank 0 picd_vmm_mc_create -> its fd to
 * the others -> picd_vmm_mc_import; every rank picd_vmm_mc_add_device; barrier; every rank picd_vmm_mc_bind_map. */
int picd_vmm_create(size_t bytes, int world, void** slab_out, void** base_out, int* fd_out, size_t* padded_bytes_out);
int picd_vmm_import_peer(void* slab, int peer_rank, int fd, void** base_out);
int picd_vmm_multicast_supported(void);
int picd_vmm_mc_create(void* slab, int world, int* fd_out);
int picd_vmm_mc_import(void* slab, int fd);
int picd_vmm_mc_add_device(void* slab);
int picd_vmm_mc_bind_map(void* slab, void** mc_base_out);
int picd_vmm_destroy(void* slab);
int picd_nvls_supported(void);   /* 1/0/-1: NVSwitch multicast available (probe only; the kernels use P2P loads/stores) */

/* ---- runtime ---------------------------------------------------------------------------------------------- */
/* reverse-order bucket plan (torch DDP Reducer semantics); host function */
int picd_plan_buckets(const int64_t* padded_sizes, int n_params, int64_t first_bytes, int64_t cap_bytes,
                      int elem_bytes, int64_t unit /* every bucket end is a multiple of this many elements */,
                      int64_t* starts, int64_t* ends, int* param_bucket, int max_buckets);

#ifdef __cplusplus
}
#endif
#endif /* PICD_B200_H */
