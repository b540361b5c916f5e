/* C ABI of libnidt_b200.so -- the sm_100a kernels of neuroimagedisttraining_b200.
 *
 * Conventions
 *   - every function returns 0 on success; otherwise nidt_last_error() (per host thread) says why;
 *   - all pointers are device pointers unless stated; `stream` is a cudaStream_t passed as void*;
 *   - nothing here allocates device memory: callers pass every buffer, including scratch
 *     (the size queries next to each function say how much);
 *   - "slab" = one flat fp32 buffer holding all parameters (then all buffers) of a model, every
 *     entry 16-byte aligned, padding kept at zero (neuroimagedisttraining_b200/runtime/slab.py);
 *   - activations are channels-last (NDHWC) bf16.
 * Citations name the reference code each entry point replaces (paths relative to the reference repo).
 */
#ifndef NIDT_B200_H_
#define NIDT_B200_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* nidt_last_error(void);
int nidt_abi_version(void);

/* ------------------------------------------------------------------ flat-slab ops (flat_ops.cu) */
int nidt_norm_scratch_floats(void);
int nidt_topk_scratch_bytes(void);
int nidt_max_agg_sources(void);

/* Per-step scalars and result pointers read from DEVICE memory by the kernels that need them (the SGD update and the head), so
 * that a captured CUDA graph of a step can be replayed with new values: a tiny kernel writes this struct in front of the replay.
 * The reference re-creates its optimizer and re-reads its python scalars every step (my_model_trainer.py:209,224-231). */
typedef struct {
    float lr, momentum, wd, clip_norm;
    int first_step, reserved;
    uint64_t seed;
    float* loss_acc;       /* training: += mean loss */
    double* counters;      /* evaluation: correct, loss_sum, total */
} nidt_dyn_t;

/* clip_grad_norm_(clip) + SGD(momentum, wd) + `w *= mask` in two launches over the whole slab.
 * ref: fedml_api/standalone/sailentgrads/my_model_trainer.py:209,224-231.  scratch: nidt_norm_scratch_floats() floats. */
int nidt_masked_sgd_step(float* w, const float* g, float* mom, const float* mask, int64_t n, float lr, float momentum,
                         float wd, float clip_norm, int first_step, float* scratch, float* norm_out, void* stream);
/* the same update with lr / momentum / wd / clip / first_step read from *dyn (device memory); mom must be NULL when unused */
int nidt_masked_sgd_step_dyn(float* w, const float* g, float* mom, const float* mask, int64_t n, float* scratch, const nidt_dyn_t* dyn,
                             void* stream);
int nidt_grad_norm(const float* g, int64_t n, float* scratch, float* norm_out, void* stream);

/* out = sum_i weights[i]*srcs[i] in list order (separate mul/add: bit-identical to the reference loop), optional
 * mask over the first n_mask elements.  srcs/weights are HOST arrays.  ref: sailentgrads_api.py:212-227, fedavg_api.py:102-117. */
int nidt_weighted_aggregate(float* out, const float* const* srcs, const float* weights, int k, const float* mask,
                            int64_t n_mask, int64_t n, void* stream);

/* scores += scale*|g*w|  (SNIP |dL/dm| at m = 1).  ref: sailentgrads/snip.py:9-16,60-71. */
int nidt_saliency_accumulate(float* scores, const float* g, const float* w, int64_t n, float scale, void* stream);

/* One cooperative kernel: sum of prunable scores, 4 radix-select passes for the k-th largest normalised score, mask
 * emission (ties kept).  flags[i]: 0 padding, 1 never pruned, 2 prunable.  scratch: nidt_topk_scratch_bytes().
 * ref: sailentgrads/snip.py:80-116. */
int nidt_global_topk_mask(const float* scores, const unsigned char* flags, float* mask, int64_t n, unsigned long long k,
                          void* scratch, void* stream);
int nidt_topk_read_result(const void* scratch, float* threshold, float* norm, unsigned int* kept, void* stream);

/* nnz / Hamming distance / squared distance without host syncs.  `out` counters are added to.
 * ref: fedml_core/trainer/model_trainer.py:49-53, fedml_api/standalone/DisPFL/slim_util.py:10-19. */
int nidt_count_nonzero(const float* a, int64_t n, unsigned long long* out, void* stream);
int nidt_hamming(const float* a, const float* b, int64_t n, unsigned long long* out, void* stream);
int nidt_sq_dist(const float* a, const float* b, int64_t n, float* scratch, float* out, void* stream);
/* dst[j] = src[idx[j]] for `rows` rows of row_bytes bytes each (any length / alignment): assembles device batches of 8-bit volumes
 * from a device-resident dataset.  idx: int64 on the device. */
int nidt_gather_rows_u8(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes, void* stream);
/* the same with at most max_blocks CTAs; src and idx may be pinned host memory (the kernel reads it over PCIe) */
int nidt_gather_rows_u8_capped(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes, int max_blocks, void* stream);
/* the same with the row numbers passed by value in the kernel parameters (idx: HOST array, rows <= 64) and, when ysrc != NULL, the
 * fp32 labels gathered by the same launch (ydst[j] = ysrc[idx[j]]): one launch per batch from a device-resident cache, no index
 * traffic.  Replaces the reference's per-batch HDF5 read + fp32 upload, sailentgrads/my_model_trainer.py:185-199. */
int nidt_gather_rows_u8_args(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes,
                             const float* ysrc, float* ydst, void* stream);
/* asynchronous host->device row copies (rows: HOST int64 array of source row numbers; host_base should be pinned) in one call;
 * runs of consecutive row numbers are merged into one copy each.  _counted also reports how many copies were issued. */
int nidt_copy_rows_h2d(const void* host_base, const int64_t* rows, int n, int64_t row_bytes, void* dst, void* stream);
int nidt_copy_rows_h2d_counted(const void* host_base, const int64_t* rows, int n, int64_t row_bytes, void* dst, void* stream, int* n_copies);
/* Dense 8-bit volumes from their zero-run-skipping transfer form (per image row: start, length, offset of the non-zero span; spans
 * packed back to back): the host sends only the spans and 6 bytes per row, this kernel writes every byte of n_vol <= 32 volumes of
 * rows x W (W <= 255).  packed_off: HOST array of each volume's byte offset in `packed`.  The reference uploads dense fp32 volumes
 * (sailentgrads/my_model_trainer.py:185-199). */
int nidt_expand_sparse_rows_u8(const void* packed, const long long* packed_off, const void* tab, const void* off, void* dst, int n_vol,
                               int rows, int W, void* stream);
/* host side of the same transfer form: span table of a volume cache ([n_vol][rows][W] bytes, W <= 255) on `threads` host threads --
 * per row (first non-zero byte, length up to the last one) into tab, the row's offset among its volume's spans into off, the packed
 * size of every volume into vol_bytes; then the spans copied back to back (vol_off = exclusive prefix of vol_bytes).
 * The reference has no counterpart (it copies dense fp32 batches, my_model_trainer.py:185-199). */
int nidt_sparse_scan_rows(const unsigned char* cache, int64_t n_vol, int rows, int W, unsigned char* tab, unsigned int* off,
                          int64_t* vol_bytes, int threads);
int nidt_sparse_copy_rows(const unsigned char* cache, int64_t n_vol, int rows, int W, const unsigned char* tab, const unsigned int* off,
                          const int64_t* vol_off, unsigned char* packed, int threads);
/* Ditto proximal pull w -= coef*(w - ref).  ref: fedml_api/standalone/ditto/my_model_trainer.py:63-64. */
int nidt_prox_pull(float* w, const float* ref, int64_t n, float coef, void* stream);

/* ------------------------------------------------- DisPFL prune / regrow on one tensor (mask_select.cu) */
/* mask (fp32 0/1) in place: mode 0 clears the k active entries with the smallest |v|, mode 1 sets the k inactive entries with the
 * largest |v| (ties at the threshold in index order: exactly min(k, #candidates) entries change).  One cooperative kernel
 * (4-pass radix select + ordered tie resolution).  ref: fedml_api/standalone/DisPFL/client.py:71-99 (torch.sort per layer). */
int nidt_mask_select_scratch_bytes(void);
int nidt_mask_select_k(const float* v, float* mask, int64_t n, int64_t k, int mode, void* scratch, void* stream);
int nidt_mask_select_read_changed(const void* scratch, unsigned int* changed, void* stream);

/* ------------------------------------------------- BatchNorm3d + ReLU + MaxPool3d, fused (bn_pool.cu) */
/* ref: fedml_api/model/cv/salient_models.py:147-168 (cuDNN BN, ATen ReLU/MaxPool3d there). */
int nidt_bn_bwd_grid(void);
/* pool 3/3 forward that also records the winning tap of every window (uint8, 255 = gated by ReLU), and the backward reductions
 * alone (c12, dgamma, dbeta): the two halves the fused first-layer weight gradient (nidt_conv_in_s2d_wgrad_fused) needs */
int nidt_bn_relu_pool3_fwd_amax(const void* y, void* a, void* ymax, void* amax, int N, int D, int H, int W, int C,
                                const float* scale, const float* shift, int relu, void* stream);
int nidt_bn_relu_pool3_bwd_stats(const void* ymax, const void* da, int N, int D, int H, int W, int C, const float* scale,
                                 const float* shift, const float* mean, const float* invstd, float* partials, float* c12,
                                 float* dgamma, float* dbeta, int relu, void* stream);
/* the three launches of nidt_bn_relu_pool_bwd one at a time (phases: 1 = reduce -> *rows_out rows of [2][C] partial sums in `partials`,
 * 4 = apply with the c12 given), with nidt_bn_bwd_finalize between them: data-parallel training of one client all-reduces the partial
 * sums over its ranks before the finalisation (ref: the SyncBN of fedml_api/model/cv/batchnorm_utils.py:150-237) */
int nidt_bn_relu_pool_bwd_phases(const void* y, const void* ymax, const void* da, void* dy, int N, int D, int H, int W, int C, int pk,
                                 int ps, const float* scale, const float* shift, const float* mean, const float* invstd,
                                 float* partials, float* c12, float* dgamma, float* dbeta, int relu, int phases, int* rows_out,
                                 void* stream);
int nidt_bn_bwd_finalize(const float* partials /*[rows][2][C]*/, int rows, int C, double count, float* c12, float* dgamma /*opt*/,
                         float* dbeta /*opt*/, void* stream);
int nidt_bn_finalize(const float* partials /*[T][2][C]*/, int T, int C, double count, const float* gamma, const float* beta,
                     float* running_mean, float* running_var, float* num_batches, float momentum, float eps, float* scale,
                     float* shift, float* mean, float* invstd, void* stream);
int nidt_bn_eval_prepare(int C, const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                         float eps, float* scale, float* shift, void* stream);
/* a = maxpool(relu(y*scale+shift)); ymax (optional, pool 3/3) = raw y at each window's arg-max, consumed by the backward. */
int nidt_bn_relu_pool_fwd(const void* y, void* a, void* ymax, int N, int D, int H, int W, int C, int pk, int ps,
                          const float* scale, const float* shift, int relu, void* stream);
/* dy, dgamma, dbeta from da.  partials: nidt_bn_bwd_grid()*2*C floats, c12: 2*C floats. */
int nidt_bn_relu_pool_bwd(const void* y, const void* ymax, const void* da, void* dy, int N, int D, int H, int W, int C, int pk,
                          int ps, const float* scale, const float* shift, const float* mean, const float* invstd,
                          float* partials, float* c12, float* dgamma, float* dbeta, int relu, void* stream);

/* ------------------------------------------------------------------ classifier head (head.cu) */
/* Dropout-Linear-ReLU-Dropout-Linear(1)-BCEWithLogits forward + loss + full backward, or eval counters
 * (incl. the reference's double-sigmoid loss).  ref: salient_models.py:171-176, my_model_trainer.py:207-222,239-274. */
int nidt_head_fwd_bwd(const void* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                      float* gW1, float* gb1, float* gW2, float* gb2, void* dfeat, float* loss_acc, double* eval_counters,
                      float* logits_out, int B, int S, int C, int Hd, float p_drop, uint64_t seed, int train,
                      int double_sigmoid, void* stream);
/* the same head with the dropout seed and the loss accumulator (train) / eval counters (eval) read from *dyn (device memory) */
int nidt_head_fwd_bwd_dyn(const void* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                          float* gW1, float* gb1, float* gW2, float* gb2, void* dfeat, float* logits_out, int B, int S, int C, int Hd,
                          float p_drop, int train, int double_sigmoid, const nidt_dyn_t* dyn, void* stream);

/* ------------------------------------------------- tcgen05/TMA implicit-GEMM 3-D convolution (conv_tc.cu) */
/* ref: nn.Conv3d forward/backward at salient_models.py:152-166 (cuDNN fp32).  Cin % 64 == 0; box = tile of output
 * positions (bw*bh*bd*bn <= 128).  dgrad = the same entry point on dY with the `wd` weights and pad' = k-1-pad. */
int nidt_conv3d_m_tiles(int N, int Do, int Ho, int Wo, int bw, int bh, int bd, int bn);
int nidt_conv3d_fprop_tc(const void* x, const void* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi,
                         int Wi, int Cin, int Do, int Ho, int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd,
                         int bn, int block_n, void* stream);
int64_t nidt_conv3d_wgrad_scratch_floats(int Cout, int Cin, int k, int splits);
/* grad: fp32 [Cout][Cin][k^3] (PyTorch layout); scratch: split-K partials, nidt_conv3d_wgrad_scratch_floats() floats. */
int nidt_conv3d_wgrad_tc(const void* dy, const void* x, float* grad, float* scratch, int N, int Di, int Hi, int Wi, int Cin,
                         int Do, int Ho, int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn,
                         int splits, void* stream);
/* nidt_conv3d_wgrad_tc with the SNIP saliency epilogue: the split-K reduction that writes dW also accumulates
 * scores[co][ci][tap] += scale*|dW*w| (w = the layer's fp32 weights).  ref: snip.py:60-71. */
int nidt_conv3d_wgrad_tc_saliency(const void* dy, const void* x, float* grad, float* scratch, int N, int Di, int Hi, int Wi, int Cin,
                                  int Do, int Ho, int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn,
                                  int splits, const float* w, float* scores, float scale, void* stream);
/* Same result as nidt_conv3d_wgrad_tc without a caller-provided scratch buffer: the split-K partials go to a 16 MiB
 * array built into the library (held by every GPU that loads it, until exit).  For callers that cannot allocate device
 * memory where the call happens (custom-call bodies inside another framework's executor, which owns the allocator;
 * deployments whose memory is partitioned once at start-up).  Refuses problems needing more than nidt_conv3d_wgrad_static_capacity_floats() floats
 * (splits*k^3*Cout*Cin).  The first call on a device looks the array's address up: make it at start-up, before any
 * stream capture.  Calls share the array: do not let two of them overlap on the GPU.
 * This entry point has no counterpart in the reference. */
int64_t nidt_conv3d_wgrad_static_capacity_floats(void);
int nidt_conv3d_wgrad_tc_static(const void* dy, const void* x, float* grad, int N, int Di, int Hi, int Wi, int Cin, int Do, int Ho,
                                int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn, int splits,
                                void* stream);
/* fp32 [Cout][Cin][taps] -> bf16 wf [Cout][taps][Cin] and wd [Cin][taps (flipped)][Cout]. */
int nidt_pack_conv_weights(const float* w, int Cout, int Cin, int taps, void* wf, void* wd, void* stream);
/* the same for up to 8 layers in one launch; the arrays are HOST arrays */
int nidt_pack_conv_weights_multi(int n, const float* const* w, const int* Cout, const int* Cin, const int* taps, void* const* wf,
                                 void* const* wd, void* stream);

/* ----------------------------- 3x3x3 / stride-1 convolution with the input halo resident in shared memory (conv_halo.cu) */
/* Same math and layouts as nidt_conv3d_fprop_tc (k = 3, stride = 1; dgrad = the `wd` weights with pad' = 2 - pad), but three
 * input depth-planes stay in shared memory while the CTA slides along the depth axis and the 27 taps are shifted descriptors
 * into them: L2 -> SM traffic per output row drops ~5-9x.  stats_partials: nidt_conv3d_halo_stats_rows() x 2 x Cout floats. */
int nidt_conv3d_halo_stats_rows(void);
int nidt_conv3d_halo_supported(int N, int Do, int Ho, int Wo, int Cin, int Cout, int pad, int block_n);
int nidt_conv3d_halo_fprop(const void* x, const void* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                           int Cin, int Do, int Ho, int Wo, int Cout, int pad, int block_n, int debug_flags, void* stream);

/* ------------------------------------------------------- first layer, Cin = 1 -> 64 (conv_in.cu) */
/* ref: nn.Conv3d(1,64,5,stride 2) at salient_models.py:147.  x is the 8-bit volume; normalisation x*scale is fused. */
int nidt_conv_in_grid(void);
int nidt_conv_in_fprop(const void* x, const float* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi,
                       int Wi, int k, int stride, int pad, float scale, void* stream);
int nidt_conv_in_wgrad(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi, int Wi,
                       int k, int stride, int pad, float scale, void* stream);

/* Same layer for the flagship shape (k = 5, stride 2, no padding) WITHOUT im2col: the 8-bit volume is staged as
 * space-to-depth lines (8 parity channels = one 16-byte UMMA core-matrix row) that tcgen05.mma reads directly through
 * descriptors whose K-chunk stride equals the position stride (conv_in_s2d.cu).  Same contracts as the two above. */
int nidt_conv_in_s2d_supported(int Di, int Hi, int Wi, int k, int stride, int pad);
int64_t nidt_conv_in_s2d_wgrad_scratch_floats(void);
int nidt_conv_in_s2d_fprop(const void* x, const float* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi,
                           int Wi, float scale, void* stream);
/* the same convolution with every 3 consecutive output columns reduced in the epilogue to the one BatchNorm + ReLU + MaxPool3d(3,3) would
 * pick (max of y where gamma >= 0, min where gamma < 0): yw bf16 [N][Do][Ho][Wp][64], aw uint16 [N][Do][Ho][Wp][8] (2 bits per channel: the
 * column), Wp = (Wo - 3) / 3 + 1; the full-resolution output is not written.  nidt_bn_relu_pool3w_fwd finishes the pooling from these
 * (a, and -- both or neither -- ymax / amax as nidt_bn_relu_pool3_fwd_amax would produce them from the full tensor). */
int nidt_conv_in_s2d_fprop_poolw(const void* x, const float* w, const float* bias, const float* gamma, void* yw, void* aw,
                                 float* stats_partials, int N, int Di, int Hi, int Wi, float scale, void* stream);
int nidt_bn_relu_pool3w_fwd(const void* yw, const void* aw, void* a, void* ymax, void* amax, int N, int D, int H, int W, int C,
                            const float* scale, const float* shift, int relu, void* stream);
int nidt_conv_in_s2d_wgrad(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi,
                           int Wi, float scale, void* stream);

/* BatchNorm backward (batch statistics) + ReLU gate + MaxPool(3,3) scatter FUSED into the first layer's weight gradient: four extra
 * warps turn every conv-output tile y into dY in shared memory (dY = cA*y + cB + [position won its window] * da*scale, `amax` from
 * nidt_bn_relu_pool3_fwd_amax, c12 from nidt_bn_relu_pool3_bwd_stats), so dY (506 MB at batch 16) is neither written nor read
 * back.  Result identical to nidt_bn_relu_pool_bwd + nidt_conv_in_s2d_wgrad.  ref: autograd of salient_models.py:147-150. */
int nidt_conv_in_s2d_wgrad_fused(const void* x, const void* y, const void* amax, const void* da, const float* bn_scale,
                                 const float* bn_mean, const float* bn_invstd, const float* c12, float* grad, float* grad_bias,
                                 float* scratch, int N, int Di, int Hi, int Wi, float scale, const float* sal_w, float* sal_scores,
                                 float sal_scale, void* stream);
int nidt_conv_in_s2d_wgrad_saliency(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi,
                                    int Wi, float scale, const float* w, float* scores, float sal_scale, void* stream);

/* ------------------------------------------------------- residual-network glue (resnet_ops.cu) */
/* ref: ResNet_l3 / BasicBlock, fedml_api/model/cv/salient_models.py:13-139 (ATen / cuDNN / cuBLAS under autograd there). */
/* mode 0: out = relu(a + b); 1: out = b > 0 ? a : 0 (ReLU gate, b = forward output); 2: out = a + b.  bf16, n % 8 == 0. */
int nidt_ew_bf16(const void* a, const void* b, void* out, int64_t n, int mode, void* stream);
/* MaxPool3d(k, s, p) with overlapping windows; idx = uint8 tap of the first maximum per output element. */
int nidt_maxpool3d_fwd(const void* x, void* out, void* idx, int N, int D, int H, int W, int C, int k, int s, int p, void* stream);
int nidt_maxpool3d_bwd(const void* dout, const void* idx, void* dx, int N, int D, int H, int W, int C, int k, int s, int p, void* stream);
/* u = dy with s-1 zeros between neighbours: dgrad of a stride-s conv = stride-1 nidt_conv3d_fprop_tc over u with the `wd` weights. */
int nidt_upsample_zero(const void* dy, void* u, int N, int D, int H, int W, int C, int Du, int Hu, int Wu, int s, void* stream);
/* AvgPool3d(k): bf16 NDHWC -> fp32 features in the NCDHW flatten order of the reference's fc weights, and its backward. */
int nidt_avgpool3d_fwd(const void* x, float* feat, int N, int D, int H, int W, int C, int k, void* stream);
int nidt_avgpool3d_bwd(const float* dfeat, void* dx, int N, int D, int H, int W, int C, int k, void* stream);
/* Linear(F,Hd) -> Linear(Hd,1) -> BCEWithLogits forward, loss / eval counters and full backward.  hidden, dhidden: B*Hd floats. */
int nidt_linear_head_fwd_bwd(const float* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                             float* gW1, float* gb1, float* gW2, float* gb2, float* dfeat, float* hidden, float* dhidden,
                             float* loss_acc, double* eval_counters, float* logits_out, int B, int F, int Hd, int train,
                             int double_sigmoid, void* stream);

/* ------------------------------------------- first-layer weight gradient from pooled tensors + cached input Gram matrices (conv_in_alg.cu) */
/* dW of Conv3d(1,64,5,stride 2) -> BN -> ReLU -> MaxPool(3,3) without reading y or writing dY: the dense part of the gradient is a
 * product with the Gram matrix of the INPUT patches (128 x 128 fp32 per sample, tap 125 = constant 1), which depends on the volume
 * only and is computed once per sample; per step only one patch per pooling window and channel is touched.  rows / slots are HOST
 * int arrays (<= nidt_conv_in_gram_max_samples()).  ref: autograd through salient_models.py:147-150. */
int nidt_conv_in_gram_floats(void);
int nidt_conv_in_gram_max_samples(void);
int nidt_conv_in_alg_grid(void);
int64_t nidt_conv_in_gram_scratch_floats(int n, int Di);
int64_t nidt_conv_in_alg_scratch_floats(void);
int nidt_conv_in_gram_samples(const void* x_u8, const int* rows, const int* slots, int n, float* cache, float* scratch, int Di, int Hi, int Wi,
                              float scale, void* stream);
int nidt_conv_in_gram_sum(const float* cache, const int* slots, int n, double* gb, void* stream);
int nidt_conv_in_alg_wgrad(const void* x_u8, const void* amax, const void* da, const float* W, const float* bias, const float* bn_scale,
                           const float* mean, const float* invstd, const float* c12, const double* gb, float* grad_w, float* grad_b,
                           float* scratch, int N, int Di, int Hi, int Wi, float scale, const float* sal_w, float* sal_scores,
                           float sal_scale, void* stream);

/* ------------------------------------------------------- native step executor (executor.cu) */
/* One call runs a whole forward / training step / evaluation batch of a conv-BN-ReLU-pool stack + the classifier head: the
 * ~45 kernel launches of a step are issued from native code instead of one Python -> C transition each (host time per step
 * 0.5 ms -> < 0.1 ms, which is what small evaluation batches and the end-to-end path are bound by).  The plan holds shapes,
 * parameter offsets (in floats) into the flat slab and the caller-allocated activation / scratch buffers; nothing is allocated. */
typedef struct {
    int cin, cout, k, stride, pad, pool_k, pool_s, relu, has_bias;
    int in_d, in_h, in_w, c_d, c_h, c_w;                  /* input extents, conv-output extents */
    int64_t off_w, off_b, off_gamma, off_beta, off_rm, off_rv, off_nbt;
    void *y, *dy, *a, *da, *ymax;                         /* conv output, its gradient, block output, its gradient, arg-max values */
    float *scale, *shift, *mean, *invstd, *c12, *bwd_partials, *stats;
    void *wf, *wd;                                        /* packed bf16 weights (blocks >= 1) */
    float* wg_scratch;
    int stats_tiles, s2d, block_n, dgrad_block_n, splits;
    int box[4], dbox[4];                                  /* M-tile boxes (bw,bh,bd,bn) of fprop/wgrad and of dgrad at max_batch */
    void* amax;                                           /* block 0, optional: winning pool taps (uint8), written by the forward when bwd0_mode != 0 */
    double* gram;                                         /* block 0, bwd0_mode 2: the batch's patch Gram matrix (nidt_conv_in_gram_sum) */
    float* alg_scratch;                                   /* block 0, bwd0_mode 2: nidt_conv_in_alg_scratch_floats() floats */
    void *yw, *aw;                                        /* block 0, optional: conv output reduced along W in the conv epilogue (nidt_conv_in_s2d_fprop_poolw)
                                                             -- used by evaluation batches and by training steps with bwd0_mode 2, which never read y */
    int bwd0_mode;                                        /* block 0 backward: 0 = dY written, then wgrad; 1 = dY formed inside the wgrad kernel
                                                             (nidt_conv_in_s2d_wgrad_fused); 2 = no pass over y / dY at all (nidt_conv_in_alg_wgrad) */
    int poolw_train;                                      /* block 0: use yw / aw in training steps too (bwd0_mode 2 only); evaluation always does */
} nidt_block_t;
typedef struct {
    int n_blocks, max_batch;
    nidt_block_t blocks[8];
    int64_t off_fc1_w, off_fc1_b, off_fc2_w, off_fc2_b;
    int head_s, head_c, head_hidden;
    float p_drop, input_scale;
    float *loss_acc, *eval_logits, *norm_scratch;
} nidt_plan_t;
/* nidt_plan_train_step / nidt_plan_eval as CUDA-graph replays: the launch sequence is captured once per distinct set of
 * pointer arguments (plan, x, y, batch, slab, grad, mom, mask, do_update / double_sigmoid) and replayed afterwards; the scalars of
 * the step (lr, momentum, wd, clip, first_step, seed) and the result pointers (loss_acc / counters) travel through `dyn`, a
 * caller-allocated DEVICE nidt_dyn_t that a one-thread kernel fills in front of every replay.  Kernel-to-kernel dependencies are
 * resolved on the device, so the ~48 launches of a step neither wait for the host nor for command fetches over a PCIe link that
 * is busy with the next client's upload.  x / y must therefore be stable buffers (the stagers gather every batch into one).
 * The reference has no counterpart (eager autograd, my_model_trainer.py:201-235). */
int nidt_plan_train_step_graph(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, float* slab, float* grad, float* mom,
                               const float* mask, int64_t n_params, float lr, float momentum, float wd, float clip_norm, int first_step,
                               uint64_t seed, int do_update, float* loss_acc, nidt_dyn_t* dyn, int* launches, void* stream);
int nidt_plan_eval_graph(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, const float* slab, double* counters,
                         int double_sigmoid, nidt_dyn_t* dyn, int* launches, void* stream);
/* drop the cached graphs of one plan (plan == NULL: all); the number of cached graphs */
int nidt_plan_graph_clear(const nidt_plan_t* plan);
int nidt_plan_graph_count(void);
/* weights of the attached slab -> bf16 packed copies (call when the slab changed outside a training step) */
int nidt_plan_pack_weights(const nidt_plan_t* plan, const float* slab, void* stream);
/* forward + head + backward (gradients into `grad`, params-first layout of the slab) and, when do_update != 0, the masked SGD
 * update + re-pack of the weights.  Returns the number of kernel launches through *launches. */
int nidt_plan_train_step(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, float* slab, float* grad, float* mom,
                         const float* mask, int64_t n_params, float lr, float momentum, float wd, float clip_norm, int first_step,
                         uint64_t seed, int do_update, int* launches, void* stream);
/* One SNIP evaluation: forward (batch statistics, dropout on) + backward with the saliency epilogues -- the conv layers'
 * scores += scale*|dW*w| leave their wgrad reductions directly, the classifier's by one flat pass over its slab segment.
 * `scores` has the slab's parameter layout.  BN running statistics of `slab` ARE updated (callers that score a model without
 * touching it save / restore that tail).  ref: snip.py:38-71. */
int nidt_plan_saliency_step(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, float* slab, float* grad, float* scores,
                            float scale, int64_t n_params, uint64_t seed, int* launches, void* stream);
/* struct layout probes for bindings: sizeof, and offsetof for a few fields (see executor.cu for the field numbers) */
int nidt_plan_sizeof(void);
int nidt_block_sizeof(void);
int nidt_plan_offsetof(int field);
/* the box of output positions (bw,bh,bd,bn; product <= 128) the executor uses for a conv output of extent WxHxD at batch N */
int nidt_choose_box(int W, int H, int D, int N, int* out);
/* eval-mode forward + eval head (counters may be null: logits only) */
int nidt_plan_eval(const nidt_plan_t* plan, const void* x_u8, const float* y, int batch, const float* slab, double* counters,
                   int double_sigmoid, int* launches, void* stream);

/* ------------------------------------- compute fused with its collective over NVLink symmetric memory (sym_agg.cu) */
/* All ranks call these with the same arguments order; the kernels wait for each other through release/acquire flags.
 * partial_ptrs / result_ptrs / flag_ptrs: HOST arrays of `world` peer-mapped device pointers (identical order on every rank);
 * flag pads hold nidt_sym_flag_words() uint32, zero before the first call; `epoch` grows by 1 per call on every rank.
 * mc_*: NVLS multicast addresses of the same buffers (or null). */
int nidt_sym_max_ranks(void);
int nidt_sym_flag_words(void);
int nidt_sym_max_local(void);          /* client slabs that can be weighted inside one nidt_sym_aggregate call */
int nidt_sym_bulk_max_ranks(void);     /* ranks the shared-memory (bulk-async) data path supports */
int nidt_sym_hist_words(void);         /* uint32 words of a W2 histogram-exchange pad */
int nidt_sym_sal_scratch_bytes(void);  /* bytes of local scratch of W2 */
/* W1: out = mask * sum_ranks sum_local wt*src.  Sample-count weighting of the rank's k_local client slabs (k_local == 0:
 * partial_ptrs[rank] already holds the rank's contribution) + reduce-scatter + SalientGrads mask + all-gather (+ write-back when
 * out_local is neither NULL nor result_ptrs[rank]), one launch per rank.  mode: 0 P2P loads/stores through registers, 1 NVLS
 * multimem.ld_reduce / multimem.st, 2 P2P through shared memory with bulk-async copies; mm_frac_q8 (0..256): share of every
 * chunk sent through the in-switch path next to the P2P path; ctas: grid size (0 = auto).
 * ref: sailentgrads_api.py:212-227 (CPU dict loop). */
int nidt_sym_aggregate(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial, void* mc_result,
                       int rank, int world, const float* const* local_srcs, const float* local_wts, int k_local, const float* mask,
                       int64_t n_mask, int64_t n, float* out_local, uint32_t epoch, int mode, int mm_frac_q8, int ctas, void* stream);
/* the first interface of W1: partial already in place, P2P (registers) or multimem */
int nidt_sym_reduce_mask_broadcast(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial,
                                   void* mc_result, int rank, int world, const float* mask, int64_t n_mask, int64_t n,
                                   float* out_local, uint32_t epoch, int use_multimem, void* stream);
/* W2: cross-rank saliency reduce-scatter fused with the global top-k threshold and the mask emission: every rank histograms only
 * its shard of the reduced scores, the 2048-bin histograms (3 radix passes) and shard sums are exchanged through hist_ptrs
 * (peer-mapped pads of nidt_sym_hist_words() uint32), the mask is all-gathered into mask_ptrs (peer-mapped, n floats).  Uses a
 * flag pad of its own (6 barriers per call).  scratch: nidt_sym_sal_scratch_bytes() bytes.  ref: snip.py:80-140. */
int nidt_sym_saliency_topk_mask(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial,
                                void* mc_result, void* const* hist_ptrs, void* const* mask_ptrs, int rank, int world, int64_t n,
                                const unsigned char* flags, float* mask, unsigned long long k, void* scratch, float* scores_out,
                                uint32_t epoch, int use_multimem, void* stream);
int nidt_sym_saliency_read_result(const void* scratch, float* threshold, float* norm, unsigned int* kept, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NIDT_B200_H_ */
