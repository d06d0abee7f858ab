/* cnx_b200.h — public C ABI of libcnx_b200.so (sm_100a only, no fallback path).
 *
 * The reference (Garhlz/image_classification) has no FFI of its own: its hot path is reached
 * through torch.nn.Module calls into the un-vendored `timm` package.  Each entry
 * point below therefore cites the reference call site / third-party op sequence it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated otherwise; the library never allocates,
 *     frees or retains device memory; all work is enqueued on `stream` (a cudaStream_t) with no
 *     internal synchronisation;
 *   - activations are NHWC bf16, i.e. row-major [M, C] with M = B*H*W tokens; parameters and
 *     their gradients are fp32 in the exact shapes timm uses;
 *   - return value: 0 on success, negative cnx_status otherwise (never throws, never exits);
 *     cnx_last_error_string() gives the reason for the calling thread;
 *   - C must be a multiple of 8 and splittable into <= 8 equal slabs of <= 128 channels
 *     (ConvNeXt-T/S/B/L/XL widths all qualify).
 */
#ifndef CNX_B200_H_
#define CNX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum cnx_status {
  CNX_OK = 0,
  CNX_ERR_BAD_SHAPE = -1,
  CNX_ERR_BAD_ARG = -2,
  CNX_ERR_WORKSPACE = -3,
  CNX_ERR_NOT_SM100 = -4,
  CNX_ERR_DRIVER = -5,
  CNX_ERR_LAUNCH = -6
} cnx_status;

typedef enum cnx_op {
  CNX_OP_DWCONV_LN_FWD = 0,
  CNX_OP_DWCONV_LN_BWD = 1,
  CNX_OP_MLP_FWD = 2,
  CNX_OP_MLP_BWD = 3,
  CNX_OP_MIX_BATCH = 4,
  CNX_OP_SOFTCE = 5,
  CNX_OP_EMA = 6
} cnx_op;

/* GEMM epilogues of cnx_gemm_bf16 (the MLP entry points below are thin presets of it). */
typedef enum cnx_epilogue {
  CNX_EPI_STORE_BF16 = 0,    /* D0 = acc                                   */
  CNX_EPI_BIAS_GELU = 1,     /* D0 = z = acc + bias ; D1 = gelu(z)         */
  CNX_EPI_BIAS_RESIDUAL = 2, /* D0 = acc + bias + R                        */
  CNX_EPI_GELU_GRAD = 3,     /* D0 = acc * gelu'(R) ; colsum += sum_m D0   */
  CNX_EPI_STORE_F32 = 4,     /* D0 (fp32, zero-filled by caller) += acc    */
  CNX_EPI_BIAS = 6,          /* D0 = acc + bias (patchify convolutions)    */
  CNX_EPI_LN_BWD = 7         /* D0 = LayerNorm backward of acc (internal preset: cnx_mlp_bwd_dconv) */
} cnx_epilogue;

/* One entry of the host-generated MixUp/CutMix plan (see cnx_mix_batch). 32 bytes. */
typedef struct cnx_mix_plan_entry {
  int32_t mode;          /* 0 passthrough, 1 mixup, 2 cutmix */
  int32_t partner;       /* index of the second sample inside the batch */
  float lam;             /* float32(lam) */
  float one_minus_lam;   /* float32(1.0 - lam), subtraction in float64 */
  int32_t bbx1, bby1, bbx2, bby2; /* cutmix box, [x1,x2) x [y1,y2) */
} cnx_mix_plan_entry;

int cnx_version(void);
const char* cnx_last_error_string(void);
/* 0 if the current CUDA device is sm_100, CNX_ERR_NOT_SM100 otherwise. */
int cnx_check_device(void);
/* 1 if a kernel gave up waiting on an mbarrier since the last call (then that launch's results
 * are garbage), 0 otherwise; clears the condition.  Call after synchronising the stream.
 * The condition is STICKY: from the moment a kernel has recorded a timeout, every cnx_* entry
 * point that launches work returns CNX_ERR_LAUNCH (cnx_last_error_string() says why) until this
 * call clears it, so a training loop cannot keep stepping on garbage. */
int cnx_debug_timeout_flag(void);
/* Test hook: set (1) or clear (0) the timeout condition by hand. */
void cnx_debug_set_timeout_flag(int value);
/* Debug / tuning: override the GEMM's tiling heuristics for the following calls (development tool
 * behind tools/gemm_sweep.py; -1 = automatic).  stages: operand-ring depth actually used (<= the
 * compiled depth); pair: 0/1 CTA-pair kernel; bn: 128/256 column tile; splits: split-K count of the
 * fp32 reduction epilogue. */
void cnx_debug_gemm_tuning(int stages, int pair, int bn, int splits);
/* Debug / tuning: epilogue warps of the bf16-output GEMM kernels for the following calls: 8, 16, or
 * 0 = automatic (16 for the GELU / GELU-gradient epilogues, 8 otherwise). */
void cnx_debug_gemm_epilogue_warps(int n);
/* Debug / tuning: 64 = every bf16 epilogue stages through 64-column chunks (the layout of the
 * epilogues with a residual operand); 0 = automatic (32-column chunks and one more operand stage for
 * the 8-warp epilogues without a residual operand). */
void cnx_debug_gemm_staging_chunk(int cols);
/* Debug: clusters of `cluster_size` GEMM CTAs the device can hold at once (occupancy probe). */
int cnx_debug_max_clusters(int cluster_size);
/* Scratch bytes an op needs for the given problem (B*H*W*C*2 for CNX_OP_DWCONV_LN_BWD, else 0). */
size_t cnx_workspace_bytes(int op, int B, int H, int W, int C);

/* Fold LN affine into fc1 and layer-scale into fc2, cast to bf16 in both orientations, and
 * transpose the depthwise taps to [49][C].  Inputs are the nine timm ConvNeXtBlock parameters
 * (conv_dw.weight [C,1,7,7], norm.weight/bias [C], mlp.fc1.weight [4C,C], mlp.fc1.bias [4C],
 * mlp.fc2.weight [C,4C], mlp.fc2.bias [C], gamma [C]); call once per weight version
 * (after optimizer.step(), /root/reference/train_advanced_v4.py:229). */
int cnx_prep_block_weights(const float* conv_dw_w, const float* norm_w, const float* norm_b, const float* fc1_w,
                           const float* fc1_b, const float* fc2_w, const float* fc2_b,
                           const float* gamma, float* wt_dw /*[49][C]*/, void* w1f /*[4C][C] bf16*/,
                           void* w1fT /*[C][4C] bf16*/, float* b1f /*[4C]*/,
                           void* w2f /*[C][4C] bf16*/, void* w2fT /*[4C][C] bf16*/,
                           float* b2f /*[C]*/,
                           float* rb /*[4C][2]: (rowsum of the bf16 w1f row, -b1f), may be NULL*/,
                           int C, void* stream);

/* Adjoint of cnx_prep_block_weights: gradients of the folded tensors -> gradients of the timm
 * parameters (d_fc1_b equals db1f and d_conv_dw_b equals dbias of cnx_dwconv_ln_bwd). */
int cnx_unfold_block_grads(const float* dw1f, const float* db1f, const float* dw2f,
                           const float* db2f, const float* dwt_dw, const float* norm_w,
                           const float* norm_b, const float* fc1_w, const float* fc2_w,
                           const float* fc2_b, const float* gamma, float* d_conv_dw_w,
                           float* d_norm_w, float* d_norm_b, float* d_fc1_w, float* d_fc2_w,
                           float* d_fc2_b, float* d_gamma, int C, void* stream);

/* conv_dw (7x7 depthwise, pad 3, +bias) -> permute -> LayerNorm(C, eps) WITHOUT the affine
 * (folded into fc1).  Replaces the first three ops of timm ConvNeXtBlock.forward reached from
 * /root/reference/train_advanced_v4.py:134-135.  x, xhat: [B,H,W,C] bf16; mean, rstd: [B*H*W]. */
int cnx_dwconv_ln_fwd(const void* x, const float* wt_dw, const float* conv_dw_b, void* xhat,
                      float* mean, float* rstd, int B, int H, int W, int C, float eps, void* stream);

/* Backward of the above plus the shortcut edge: dx = dy + dwconv^T(LN^T(dxhat)); dwt [49][C],
 * dbias [C] and dy_colsum [C] (= gradient of the folded fc2 bias) are overwritten. */
int cnx_dwconv_ln_bwd(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                      const void* dy, const float* wt_dw, void* dx, float* dwt, float* dbias,
                      float* dy_colsum, int B, int H, int W, int C, void* workspace,
                      size_t workspace_bytes, void* stream);

/* Same, but dwt / dbias / dy_colsum are ACCUMULATED into (+=): used inside a gradient-accumulation
 * window (/root/reference/train_advanced_v4.py:225, accumulate 2) so that the adjoint of the
 * weight folding runs once per optimizer step instead of once per micro-batch. */
int cnx_dwconv_ln_bwd_acc(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                          const void* dy, const float* wt_dw, void* dx, float* dwt, float* dbias,
                          float* dy_colsum, int B, int H, int W, int C, void* workspace,
                          size_t workspace_bytes, void* stream);

/* timm LayerNorm2d (channels-last LayerNorm WITH affine) of the stem and of every stage's
 * `downsample` (un-vendored timm; reached from /root/reference/train_advanced_v4.py:131,134-135;
 * SURVEY.md §8 f-1): x, y [M, C] bf16 rows (NHWC), weight / bias [C] fp32, mean / rstd [M] fp32
 * saved for the backward.  C multiple of 8, <= 1024. */
int cnx_layernorm_fwd(const void* x, const float* weight, const float* bias, void* y, float* mean,
                      float* rstd, int64_t M, int C, float eps, void* stream);
/* Its backward: dx [M, C] bf16; dweight, dbias [C] fp32 are overwritten. */
int cnx_layernorm_bwd(const void* dy, const void* x, const float* weight, const float* mean,
                      const float* rstd, void* dx, float* dweight, float* dbias, int64_t M, int C,
                      void* stream);
/* The k x k stride-k ("patchify") convolutions of the stem and the downsample layers as plain GEMMs
 * (SURVEY.md 8 f-1): their im2col matrix is a permutation of the input, [B*H/k*W/k, k*k*C] with
 * inner order (kh, kw, c) - the layout of a channels_last conv weight [Cout][kh][kw][C].
 * cnx_layernorm_patch_fwd: LayerNorm2d whose OUTPUT rows are written patch-major (downsample:
 *   LayerNorm2d -> 2x2 s2 conv); mean / rstd stay indexed by input pixel.
 * cnx_layernorm_patch_bwd: its backward; dy_patch is the patch-major gradient the conv's data-
 *   gradient GEMM produced, dx is in pixel order.  dweight / dbias are overwritten.
 * cnx_patchify_nchw_f32: stem input [B,Cin,H,W] fp32 -> patch-major bf16 (k*k*Cin % 8 == 0). */
int cnx_layernorm_patch_fwd(const void* x, const float* weight, const float* bias, void* y_patch,
                            float* mean, float* rstd, int B, int H, int W, int C, int patch,
                            float eps, void* stream);
int cnx_layernorm_patch_bwd(const void* dy_patch, const void* x, const float* weight,
                            const float* mean, const float* rstd, void* dx, float* dweight,
                            float* dbias, int B, int H, int W, int C, int patch, void* stream);
int cnx_patchify_nchw_f32(const float* x, void* P, int B, int Cin, int H, int W, int patch,
                          void* stream);
/* out[c] = sum_m x[m,c] for a bf16 [M,C] matrix (fp32 result, overwritten): the bias gradient of the
 * stem / downsample convolutions (sum of grad_output over batch and space in NHWC). */
int cnx_colsum_bf16(const void* x, float* out, int64_t M, int C, void* stream);
/* Depthwise k x k convolution, k in {3, 5}, stride in {1, 2}, padding k/2, no bias, NHWC bf16 with
 * fp32 accumulation: the `conv_dw` of timm efficientnet_b2's DepthwiseSeparableConv /
 * InvertedResidual blocks (model built at /root/reference/previous/V3-efb2/train_advanced_v3.0.py:330;
 * SURVEY.md §8 f-4, BASELINE config 5).  Replaces F.conv2d(x, w, groups=C) and its autograd backward.
 *   x  [B,H,W,C] bf16;  wt [k*k, C] fp32 tap-major (wt[kh*k+kw, c] = conv.weight[c, 0, kh, kw]);
 *   y / dy [B,Ho,Wo,C] bf16 with Ho = (H + 2*(k/2) - k) / stride + 1.
 * cnx_dwconv_kxk_bwd: dx [B,H,W,C] bf16 (may be NULL), dwt [k*k, C] fp32 (may be NULL; overwritten,
 * or += when accumulate != 0).  C % 8 == 0. */
int cnx_dwconv_kxk_fwd(const void* x, const float* wt, void* y, int B, int H, int W, int C, int k,
                       int stride, void* stream);
int cnx_dwconv_kxk_bwd(const void* dy, const void* x, const float* wt, void* dx, float* dwt, int B,
                       int H, int W, int C, int k, int stride, int accumulate, void* stream);
/* Global average pool over the pixels of an NHWC bf16 map: y[b,c] = mean_p x[b,p,c] (fp32 sum,
 * bf16 result) and its backward dx[b,p,c] = g[b,c] / HW.  Replaces `nn.AdaptiveAvgPool2d(1)` of
 * head.global_pool (/root/reference/train_advanced_v4.py:140) and of the auxiliary classifiers
 * (:119-121) on the channels-last stream.  C % 8 == 0. */
int cnx_avgpool_nhwc_fwd(const void* x, void* y, int B, int HW, int C, void* stream);
int cnx_avgpool_nhwc_bwd(const void* g, void* dx, int B, int HW, int C, void* stream);

/* BatchNorm2d with batch statistics (+ optional SiLU) over an NHWC bf16 map, the `bn1 / bn2 / bn3`
 * (timm BatchNormAct2d) of efficientnet_b2's MBConv blocks (model built at
 * /root/reference/previous/V3-efb2/train_advanced_v3.0.py:330; forward / backward :100-104;
 * SURVEY.md §8 f-4).  Replaces F.batch_norm(training) -> F.silu and their autograd backward.
 *   cnx_bn_stats   : sums[0..C) = sum_m d[m,c], sums[C..2C) = sum_m d[m,c]^2 with d = x - x[0,c]
 *                    (fp32, overwritten; shifted sums: no cancellation for channels whose mean is
 *                    far from zero).  cnx_bn_act_fwd must be given the same x, and y must not alias x.
 *   cnx_bn_act_fwd : y = act((x - mean) * rstd * weight + bias), act 0 = identity, 1 = SiLU.
 *                    sums != NULL: training (mean / biased var from sums over M = B*HW rows;
 *                    running_mean / running_var, when given, are updated with `momentum` and the
 *                    unbiased variance like nn.BatchNorm2d).  sums == NULL: eval, statistics are
 *                    running_mean / running_var.  mean_out / rstd_out [C] are saved for backward.
 *                    pool_sum [B,C] fp32 (may be NULL; overwritten) = sum over the HW pixels of
 *                    image b of y: the squeeze-excite average pool without another pass.
 *   cnx_bn_act_bwd : dx [M,C] bf16, dweight [C], dbias [C] fp32 (overwritten) from g = dL/dy and
 *                    the saved x, mean, rstd (u = xhat*w + b is recomputed for SiLU').
 * C % 8 == 0. */
int cnx_bn_stats(const void* x, float* sums, int64_t M, int C, void* stream);
int cnx_bn_act_fwd(const void* x, const float* sums, const float* weight, const float* bias,
                   float* running_mean, float* running_var, float momentum, float eps, int act,
                   void* y, float* mean_out, float* rstd_out, float* pool_sum, int B, int HW, int C,
                   void* stream);
int cnx_bn_act_bwd(const void* g, const void* x, const float* weight, const float* bias,
                   const float* mean, const float* rstd, int act, void* dx, float* dweight,
                   float* dbias, int64_t M, int C, void* stream);
/* Squeeze-excite scaling of an NHWC bf16 map by a per-image, per-channel gate (timm SqueezeExcite /
 * torchvision SqueezeExcitation: x * sigmoid(fc2(silu(fc1(mean_p x))))); the two tiny [B,C] layers
 * stay with the caller.
 *   cnx_se_scale_fwd        : y[b,p,c] = x[b,p,c] * gate[b,c]              (gate fp32 [B,C])
 *   cnx_se_scale_bwd_reduce : dgate[b,c] = sum_p g[b,p,c] * x[b,p,c]       (fp32, overwritten)
 *   cnx_se_scale_bwd_apply  : dx[b,p,c] = g[b,p,c] * gate[b,c] + dpool[b,c] / HW, where dpool is
 *                             the gradient that came back through the gate layers to the pooled
 *                             mean: scaling path and pooling path in one pass.            */
int cnx_se_scale_fwd(const void* x, const float* gate, void* y, int B, int HW, int C, void* stream);
int cnx_se_scale_bwd_reduce(const void* g, const void* x, float* dgate, int B, int HW, int C,
                            void* stream);
int cnx_se_scale_bwd_apply(const void* g, const float* gate, const float* dpool, void* dx, int B,
                           int HW, int C, void* stream);
/* The two gate layers of the squeeze-excite block on the pooled [B,C] vector (fp32):
 *   pooled = pool_sum * pool_scale;  hidden = w1 pooled + b1  ([B,Cs], saved);
 *   gate = sigmoid(w2 silu(hidden) + b2)  ([B,C]).
 * w1 [Cs,C] is the reduce convolution's weight, w2t [Cs,C] the expand convolution's weight
 * TRANSPOSED (so that all reads are coalesced along C).
 * cnx_se_gate_bwd: from dgate [B,C] (cnx_se_scale_bwd_reduce) -> dpool [B,C] (gradient at the
 * pooled vector, for cnx_se_scale_bwd_apply), dw1 [Cs,C], db1 [Cs], dw2t [Cs,C], db2 [C] (all
 * overwritten); dz2 [B,C] and dh [B,Cs] are caller-owned scratch.  Cs * 32 B and B * 64 B of
 * shared memory must fit 48 KB. */
int cnx_se_gate_fwd(const float* pool_sum, float pool_scale, const float* w1, const float* b1,
                    const float* w2t, const float* b2, float* hidden, float* gate, int B, int C,
                    int Cs, void* stream);
int cnx_se_gate_bwd(const float* dgate, const float* gate, const float* hidden,
                    const float* pool_sum, float pool_scale, const float* w1, const float* w2t,
                    float* dz2, float* dh, float* dpool, float* dw1, float* db1, float* dw2t,
                    float* db2, int B, int C, int Cs, void* stream);

/* Generic tensor-core GEMM: D[M,N] = epilogue(A[M,K] * B[N,K]^T), bf16 in, fp32 accumulate.
 * a_mn_major / b_mn_major != 0: the operand is stored [K, rows] (row-major) instead. */
int cnx_gemm_bf16(const void* A, int64_t lda, int a_mn_major, const void* B, int64_t ldb,
                  int b_mn_major, int M, int N, int K, int epilogue, void* D0, int64_t ldd0,
                  void* D1, int64_t ldd1, const float* bias, const void* R, int64_t ldr,
                  float* colsum, void* stream);

/* mlp.fc1 + GELU(erf): z = xhat @ w1f^T + b1f (bf16), h = gelu(z) (bf16); [M,4C] each. */
int cnx_fc1_gelu_fwd(const void* xhat, const void* w1f, const float* b1f, void* z, void* h, int M,
                     int C, void* stream);
/* mlp.fc2 * gamma + shortcut: y = x + h @ w2f^T + b2f   ([M,C] bf16). */
int cnx_fc2_residual_fwd(const void* h, const void* w2f, const float* b2f, const void* x, void* y,
                         int M, int C, void* stream);
/* dz = (dy @ w2f) * gelu'(z) ([M,4C] bf16); db1f[4C] (zero-filled by caller) += column sums. */
int cnx_mlp_bwd_dz(const void* dy, const void* w2fT, const void* z, void* dz, float* db1f, int M,
                   int C, void* stream);
/* LayerNorm backward WITHOUT a pass over dxhat (round 2).  dxhat = dz @ w1f and z = xhat @ w1f^T +
 * b1f, hence per token m
 *     s1[m] = sum_c dxhat[m,c]             = dz[m,:] . rowsum(w1f)
 *     s2[m] = sum_c dxhat[m,c] * xhat[m,c] = dz[m,:] . (z[m,:] - b1f)
 * cnx_mlp_bwd_dz_stats = cnx_mlp_bwd_dz that also accumulates rowstats[m] = (s1, s2) ([M][2] fp32,
 * zero-filled by the caller) in its epilogue; `rb` [4C][2] fp32 = (rowsum of the bf16 w1f row, -b1f),
 * written by cnx_prep_block_weights.  cnx_mlp_bwd_dconv = the fc1 data-gradient GEMM whose epilogue
 * applies  dconv = rstd * (dxhat - s1/C - xhat * s2/C)  ([M,C] bf16): the gradient of the depthwise
 * conv output, ready for cnx_dwconv_bwd.  Replaces cnx_mlp_bwd_dxhat + the LayerNorm-backward half
 * of cnx_dwconv_ln_bwd for every channel count. */
int cnx_mlp_bwd_dz_stats(const void* dy, const void* w2fT, const void* z, const float* rb, void* dz,
                         float* db1f, float* rowstats, int M, int C, void* stream);
int cnx_mlp_bwd_dconv(const void* dz, const void* w1fT, const void* xhat, const float* rowstats,
                      const float* rstd, void* dconv, int M, int C, void* stream);
/* Depthwise 7x7 conv backward alone (the conv half of cnx_dwconv_ln_bwd): dx = dy + conv^T(dconv),
 * dwt[49,C], dbias[C], dy_colsum[C] overwritten (accumulate = 0) or += (accumulate != 0). */
int cnx_dwconv_bwd(const void* dconv, const void* x, const void* dy, const float* wt_dw, void* dx,
                   float* dwt, float* dbias, float* dy_colsum, int B, int H, int W, int C,
                   int accumulate, void* stream);
/* dxhat = dz @ w1f   ([M,C] bf16). */
int cnx_mlp_bwd_dxhat(const void* dz, const void* w1fT, void* dxhat, int M, int C, void* stream);
/* dw2f[C,4C] += dy^T @ h ; dw1f[4C,C] += dz^T @ xhat   (fp32; always accumulated: zero-fill or keep
 * accumulating across micro-batches). */
int cnx_mlp_bwd_wgrad(const void* dy, const void* h, const void* dz, const void* xhat, float* dw2f,
                      float* dw1f, int M, int C, void* stream);

/* Batch-level MixUp/CutMix (MixUpCutMixDataset, /root/reference/dataset.py:81-190) from a host
 * plan: x [B,CH,H,W] fp32 NCHW -> out (same shape, must not alias x); labels [B] int64 ->
 * targets [B,num_classes] fp32.  Pixels and targets are bit-identical to the reference. */
int cnx_mix_batch(const float* x, const int64_t* labels, const cnx_mix_plan_entry* plan,
                  float* out, float* targets, int B, int CH, int H, int W, int num_classes,
                  void* stream);
/* Same with label-smoothed one-hot rows (timm Mixup semantics; constructed at
 * /root/reference/previous/V2-convbase/dataset.py:119-129): a target row is
 * y(label)*lam + y(partner label)*(1-lam) with y = on_value at the label, off_value elsewhere. */
int cnx_mix_batch_smooth(const float* x, const int64_t* labels, const cnx_mix_plan_entry* plan,
                         float* out, float* targets, int B, int CH, int H, int W, int num_classes,
                         float on_value, float off_value, void* stream);

/* nn.CrossEntropyLoss(label_smoothing) on probability targets
 * (/root/reference/train_advanced_v4.py:626,221): loss (scalar, mean over B) and, if dlogits is
 * non-NULL, d loss / d logits.  logits, targets, dlogits: [B,NC] fp32. */
int cnx_softce_fwd_bwd(const float* logits, const float* targets, float* loss, float* dlogits,
                       int B, int NC, float label_smoothing, void* stream);

/* EMA.update (/root/reference/train_advanced_v4.py:75-80) over all tensors in one launch:
 * shadow[t] = one_minus_decay * param[t] + decay * shadow[t].  All four tables are DEVICE arrays;
 * chunk_prefix[t] = number of 8192-element chunks before tensor t, [num_tensors + 1]. */
int cnx_ema_lerp_multi(const void* const* param_ptrs, void* const* shadow_ptrs,
                       const int64_t* numels, const int64_t* chunk_prefix, int num_tensors,
                       int64_t total_chunks, float one_minus_decay, float decay, void* stream);
int cnx_ema_chunk_elems(void);

/* The optimizer step of the loop (/root/reference/train_advanced_v4.py:226-237) as two launches over
 * device pointer tables (same chunking as cnx_ema_lerp_multi; fp32 tensors only).
 * cnx_optim_grad_norm  = GradScaler.unscale_ (inf/nan check) + the norm of clip_grad_norm_:
 *   reads every gradient once.  scale: device float (GradScaler's scale) or NULL (= 1);
 *   partials: device float[total_chunks] scratch; state: device float[8], zero-initialised ONCE by
 *   the caller and then owned by these two calls:
 *     [0] total gradient norm (unscaled)   [1] gradient multiplier = clip_coef / scale
 *     [2] found_inf (1.0 when any gradient is non-finite: the AdamW update is skipped)
 *     [3] AdamW step count (incremented here unless found_inf)   [4] internal   [5] clip_coef
 * cnx_optim_adamw_ema  = AdamW.step (torch.optim.AdamW arithmetic, one parameter group) with the
 *   multiplier applied to the gradient on the fly, then EMA.update of the new parameter
 *   (shadow_ptrs may be NULL: no EMA).  Must follow cnx_optim_grad_norm on the same stream. */
int cnx_optim_grad_norm(const void* const* grad_ptrs, const int64_t* numels,
                        const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                        const float* scale, float max_norm, float* partials, float* state,
                        void* stream);
/* cnx_optim_grad_norm without a caller-provided `partials` buffer and without any allocation.
 * NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  The per-chunk partial sums go to a file-scope `static __device__ __align__(256)` array of
 * cnx_optim_grad_norm_noalloc_max_chunks() = 65536 floats (256 KiB), which every GPU that loads the
 * library holds until the process exits; its device address is resolved with cudaGetSymbolAddress
 * once per device on the first call (that device current).  The call performs no cudaMalloc /
 * cudaFree / allocating library routine.  total_chunks > 65536 (more than 536 870 912 gradient
 * elements incl. the per-tensor round-up to 8192) -> CNX_ERR_BAD_ARG with a message naming the
 * limit; use cnx_optim_grad_norm then.  All calls on a device SHARE the array and it keeps the
 * last call's contents: every slot a call reads is written by that call first, so nothing needs
 * clearing, but two calls that overlap on the GPU (different streams without an ordering edge)
 * race — keep them on one stream or order them with events.  Everything else (arguments, `state`
 * contract, results bit for bit) is cnx_optim_grad_norm's. */
int cnx_optim_grad_norm_noalloc(const void* const* grad_ptrs, const int64_t* numels,
                                const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                                const float* scale, float max_norm, float* state, void* stream);
int cnx_optim_grad_norm_noalloc_max_chunks(void);
int cnx_optim_adamw_ema(void* const* param_ptrs, const void* const* grad_ptrs, void* const* m_ptrs,
                        void* const* v_ptrs, void* const* shadow_ptrs, const int64_t* numels,
                        const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                        float lr, float beta1, float beta2, float eps, float weight_decay,
                        float ema_one_minus_decay, float ema_decay, const float* state,
                        void* stream);

/* Module-boundary layout conversion (timm blocks are NCHW): [B,C,H,W] fp32 <-> [B,H,W,C] bf16. */
int cnx_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int H, int W, void* stream);
int cnx_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int H, int W, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CNX_B200_H_ */
