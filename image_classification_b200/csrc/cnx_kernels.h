// cnx_kernels.h — internal host-side launchers (C++), one per kernel family.
// The public C ABI in include/cnx_b200.h forwards to these after argument checks.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cnx {

// ---- dwconv7.cu (conv engine) + ln_rows.cu (row LayerNorm)
int dwconv_ln_fwd(const void* x, const float* wt, const float* bias, void* xhat, float* mean,
                  float* rstd, int B, int H, int W, int C, float eps, cudaStream_t stream);

size_t dwconv_ln_bwd_workspace_bytes(int B, int H, int W, int C);
int dwconv_ln_bwd(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                  const void* dy, const float* wt, void* dx, float* dwt, float* dbias,
                  float* dy_colsum, int B, int H, int W, int C, void* workspace, size_t ws_bytes,
                  int accumulate, cudaStream_t stream);
// pure depthwise-conv backward: dconv = dL/d(conv output) already formed (by ln_rows_bwd or by the
// LN-backward epilogue of the fc1 data-gradient GEMM)
int dwconv_bwd(const void* dconv, const void* x, const void* dy, const float* wt, void* dx,
               float* dwt, float* dbias, float* dy_colsum, int B, int H, int W, int C,
               int accumulate, cudaStream_t stream);

// ---- dwconvk.cu : depthwise k x k (3 / 5), stride 1 / 2, NHWC bf16 (EfficientNet path, §8 f-4)
int dwconv_kxk_fwd(const void* x, const float* wt, void* y, int B, int H, int W, int C, int k,
                   int stride, cudaStream_t stream);
int dwconv_kxk_bwd(const void* dy, const void* x, const float* wt, void* dx, float* dwt, int B,
                   int H, int W, int C, int k, int stride, int accumulate, cudaStream_t stream);

// ---- bnse.cu : BatchNorm2d (batch statistics) + SiLU, squeeze-excite scaling (EfficientNet path)
int bn_stats(const void* x, float* sums, int64_t M, int C, cudaStream_t stream);
int bn_act_fwd(const void* x, const float* sums, const float* weight, const float* bias,
               float* running_mean, float* running_var, float momentum, float eps, int act, void* y,
               float* mean_out, float* rstd_out, float* pool_sum, int B, int HW, int C,
               cudaStream_t stream);
int bn_act_bwd(const void* g, const void* x, const float* weight, const float* bias,
               const float* mean, const float* rstd, int act, void* dx, float* dweight,
               float* dbias, int64_t M, int C, cudaStream_t stream);
int se_scale_fwd(const void* x, const float* gate, void* y, int B, int HW, int C,
                 cudaStream_t stream);
int se_scale_bwd_reduce(const void* g, const void* x, float* dgate, int B, int HW, int C,
                        cudaStream_t stream);
int se_scale_bwd_apply(const void* g, const float* gate, const float* dpool, void* dx, int B,
                       int HW, int C, cudaStream_t stream);

int se_gate_fwd(const float* pool_sum, float pool_scale, const float* w1, const float* b1,
                const float* w2t, const float* b2, float* hidden, float* gate, int B, int C, int Cs,
                cudaStream_t stream);
int se_gate_bwd(const float* dgate, const float* gate, const float* hidden, const float* pool_sum,
                float pool_scale, const float* w1, const float* w2t, float* dz2, float* dh,
                float* dpool, float* dw1, float* db1, float* dw2t, float* db2, int B, int C, int Cs,
                cudaStream_t stream);

// ---- ln_rows.cu : LayerNorm with affine over [M, C] bf16 rows (stem / downsample LayerNorm2d)
int ln_affine_fwd(const void* x, const float* w, const float* b, void* y, float* mean, float* rstd,
                  int64_t M, int C, float eps, int pH, int pW, int pk, cudaStream_t stream);
int patchify_nchw_f32(const float* x, void* P, int B, int Cin, int H, int W, int k,
                      cudaStream_t stream);   // layout.cu
int colsum_bf16(const void* x, float* out, int64_t M, int C, cudaStream_t stream);   // ln_rows.cu
int avgpool_nhwc_fwd(const void* x, void* y, int B, int HW, int C, cudaStream_t stream);
int avgpool_nhwc_bwd(const void* g, void* dx, int B, int HW, int C, cudaStream_t stream);
int ln_affine_bwd(const void* g, const void* x, const float* w, const float* mean,
                  const float* rstd, void* dx, float* dw, float* db, int64_t M, int C,
                  int pH, int pW, int pk, cudaStream_t stream);

// ---- mlp_gemm.cu  (tcgen05 / TMEM / TMA)
enum GemmEpilogue : int {
  EPI_STORE_BF16 = 0,       // D = acc                                  (bf16 out)
  EPI_BIAS_GELU = 1,        // z = acc + bias[n]; D0 = z, D1 = gelu(z)  (both bf16)
  EPI_BIAS_RESIDUAL = 2,    // D = acc + bias[n] + R[m,n]               (bf16 out, R bf16)
  EPI_GELU_GRAD = 3,        // D = acc * gelu'(Z[m,n]); colsum[n] += D  (bf16 out, Z bf16)
  EPI_STORE_F32 = 4,        // D = acc                                  (fp32 out; weight grads)
  EPI_BIAS = 6,             // D = acc + bias[n]                        (bf16 out; 5 is internal)
  EPI_LN_BWD = 7,           // D = rstd[m] * (acc - s1[m]/N - R[m,n] * s2[m]/N): LayerNorm backward of
                            // the fc1 data gradient (R = xhat, (s1, s2) = rowstats[m])   (bf16 out)
};
struct GemmArgs {
  // D[M,N] = A[M,K] * B[N,K]^T.  a_mn_major / b_mn_major: operand stored as [K,M] / [K,N].
  const void* A; int64_t lda;   // elements between consecutive rows of the STORED matrix
  const void* B; int64_t ldb;
  int M, N, K;
  int a_mn_major, b_mn_major;
  int epilogue;
  void* D0; int64_t ldd0;
  void* D1; int64_t ldd1;       // second output (EPI_BIAS_GELU)
  const float* bias;            // [N]
  const void* R; int64_t ldr;   // residual (EPI_BIAS_RESIDUAL) or Z (EPI_GELU_GRAD), bf16
  float* colsum;                // [N] fp32, atomically accumulated (EPI_GELU_GRAD), may be null
  // LayerNorm backward by linearity (see gemm2.cu): EPI_GELU_GRAD also accumulates, per row m,
  //   rowstats[m] += ( sum_n D[m,n] * rowvec2[n].x ,  sum_n D[m,n] * (Z[m,n] + rowvec2[n].y) )
  // when rowvec2 / rowstats are given; EPI_LN_BWD consumes rowstats and rstd.
  const float* rowvec2;         // [N][2] fp32
  float* rowstats;              // [M][2] fp32
  const float* rstd;            // [M] fp32 (EPI_LN_BWD)
};
int gemm_bf16(const GemmArgs& a, cudaStream_t stream);      // gemm2.cu (persistent, warp-specialised)
int gemm_debug_max_clusters(int cluster_size);                      // gemm2.cu, occupancy probe
void gemm_debug_tuning(int stages, int pair, int bn, int splits);   // gemm2.cu, -1 = automatic
void gemm_debug_epilogue_warps(int n);
void gemm_debug_staging_chunk(int cols);   // 64 = wide staging chunks everywhere, 0 = automatic                              // 8 / 16, 0 = automatic

// ---- elementwise.cu
struct MixPlanEntry {          // one per sample; integers are bit-exact vs the reference RNG order
  int32_t mode;                // 0 passthrough, 1 mixup, 2 cutmix
  int32_t partner;             // index into the batch
  float lam;                   // float32(lam)
  float one_minus_lam;         // float32(1 - lam) with the subtraction done in float64
  int32_t bbx1, bby1, bbx2, bby2;
};
int mix_batch(const float* x, const int64_t* labels, const MixPlanEntry* plan, float* out,
              float* targets, int B, int CH, int H, int W, int num_classes, float on_value,
              float off_value, cudaStream_t stream);
int softce_fwd_bwd(const float* logits, const float* targets, float* loss, float* dlogits, int B,
                   int NC, float label_smoothing, cudaStream_t stream);
int ema_lerp_multi(const void* const* param_ptrs, void* const* shadow_ptrs,
                   const int64_t* numels, const int64_t* chunk_prefix, int num_tensors,
                   int64_t total_chunks, float one_minus_decay, float decay, cudaStream_t stream);
constexpr int kEmaChunkElems = 8192;

// ---- optim.cu : unscale + clip-norm + AdamW + EMA over pointer tables
int optim_grad_norm(const void* const* grad_ptrs, const int64_t* numels,
                    const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                    const float* scale, float max_norm, float* partials, float* state,
                    cudaStream_t stream);
// Allocation-free variant: the per-chunk partial sums live in a file-scope __device__ array of
// kStaticNormChunks floats (optim.cu) instead of a caller buffer.
constexpr int kStaticNormChunks = 65536;      // 256 KiB; 65536 chunks x 8192 = 536 870 912 elements
int optim_grad_norm_static(const void* const* grad_ptrs, const int64_t* numels,
                           const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                           const float* scale, float max_norm, float* state, cudaStream_t stream);
int optim_adamw_ema(void* const* param_ptrs, const void* const* grad_ptrs, void* const* m_ptrs,
                    void* const* v_ptrs, void* const* shadow_ptrs, const int64_t* numels,
                    const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks, float lr,
                    float beta1, float beta2, float eps, float weight_decay,
                    float ema_one_minus_decay, float ema_decay, const float* state,
                    cudaStream_t stream);

// ---- prep.cu : weight folding before the GEMMs and its adjoint after them
int prep_block_weights(const float* dw_w, const float* ln_w, const float* ln_b, const float* fc1_w,
                       const float* fc1_b, const float* fc2_w, const float* fc2_b,
                       const float* ls_gamma, float* wt_dw, void* w1f, void* w1fT, float* b1f,
                       void* w2f, void* w2fT, float* b2f, float* rb, int C, cudaStream_t stream);
int unfold_block_grads(const float* dw1f, const float* db1f, const float* dw2f, const float* db2f,
                       const float* dwt_dw, const float* ln_w, const float* ln_b,
                       const float* fc1_w, const float* fc2_w, const float* fc2_b,
                       const float* ls_gamma, float* d_dw_w, float* d_ln_w, float* d_ln_b,
                       float* d_fc1_w, float* d_fc2_w, float* d_fc2_b, float* d_ls_gamma, int C,
                       cudaStream_t stream);

// ---- layout.cu : NCHW fp32 <-> NHWC bf16 at the module boundary
int nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int H, int W,
                          cudaStream_t stream);
int nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int H, int W,
                          cudaStream_t stream);

}  // namespace cnx
