// cnx_api.cu — the extern "C" surface declared in include/cnx_b200.h.
// Every entry point: device check (sm_100 or error, never a fallback) -> argument checks ->
// launcher in namespace cnx.  No exceptions cross this boundary.
#include "../../include/cnx_b200.h"
#include "cnx_common.cuh"
#include "cnx_kernels.h"
#include <stdlib.h>

using namespace cnx;

#define CNX_REQUIRE_DEVICE()              \
  do {                                    \
    int rc__ = check_device_sm100();      \
    if (rc__ != CNX_OK) return rc__;      \
    rc__ = check_sticky_timeout();        \
    if (rc__ != CNX_OK) return rc__;      \
  } while (0)

#define CNX_REQUIRE_PTR(p)                                         \
  do {                                                             \
    if ((p) == nullptr) {                                          \
      set_last_error("%s: argument `%s` is NULL", __func__, #p);   \
      return CNX_ERR_BAD_ARG;                                      \
    }                                                              \
  } while (0)

static inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

extern "C" {

int cnx_version(void) { return 100; }

const char* cnx_last_error_string(void) { return last_error(); }

int cnx_check_device(void) { return check_device_sm100(); }

int cnx_debug_timeout_flag(void) { return read_and_clear_timeout_flag(); }
void cnx_debug_set_timeout_flag(int value) { debug_set_timeout_flag(value); }
int cnx_debug_max_clusters(int cluster_size) { return gemm_debug_max_clusters(cluster_size); }
void cnx_debug_gemm_tuning(int stages, int pair, int bn, int splits) {
  gemm_debug_tuning(stages, pair, bn, splits);
}
void cnx_debug_gemm_epilogue_warps(int n) { gemm_debug_epilogue_warps(n); }
void cnx_debug_gemm_staging_chunk(int cols) { gemm_debug_staging_chunk(cols); }

size_t cnx_workspace_bytes(int op, int B, int H, int W, int C) {
  if (op == CNX_OP_DWCONV_LN_BWD) return dwconv_ln_bwd_workspace_bytes(B, H, W, C);
  return 0;
}

int cnx_prep_block_weights(const float* conv_dw_w, const float* norm_w, const float* norm_b,
                           const float* fc1_w, const float* fc1_b, const float* fc2_w,
                           const float* fc2_b, const float* gamma, float* wt_dw, void* w1f,
                           void* w1fT, float* b1f, void* w2f, void* w2fT, float* b2f, float* rb,
                           int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(conv_dw_w); CNX_REQUIRE_PTR(norm_w); CNX_REQUIRE_PTR(norm_b);
  CNX_REQUIRE_PTR(fc1_w); CNX_REQUIRE_PTR(fc1_b); CNX_REQUIRE_PTR(fc2_w); CNX_REQUIRE_PTR(fc2_b);
  CNX_REQUIRE_PTR(gamma); CNX_REQUIRE_PTR(wt_dw); CNX_REQUIRE_PTR(w1f); CNX_REQUIRE_PTR(w1fT);
  CNX_REQUIRE_PTR(b1f); CNX_REQUIRE_PTR(w2f); CNX_REQUIRE_PTR(w2fT); CNX_REQUIRE_PTR(b2f);
  return prep_block_weights(conv_dw_w, norm_w, norm_b, fc1_w, fc1_b, fc2_w, fc2_b, gamma, wt_dw,
                            w1f, w1fT, b1f, w2f, w2fT, b2f, rb, C, as_stream(stream));
}

int cnx_unfold_block_grads(const float* dw1f, const float* db1f, const float* dw2f,
                           const float* db2f, const float* dwt_dw, const float* norm_w,
                           const float* norm_b, const float* fc1_w, const float* fc2_w,
                           const float* fc2_b, const float* gamma, float* d_conv_dw_w,
                           float* d_norm_w, float* d_norm_b, float* d_fc1_w, float* d_fc2_w,
                           float* d_fc2_b, float* d_gamma, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dw1f); CNX_REQUIRE_PTR(db1f); CNX_REQUIRE_PTR(dw2f); CNX_REQUIRE_PTR(db2f);
  CNX_REQUIRE_PTR(dwt_dw); CNX_REQUIRE_PTR(norm_w); CNX_REQUIRE_PTR(norm_b); CNX_REQUIRE_PTR(fc1_w);
  CNX_REQUIRE_PTR(fc2_w); CNX_REQUIRE_PTR(fc2_b); CNX_REQUIRE_PTR(gamma);
  CNX_REQUIRE_PTR(d_conv_dw_w); CNX_REQUIRE_PTR(d_norm_w); CNX_REQUIRE_PTR(d_norm_b);
  CNX_REQUIRE_PTR(d_fc1_w); CNX_REQUIRE_PTR(d_fc2_w); CNX_REQUIRE_PTR(d_fc2_b);
  CNX_REQUIRE_PTR(d_gamma);
  return unfold_block_grads(dw1f, db1f, dw2f, db2f, dwt_dw, norm_w, norm_b, fc1_w, fc2_w, fc2_b,
                            gamma, d_conv_dw_w, d_norm_w, d_norm_b, d_fc1_w, d_fc2_w, d_fc2_b,
                            d_gamma, C, as_stream(stream));
}

int cnx_dwconv_ln_fwd(const void* x, const float* wt_dw, const float* conv_dw_b, void* xhat,
                      float* mean, float* rstd, int B, int H, int W, int C, float eps,
                      void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(wt_dw); CNX_REQUIRE_PTR(conv_dw_b); CNX_REQUIRE_PTR(xhat);
  CNX_REQUIRE_PTR(mean); CNX_REQUIRE_PTR(rstd);
  return dwconv_ln_fwd(x, wt_dw, conv_dw_b, xhat, mean, rstd, B, H, W, C, eps, as_stream(stream));
}

int cnx_dwconv_ln_bwd(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                      const void* dy, const float* wt_dw, void* dx, float* dwt, float* dbias,
                      float* dy_colsum, int B, int H, int W, int C, void* workspace,
                      size_t workspace_bytes, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dxhat); CNX_REQUIRE_PTR(xhat); CNX_REQUIRE_PTR(rstd); CNX_REQUIRE_PTR(x);
  CNX_REQUIRE_PTR(dy); CNX_REQUIRE_PTR(wt_dw); CNX_REQUIRE_PTR(dx); CNX_REQUIRE_PTR(dwt);
  CNX_REQUIRE_PTR(dbias); CNX_REQUIRE_PTR(dy_colsum);
  return dwconv_ln_bwd(dxhat, xhat, rstd, x, dy, wt_dw, dx, dwt, dbias, dy_colsum, B, H, W, C,
                       workspace, workspace_bytes, 0, as_stream(stream));
}

int cnx_dwconv_ln_bwd_acc(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                      const void* dy, const float* wt_dw, void* dx, float* dwt, float* dbias,
                      float* dy_colsum, int B, int H, int W, int C, void* workspace,
                      size_t workspace_bytes, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dxhat); CNX_REQUIRE_PTR(xhat); CNX_REQUIRE_PTR(rstd); CNX_REQUIRE_PTR(x);
  CNX_REQUIRE_PTR(dy); CNX_REQUIRE_PTR(wt_dw); CNX_REQUIRE_PTR(dx); CNX_REQUIRE_PTR(dwt);
  CNX_REQUIRE_PTR(dbias); CNX_REQUIRE_PTR(dy_colsum);
  return dwconv_ln_bwd(dxhat, xhat, rstd, x, dy, wt_dw, dx, dwt, dbias, dy_colsum, B, H, W, C,
                       workspace, workspace_bytes, 1, as_stream(stream));
}

int cnx_layernorm_fwd(const void* x, const float* weight, const float* bias, void* y, float* mean,
                      float* rstd, int64_t M, int C, float eps, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(weight); CNX_REQUIRE_PTR(bias); CNX_REQUIRE_PTR(y);
  CNX_REQUIRE_PTR(mean); CNX_REQUIRE_PTR(rstd);
  return ln_affine_fwd(x, weight, bias, y, mean, rstd, M, C, eps, 0, 0, 0, as_stream(stream));
}

int cnx_layernorm_patch_fwd(const void* x, const float* weight, const float* bias, void* y_patch,
                            float* mean, float* rstd, int B, int H, int W, int C, int patch,
                            float eps, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(weight); CNX_REQUIRE_PTR(bias); CNX_REQUIRE_PTR(y_patch);
  CNX_REQUIRE_PTR(mean); CNX_REQUIRE_PTR(rstd);
  return ln_affine_fwd(x, weight, bias, y_patch, mean, rstd, static_cast<int64_t>(B) * H * W, C,
                       eps, H, W, patch, as_stream(stream));
}

int cnx_layernorm_patch_bwd(const void* dy_patch, const void* x, const float* weight,
                            const float* mean, const float* rstd, void* dx, float* dweight,
                            float* dbias, int B, int H, int W, int C, int patch, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dy_patch); CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(weight); CNX_REQUIRE_PTR(mean);
  CNX_REQUIRE_PTR(rstd); CNX_REQUIRE_PTR(dx); CNX_REQUIRE_PTR(dweight); CNX_REQUIRE_PTR(dbias);
  return ln_affine_bwd(dy_patch, x, weight, mean, rstd, dx, dweight, dbias,
                       static_cast<int64_t>(B) * H * W, C, H, W, patch, as_stream(stream));
}

int cnx_patchify_nchw_f32(const float* x, void* P, int B, int Cin, int H, int W, int patch,
                          void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(P);
  return patchify_nchw_f32(x, P, B, Cin, H, W, patch, as_stream(stream));
}

int cnx_colsum_bf16(const void* x, float* out, int64_t M, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(out);
  return colsum_bf16(x, out, M, C, as_stream(stream));
}

int cnx_dwconv_kxk_fwd(const void* x, const float* wt, void* y, int B, int H, int W, int C, int k,
                       int stride, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(wt); CNX_REQUIRE_PTR(y);
  return dwconv_kxk_fwd(x, wt, y, B, H, W, C, k, stride, as_stream(stream));
}

int cnx_dwconv_kxk_bwd(const void* dy, const void* x, const float* wt, void* dx, float* dwt, int B,
                       int H, int W, int C, int k, int stride, int accumulate, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dy); CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(wt);
  if (dx == nullptr && dwt == nullptr) {
    set_last_error("cnx_dwconv_kxk_bwd: both dx and dwt are NULL, nothing to compute");
    return CNX_ERR_BAD_ARG;
  }
  return dwconv_kxk_bwd(dy, x, wt, dx, dwt, B, H, W, C, k, stride, accumulate, as_stream(stream));
}

int cnx_avgpool_nhwc_fwd(const void* x, void* y, int B, int HW, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(y);
  return avgpool_nhwc_fwd(x, y, B, HW, C, as_stream(stream));
}

int cnx_avgpool_nhwc_bwd(const void* g, void* dx, int B, int HW, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(g); CNX_REQUIRE_PTR(dx);
  return avgpool_nhwc_bwd(g, dx, B, HW, C, as_stream(stream));
}

int cnx_bn_stats(const void* x, float* sums, int64_t M, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(sums);
  return bn_stats(x, sums, M, C, as_stream(stream));
}

int cnx_bn_act_fwd(const void* x, const float* sums, const float* weight, const float* bias,
                   float* running_mean, float* running_var, float momentum, float eps, int act,
                   void* y, float* mean_out, float* rstd_out, float* pool_sum, int B, int HW, int C,
                   void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(weight); CNX_REQUIRE_PTR(bias); CNX_REQUIRE_PTR(y);
  CNX_REQUIRE_PTR(mean_out); CNX_REQUIRE_PTR(rstd_out);
  return bn_act_fwd(x, sums, weight, bias, running_mean, running_var, momentum, eps, act, y,
                    mean_out, rstd_out, pool_sum, B, HW, C, as_stream(stream));
}

int cnx_bn_act_bwd(const void* g, const void* x, const float* weight, const float* bias,
                   const float* mean, const float* rstd, int act, void* dx, float* dweight,
                   float* dbias, int64_t M, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(g); CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(weight); CNX_REQUIRE_PTR(bias);
  CNX_REQUIRE_PTR(mean); CNX_REQUIRE_PTR(rstd); CNX_REQUIRE_PTR(dx); CNX_REQUIRE_PTR(dweight);
  CNX_REQUIRE_PTR(dbias);
  return bn_act_bwd(g, x, weight, bias, mean, rstd, act, dx, dweight, dbias, M, C,
                    as_stream(stream));
}

int cnx_se_scale_fwd(const void* x, const float* gate, void* y, int B, int HW, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(gate); CNX_REQUIRE_PTR(y);
  return se_scale_fwd(x, gate, y, B, HW, C, as_stream(stream));
}

int cnx_se_scale_bwd_reduce(const void* g, const void* x, float* dgate, int B, int HW, int C,
                            void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(g); CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(dgate);
  return se_scale_bwd_reduce(g, x, dgate, B, HW, C, as_stream(stream));
}

int cnx_se_scale_bwd_apply(const void* g, const float* gate, const float* dpool, void* dx, int B,
                           int HW, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(g); CNX_REQUIRE_PTR(gate); CNX_REQUIRE_PTR(dpool); CNX_REQUIRE_PTR(dx);
  return se_scale_bwd_apply(g, gate, dpool, dx, B, HW, C, as_stream(stream));
}

int cnx_se_gate_fwd(const float* pool_sum, float pool_scale, const float* w1, const float* b1,
                    const float* w2t, const float* b2, float* hidden, float* gate, int B, int C,
                    int Cs, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(pool_sum); CNX_REQUIRE_PTR(w1); CNX_REQUIRE_PTR(b1); CNX_REQUIRE_PTR(w2t);
  CNX_REQUIRE_PTR(b2); CNX_REQUIRE_PTR(hidden); CNX_REQUIRE_PTR(gate);
  return se_gate_fwd(pool_sum, pool_scale, w1, b1, w2t, b2, hidden, gate, B, C, Cs,
                     as_stream(stream));
}

int cnx_se_gate_bwd(const float* dgate, const float* gate, const float* hidden,
                    const float* pool_sum, float pool_scale, const float* w1, const float* w2t,
                    float* dz2, float* dh, float* dpool, float* dw1, float* db1, float* dw2t,
                    float* db2, int B, int C, int Cs, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dgate); CNX_REQUIRE_PTR(gate); CNX_REQUIRE_PTR(hidden); CNX_REQUIRE_PTR(pool_sum);
  CNX_REQUIRE_PTR(w1); CNX_REQUIRE_PTR(w2t); CNX_REQUIRE_PTR(dz2); CNX_REQUIRE_PTR(dh);
  CNX_REQUIRE_PTR(dpool); CNX_REQUIRE_PTR(dw1); CNX_REQUIRE_PTR(db1); CNX_REQUIRE_PTR(dw2t);
  CNX_REQUIRE_PTR(db2);
  return se_gate_bwd(dgate, gate, hidden, pool_sum, pool_scale, w1, w2t, dz2, dh, dpool, dw1, db1,
                     dw2t, db2, B, C, Cs, as_stream(stream));
}

int cnx_layernorm_bwd(const void* dy, const void* x, const float* weight, const float* mean,
                      const float* rstd, void* dx, float* dweight, float* dbias, int64_t M, int C,
                      void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dy); CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(weight); CNX_REQUIRE_PTR(mean);
  CNX_REQUIRE_PTR(rstd); CNX_REQUIRE_PTR(dx); CNX_REQUIRE_PTR(dweight); CNX_REQUIRE_PTR(dbias);
  return ln_affine_bwd(dy, x, weight, mean, rstd, dx, dweight, dbias, M, C, 0, 0, 0,
                       as_stream(stream));
}

int cnx_gemm_bf16(const void* A, int64_t lda, int a_mn_major, const void* B, int64_t ldb,
                  int b_mn_major, int M, int N, int K, int epilogue, void* D0, int64_t ldd0,
                  void* D1, int64_t ldd1, const float* bias, const void* R, int64_t ldr,
                  float* colsum, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(A); CNX_REQUIRE_PTR(B); CNX_REQUIRE_PTR(D0);
  GemmArgs a;
  a.A = A; a.lda = lda; a.B = B; a.ldb = ldb;
  a.M = M; a.N = N; a.K = K;
  a.a_mn_major = a_mn_major; a.b_mn_major = b_mn_major;
  a.epilogue = epilogue;
  a.D0 = D0; a.ldd0 = ldd0; a.D1 = D1; a.ldd1 = ldd1;
  a.bias = bias; a.R = R; a.ldr = ldr; a.colsum = colsum;
  a.rowvec2 = nullptr; a.rowstats = nullptr; a.rstd = nullptr;
  return gemm_bf16(a, as_stream(stream));
}

int cnx_fc1_gelu_fwd(const void* xhat, const void* w1f, const float* b1f, void* z, void* h, int M,
                     int C, void* stream) {
  return cnx_gemm_bf16(xhat, C, 0, w1f, C, 0, M, 4 * C, C, CNX_EPI_BIAS_GELU, z, 4 * C, h, 4 * C,
                       b1f, nullptr, 0, nullptr, stream);
}

int cnx_fc2_residual_fwd(const void* h, const void* w2f, const float* b2f, const void* x, void* y,
                         int M, int C, void* stream) {
  return cnx_gemm_bf16(h, 4 * C, 0, w2f, 4 * C, 0, M, C, 4 * C, CNX_EPI_BIAS_RESIDUAL, y, C,
                       nullptr, 0, b2f, x, C, nullptr, stream);
}

int cnx_mlp_bwd_dz(const void* dy, const void* w2fT, const void* z, void* dz, float* db1f, int M,
                   int C, void* stream) {
  return cnx_gemm_bf16(dy, C, 0, w2fT, C, 0, M, 4 * C, C, CNX_EPI_GELU_GRAD, dz, 4 * C, nullptr, 0,
                       nullptr, z, 4 * C, db1f, stream);
}

int cnx_mlp_bwd_dz_stats(const void* dy, const void* w2fT, const void* z, const float* rb, void* dz,
                         float* db1f, float* rowstats, int M, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dy); CNX_REQUIRE_PTR(w2fT); CNX_REQUIRE_PTR(z); CNX_REQUIRE_PTR(rb);
  CNX_REQUIRE_PTR(dz); CNX_REQUIRE_PTR(rowstats);
  GemmArgs a{};
  a.A = dy; a.lda = C; a.B = w2fT; a.ldb = C;
  a.M = M; a.N = 4 * C; a.K = C;
  a.epilogue = EPI_GELU_GRAD;
  a.D0 = dz; a.ldd0 = 4 * C;
  a.R = z; a.ldr = 4 * C; a.colsum = db1f;
  a.rowvec2 = rb; a.rowstats = rowstats;
  return gemm_bf16(a, as_stream(stream));
}

int cnx_mlp_bwd_dconv(const void* dz, const void* w1fT, const void* xhat, const float* rowstats,
                      const float* rstd, void* dconv, int M, int C, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dz); CNX_REQUIRE_PTR(w1fT); CNX_REQUIRE_PTR(xhat); CNX_REQUIRE_PTR(rowstats);
  CNX_REQUIRE_PTR(rstd); CNX_REQUIRE_PTR(dconv);
  GemmArgs a{};
  a.A = dz; a.lda = 4 * C; a.B = w1fT; a.ldb = 4 * C;
  a.M = M; a.N = C; a.K = 4 * C;
  a.epilogue = EPI_LN_BWD;
  a.D0 = dconv; a.ldd0 = C;
  a.R = xhat; a.ldr = C;
  a.rowstats = const_cast<float*>(rowstats); a.rstd = rstd;
  return gemm_bf16(a, as_stream(stream));
}

int cnx_dwconv_bwd(const void* dconv, const void* x, const void* dy, const float* wt_dw, void* dx,
                   float* dwt, float* dbias, float* dy_colsum, int B, int H, int W, int C,
                   int accumulate, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(dconv); CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(dy); CNX_REQUIRE_PTR(wt_dw);
  CNX_REQUIRE_PTR(dx); CNX_REQUIRE_PTR(dwt); CNX_REQUIRE_PTR(dbias); CNX_REQUIRE_PTR(dy_colsum);
  return dwconv_bwd(dconv, x, dy, wt_dw, dx, dwt, dbias, dy_colsum, B, H, W, C, accumulate,
                    as_stream(stream));
}

int cnx_mlp_bwd_dxhat(const void* dz, const void* w1fT, void* dxhat, int M, int C, void* stream) {
  return cnx_gemm_bf16(dz, 4 * C, 0, w1fT, 4 * C, 0, M, C, 4 * C, CNX_EPI_STORE_BF16, dxhat, C,
                       nullptr, 0, nullptr, nullptr, 0, nullptr, stream);
}

int cnx_mlp_bwd_wgrad(const void* dy, const void* h, const void* dz, const void* xhat, float* dw2f,
                      float* dw1f, int M, int C, void* stream) {
  // dw2f[C,4C] = dy^T[C,M] * h[M,4C]   : A = dy stored [K=M, C] (MN-major), B = h stored [K=M, 4C]
  int rc = cnx_gemm_bf16(dy, C, 1, h, 4 * C, 1, C, 4 * C, M, CNX_EPI_STORE_F32, dw2f, 4 * C,
                         nullptr, 0, nullptr, nullptr, 0, nullptr, stream);
  if (rc != CNX_OK) return rc;
  // dw1f[4C,C] = dz^T[4C,M] * xhat[M,C]
  return cnx_gemm_bf16(dz, 4 * C, 1, xhat, C, 1, 4 * C, C, M, CNX_EPI_STORE_F32, dw1f, C, nullptr,
                       0, nullptr, nullptr, 0, nullptr, stream);
}

int cnx_mix_batch(const float* x, const int64_t* labels, const cnx_mix_plan_entry* plan,
                  float* out, float* targets, int B, int CH, int H, int W, int num_classes,
                  void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(labels); CNX_REQUIRE_PTR(plan); CNX_REQUIRE_PTR(out);
  CNX_REQUIRE_PTR(targets);
  static_assert(sizeof(cnx_mix_plan_entry) == sizeof(MixPlanEntry), "plan entry layout");
  return mix_batch(x, labels, reinterpret_cast<const MixPlanEntry*>(plan), out, targets, B, CH, H,
                   W, num_classes, 1.0f, 0.0f, as_stream(stream));
}

int cnx_mix_batch_smooth(const float* x, const int64_t* labels, const cnx_mix_plan_entry* plan,
                         float* out, float* targets, int B, int CH, int H, int W, int num_classes,
                         float on_value, float off_value, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(x); CNX_REQUIRE_PTR(labels); CNX_REQUIRE_PTR(plan); CNX_REQUIRE_PTR(out);
  CNX_REQUIRE_PTR(targets);
  return mix_batch(x, labels, reinterpret_cast<const MixPlanEntry*>(plan), out, targets, B, CH, H,
                   W, num_classes, on_value, off_value, as_stream(stream));
}

int cnx_softce_fwd_bwd(const float* logits, const float* targets, float* loss, float* dlogits,
                       int B, int NC, float label_smoothing, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(logits); CNX_REQUIRE_PTR(targets); CNX_REQUIRE_PTR(loss);
  return softce_fwd_bwd(logits, targets, loss, dlogits, B, NC, label_smoothing, as_stream(stream));
}

int cnx_ema_lerp_multi(const void* const* param_ptrs, void* const* shadow_ptrs,
                       const int64_t* numels, const int64_t* chunk_prefix, int num_tensors,
                       int64_t total_chunks, float one_minus_decay, float decay, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(param_ptrs); CNX_REQUIRE_PTR(shadow_ptrs); CNX_REQUIRE_PTR(numels);
  CNX_REQUIRE_PTR(chunk_prefix);
  return ema_lerp_multi(param_ptrs, shadow_ptrs, numels, chunk_prefix, num_tensors, total_chunks,
                        one_minus_decay, decay, as_stream(stream));
}

int cnx_ema_chunk_elems(void) { return kEmaChunkElems; }

int cnx_optim_grad_norm(const void* const* grad_ptrs, const int64_t* numels,
                        const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                        const float* scale, float max_norm, float* partials, float* state,
                        void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(grad_ptrs); CNX_REQUIRE_PTR(numels); CNX_REQUIRE_PTR(chunk_prefix);
  CNX_REQUIRE_PTR(partials); CNX_REQUIRE_PTR(state);
  return optim_grad_norm(grad_ptrs, numels, chunk_prefix, num_tensors, total_chunks, scale,
                         max_norm, partials, state, as_stream(stream));
}

int cnx_optim_grad_norm_noalloc(const void* const* grad_ptrs, const int64_t* numels,
                                const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                                const float* scale, float max_norm, float* state, void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(grad_ptrs); CNX_REQUIRE_PTR(numels); CNX_REQUIRE_PTR(chunk_prefix);
  CNX_REQUIRE_PTR(state);
  return optim_grad_norm_static(grad_ptrs, numels, chunk_prefix, num_tensors, total_chunks, scale,
                                max_norm, state, as_stream(stream));
}

int cnx_optim_grad_norm_noalloc_max_chunks(void) { return kStaticNormChunks; }

int cnx_optim_adamw_ema(void* const* param_ptrs, const void* const* grad_ptrs, void* const* m_ptrs,
                        void* const* v_ptrs, void* const* shadow_ptrs, const int64_t* numels,
                        const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                        float lr, float beta1, float beta2, float eps, float weight_decay,
                        float ema_one_minus_decay, float ema_decay, const float* state,
                        void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(param_ptrs); CNX_REQUIRE_PTR(grad_ptrs); CNX_REQUIRE_PTR(m_ptrs);
  CNX_REQUIRE_PTR(v_ptrs); CNX_REQUIRE_PTR(numels); CNX_REQUIRE_PTR(chunk_prefix);
  CNX_REQUIRE_PTR(state);
  return optim_adamw_ema(param_ptrs, grad_ptrs, m_ptrs, v_ptrs, shadow_ptrs, numels, chunk_prefix,
                         num_tensors, total_chunks, lr, beta1, beta2, eps, weight_decay,
                         ema_one_minus_decay, ema_decay, state, as_stream(stream));
}

int cnx_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int H, int W,
                              void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(src); CNX_REQUIRE_PTR(dst);
  return nchw_f32_to_nhwc_bf16(src, dst, B, C, H, W, as_stream(stream));
}

int cnx_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int H, int W,
                              void* stream) {
  CNX_REQUIRE_DEVICE();
  CNX_REQUIRE_PTR(src); CNX_REQUIRE_PTR(dst);
  return nhwc_bf16_to_nchw_f32(src, dst, B, C, H, W, as_stream(stream));
}

}  // extern "C"
