// ln_rows.cu — channels-last LayerNorm over [M, C] bf16 rows, forward and backward, without the
// affine (norm.weight / norm.bias are folded into fc1 by prep.cu).
//
// Replaces  permute -> norm  of timm's ConvNeXtBlock.forward and its autograd backward
// (/root/reference/train_advanced_v4.py:134-135, :223).  The rows are the conv
// output written by dwconv7.cu; for every ConvNeXt stage but the first the tensor is L2-resident
// when these kernels run, so they cost an L2 pass, not an HBM pass.
//
//   forward : xhat = (u - mean) * rstd  in place, mean / rstd [M] fp32 (two-pass statistics in
//             registers: mean first, then the centred second moment — no E[x^2]-E[x]^2)
//   backward: dconv = rstd * (dxhat - mean_c(dxhat) - xhat * mean_c(dxhat * xhat))
//
// Mapping: LPR lanes share a row (16-byte vectors, NV per lane); reductions are xor-shuffles
// inside the lane group.  Pure streaming: HBM/L2-bound.
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int kRowThreads = 256;

__device__ __forceinline__ uint4 ld_v4(const void* p) {
  return *reinterpret_cast<const uint4*>(p);
}
__device__ __forceinline__ void st_v4(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }

template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = bf16x2_to_f32x2(w[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  v.x = f32x2_to_bf16x2(f[0], f[1]);
  v.y = f32x2_to_bf16x2(f[2], f[3]);
  v.z = f32x2_to_bf16x2(f[4], f[5]);
  v.w = f32x2_to_bf16x2(f[6], f[7]);
  return v;
}

template <int LPR, int NV>
__global__ void __launch_bounds__(kRowThreads)
ln_rows_fwd_kernel(__nv_bfloat16* __restrict__ u, float* __restrict__ mean_out,
                   float* __restrict__ rstd_out, int64_t M, int C, float eps) {
  const int lane_in_row = threadIdx.x % LPR;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (kRowThreads / LPR) + threadIdx.x / LPR;
  const bool row_ok = row < M;
  const int nvec = C >> 3;
  __nv_bfloat16* rp = u + (row_ok ? row : 0) * C;
  float f[NV][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (row_ok && v < nvec) {
      unpack8(ld_v4(rp + v * 8), f[i]);
#pragma unroll
      for (int k = 0; k < 8; ++k) s += f[i][k];
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) f[i][k] = 0.f;
    }
  }
  const float inv_c = 1.0f / static_cast<float>(C);
  const float mean = group_sum<LPR>(s) * inv_c;
  float m2 = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (v < nvec) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float d = f[i][k] - mean;
        m2 = fmaf(d, d, m2);
      }
    }
  }
  const float rstd = rsqrtf(group_sum<LPR>(m2) * inv_c + eps);
  if (!row_ok) return;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (v < nvec) {
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = (f[i][k] - mean) * rstd;
      st_v4(rp + v * 8, pack8(o));
    }
  }
  if (lane_in_row == 0) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
}

template <int LPR, int NV>
__global__ void __launch_bounds__(kRowThreads)
ln_rows_bwd_kernel(const __nv_bfloat16* __restrict__ dxhat, const __nv_bfloat16* __restrict__ xhat,
                   const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dconv, int64_t M,
                   int C) {
  const int lane_in_row = threadIdx.x % LPR;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (kRowThreads / LPR) + threadIdx.x / LPR;
  const bool row_ok = row < M;
  const int nvec = C >> 3;
  const int64_t off = (row_ok ? row : 0) * C;
  float d[NV][8], xh[NV][8];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (row_ok && v < nvec) {
      unpack8(ld_v4(dxhat + off + v * 8), d[i]);
      unpack8(ld_v4(xhat + off + v * 8), xh[i]);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s1 += d[i][k];
        s2 = fmaf(d[i][k], xh[i][k], s2);
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) d[i][k] = xh[i][k] = 0.f;
    }
  }
  const float inv_c = 1.0f / static_cast<float>(C);
  const float m1 = group_sum<LPR>(s1) * inv_c;
  const float m2 = group_sum<LPR>(s2) * inv_c;
  if (!row_ok) return;
  const float rs = rstd_in[row];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (v < nvec) {
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = rs * (d[i][k] - m1 - xh[i][k] * m2);
      st_v4(dconv + off + v * 8, pack8(o));
    }
  }
}

struct RowCfg { int lpr, nv; };
bool pick_cfg(int C, RowCfg* cfg) {
  const int nvec = C / 8;
  if (nvec <= 8) *cfg = {8, 1};
  else if (nvec <= 16) *cfg = {16, 1};
  else if (nvec <= 32) *cfg = {32, 1};
  else if (nvec <= 64) *cfg = {32, 2};
  else if (nvec <= 128) *cfg = {32, 4};
  else return false;
  return true;
}

int check_rows(const char* who, int64_t M, int C, RowCfg* cfg) {
  if (M <= 0 || C <= 0 || (C % 8) != 0 || !pick_cfg(C, cfg)) {
    set_last_error("%s: bad shape M=%lld C=%d (C must be a multiple of 8, <= 1024)", who,
                   (long long)M, C);
    return CNX_ERR_BAD_SHAPE;
  }
  return CNX_OK;
}

int launch_status(const char* who) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("%s launch: %s", who, cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace

#define CNX_ROW_DISPATCH(KERNEL, ...)                                                          \
  do {                                                                                         \
    const unsigned grid = static_cast<unsigned>((M + (kRowThreads / cfg.lpr) - 1) /            \
                                                (kRowThreads / cfg.lpr));                      \
    if (cfg.lpr == 8) KERNEL<8, 1><<<grid, kRowThreads, 0, stream>>>(__VA_ARGS__);             \
    else if (cfg.lpr == 16) KERNEL<16, 1><<<grid, kRowThreads, 0, stream>>>(__VA_ARGS__);      \
    else if (cfg.nv == 1) KERNEL<32, 1><<<grid, kRowThreads, 0, stream>>>(__VA_ARGS__);        \
    else if (cfg.nv == 2) KERNEL<32, 2><<<grid, kRowThreads, 0, stream>>>(__VA_ARGS__);        \
    else KERNEL<32, 4><<<grid, kRowThreads, 0, stream>>>(__VA_ARGS__);                         \
  } while (0)

int ln_rows_fwd(void* u_inout, float* mean, float* rstd, int64_t M, int C, float eps,
                cudaStream_t stream) {
  RowCfg cfg;
  int rc = check_rows("ln_rows_fwd", M, C, &cfg);
  if (rc != CNX_OK) return rc;
  CNX_ROW_DISPATCH(ln_rows_fwd_kernel, static_cast<__nv_bfloat16*>(u_inout), mean, rstd, M, C, eps);
  return launch_status("ln_rows_fwd");
}

int ln_rows_bwd(const void* dxhat, const void* xhat, const float* rstd, void* dconv, int64_t M,
                int C, cudaStream_t stream) {
  RowCfg cfg;
  int rc = check_rows("ln_rows_bwd", M, C, &cfg);
  if (rc != CNX_OK) return rc;
  CNX_ROW_DISPATCH(ln_rows_bwd_kernel, static_cast<const __nv_bfloat16*>(dxhat),
                   static_cast<const __nv_bfloat16*>(xhat), rstd,
                   static_cast<__nv_bfloat16*>(dconv), M, C);
  return launch_status("ln_rows_bwd");
}

}  // namespace cnx
