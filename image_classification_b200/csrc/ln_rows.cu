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
  pdl_wait();                 // launched with launch_pdl: scheduled while the conv kernel drains
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
  pdl_wait();
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


// ---- LayerNorm WITH affine over [M, C] bf16 rows (timm LayerNorm2d of the stem and of the stage
// `downsample` modules): y = (x - mean) * rstd * w + b; mean / rstd saved.
// Row of the "patch-major" matrix a k x k stride-k convolution consumes as a plain GEMM operand:
// pixel (b, h, w) of an [B, H, W, C] tensor -> row ((b*H/k + h/k)*W/k + w/k)*k*k + (h%k)*k + w%k of
// an [B*H*W, C] matrix, i.e. [B*H/k*W/k, k*k*C] holds one im2col row per output pixel.  pk <= 1: identity.
__device__ __forceinline__ int64_t patch_row(int64_t m, int pH, int pW, int pk) {
  if (pk <= 1) return m;
  // 32-bit arithmetic (the host checks M < 2^31): 64-bit divisions would dominate the row loop
  const unsigned um = static_cast<unsigned>(m), uW = static_cast<unsigned>(pW),
                 uH = static_cast<unsigned>(pH), uk = static_cast<unsigned>(pk);
  const unsigned t = um / uW, w = um - t * uW;
  const unsigned b = t / uH, h = t - b * uH;
  const unsigned oh = h / uk, kh = h - oh * uk, ow = w / uk, kw = w - ow * uk;
  return static_cast<int64_t>(((b * (uH / uk) + oh) * (uW / uk) + ow) * (uk * uk) + kh * uk + kw);
}

template <int LPR, int NV>
__global__ void __launch_bounds__(kRowThreads)
ln_affine_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                     const float* __restrict__ b, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int64_t M, int C,
                     float eps, int pH, int pW, int pk) {
  const int lane_in_row = threadIdx.x % LPR;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (kRowThreads / LPR) + threadIdx.x / LPR;
  const bool row_ok = row < M;
  const int nvec = C >> 3;
  const int64_t off = (row_ok ? row : 0) * C;
  const int64_t yoff = patch_row(row_ok ? row : 0, pH, pW, pk) * C;     // output row (patch-major)
  float f[NV][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (row_ok && v < nvec) {
      unpack8(ld_v4(x + off + v * 8), f[i]);
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
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + v * 8));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + v * 8 + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(b + v * 8));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(b + v * 8 + 4));
      const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) o[k] = fmaf((f[i][k] - mean) * rstd, wv[k], bv[k]);
      st_v4(y + yoff + v * 8, pack8(o));
    }
  }
  if (lane_in_row == 0) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
}

// backward: dx = rstd * (g*w - mean_c(g*w) - xhat * mean_c(g*w*xhat)), dw += sum_rows g*xhat,
// db += sum_rows g.  A CTA walks `rows_per_cta` rows keeping the per-channel sums in registers,
// then reduces them through shared-memory atomics and issues one global atomic per channel.
template <int LPR, int NV>
__global__ void __launch_bounds__(kRowThreads)
ln_affine_bwd_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                     const float* __restrict__ w, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dx,
                     float* __restrict__ dw, float* __restrict__ db, int64_t M, int C,
                     int rows_per_cta, int pH, int pW, int pk) {
  __shared__ float s_dw[1024];
  __shared__ float s_db[1024];
  for (int i = threadIdx.x; i < C; i += kRowThreads) {
    s_dw[i] = 0.f;
    s_db[i] = 0.f;
  }
  __syncthreads();
  const int lane_in_row = threadIdx.x % LPR;
  const int slot = threadIdx.x / LPR;
  constexpr int kSlots = kRowThreads / LPR;
  const int nvec = C >> 3;
  const float inv_c = 1.0f / static_cast<float>(C);
  float wv[NV][8], dwa[NV][8], dba[NV][8];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      wv[i][k] = (v < nvec) ? __ldg(w + v * 8 + k) : 0.f;
      dwa[i][k] = 0.f;
      dba[i][k] = 0.f;
    }
  }
  const int64_t row_begin = static_cast<int64_t>(blockIdx.x) * rows_per_cta;
  for (int r = slot; r < rows_per_cta; r += kSlots) {
    const int64_t row = row_begin + r;
    const bool row_ok = row < M;            // uniform across the lanes sharing the row
    const int64_t off = (row_ok ? row : 0) * C;
    const int64_t goff = patch_row(row_ok ? row : 0, pH, pW, pk) * C;   // gradient row (patch-major)
    const float mu = row_ok ? mean_in[row] : 0.f;
    const float rs = row_ok ? rstd_in[row] : 0.f;
    float gw[NV][8], xh[NV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int v = lane_in_row + i * LPR;
      if (row_ok && v < nvec) {
        float gg[8], xx[8];
        unpack8(ld_v4(g + goff + v * 8), gg);
        unpack8(ld_v4(x + off + v * 8), xx);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          xh[i][k] = (xx[k] - mu) * rs;
          dwa[i][k] = fmaf(gg[k], xh[i][k], dwa[i][k]);
          dba[i][k] += gg[k];
          gw[i][k] = gg[k] * wv[i][k];
          s1 += gw[i][k];
          s2 = fmaf(gw[i][k], xh[i][k], s2);
        }
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) gw[i][k] = xh[i][k] = 0.f;
      }
    }
    const float m1 = group_sum<LPR>(s1) * inv_c;
    const float m2 = group_sum<LPR>(s2) * inv_c;
    if (row_ok) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int v = lane_in_row + i * LPR;
        if (v < nvec) {
          float o[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] = rs * (gw[i][k] - m1 - xh[i][k] * m2);
          st_v4(dx + off + v * 8, pack8(o));
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int v = lane_in_row + i * LPR;
    if (v < nvec) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        atomicAdd(&s_dw[v * 8 + k], dwa[i][k]);
        atomicAdd(&s_db[v * 8 + k], dba[i][k]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += kRowThreads) {
    atomicAdd(dw + i, s_dw[i]);
    atomicAdd(db + i, s_db[i]);
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

// Column sums of an [M, C] bf16 matrix (the bias gradient of the stem / downsample convolutions:
// sum of grad_output over batch and space).  A thread owns one 8-channel vector of the row and
// walks rows; lanes of different rows meet in shared memory, one atomic per channel per CTA.
constexpr int kColsumThreads = 256;
__global__ void __launch_bounds__(kColsumThreads)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int64_t M, int C,
                   int rows_per_cta) {
  extern __shared__ float s_cs[];                       // [C]
  const int nvec = C >> 3;
  const int rows_in_flight = kColsumThreads / nvec;     // host guarantees nvec <= 256
  const int v = threadIdx.x % nvec, rl = threadIdx.x / nvec;
  for (int i = threadIdx.x; i < C; i += kColsumThreads) s_cs[i] = 0.f;
  __syncthreads();
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const int64_t r0 = static_cast<int64_t>(blockIdx.x) * rows_per_cta;
  const int64_t r1 = min(r0 + rows_per_cta, M);
  if (rl < rows_in_flight) {
    for (int64_t r = r0 + rl; r < r1; r += rows_in_flight) {
      float f[8];
      unpack8(ld_v4(x + r * C + v * 8), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) atomicAdd(&s_cs[v * 8 + k], acc[k]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += kColsumThreads) atomicAdd(out + i, s_cs[i]);
}

// Global average pool of an NHWC bf16 map, [B, HW, C] -> [B, C] (head.global_pool and the three
// auxiliary classifiers of the reference's DeepSupervisionModel).  A CTA owns one image and a
// 64-channel slab: 8 vector lanes x 32 row lanes, row lanes meet in shared memory in a fixed order
// (deterministic).  Backward is the broadcast dx[b, p, c] = g[b, c] / HW.
constexpr int kPoolThreads = 256;
__global__ void __launch_bounds__(kPoolThreads)
avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int HW, int C,
                   float inv_hw) {
  __shared__ float s_part[32][65];
  const int v = threadIdx.x & 7, rl = threadIdx.x >> 3;
  const int c0 = blockIdx.x * 64 + v * 8;
  const __nv_bfloat16* xb = x + static_cast<int64_t>(blockIdx.y) * HW * C;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < C) {
    for (int r = rl; r < HW; r += 32) {
      float f[8];
      unpack8(ld_v4(xb + static_cast<int64_t>(r) * C + c0), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) s_part[rl][v * 8 + k] = acc[k];
  __syncthreads();
  if (threadIdx.x < 64) {
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c < C) {
      float t = 0.f;
#pragma unroll
      for (int r = 0; r < 32; ++r) t += s_part[r][threadIdx.x];
      y[static_cast<int64_t>(blockIdx.y) * C + c] = __float2bfloat16_rn(t * inv_hw);
    }
  }
}

__global__ void __launch_bounds__(kPoolThreads)
avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ dx, int HW,
                   int nvec, int64_t total_vec, float inv_hw) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * kPoolThreads;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * kPoolThreads + threadIdx.x; i < total_vec;
       i += stride) {
    const int v = static_cast<int>(i % nvec);
    const int64_t b = i / nvec / HW;
    float f[8];
    unpack8(ld_v4(g + (b * nvec + v) * 8), f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] *= inv_hw;
    st_v4(dx + i * 8, pack8(f));
  }
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

// same, through launch_pdl (kernels that begin with pdl_wait())
#define CNX_ROW_DISPATCH_PDL(KERNEL, ...)                                                      \
  do {                                                                                         \
    const unsigned grid = static_cast<unsigned>((M + (kRowThreads / cfg.lpr) - 1) /            \
                                                (kRowThreads / cfg.lpr));                      \
    if (cfg.lpr == 8) launch_pdl(KERNEL<8, 1>, grid, kRowThreads, 0, stream, 1u, __VA_ARGS__);  \
    else if (cfg.lpr == 16) launch_pdl(KERNEL<16, 1>, grid, kRowThreads, 0, stream, 1u, __VA_ARGS__); \
    else if (cfg.nv == 1) launch_pdl(KERNEL<32, 1>, grid, kRowThreads, 0, stream, 1u, __VA_ARGS__); \
    else if (cfg.nv == 2) launch_pdl(KERNEL<32, 2>, grid, kRowThreads, 0, stream, 1u, __VA_ARGS__); \
    else launch_pdl(KERNEL<32, 4>, grid, kRowThreads, 0, stream, 1u, __VA_ARGS__);              \
  } while (0)

int ln_rows_fwd(void* u_inout, float* mean, float* rstd, int64_t M, int C, float eps,
                cudaStream_t stream) {
  RowCfg cfg;
  int rc = check_rows("ln_rows_fwd", M, C, &cfg);
  if (rc != CNX_OK) return rc;
  CNX_ROW_DISPATCH_PDL(ln_rows_fwd_kernel, static_cast<__nv_bfloat16*>(u_inout), mean, rstd, M, C, eps);
  return launch_status("ln_rows_fwd");
}

int ln_rows_bwd(const void* dxhat, const void* xhat, const float* rstd, void* dconv, int64_t M,
                int C, cudaStream_t stream) {
  RowCfg cfg;
  int rc = check_rows("ln_rows_bwd", M, C, &cfg);
  if (rc != CNX_OK) return rc;
  CNX_ROW_DISPATCH_PDL(ln_rows_bwd_kernel, static_cast<const __nv_bfloat16*>(dxhat),
                   static_cast<const __nv_bfloat16*>(xhat), rstd,
                   static_cast<__nv_bfloat16*>(dconv), M, C);
  return launch_status("ln_rows_bwd");
}

int check_patch(const char* who, int64_t M, int pH, int pW, int pk) {
  if (pk <= 1) return CNX_OK;
  if (pH <= 0 || pW <= 0 || (pH % pk) != 0 || (pW % pk) != 0 || M >= (1LL << 31) ||
      (M % (static_cast<int64_t>(pH) * pW)) != 0) {
    set_last_error("%s: patch size %d must divide H=%d and W=%d, and M=%lld must be B*H*W", who, pk,
                   pH, pW, (long long)M);
    return CNX_ERR_BAD_SHAPE;
  }
  return CNX_OK;
}

int ln_affine_fwd(const void* x, const float* w, const float* b, void* y, float* mean, float* rstd,
                  int64_t M, int C, float eps, int pH, int pW, int pk, cudaStream_t stream) {
  RowCfg cfg;
  int rc = check_rows("ln_affine_fwd", M, C, &cfg);
  if (rc != CNX_OK) return rc;
  rc = check_patch("ln_affine_fwd", M, pH, pW, pk);
  if (rc != CNX_OK) return rc;
  CNX_ROW_DISPATCH(ln_affine_fwd_kernel, static_cast<const __nv_bfloat16*>(x), w, b,
                   static_cast<__nv_bfloat16*>(y), mean, rstd, M, C, eps, pH, pW, pk);
  return launch_status("ln_affine_fwd");
}

int ln_affine_bwd(const void* g, const void* x, const float* w, const float* mean,
                  const float* rstd, void* dx, float* dw, float* db, int64_t M, int C,
                  int pH, int pW, int pk, cudaStream_t stream) {
  RowCfg cfg;
  int rc = check_rows("ln_affine_bwd", M, C, &cfg);
  if (rc != CNX_OK) return rc;
  rc = check_patch("ln_affine_bwd", M, pH, pW, pk);
  if (rc != CNX_OK) return rc;
  if ((rc = zero_async(dw, sizeof(float) * C, stream, "ln_affine_bwd")) != CNX_OK) return rc;
  if ((rc = zero_async(db, sizeof(float) * C, stream, "ln_affine_bwd")) != CNX_OK) return rc;
  // ~4 CTAs per SM, each walking a contiguous run of rows (multiple of the rows in flight)
  const int slots = kRowThreads / cfg.lpr;
  int64_t rpc = (M + 4 * sm_count() - 1) / (4 * sm_count());
  rpc = (rpc + slots - 1) / slots * slots;
  if (rpc < slots) rpc = slots;
  const int rows_per_cta = static_cast<int>(rpc);
  const unsigned grid = static_cast<unsigned>((M + rpc - 1) / rpc);
#define CNX_LNB(LPR_, NV_)                                                                      \
  ln_affine_bwd_kernel<LPR_, NV_><<<grid, kRowThreads, 0, stream>>>(                            \
      static_cast<const __nv_bfloat16*>(g), static_cast<const __nv_bfloat16*>(x), w, mean, rstd, \
      static_cast<__nv_bfloat16*>(dx), dw, db, M, C, rows_per_cta, pH, pW, pk)
  if (cfg.lpr == 8) CNX_LNB(8, 1);
  else if (cfg.lpr == 16) CNX_LNB(16, 1);
  else if (cfg.nv == 1) CNX_LNB(32, 1);
  else if (cfg.nv == 2) CNX_LNB(32, 2);
  else CNX_LNB(32, 4);
#undef CNX_LNB
  return launch_status("ln_affine_bwd");
}

int colsum_bf16(const void* x, float* out, int64_t M, int C, cudaStream_t stream) {
  if (M <= 0 || C <= 0 || (C % 8) != 0 || C > 2048) {
    set_last_error("colsum_bf16: bad shape M=%lld C=%d (C must be a multiple of 8, <= 2048)",
                   (long long)M, C);
    return CNX_ERR_BAD_SHAPE;
  }
  if (int rc = zero_async(out, sizeof(float) * C, stream, "colsum_bf16")) return rc;
  const int rows_in_flight = kColsumThreads / (C >> 3);
  int64_t rpc = (M + 8 * sm_count() - 1) / (8 * sm_count());     // ~8 CTAs per SM
  rpc = (rpc + rows_in_flight - 1) / rows_in_flight * rows_in_flight;
  const unsigned grid = static_cast<unsigned>((M + rpc - 1) / rpc);
  colsum_bf16_kernel<<<grid, kColsumThreads, sizeof(float) * C, stream>>>(
      static_cast<const __nv_bfloat16*>(x), out, M, C, static_cast<int>(rpc));
  return launch_status("colsum_bf16");
}

int avgpool_nhwc_fwd(const void* x, void* y, int B, int HW, int C, cudaStream_t stream) {
  if (B <= 0 || HW <= 0 || C <= 0 || (C % 8) != 0 || B > 65535) {
    set_last_error("avgpool_nhwc_fwd: bad shape B=%d HW=%d C=%d (C must be a multiple of 8)", B, HW, C);
    return CNX_ERR_BAD_SHAPE;
  }
  dim3 grid((C + 63) / 64, B);
  avgpool_fwd_kernel<<<grid, kPoolThreads, 0, stream>>>(static_cast<const __nv_bfloat16*>(x),
                                                         static_cast<__nv_bfloat16*>(y), HW, C,
                                                         1.0f / static_cast<float>(HW));
  return launch_status("avgpool_nhwc_fwd");
}

int avgpool_nhwc_bwd(const void* g, void* dx, int B, int HW, int C, cudaStream_t stream) {
  if (B <= 0 || HW <= 0 || C <= 0 || (C % 8) != 0) {
    set_last_error("avgpool_nhwc_bwd: bad shape B=%d HW=%d C=%d (C must be a multiple of 8)", B, HW, C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int64_t total_vec = static_cast<int64_t>(B) * HW * (C >> 3);
  int64_t grid = (total_vec + kPoolThreads - 1) / kPoolThreads;
  if (grid > 16 * sm_count()) grid = 16 * sm_count();
  avgpool_bwd_kernel<<<static_cast<unsigned>(grid), kPoolThreads, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(g), static_cast<__nv_bfloat16*>(dx), HW, C >> 3, total_vec,
      1.0f / static_cast<float>(HW));
  return launch_status("avgpool_nhwc_bwd");
}

}  // namespace cnx
