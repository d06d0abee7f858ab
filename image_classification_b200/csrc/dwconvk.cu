// dwconvk.cu — depthwise k x k convolution (k = 3 or 5, stride 1 or 2, padding k/2, no bias) on
// NHWC bf16, forward and backward, fp32 accumulation.
//
// SURVEY.md §8 f-4 (BASELINE config 5): the depthwise half of the depthwise-separable blocks of
// timm `efficientnet_b2` that /root/reference/previous/V3-efb2/train_advanced_v3.0.py:330 builds
// (`conv_dw` of DepthwiseSeparableConv / InvertedResidual: 3x3 and 5x5, stride 1 and 2, channels
// 32 .. 2112, maps 130^2 .. 9^2).  The pointwise half runs on the tcgen05 GEMM (gemm2.cu).
//
// 9 or 25 FMAs per element against 4 bytes of HBM traffic: these are HBM / L1 bound, not FMA
// bound like the 7x7 kernels, so the design is a streaming one:
//   * a thread owns VEC channels (one 16-byte or 8-byte vector; a warp reads 512 / 256 contiguous
//     bytes of a pixel) and a strip of TW = 4 consecutive output columns;
//   * per kernel row the strip's (TW-1)*S + K input columns are loaded once into registers and
//     reused by every tap that touches them (sliding window; all indices are compile-time after
//     unrolling), 2.5-4x fewer L1 requests than one load per tap;
//   * padding is a predicate (zero contribution), weights are tap-major fp32 [K*K, C] read
//     through the read-only path (<= 211 KB, L1/L2 resident);
//   * the filter gradient keeps K*K*VEC accumulators in registers over all the strips a thread
//     walks, then meets the other threads of the CTA in shared memory: one atomic per
//     (tap, channel) per CTA.
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int kThreads = 256;
constexpr int TW = 4;                     // output (fwd, bwd-weight) / input (bwd-data) columns per strip

__host__ __device__ constexpr int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
__host__ __device__ constexpr int ceil_div(int a, int b) { return -floor_div(-a, b); }

template <int VEC> struct Vec;
template <> struct Vec<8> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&f)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = bf16x2_to_f32x2(w[i]);
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&f)[8]) {
    uint4 v;
    v.x = f32x2_to_bf16x2(f[0], f[1]); v.y = f32x2_to_bf16x2(f[2], f[3]);
    v.z = f32x2_to_bf16x2(f[4], f[5]); v.w = f32x2_to_bf16x2(f[6], f[7]);
    *reinterpret_cast<uint4*>(p) = v;
  }
};
template <> struct Vec<4> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&f)[4]) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = bf16x2_to_f32x2(v.x), b = bf16x2_to_f32x2(v.y);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
  }
};

// packed 16-byte load (zeros when the pixel is padding) and its unpack: the conv kernels issue all
// loads of a kernel row before the first FMA (a load per tap interleaved with its FMAs left one or
// two loads in flight per thread)
__device__ __forceinline__ uint4 ldg_or_zero(const __nv_bfloat16* p, bool ok) {
  return ok ? __ldg(reinterpret_cast<const uint4*>(p)) : make_uint4(0u, 0u, 0u, 0u);
}
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = bf16x2_to_f32x2(w[i]);
    f[2 * i] = t.x; f[2 * i + 1] = t.y;
  }
}

template <int VEC>
__device__ __forceinline__ void load_w(const float* __restrict__ p, float (&w)[VEC]) {
#pragma unroll
  for (int i = 0; i < VEC; i += 4) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p + i));
    w[i] = t.x; w[i + 1] = t.y; w[i + 2] = t.z; w[i + 3] = t.w;
  }
}

struct Geo {
  int B, H, W, C, Ho, Wo;
};

// ------------------------------------------------------------------------------------ forward
// y[b, oy, ox, c] = sum_{kh, kw} x[b, oy*S - P + kh, ox*S - P + kw, c] * w[kh*K + kw, c]
template <int K, int S>
__global__ void __launch_bounds__(kThreads)
dwk_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ wt,
               __nv_bfloat16* __restrict__ y, Geo g, long long total) {
  constexpr int P = K / 2, NC = (TW - 1) * S + K;
  const int nvec = g.C >> 3;
  const int strips = (g.Wo + TW - 1) / TW;
  // 32-bit work-item arithmetic (the launcher rejects total >= 2^31): six 64-bit div / mod per
  // strip cost more instructions than the strip's FMAs
  const unsigned utotal = static_cast<unsigned>(total), ustep = gridDim.x * kThreads;
  for (unsigned idx = blockIdx.x * kThreads + threadIdx.x; idx < utotal; idx += ustep) {
    const int v = static_cast<int>(idx % nvec);
    unsigned r = idx / nvec;
    const int strip = static_cast<int>(r % strips); r /= strips;
    const int oy = static_cast<int>(r % g.Ho);
    const int b = static_cast<int>(r / g.Ho);
    const int ox0 = strip * TW, c0 = v * 8;
    const int ix0 = ox0 * S - P;
    float acc[TW][8];
#pragma unroll
    for (int t = 0; t < TW; ++t)
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[t][i] = 0.f;
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int iy = oy * S - P + kh;
      if (iy < 0 || iy >= g.H) continue;
      const __nv_bfloat16* xrow = x + (static_cast<long long>(b) * g.H + iy) * g.W * g.C + c0;
      float w[K][8];
#pragma unroll
      for (int kw = 0; kw < K; ++kw) load_w<8>(wt + (kh * K + kw) * g.C + c0, w[kw]);
      uint4 xr[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const int ix = ix0 + c;
        xr[c] = ldg_or_zero(xrow + static_cast<long long>(ix) * g.C, ix >= 0 && ix < g.W);
      }
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        float xv[8];
        unpack8(xr[c], xv);
#pragma unroll
        for (int t = 0; t < TW; ++t) {
          const int kw = c - t * S;                      // compile-time after unrolling
          if (kw >= 0 && kw < K) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[t][i] = fmaf(xv[i], w[kw][i], acc[t][i]);
          }
        }
      }
    }
    __nv_bfloat16* yrow = y + ((static_cast<long long>(b) * g.Ho + oy) * g.Wo + ox0) * g.C + c0;
#pragma unroll
    for (int t = 0; t < TW; ++t)
      if (ox0 + t < g.Wo) Vec<8>::store(yrow + static_cast<long long>(t) * g.C, acc[t]);
  }
}

// ------------------------------------------------------------------------------ backward, data
// dx[b, iy, ix, c] = sum_{kh, kw : (iy+P-kh) % S == 0, (ix+P-kw) % S == 0}
//                        dy[b, (iy+P-kh)/S, (ix+P-kw)/S, c] * w[kh*K + kw, c]
template <int K, int S>
__global__ void __launch_bounds__(kThreads)
dwk_bwd_data_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ wt,
                    __nv_bfloat16* __restrict__ dx, Geo g, long long total) {
  constexpr int P = K / 2;
  constexpr int LO = ceil_div(P - K + 1, S), HI = floor_div(TW - 1 + P, S), NC = HI - LO + 1;
  const int nvec = g.C >> 3;
  const int strips = (g.W + TW - 1) / TW;
  const unsigned utotal = static_cast<unsigned>(total), ustep = gridDim.x * kThreads;
  for (unsigned idx = blockIdx.x * kThreads + threadIdx.x; idx < utotal; idx += ustep) {
    const int v = static_cast<int>(idx % nvec);
    unsigned r = idx / nvec;
    const int strip = static_cast<int>(r % strips); r /= strips;
    const int iy = static_cast<int>(r % g.H);
    const int b = static_cast<int>(r / g.H);
    const int ix0 = strip * TW, c0 = v * 8;      // ix0 % S == 0 (TW is a multiple of S)
    const int oxb = ix0 / S + LO;                // dy column of window slot 0
    float acc[TW][8];
#pragma unroll
    for (int t = 0; t < TW; ++t)
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[t][i] = 0.f;
#pragma unroll
    for (int kh = 0; kh < K; ++kh) {
      const int num = iy + P - kh;
      if (num < 0 || (num % S) != 0) continue;
      const int oy = num / S;
      if (oy >= g.Ho) continue;
      const __nv_bfloat16* drow = dy + (static_cast<long long>(b) * g.Ho + oy) * g.Wo * g.C + c0;
      float w[K][8];
#pragma unroll
      for (int kw = 0; kw < K; ++kw) load_w<8>(wt + (kh * K + kw) * g.C + c0, w[kw]);
      uint4 dr[NC];
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        const int ox = oxb + j;
        dr[j] = ldg_or_zero(drow + static_cast<long long>(ox) * g.C, ox >= 0 && ox < g.Wo);
      }
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        float dv[8];
        unpack8(dr[j], dv);
#pragma unroll
        for (int t = 0; t < TW; ++t) {
          const int kw = t + P - (LO + j) * S;           // compile-time
          if (kw >= 0 && kw < K) {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[t][i] = fmaf(dv[i], w[kw][i], acc[t][i]);
          }
        }
      }
    }
    __nv_bfloat16* xrow = dx + ((static_cast<long long>(b) * g.H + iy) * g.W + ix0) * g.C + c0;
#pragma unroll
    for (int t = 0; t < TW; ++t)
      if (ix0 + t < g.W) Vec<8>::store(xrow + static_cast<long long>(t) * g.C, acc[t]);
  }
}

// ---------------------------------------------------------------------------- backward, filter
// dw[kh*K + kw, c] (+)= sum_{b, oy, ox} dy[b, oy, ox, c] * x[b, oy*S - P + kh, ox*S - P + kw, c]
// Block = vpb channel vectors x (kThreads / vpb) strip lanes; grid.x = channel blocks,
// grid.y = strip chunks, grid.z = kernel row kh.  One kernel row per CTA keeps K*VEC accumulators
// per thread (40 registers at most) instead of K*K*VEC: the all-taps form needed 240 registers
// (one CTA of 8 warps per SM, loads issued one by one) and ran at 0.2-0.3 TB/s on L2-resident
// maps — a quarter of the EfficientNet-B2 step.  Here a strip costs TW dy loads and NC x loads,
// all issued before the first FMA; dy and the x rows are re-read by the K rows' CTAs out of L2.
template <int K, int S, int VEC>
__global__ void __launch_bounds__(kThreads)
dwk_bwd_weight_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                      float* __restrict__ dwt, Geo g, int vpb, long long strips_total,
                      long long strips_per_cta) {
  constexpr int P = K / 2, NC = (TW - 1) * S + K;
  extern __shared__ float s_dw[];                       // [K][vpb * VEC]
  const int nvec = g.C / VEC;
  const int vl = threadIdx.x % vpb, pl = threadIdx.x / vpb, npl = kThreads / vpb;
  const int v = blockIdx.x * vpb + vl;
  const int cw = vpb * VEC;
  const int kh = blockIdx.z;
  for (int i = threadIdx.x; i < K * cw; i += kThreads) s_dw[i] = 0.f;
  __syncthreads();
  const bool active = pl < npl && v < nvec;
  float acc[K][VEC];
#pragma unroll
  for (int kw = 0; kw < K; ++kw)
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[kw][i] = 0.f;
  if (active) {
    const int c0 = v * VEC;
    const int strips = (g.Wo + TW - 1) / TW;
    const unsigned s_begin = blockIdx.y * static_cast<unsigned>(strips_per_cta);
    const unsigned s_end = min(s_begin + static_cast<unsigned>(strips_per_cta),
                               static_cast<unsigned>(strips_total));
    for (unsigned sidx = s_begin + pl; sidx < s_end; sidx += npl) {
      const int strip = static_cast<int>(sidx % strips);
      const unsigned r = sidx / strips;
      const int oy = static_cast<int>(r % g.Ho);
      const int b = static_cast<int>(r / g.Ho);
      const int iy = oy * S - P + kh;
      if (iy < 0 || iy >= g.H) continue;                 // this kernel row sees padding only
      const int ox0 = strip * TW, ix0 = ox0 * S - P;
      float d[TW][VEC], xv[NC][VEC];
      const __nv_bfloat16* drow = dy + ((static_cast<long long>(b) * g.Ho + oy) * g.Wo + ox0) * g.C + c0;
      const __nv_bfloat16* xrow = x + ((static_cast<long long>(b) * g.H + iy) * g.W + ix0) * g.C + c0;
      if (ox0 + TW <= g.Wo && ix0 >= 0 && ix0 + NC <= g.W) {
        // interior strip: no padding column, no ragged end — plain loads, no predicates, no
        // zero-fill moves (two thirds of the issued instructions were integer / move work)
#pragma unroll
        for (int t = 0; t < TW; ++t) Vec<VEC>::load(drow + t * g.C, d[t]);
#pragma unroll
        for (int c = 0; c < NC; ++c) Vec<VEC>::load(xrow + c * g.C, xv[c]);
      } else {
#pragma unroll
        for (int t = 0; t < TW; ++t) {
          if (ox0 + t < g.Wo) {
            Vec<VEC>::load(drow + t * g.C, d[t]);
          } else {
#pragma unroll
            for (int i = 0; i < VEC; ++i) d[t][i] = 0.f;
          }
        }
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const int ix = ix0 + c;
          if (ix >= 0 && ix < g.W) {
            Vec<VEC>::load(xrow + c * g.C, xv[c]);
          } else {
#pragma unroll
            for (int i = 0; i < VEC; ++i) xv[c][i] = 0.f;
          }
        }
      }
#pragma unroll
      for (int c = 0; c < NC; ++c) {
#pragma unroll
        for (int t = 0; t < TW; ++t) {
          const int kw = c - t * S;                      // compile-time
          if (kw >= 0 && kw < K) {
#pragma unroll
            for (int i = 0; i < VEC; ++i) acc[kw][i] = fmaf(d[t][i], xv[c][i], acc[kw][i]);
          }
        }
      }
    }
#pragma unroll
    for (int kw = 0; kw < K; ++kw)
#pragma unroll
      for (int i = 0; i < VEC; ++i) atomicAdd(&s_dw[kw * cw + vl * VEC + i], acc[kw][i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K * cw; i += kThreads) {
    const int kw = i / cw, cc = i % cw;
    const int c = blockIdx.x * cw + cc;
    if (c < g.C) atomicAdd(dwt + (kh * K + kw) * g.C + c, s_dw[i]);
  }
}

int check_geo(const char* who, int B, int H, int W, int C, int k, int stride) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (C % 8) != 0 || (k != 3 && k != 5) ||
      (stride != 1 && stride != 2)) {
    set_last_error("%s: bad shape B=%d H=%d W=%d C=%d k=%d stride=%d (C %% 8 == 0, k in {3,5}, "
                   "stride in {1,2})", who, B, H, W, C, k, stride);
    return CNX_ERR_BAD_SHAPE;
  }
  return CNX_OK;
}

Geo make_geo(int B, int H, int W, int C, int k, int stride) {
  Geo g;
  g.B = B; g.H = H; g.W = W; g.C = C;
  g.Ho = (H + 2 * (k / 2) - k) / stride + 1;
  g.Wo = (W + 2 * (k / 2) - k) / stride + 1;
  return g;
}

unsigned grid_for(long long total) {
  long long grid = (total + kThreads - 1) / kThreads;
  const long long cap = 32LL * sm_count();
  return static_cast<unsigned>(grid < cap ? (grid > 0 ? grid : 1) : cap);
}

int status(const char* who) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("%s launch: %s", who, cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

template <int K, int S, int VEC>
int launch_weight(const __nv_bfloat16* dy, const __nv_bfloat16* x, float* dwt, const Geo& g,
                  cudaStream_t stream) {
  const int nvec = g.C / VEC;
  const int vpb = nvec < 32 ? nvec : 32;
  const int npl = kThreads / vpb;
  const long long strips_total = static_cast<long long>(g.B) * g.Ho * ((g.Wo + TW - 1) / TW);
  if (strips_total >= (1LL << 31) - (1LL << 24)) {
    set_last_error("dwconv_kxk_bwd: %lld strips exceed the 32-bit index range", strips_total);
    return CNX_ERR_BAD_SHAPE;
  }
  const int gx = (nvec + vpb - 1) / vpb;
  long long gy = (12LL * sm_count() + gx * K - 1) / (gx * K);       // ~12 CTAs per SM in total
  const long long max_gy = (strips_total + npl - 1) / npl;
  if (gy > max_gy) gy = max_gy;
  if (gy < 1) gy = 1;
  if (gy > 65535) gy = 65535;
  const long long per = (strips_total + gy - 1) / gy;
  const size_t smem = sizeof(float) * K * vpb * VEC;
  dwk_bwd_weight_kernel<K, S, VEC><<<dim3(gx, static_cast<unsigned>(gy), K), kThreads, smem, stream>>>(
      dy, x, dwt, g, vpb, strips_total, per);
  return status("dwconv_kxk_bwd (filter)");
}

}  // namespace

int dwconv_kxk_fwd(const void* x, const float* wt, void* y, int B, int H, int W, int C, int k,
                   int stride, cudaStream_t stream) {
  if (int rc = check_geo("dwconv_kxk_fwd", B, H, W, C, k, stride)) return rc;
  const Geo g = make_geo(B, H, W, C, k, stride);
  const long long total = static_cast<long long>(B) * g.Ho * ((g.Wo + TW - 1) / TW) * (C >> 3);
  if (total >= (1LL << 31) - (1LL << 24)) {
    set_last_error("dwconv_kxk_fwd: %lld work items exceed the 32-bit index range", total);
    return CNX_ERR_BAD_SHAPE;
  }
  const unsigned grid = grid_for(total);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yp = static_cast<__nv_bfloat16*>(y);
#define CNX_DWK_FWD(KK, SS) dwk_fwd_kernel<KK, SS><<<grid, kThreads, 0, stream>>>(xp, wt, yp, g, total)
  if (k == 3 && stride == 1) CNX_DWK_FWD(3, 1);
  else if (k == 3) CNX_DWK_FWD(3, 2);
  else if (stride == 1) CNX_DWK_FWD(5, 1);
  else CNX_DWK_FWD(5, 2);
#undef CNX_DWK_FWD
  return status("dwconv_kxk_fwd");
}

int dwconv_kxk_bwd(const void* dy, const void* x, const float* wt, void* dx, float* dwt, int B,
                   int H, int W, int C, int k, int stride, int accumulate, cudaStream_t stream) {
  if (int rc = check_geo("dwconv_kxk_bwd", B, H, W, C, k, stride)) return rc;
  const Geo g = make_geo(B, H, W, C, k, stride);
  const __nv_bfloat16* dyp = static_cast<const __nv_bfloat16*>(dy);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  if (dx != nullptr) {
    const long long total = static_cast<long long>(B) * H * ((W + TW - 1) / TW) * (C >> 3);
    if (total >= (1LL << 31) - (1LL << 24)) {
      set_last_error("dwconv_kxk_bwd: %lld work items exceed the 32-bit index range", total);
      return CNX_ERR_BAD_SHAPE;
    }
    const unsigned grid = grid_for(total);
    __nv_bfloat16* dxp = static_cast<__nv_bfloat16*>(dx);
#define CNX_DWK_BD(KK, SS) dwk_bwd_data_kernel<KK, SS><<<grid, kThreads, 0, stream>>>(dyp, wt, dxp, g, total)
    if (k == 3 && stride == 1) CNX_DWK_BD(3, 1);
    else if (k == 3) CNX_DWK_BD(3, 2);
    else if (stride == 1) CNX_DWK_BD(5, 1);
    else CNX_DWK_BD(5, 2);
#undef CNX_DWK_BD
    if (int rc = status("dwconv_kxk_bwd (data)")) return rc;
  }
  if (dwt != nullptr) {
    if (!accumulate)
      if (int rc = zero_async(dwt, sizeof(float) * k * k * C, stream, "dwconv_kxk_bwd")) return rc;
    if (k == 3 && stride == 1) return launch_weight<3, 1, 4>(dyp, xp, dwt, g, stream);
    if (k == 3) return launch_weight<3, 2, 4>(dyp, xp, dwt, g, stream);
    if (stride == 1) return launch_weight<5, 1, 4>(dyp, xp, dwt, g, stream);
    return launch_weight<5, 2, 4>(dyp, xp, dwt, g, stream);
  }
  return CNX_OK;
}

}  // namespace cnx
