// layout.cu — NCHW fp32 <-> NHWC bf16 conversion at the nn.Module boundary.
//
// timm blocks take and return NCHW tensors; the kernels of this library work
// on NHWC bf16.  A model whose residual stream already is channels_last bf16 never calls these;
// they exist so ConvNeXtBlockB200 is a drop-in when handed a contiguous fp32 NCHW activation.
// Per image this is a [C][HW] <-> [HW][C] transpose: 32x32 tiles through padded shared memory,
// coalesced on both sides.
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int C, int HW) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const size_t img = static_cast<size_t>(blockIdx.z) * C * HW;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, p = p0 + tx;
    tile[ty + i][tx] = (c < C && p < HW) ? src[img + static_cast<size_t>(c) * HW + p] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int p = p0 + ty + i, c = c0 + tx;
    if (p < HW && c < C)
      dst[img + static_cast<size_t>(p) * C + c] = __float2bfloat16_rn(tile[tx][ty + i]);
  }
}

__global__ void __launch_bounds__(256)
nhwc_to_nchw_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, int C, int HW) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const size_t img = static_cast<size_t>(blockIdx.z) * C * HW;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int p = p0 + ty + i, c = c0 + tx;
    tile[ty + i][tx] =
        (c < C && p < HW) ? __bfloat162float(src[img + static_cast<size_t>(p) * C + c]) : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c = c0 + ty + i, p = p0 + tx;
    if (p < HW && c < C) dst[img + static_cast<size_t>(c) * HW + p] = tile[tx][ty + i];
  }
}

// Stem patchify: x [B, Cin, H, W] fp32 (NCHW, what the batch mixer emits) -> P [B*H/k*W/k, k*k*Cin]
// bf16, inner order (kh, kw, c): one im2col row per output pixel of a k x k stride-k convolution
// (no overlap, so this is a permutation + cast).  One CTA per (image, row of patches): coalesced
// reads along W into shared memory, one contiguous run of W/k*k*k*Cin bf16 written out.
// CIN / KK > 0: compile-time channel count and patch size (divisions become shifts / multiplies)
template <int CIN, int KK>
__global__ void __launch_bounds__(256)
patchify_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ P, int Cin_rt, int H, int W,
                int k_rt) {
  extern __shared__ float s_in[];                 // [Cin][k][W + 1]  (+1: conflict-free gathers)
  const int Cin = CIN > 0 ? CIN : Cin_rt, k = KK > 0 ? KK : k_rt;
  const int oh = blockIdx.x, b = blockIdx.y;
  const int WP = W + 1;
  for (int r = threadIdx.x / 32; r < Cin * k; r += 8) {      // one warp per input row: coalesced
    const int kh = r % k, c = r / k;
    const float* src = x + ((static_cast<size_t>(b) * Cin + c) * H + oh * k + kh) * W;
    for (int w = threadIdx.x & 31; w < W; w += 32) s_in[r * WP + w] = src[w];
  }
  __syncthreads();
  const int OW = W / k, K = k * k * Cin;
  __nv_bfloat16* dst = P + (static_cast<size_t>(b) * (H / k) + oh) * OW * K;
  for (int e = 2 * threadIdx.x; e < OW * K; e += 512) {      // K is even: two outputs per thread
    float v[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int ee = e + u;
      const int ow = ee / K, r = ee - ow * K;
      const int c = r % Cin, t = r / Cin, kw = t % k, kh = t / k;
      v[u] = s_in[(c * k + kh) * WP + ow * k + kw];
    }
    *reinterpret_cast<__nv_bfloat162*>(dst + e) = __floats2bfloat162_rn(v[0], v[1]);
  }
}

}  // namespace

int patchify_nchw_f32(const float* x, void* P, int B, int Cin, int H, int W, int k,
                      cudaStream_t stream) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || k <= 0 || (H % k) != 0 || (W % k) != 0 ||
      B > 65535 || ((k * k * Cin) % 8) != 0 || static_cast<size_t>(Cin) * k * (W + 1) * 4 > 160 * 1024) {
    set_last_error("patchify: bad shape B=%d Cin=%d H=%d W=%d k=%d (k must divide H and W, "
                   "k*k*Cin must be a multiple of 8)", B, Cin, H, W, k);
    return CNX_ERR_BAD_SHAPE;
  }
  const size_t smem = static_cast<size_t>(Cin) * k * (W + 1) * 4;
  auto kern = (Cin == 3 && k == 4) ? patchify_kernel<3, 4> : patchify_kernel<0, 0>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) {
      set_last_error("patchify: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return CNX_ERR_LAUNCH;
    }
  }
  kern<<<dim3(H / k, B), 256, smem, stream>>>(x, static_cast<__nv_bfloat16*>(P), Cin, H, W, k);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("patchify launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int H, int W,
                          cudaStream_t stream) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || B > 65535) {
    set_last_error("nchw_f32_to_nhwc_bf16: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    return CNX_ERR_BAD_SHAPE;
  }
  const int HW = H * W;
  dim3 grid((HW + 31) / 32, (C + 31) / 32, B);
  nchw_to_nhwc_kernel<<<grid, 256, 0, stream>>>(src, static_cast<__nv_bfloat16*>(dst), C, HW);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("nchw_f32_to_nhwc_bf16 launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int H, int W,
                          cudaStream_t stream) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || B > 65535) {
    set_last_error("nhwc_bf16_to_nchw_f32: bad shape B=%d C=%d H=%d W=%d", B, C, H, W);
    return CNX_ERR_BAD_SHAPE;
  }
  const int HW = H * W;
  dim3 grid((HW + 31) / 32, (C + 31) / 32, B);
  nhwc_to_nchw_kernel<<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(src), dst, C, HW);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("nhwc_bf16_to_nchw_f32 launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace cnx
