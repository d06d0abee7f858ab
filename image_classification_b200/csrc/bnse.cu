// bnse.cu — BatchNorm2d (batch statistics) + SiLU and squeeze-excite scaling over NHWC bf16 maps:
// the memory-bound glue between the convolutions of an EfficientNet MBConv block
// (BASELINE config 5).
//
// Replaces, in timm's / torchvision's efficientnet_b2 that the older reference trains
// (/root/reference/previous/V3-efb2/train_advanced_v3.0.py:330 model, :100-104 forward/backward):
//   conv -> BatchNorm2d -> SiLU            => bn_stats + bn_act_fwd          (2 passes over x)
//   its autograd backward (5 ATen kernels) => bn_act_bwd_reduce + _apply     (2 passes over g, x)
//   x * sigmoid(fc2(silu(fc1(avgpool(x))))) => pooled sums come out of bn_act_fwd for free, the
//                                             two [B,C] gate layers are se_rowdot / se_coldot /
//                                             se_gate_wgrad, the scaling se_scale_fwd; backward
//                                             se_bwd_reduce + se_bwd_apply
//
// One thread mapping for every kernel: a CTA owns a channel slab of LW 16-byte vectors (LW = 1, 2,
// 4 or 8, i.e. 8..64 channels) and a range of rows of ONE image; its 256 threads are LW vector
// lanes x 256/LW row lanes.  A thread's channels never change, so the per-channel constants
// (scale/shift, mean/rstd) live in registers; row lanes meet in shared memory once per CTA and
// leave one atomic per channel.  Pure streaming, HBM-bound: algorithmic bytes per element
// (E = M*C, bf16) are 2+2 (stats: read; apply: read + write = 3*E*2 B in total) forward and
// 5*E*2 B backward (g and x read twice, dx written once).
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ uint4 ldv(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void stv(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
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
// ex2.approx + rcp.approx (2 MUFU ops, ~2 ulp): the IEEE division costs ~8 more instructions per
// element and made the SiLU passes issue-bound instead of HBM-bound
__device__ __forceinline__ float sigmoidf_(float u) { return __fdividef(1.0f, 1.0f + __expf(-u)); }

// Where this thread sits: vector lane / row lane, first channel, row range inside image blockIdx.z.
struct Slot {
  int lane, rl, nrl, c0;
  bool live;
  int64_t row0;        // first row of the image in the [M, C] matrix
  int r_begin, r_end;  // rows of that image owned by the CTA
};
__device__ __forceinline__ Slot make_slot(int lw, int C, int HW, int rows_per_cta) {
  Slot s;
  s.lane = threadIdx.x % lw;
  s.rl = threadIdx.x / lw;
  s.nrl = kThreads / lw;
  s.c0 = (blockIdx.x * lw + s.lane) * 8;
  s.live = s.c0 < C;
  s.row0 = static_cast<int64_t>(blockIdx.z) * HW;
  s.r_begin = blockIdx.y * rows_per_cta;
  s.r_end = min(s.r_begin + rows_per_cta, HW);
  return s;
}

// Rows are walked kBatch at a time with every 16-byte load of the batch issued before the first
// use: left to itself the compiler interleaves load and arithmetic and a thread has one or two
// loads in flight, which showed as ~3 TB/s in ncu where the same traffic pattern reaches 5.5 TB/s.
constexpr int kBatch = 4;
template <int NB>
__device__ __forceinline__ void load_rows(uint4 (&v)[NB], const __nv_bfloat16* base, int r,
                                          const Slot& s, int C) {
  const __nv_bfloat16* p = base + static_cast<int64_t>(r) * C;
  const int64_t step = static_cast<int64_t>(s.nrl) * C;
  if (r + (NB - 1) * s.nrl < s.r_end) {          // whole batch inside the range: plain loads
#pragma unroll
    for (int j = 0; j < NB; ++j) v[j] = ldv(p + j * step);
  } else {
#pragma unroll
    for (int j = 0; j < NB; ++j)
      v[j] = r + j * s.nrl < s.r_end ? ldv(p + j * step) : make_uint4(0u, 0u, 0u, 0u);
  }
}

// Sum the per-thread partials acc[8] over the row lanes of the CTA and atomically add the slab's
// totals to out[c] (out points at channel 0).  s_red: kThreads * 8 floats.
__device__ __forceinline__ void slab_reduce_atomic(const float (&acc)[8], float* s_red, int lw,
                                                   int C, float* out) {
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 8; ++k) s_red[threadIdx.x * 8 + k] = acc[k];
  __syncthreads();
  const int width = lw * 8;
  if (threadIdx.x < width) {
    const int c = blockIdx.x * width + threadIdx.x;
    if (c < C) {
      float t = 0.f;
      for (int r = 0; r < kThreads / lw; ++r) t += s_red[r * width + threadIdx.x];
      atomicAdd(out + c, t);
    }
  }
}

// ------------------------------------------------------------------ BatchNorm statistics
// sums[0..C) += sum_m d[m,c], sums[C..2C) += sum_m d[m,c]^2 with d = x - x[0,c]   (sums zeroed
// by the launcher)
__global__ void __launch_bounds__(kThreads)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums, int C, int HW,
                int rows_per_cta, int lw) {
  __shared__ float s_red[kThreads * 8];
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  float a1[8], a2[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a1[k] = a2[k] = 0.f;
  if (s.live) {
    // sums of (x - pivot), pivot = row 0 of the matrix: E[d^2] - E[d]^2 then cancels only if the
    // mean is many standard deviations away from a SAMPLE of the channel, not from zero
    float pv[8];
    unpack8(ldv(x + s.c0), pv);
    const __nv_bfloat16* xp = x + s.row0 * C + s.c0;
    for (int r = s.r_begin + s.rl; r < s.r_end; r += kBatch * s.nrl) {
      uint4 v[kBatch];
      load_rows(v, xp, r, s, C);
#pragma unroll
      for (int j = 0; j < kBatch; ++j) {
        if (r + j * s.nrl >= s.r_end) break;      // rows past the end loaded zeros, not the pivot
        float f[8];
        unpack8(v[j], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float d = f[k] - pv[k];
          a1[k] += d;
          a2[k] = fmaf(d, d, a2[k]);
        }
      }
    }
  }
  slab_reduce_atomic(a1, s_red, lw, C, sums);
  slab_reduce_atomic(a2, s_red, lw, C, sums + C);
}

// Per-channel scale / shift of this thread's 8 channels.  Training: from the batch sums (biased
// variance, as F.batch_norm normalises with); the CTAs of row-chunk 0 / image 0 also publish
// mean / rstd for backward and update the running statistics (unbiased variance, momentum) the
// way nn.BatchNorm2d does.  Eval: from the running statistics.
struct BnArgs {
  const float* sums;          // [2C] or nullptr (eval)
  const __nv_bfloat16* pivot; // row 0 of x: the shift bn_stats subtracted
  const float* weight;        // [C]
  const float* bias;          // [C]
  float* running_mean;        // [C] or nullptr
  float* running_var;         // [C] or nullptr
  float* mean_out;            // [C]
  float* rstd_out;            // [C]
  float inv_m, unbias, momentum, eps;
};
__device__ __forceinline__ void bn_channel_consts(const BnArgs& p, int C, int c0, float (&sc)[8],
                                                  float (&sh)[8], bool publish) {
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = c0 + k;
    float mean, var;
    if (p.sums != nullptr) {
      const float dm = p.sums[c] * p.inv_m;
      mean = __bfloat162float(p.pivot[c]) + dm;
      var = fmaxf(p.sums[C + c] * p.inv_m - dm * dm, 0.f);
    } else {
      mean = p.running_mean[c];
      var = p.running_var[c];
    }
    const float rstd = rsqrtf(var + p.eps);
    const float w = p.weight[c];
    sc[k] = w * rstd;
    sh[k] = p.bias[c] - mean * w * rstd;
    if (publish) {
      p.mean_out[c] = mean;
      p.rstd_out[c] = rstd;
      if (p.sums != nullptr && p.running_mean != nullptr) {
        p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * mean;
        p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * var * p.unbias;
      }
    }
  }
}

// y = act(x * scale + shift)  (act: 0 identity, 1 SiLU), bf16 out.  pool_sum [B, C] fp32 (optional,
// zeroed by the launcher) receives sum over the image's rows of the ROUNDED y: the squeeze-excite
// average pool that follows the depthwise stage costs no extra pass.
template <int ACT>
__global__ void __launch_bounds__(kThreads)
bn_act_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, BnArgs p,
                  float* __restrict__ pool_sum, int C, int HW, int rows_per_cta, int lw) {
  __shared__ float s_red[kThreads * 8];
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  if (s.live) {
    float sc[8], sh[8];
    bn_channel_consts(p, C, s.c0, sc, sh, s.rl == 0 && blockIdx.y == 0 && blockIdx.z == 0);
    const __nv_bfloat16* xp = x + s.row0 * C + s.c0;
    __nv_bfloat16* yp = y + s.row0 * C + s.c0;
    for (int r = s.r_begin + s.rl; r < s.r_end; r += kBatch * s.nrl) {
      uint4 v[kBatch];
      load_rows(v, xp, r, s, C);
#pragma unroll
      for (int j = 0; j < kBatch; ++j) {
        const int rr = r + j * s.nrl;
        if (rr >= s.r_end) break;
        float f[8];
        unpack8(v[j], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float u = fmaf(f[k], sc[k], sh[k]);
          f[k] = ACT == 1 ? u * sigmoidf_(u) : u;
        }
        const uint4 o = pack8(f);
        stv(yp + static_cast<int64_t>(rr) * C, o);
        if (pool_sum != nullptr) {
          unpack8(o, f);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += f[k];
        }
      }
    }
  }
  if (pool_sum != nullptr)
    slab_reduce_atomic(acc, s_red, lw, C, pool_sum + static_cast<int64_t>(blockIdx.z) * C);
}

// du = g * act'(u) with u = xhat * w + b recomputed from x.  Pass 1: dbias += sum du,
// dweight += sum du * xhat.
template <int ACT>
__device__ __forceinline__ float act_grad(float g, float u) {
  if (ACT == 1) {
    const float sg = sigmoidf_(u);
    return g * sg * fmaf(u, 1.f - sg, 1.f);
  }
  return g;
}
struct BnBwdArgs {
  const float* weight;   // [C]
  const float* bias;     // [C]
  const float* mean;     // [C]
  const float* rstd;     // [C]
  float* dweight;        // [C]
  float* dbias;          // [C]
  float inv_m;
};
template <int ACT>
__global__ void __launch_bounds__(kThreads)
bn_act_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                         BnBwdArgs p, int C, int HW, int rows_per_cta, int lw) {
  __shared__ float s_red[kThreads * 8];
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  float a1[8], a2[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a1[k] = a2[k] = 0.f;
  if (s.live) {
    // per channel: xhat = x * rs + nmr, u = x * A + Bc  (one FMA each, four constants instead of
    // mean / rstd / weight / bias kept apart)
    float rs[8], nmr[8], A[8], Bc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      rs[k] = p.rstd[s.c0 + k];
      nmr[k] = -p.mean[s.c0 + k] * rs[k];
      A[k] = p.weight[s.c0 + k] * rs[k];
      Bc[k] = fmaf(nmr[k], p.weight[s.c0 + k], p.bias[s.c0 + k]);
    }
    const int64_t base = s.row0 * C + s.c0;
    for (int r = s.r_begin + s.rl; r < s.r_end; r += 2 * s.nrl) {
      uint4 vg[2], vx[2];
      load_rows(vg, g + base, r, s, C);           // g = 0 past the end: du = 0, nothing added
      load_rows(vx, x + base, r, s, C);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float fg[8], fx[8];
        unpack8(vg[j], fg);
        unpack8(vx[j], fx);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float xh = fmaf(fx[k], rs[k], nmr[k]);
          const float du = act_grad<ACT>(fg[k], fmaf(fx[k], A[k], Bc[k]));
          a1[k] += du;
          a2[k] = fmaf(du, xh, a2[k]);
        }
      }
    }
  }
  slab_reduce_atomic(a1, s_red, lw, C, p.dbias);
  slab_reduce_atomic(a2, s_red, lw, C, p.dweight);
}

// Pass 2: dx = w * rstd * (du - dbias / M - xhat * dweight / M)
template <int ACT>
__global__ void __launch_bounds__(kThreads)
bn_act_bwd_apply_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                        __nv_bfloat16* __restrict__ dx, BnBwdArgs p, int C, int HW,
                        int rows_per_cta, int lw) {
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  if (!s.live) return;
  // dx = A * du + x * P + Q with A = w * rstd, P = -A * k2 * rstd, Q = -A * (k1 + k2 * nmr),
  // k1 = dbias / M, k2 = dweight / M, nmr = -mean * rstd; u = x * A + Bc as in the reduce pass
  float A[8], Bc[8], P[8], Q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float rs = p.rstd[s.c0 + k], w = p.weight[s.c0 + k];
    const float nmr = -p.mean[s.c0 + k] * rs;
    const float k1 = p.dbias[s.c0 + k] * p.inv_m, k2 = p.dweight[s.c0 + k] * p.inv_m;
    A[k] = w * rs;
    Bc[k] = fmaf(nmr, w, p.bias[s.c0 + k]);
    P[k] = -A[k] * k2 * rs;
    Q[k] = -A[k] * fmaf(k2, nmr, k1);
  }
  const int64_t base = s.row0 * C + s.c0;
  for (int r = s.r_begin + s.rl; r < s.r_end; r += 2 * s.nrl) {
    uint4 vg[2], vx[2];
    load_rows(vg, g + base, r, s, C);
    load_rows(vx, x + base, r, s, C);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int rr = r + j * s.nrl;
      if (rr >= s.r_end) break;
      float fg[8], fx[8];
      unpack8(vg[j], fg);
      unpack8(vx[j], fx);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float du = act_grad<ACT>(fg[k], fmaf(fx[k], A[k], Bc[k]));
        fx[k] = fmaf(A[k], du, fmaf(fx[k], P[k], Q[k]));
      }
      stv(dx + base + static_cast<int64_t>(rr) * C, pack8(fx));
    }
  }
}

// ------------------------------------------------------------------ squeeze-excite scaling
// y[b, p, c] = x[b, p, c] * s[b, c]   (s = sigmoid gate, fp32 [B, C])
__global__ void __launch_bounds__(kThreads)
se_scale_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gate,
                    __nv_bfloat16* __restrict__ y, int C, int HW, int rows_per_cta, int lw) {
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  if (!s.live) return;
  float sc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) sc[k] = gate[static_cast<int64_t>(blockIdx.z) * C + s.c0 + k];
  const int64_t base = s.row0 * C + s.c0;
  for (int r = s.r_begin + s.rl; r < s.r_end; r += kBatch * s.nrl) {
    uint4 v[kBatch];
    load_rows(v, x + base, r, s, C);
#pragma unroll
    for (int j = 0; j < kBatch; ++j) {
      const int rr = r + j * s.nrl;
      if (rr >= s.r_end) break;
      float f[8];
      unpack8(v[j], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] *= sc[k];
      stv(y + base + static_cast<int64_t>(rr) * C, pack8(f));
    }
  }
}

// dgate[b, c] += sum_p g[b, p, c] * x[b, p, c]        (dgate zeroed by the launcher)
__global__ void __launch_bounds__(kThreads)
se_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                     float* __restrict__ dgate, int C, int HW, int rows_per_cta, int lw) {
  __shared__ float s_red[kThreads * 8];
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  if (s.live) {
    const int64_t base = s.row0 * C + s.c0;
    for (int r = s.r_begin + s.rl; r < s.r_end; r += 2 * s.nrl) {
      uint4 vg[2], vx[2];
      load_rows(vg, g + base, r, s, C);
      load_rows(vx, x + base, r, s, C);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float fg[8], fx[8];
        unpack8(vg[j], fg);
        unpack8(vx[j], fx);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(fg[k], fx[k], acc[k]);
      }
    }
  }
  slab_reduce_atomic(acc, s_red, lw, C, dgate + static_cast<int64_t>(blockIdx.z) * C);
}

// dx[b, p, c] = g[b, p, c] * s[b, c] + dpool[b, c] / HW   (the gate path's gradient through the
// average pool joins the scaling path's in the same pass)
__global__ void __launch_bounds__(kThreads)
se_bwd_apply_kernel(const __nv_bfloat16* __restrict__ g, const float* __restrict__ gate,
                    const float* __restrict__ dpool, __nv_bfloat16* __restrict__ dx, int C, int HW,
                    int rows_per_cta, int lw, float inv_hw) {
  const Slot s = make_slot(lw, C, HW, rows_per_cta);
  if (!s.live) return;
  float sc[8], ad[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    sc[k] = gate[static_cast<int64_t>(blockIdx.z) * C + s.c0 + k];
    ad[k] = dpool[static_cast<int64_t>(blockIdx.z) * C + s.c0 + k] * inv_hw;
  }
  const int64_t base = s.row0 * C + s.c0;
  for (int r = s.r_begin + s.rl; r < s.r_end; r += kBatch * s.nrl) {
    uint4 v[kBatch];
    load_rows(v, g + base, r, s, C);
#pragma unroll
    for (int j = 0; j < kBatch; ++j) {
      const int rr = r + j * s.nrl;
      if (rr >= s.r_end) break;
      float f[8];
      unpack8(v[j], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], sc[k], ad[k]);
      stv(dx + base + static_cast<int64_t>(rr) * C, pack8(f));
    }
  }
}

// ------------------------------------------------------------------ squeeze-excite gate layers
// gate[b,c] = sigmoid(b2[c] + sum_j w2t[j,c] * silu(b1[j] + sum_c' w1[j,c'] * pooled[b,c'])),
// pooled = pool_sum * pool_scale.  Two [B,C]-sized layers (Cs <= C/4 hidden units): as PyTorch ops
// they were ~30 launches per block (SIMT sgemm + bias + activations, forward / recompute /
// backward), 2 ms of a 15.5 ms EfficientNet-B2 step.  w2t is the expand weight transposed to
// [Cs, C] so that every global read is coalesced along c.  Two kernel shapes serve forward and
// backward; both tile (hidden units or channels) x kSeB images per CTA, so a weight element is
// read once per kSeB images (one CTA per image re-read both matrices per image and was bound by
// one SM's L2 port: 26-31 us per launch):
//   rowdot: out[b,j] = f(sum_c W[j,c] * V[b,c])          warp = image, lanes stride c
//           FWD: V = pool_sum*scale, f = +b1 (hidden);   BWD: V = dgate*gate*(1-gate) (also written
//           out as dz2), f = * silu'(hidden[b,j]) (dh)
//   coldot: out[b,c] = f(sum_j W[j,c] * U[b,j])          thread = channel
//           FWD: U = silu(hidden), f = sigmoid(. + b2);  BWD: U = dh, f = identity (dpool)
constexpr int kSeB = 8;     // images per CTA (= warps per CTA in rowdot)
constexpr int kSeJ = 8;     // hidden units per CTA in rowdot / weight-gradient kernels

constexpr int kSeChunk = 1024;   // channels of the CTA's kSeJ weight rows staged per pass (32 KB)
template <bool BWD>
__global__ void __launch_bounds__(kThreads)
se_rowdot_kernel(const float* __restrict__ W, const float* __restrict__ v0,
                 const float* __restrict__ v1, float v_scale, const float* __restrict__ bias,
                 const float* __restrict__ hidden_in, float* __restrict__ out,
                 float* __restrict__ dz2_out, int B, int C, int Cs) {
  __shared__ __align__(16) float s_w[kSeJ][kSeChunk];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y * kSeB + warp, j0 = blockIdx.x * kSeJ;
  const bool live = b < B;
  float acc[kSeJ];
#pragma unroll
  for (int k = 0; k < kSeJ; ++k) acc[k] = 0.f;
  const int64_t vb = static_cast<int64_t>(live ? b : 0) * C;
  for (int c0 = 0; c0 < C; c0 += kSeChunk) {
    const int n = min(kSeChunk, C - c0);          // multiple of 8 (C % 8 == 0)
    __syncthreads();
    for (int i = threadIdx.x * 4; i < kSeJ * n; i += kThreads * 4) {
      const int k = i / n, c = i % n;
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (j0 + k < Cs) w = *reinterpret_cast<const float4*>(W + static_cast<int64_t>(j0 + k) * C + c0 + c);
      *reinterpret_cast<float4*>(&s_w[k][c]) = w;
    }
    __syncthreads();
    if (live) {
#pragma unroll 4
      for (int c = lane; c < n; c += 32) {
        float v;
        if (BWD) {
          const float g = v1[vb + c0 + c];
          v = v0[vb + c0 + c] * g * (1.f - g);
          if (blockIdx.x == 0) dz2_out[vb + c0 + c] = v;
        } else {
          v = v0[vb + c0 + c] * v_scale;
        }
#pragma unroll
        for (int k = 0; k < kSeJ; ++k) acc[k] = fmaf(s_w[k][c], v, acc[k]);
      }
    }
  }
  if (!live) return;
#pragma unroll
  for (int k = 0; k < kSeJ; ++k) acc[k] = warp_sum(acc[k]);
  if (lane < kSeJ && j0 + lane < Cs) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < kSeJ; ++k) t = lane == k ? acc[k] : t;
    const int64_t o = static_cast<int64_t>(b) * Cs + j0 + lane;
    if (BWD) {
      const float h = hidden_in[o];
      const float sg = sigmoidf_(h);
      out[o] = t * sg * fmaf(h, 1.f - sg, 1.f);
    } else {
      out[o] = t + bias[j0 + lane];
    }
  }
}

template <bool BWD>
__global__ void __launch_bounds__(kThreads)
se_coldot_kernel(const float* __restrict__ W, const float* __restrict__ u,
                 const float* __restrict__ bias, float* __restrict__ out, int B, int C, int Cs) {
  extern __shared__ float s_se[];                 // [kSeB][Cs]: silu(hidden) or dh of the CTA's images
  const int b0 = blockIdx.y * kSeB;
  for (int i = threadIdx.x; i < kSeB * Cs; i += kThreads) {
    const int bb = b0 + i / Cs;
    float t = 0.f;
    if (bb < B) {
      t = u[static_cast<int64_t>(bb) * Cs + i % Cs];
      if (!BWD) t = t * sigmoidf_(t);
    }
    s_se[i] = t;
  }
  __syncthreads();
  const int c = blockIdx.x * kThreads + threadIdx.x;
  if (c >= C) return;
  float acc[kSeB];
#pragma unroll
  for (int k = 0; k < kSeB; ++k) acc[k] = 0.f;
#pragma unroll 4
  for (int j = 0; j < Cs; ++j) {
    const float w = W[static_cast<int64_t>(j) * C + c];
#pragma unroll
    for (int k = 0; k < kSeB; ++k) acc[k] = fmaf(w, s_se[k * Cs + j], acc[k]);
  }
  const float bc = BWD ? 0.f : bias[c];
#pragma unroll
  for (int k = 0; k < kSeB; ++k)
    if (b0 + k < B)
      out[static_cast<int64_t>(b0 + k) * C + c] = BWD ? acc[k] : sigmoidf_(acc[k] + bc);
}

// dw1[j,c] = sum_b dh[b,j] * pooled[b,c];  dw2t[j,c] = sum_b silu(hidden[b,j]) * dz2[b,c];
// db1[j] = sum_b dh[b,j];  db2[c] = sum_b dz2[b,c].   grid = (C / 256, Cs / kSeJ); a thread owns one
// channel and kSeJ hidden units, the per-image [kSeJ] vectors sit in shared memory.
__global__ void __launch_bounds__(kThreads)
se_gate_wgrad_kernel(const float* __restrict__ pool_sum, float pool_scale,
                     const float* __restrict__ hidden, const float* __restrict__ dz2,
                     const float* __restrict__ dh, float* __restrict__ dw1, float* __restrict__ db1,
                     float* __restrict__ dw2t, float* __restrict__ db2, int B, int C, int Cs) {
  extern __shared__ float s_se[];                 // [B][kSeJ] dh, [B][kSeJ] silu(hidden)
  float* s_dh = s_se;
  float* s_a = s_se + B * kSeJ;
  const int j0 = blockIdx.y * kSeJ;
  for (int i = threadIdx.x; i < B * kSeJ; i += kThreads) {
    const int bb = i / kSeJ, j = j0 + i % kSeJ;
    float d = 0.f, a = 0.f;
    if (j < Cs) {
      d = dh[static_cast<int64_t>(bb) * Cs + j];
      const float h = hidden[static_cast<int64_t>(bb) * Cs + j];
      a = h * sigmoidf_(h);
    }
    s_dh[i] = d;
    s_a[i] = a;
  }
  __syncthreads();
  const int c = blockIdx.x * kThreads + threadIdx.x;
  if (c < C) {
    float a1[kSeJ], a2[kSeJ], sb2 = 0.f;
#pragma unroll
    for (int k = 0; k < kSeJ; ++k) a1[k] = a2[k] = 0.f;
    for (int bb = 0; bb < B; ++bb) {
      const float p = pool_sum[static_cast<int64_t>(bb) * C + c] * pool_scale;
      const float d = dz2[static_cast<int64_t>(bb) * C + c];
      sb2 += d;
#pragma unroll
      for (int k = 0; k < kSeJ; ++k) {
        a1[k] = fmaf(s_dh[bb * kSeJ + k], p, a1[k]);
        a2[k] = fmaf(s_a[bb * kSeJ + k], d, a2[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < kSeJ; ++k) {
      if (j0 + k < Cs) {
        dw1[static_cast<int64_t>(j0 + k) * C + c] = a1[k];
        dw2t[static_cast<int64_t>(j0 + k) * C + c] = a2[k];
      }
    }
    if (blockIdx.y == 0) db2[c] = sb2;
  }
  if (blockIdx.x == 0 && threadIdx.x < kSeJ && j0 + threadIdx.x < Cs) {
    float t = 0.f;
    for (int bb = 0; bb < B; ++bb) t += s_dh[bb * kSeJ + threadIdx.x];
    db1[j0 + threadIdx.x] = t;
  }
}

// ------------------------------------------------------------------ launch geometry
struct Geo {
  dim3 grid;
  int rows_per_cta, lw;
};
// images x rows-per-image x C.  lw = vector lanes per row (power of two <= 8); the row range of a
// CTA is sized for ~8 CTAs per SM and rounded to whole sweeps of the row lanes.
bool make_geo(const char* who, int images, int HW, int C, Geo* g) {
  if (images <= 0 || HW <= 0 || C <= 0 || (C % 8) != 0 || images > 65535) {
    set_last_error("%s: bad shape images=%d rows=%d C=%d (C must be a multiple of 8)", who, images,
                   HW, C);
    return false;
  }
  const int nvec = C >> 3;
  int lw = 8;
  while (lw > 1 && (lw >> 1) >= nvec) lw >>= 1;
  const int nrl = kThreads / lw;
  const int slabs = (nvec + lw - 1) / lw;
  const int64_t want = 8LL * sm_count();
  int64_t chunks = (want + static_cast<int64_t>(slabs) * images - 1) / (static_cast<int64_t>(slabs) * images);
  const int64_t max_chunks = (HW + nrl - 1) / nrl;
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  if (chunks > 65535) chunks = 65535;
  int64_t rpc = (HW + chunks - 1) / chunks;
  rpc = (rpc + nrl - 1) / nrl * nrl;
  chunks = (HW + rpc - 1) / rpc;
  g->grid = dim3(static_cast<unsigned>(slabs), static_cast<unsigned>(chunks),
                 static_cast<unsigned>(images));
  g->rows_per_cta = static_cast<int>(rpc);
  g->lw = lw;
  return true;
}
// A [M, C] matrix with no per-image output: split M into pseudo-images so that grid.z carries the
// parallelism when M is huge (rows per pseudo-image must divide M: use 1 image unless M > 2^30).
bool make_geo_flat(const char* who, int64_t M, int C, Geo* g) {
  if (M <= 0 || M > 0x7fffffffLL) {
    set_last_error("%s: bad row count M=%lld", who, (long long)M);
    return false;
  }
  return make_geo(who, 1, static_cast<int>(M), C, g);
}

int status(const char* who) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("%s launch: %s", who, cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace

int bn_stats(const void* x, float* sums, int64_t M, int C, cudaStream_t stream) {
  Geo g;
  if (!make_geo_flat("bn_stats", M, C, &g)) return CNX_ERR_BAD_SHAPE;
  if (int rc = zero_async(sums, sizeof(float) * 2 * C, stream, "bn_stats")) return rc;
  bn_stats_kernel<<<g.grid, kThreads, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), sums, C,
                                                   static_cast<int>(M), g.rows_per_cta, g.lw);
  return status("bn_stats");
}

int bn_act_fwd(const void* x, const float* sums, const float* weight, const float* bias,
               float* running_mean, float* running_var, float momentum, float eps, int act, void* y,
               float* mean_out, float* rstd_out, float* pool_sum, int B, int HW, int C,
               cudaStream_t stream) {
  if (act != 0 && act != 1) {
    set_last_error("bn_act_fwd: act must be 0 (identity) or 1 (SiLU), got %d", act);
    return CNX_ERR_BAD_ARG;
  }
  if (sums == nullptr && (running_mean == nullptr || running_var == nullptr)) {
    set_last_error("bn_act_fwd: eval mode (sums == NULL) needs running_mean and running_var");
    return CNX_ERR_BAD_ARG;
  }
  const int64_t M = static_cast<int64_t>(B) * HW;
  Geo g;
  const bool ok = pool_sum != nullptr ? make_geo("bn_act_fwd", B, HW, C, &g)
                                      : make_geo_flat("bn_act_fwd", M, C, &g);
  if (!ok) return CNX_ERR_BAD_SHAPE;
  if (pool_sum != nullptr)
    if (int rc = zero_async(pool_sum, sizeof(float) * B * C, stream, "bn_act_fwd")) return rc;
  BnArgs p;
  p.sums = sums; p.pivot = static_cast<const __nv_bfloat16*>(x); p.weight = weight; p.bias = bias;
  p.running_mean = running_mean; p.running_var = running_var;
  p.mean_out = mean_out; p.rstd_out = rstd_out;
  p.inv_m = 1.0f / static_cast<float>(M);
  p.unbias = M > 1 ? static_cast<float>(M) / static_cast<float>(M - 1) : 1.0f;
  p.momentum = momentum; p.eps = eps;
  const int hw = pool_sum != nullptr ? HW : static_cast<int>(M);
  auto* xp = static_cast<const __nv_bfloat16*>(x);
  auto* yp = static_cast<__nv_bfloat16*>(y);
  if (act == 1)
    bn_act_fwd_kernel<1><<<g.grid, kThreads, 0, stream>>>(xp, yp, p, pool_sum, C, hw, g.rows_per_cta, g.lw);
  else
    bn_act_fwd_kernel<0><<<g.grid, kThreads, 0, stream>>>(xp, yp, p, pool_sum, C, hw, g.rows_per_cta, g.lw);
  return status("bn_act_fwd");
}

int bn_act_bwd(const void* g, const void* x, const float* weight, const float* bias,
               const float* mean, const float* rstd, int act, void* dx, float* dweight,
               float* dbias, int64_t M, int C, cudaStream_t stream) {
  if (act != 0 && act != 1) {
    set_last_error("bn_act_bwd: act must be 0 (identity) or 1 (SiLU), got %d", act);
    return CNX_ERR_BAD_ARG;
  }
  Geo geo;
  if (!make_geo_flat("bn_act_bwd", M, C, &geo)) return CNX_ERR_BAD_SHAPE;
  if (int rc = zero_async(dweight, sizeof(float) * C, stream, "bn_act_bwd")) return rc;
  if (int rc = zero_async(dbias, sizeof(float) * C, stream, "bn_act_bwd")) return rc;
  BnBwdArgs p;
  p.weight = weight; p.bias = bias; p.mean = mean; p.rstd = rstd;
  p.dweight = dweight; p.dbias = dbias;
  p.inv_m = 1.0f / static_cast<float>(M);
  auto* gp = static_cast<const __nv_bfloat16*>(g);
  auto* xp = static_cast<const __nv_bfloat16*>(x);
  auto* dp = static_cast<__nv_bfloat16*>(dx);
  const int hw = static_cast<int>(M);
  if (act == 1) {
    bn_act_bwd_reduce_kernel<1><<<geo.grid, kThreads, 0, stream>>>(gp, xp, p, C, hw, geo.rows_per_cta, geo.lw);
    bn_act_bwd_apply_kernel<1><<<geo.grid, kThreads, 0, stream>>>(gp, xp, dp, p, C, hw, geo.rows_per_cta, geo.lw);
  } else {
    bn_act_bwd_reduce_kernel<0><<<geo.grid, kThreads, 0, stream>>>(gp, xp, p, C, hw, geo.rows_per_cta, geo.lw);
    bn_act_bwd_apply_kernel<0><<<geo.grid, kThreads, 0, stream>>>(gp, xp, dp, p, C, hw, geo.rows_per_cta, geo.lw);
  }
  return status("bn_act_bwd");
}

int se_scale_fwd(const void* x, const float* gate, void* y, int B, int HW, int C,
                 cudaStream_t stream) {
  Geo g;
  if (!make_geo("se_scale_fwd", B, HW, C, &g)) return CNX_ERR_BAD_SHAPE;
  se_scale_fwd_kernel<<<g.grid, kThreads, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), gate,
                                                       static_cast<__nv_bfloat16*>(y), C, HW,
                                                       g.rows_per_cta, g.lw);
  return status("se_scale_fwd");
}

int se_scale_bwd_reduce(const void* g, const void* x, float* dgate, int B, int HW, int C,
                        cudaStream_t stream) {
  Geo geo;
  if (!make_geo("se_scale_bwd_reduce", B, HW, C, &geo)) return CNX_ERR_BAD_SHAPE;
  if (int rc = zero_async(dgate, sizeof(float) * B * C, stream, "se_scale_bwd_reduce")) return rc;
  se_bwd_reduce_kernel<<<geo.grid, kThreads, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(g), static_cast<const __nv_bfloat16*>(x), dgate, C, HW,
      geo.rows_per_cta, geo.lw);
  return status("se_scale_bwd_reduce");
}

int se_scale_bwd_apply(const void* g, const float* gate, const float* dpool, void* dx, int B,
                       int HW, int C, cudaStream_t stream) {
  Geo geo;
  if (!make_geo("se_scale_bwd_apply", B, HW, C, &geo)) return CNX_ERR_BAD_SHAPE;
  se_bwd_apply_kernel<<<geo.grid, kThreads, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(g), gate, dpool, static_cast<__nv_bfloat16*>(dx), C, HW,
      geo.rows_per_cta, geo.lw, 1.0f / static_cast<float>(HW));
  return status("se_scale_bwd_apply");
}

namespace {
bool se_gate_shape_ok(const char* who, int B, int C, int Cs, size_t smem) {
  if (B <= 0 || C <= 0 || Cs <= 0 || B > 65535 || smem > 48 * 1024) {
    set_last_error("%s: bad shape B=%d C=%d Cs=%d (shared memory %zu B, limit 48 KB)", who, B, C, Cs,
                   smem);
    return false;
  }
  return true;
}
}  // namespace

int se_gate_fwd(const float* pool_sum, float pool_scale, const float* w1, const float* b1,
                const float* w2t, const float* b2, float* hidden, float* gate, int B, int C, int Cs,
                cudaStream_t stream) {
  const size_t smem = sizeof(float) * kSeB * static_cast<size_t>(Cs);
  if (!se_gate_shape_ok("se_gate_fwd", B, C, Cs, smem)) return CNX_ERR_BAD_SHAPE;
  const unsigned gb = (B + kSeB - 1) / kSeB;
  se_rowdot_kernel<false><<<dim3((Cs + kSeJ - 1) / kSeJ, gb), kThreads, 0, stream>>>(
      w1, pool_sum, nullptr, pool_scale, b1, nullptr, hidden, nullptr, B, C, Cs);
  se_coldot_kernel<false><<<dim3((C + kThreads - 1) / kThreads, gb), kThreads, smem, stream>>>(
      w2t, hidden, b2, gate, B, C, Cs);
  return status("se_gate_fwd");
}

int se_gate_bwd(const float* dgate, const float* gate, const float* hidden, const float* pool_sum,
                float pool_scale, const float* w1, const float* w2t, float* dz2, float* dh,
                float* dpool, float* dw1, float* db1, float* dw2t, float* db2, int B, int C, int Cs,
                cudaStream_t stream) {
  const size_t smem = sizeof(float) * kSeB * static_cast<size_t>(Cs);
  const size_t smem_w = sizeof(float) * 2 * static_cast<size_t>(B) * kSeJ;
  if (!se_gate_shape_ok("se_gate_bwd", B, C, Cs, smem > smem_w ? smem : smem_w)) return CNX_ERR_BAD_SHAPE;
  const unsigned gb = (B + kSeB - 1) / kSeB;
  se_rowdot_kernel<true><<<dim3((Cs + kSeJ - 1) / kSeJ, gb), kThreads, 0, stream>>>(
      w2t, dgate, gate, 1.f, nullptr, hidden, dh, dz2, B, C, Cs);
  se_coldot_kernel<true><<<dim3((C + kThreads - 1) / kThreads, gb), kThreads, smem, stream>>>(
      w1, dh, nullptr, dpool, B, C, Cs);
  dim3 grid((C + kThreads - 1) / kThreads, (Cs + kSeJ - 1) / kSeJ);
  se_gate_wgrad_kernel<<<grid, kThreads, smem_w, stream>>>(pool_sum, pool_scale, hidden, dz2, dh, dw1,
                                                           db1, dw2t, db2, B, C, Cs);
  return status("se_gate_bwd");
}

}  // namespace cnx
