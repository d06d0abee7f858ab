// optim.cu — the optimizer step of the reference loop as two multi-tensor launches.
//
// Reference (/root/reference/train_advanced_v4.py:226-237): GradScaler.unscale_ (+ inf check),
// clip_grad_norm_(1.0), AdamW.step (skipped when a gradient is non-finite), zero_grad,
// EMA.update — in eager PyTorch about 40 multi-tensor launches and 60 B of HBM traffic per
// parameter.  Here:
//   launch 1  grad_norm_kernel   4 B/param : sum of squares of the UNSCALED gradients, non-finite
//                                           flag; the last CTA folds the per-chunk partials in a
//                                           fixed order (deterministic) and publishes the clip
//                                           coefficient, the skip flag and the step count
//   launch 2  adamw_ema_kernel  36 B/param : unscale * clip, decoupled weight decay, moments,
//                                           bias-corrected update (torch's fused AdamW
//                                           arithmetic), EMA lerp of the NEW parameter
// Tensors are addressed through device pointer tables cut into 8192-element chunks, like
// ema_lerp_multi (elementwise.cu).
#include <mutex>

#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int kOptThreads = 256;

struct ChunkRef { int t; int64_t start, end; };

__device__ __forceinline__ ChunkRef find_chunk(const int64_t* __restrict__ numels,
                                               const int64_t* __restrict__ prefix, int T,
                                               int64_t chunk) {
  int lo = 0, hi = T;          // prefix[lo] <= chunk < prefix[lo + 1]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (prefix[mid] <= chunk) lo = mid; else hi = mid;
  }
  ChunkRef r;
  r.t = lo;
  r.start = (chunk - prefix[lo]) * kEmaChunkElems;
  r.end = min(r.start + static_cast<int64_t>(kEmaChunkElems), numels[lo]);
  return r;
}

__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float s = 0.f;
  if (threadIdx.x < kOptThreads / 32) s = red[threadIdx.x];
  if (warp == 0) {
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  }
  return s;          // valid in thread 0
}

// state (device float[8], see cnx_b200.h): [0] total_norm  [1] grad multiplier  [2] found_inf
// [3] step count (persistent)  [4] CTA ticket (uint32)  [5] clip coefficient
__global__ void __launch_bounds__(kOptThreads)
grad_norm_kernel(const void* const* __restrict__ gptr, const int64_t* __restrict__ numels,
                 const int64_t* __restrict__ prefix, int T, const float* __restrict__ scale,
                 float max_norm, float* __restrict__ partials, float* __restrict__ state) {
  __shared__ float red[kOptThreads / 32];
  __shared__ bool is_last;
  const ChunkRef c = find_chunk(numels, prefix, T, blockIdx.x);
  const float* __restrict__ g = static_cast<const float*>(gptr[c.t]);
  const float inv_scale = scale ? 1.0f / __ldg(scale) : 1.0f;
  float ss = 0.f;
  bool bad = false;
  if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
    const int64_t vend = c.start + ((c.end - c.start) & ~static_cast<int64_t>(3));
    for (int64_t i = c.start + threadIdx.x * 4; i < vend; i += kOptThreads * 4) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(g + i));
      const float a = v.x * inv_scale, b = v.y * inv_scale, cc = v.z * inv_scale, d = v.w * inv_scale;
      bad |= !(isfinite(a) && isfinite(b) && isfinite(cc) && isfinite(d));
      ss += a * a + b * b + cc * cc + d * d;
    }
    for (int64_t i = vend + threadIdx.x; i < c.end; i += kOptThreads) {
      const float a = g[i] * inv_scale;
      bad |= !isfinite(a);
      ss += a * a;
    }
  } else {
    for (int64_t i = c.start + threadIdx.x; i < c.end; i += kOptThreads) {
      const float a = g[i] * inv_scale;
      bad |= !isfinite(a);
      ss += a * a;
    }
  }
  const float tot = block_sum(ss, red);
  const int any_bad = __syncthreads_or(bad ? 1 : 0);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(state + 4);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = any_bad ? __int_as_float(0x7f800000) : tot;   // +inf marks the chunk
    __threadfence();
    is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  // ---- last CTA: fixed-order fold of the partials
  __threadfence();
  float s = 0.f;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += kOptThreads) s += __ldcg(partials + i);
  const float total = block_sum(s, red);
  if (threadIdx.x == 0) {
    const bool found_inf = !isfinite(total);
    const float norm = sqrtf(total);
    float coef = 1.0f;
    if (max_norm > 0.f) coef = fminf(max_norm / (norm + 1e-6f), 1.0f);   // torch clip_grad_norm_
    state[0] = norm;
    state[5] = coef;
    state[1] = inv_scale * coef;
    state[2] = found_inf ? 1.0f : 0.0f;
    if (!found_inf) state[3] += 1.0f;      // GradScaler.step skips the optimizer otherwise
    *ticket = 0u;
  }
}

struct AdamWArgs {
  float lr, beta1, beta2, eps, weight_decay;
  float ema_omd, ema_d;
};

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, const AdamWArgs& a,
                                          float mult, float step_size, float bc2_sqrt) {
  g *= mult;
  p -= a.lr * a.weight_decay * p;                    // decoupled decay
  m = m + (1.0f - a.beta1) * (g - m);                // lerp(m, g, 1 - beta1)
  v = a.beta2 * v + (1.0f - a.beta2) * g * g;
  const float denom = sqrtf(v) / bc2_sqrt + a.eps;
  p -= step_size * m / denom;
}

template <bool EMA>
__global__ void __launch_bounds__(kOptThreads)
adamw_ema_kernel(void* const* __restrict__ pptr, const void* const* __restrict__ gptr,
                 void* const* __restrict__ mptr, void* const* __restrict__ vptr,
                 void* const* __restrict__ sptr, const int64_t* __restrict__ numels,
                 const int64_t* __restrict__ prefix, int T, const AdamWArgs a,
                 const float* __restrict__ state) {
  const ChunkRef c = find_chunk(numels, prefix, T, blockIdx.x);
  float* __restrict__ p = static_cast<float*>(pptr[c.t]);
  const float* __restrict__ g = static_cast<const float*>(gptr[c.t]);
  float* __restrict__ m = static_cast<float*>(mptr[c.t]);
  float* __restrict__ v = static_cast<float*>(vptr[c.t]);
  float* __restrict__ s = EMA ? static_cast<float*>(sptr[c.t]) : nullptr;
  const float mult = state[1];
  const bool skip = state[2] != 0.0f;
  const double step = static_cast<double>(state[3]);
  // bias corrections in double, like torch's fused kernel (pow on double betas)
  const float bc1 = static_cast<float>(1.0 - pow(static_cast<double>(a.beta1), step));
  const float bc2 = static_cast<float>(1.0 - pow(static_cast<double>(a.beta2), step));
  const float step_size = a.lr / bc1;
  const float bc2_sqrt = sqrtf(bc2);
  uintptr_t al = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                 reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v);
  if (EMA) al |= reinterpret_cast<uintptr_t>(s);
  const bool aligned = (al & 15) == 0;
  const int64_t vend = aligned ? c.start + ((c.end - c.start) & ~static_cast<int64_t>(3)) : c.start;
  for (int64_t i = c.start + threadIdx.x * 4; i < vend; i += kOptThreads * 4) {
    float4 pv = *reinterpret_cast<const float4*>(p + i);
    if (!skip) {
      const float4 gv = __ldg(reinterpret_cast<const float4*>(g + i));
      float4 mv = *reinterpret_cast<const float4*>(m + i);
      float4 vv = *reinterpret_cast<const float4*>(v + i);
      adamw_one(pv.x, gv.x, mv.x, vv.x, a, mult, step_size, bc2_sqrt);
      adamw_one(pv.y, gv.y, mv.y, vv.y, a, mult, step_size, bc2_sqrt);
      adamw_one(pv.z, gv.z, mv.z, vv.z, a, mult, step_size, bc2_sqrt);
      adamw_one(pv.w, gv.w, mv.w, vv.w, a, mult, step_size, bc2_sqrt);
      *reinterpret_cast<float4*>(p + i) = pv;
      *reinterpret_cast<float4*>(m + i) = mv;
      *reinterpret_cast<float4*>(v + i) = vv;
    }
    if (EMA) {
      float4 sv = *reinterpret_cast<const float4*>(s + i);
      sv.x = __fadd_rn(__fmul_rn(a.ema_omd, pv.x), __fmul_rn(a.ema_d, sv.x));
      sv.y = __fadd_rn(__fmul_rn(a.ema_omd, pv.y), __fmul_rn(a.ema_d, sv.y));
      sv.z = __fadd_rn(__fmul_rn(a.ema_omd, pv.z), __fmul_rn(a.ema_d, sv.z));
      sv.w = __fadd_rn(__fmul_rn(a.ema_omd, pv.w), __fmul_rn(a.ema_d, sv.w));
      *reinterpret_cast<float4*>(s + i) = sv;
    }
  }
  for (int64_t i = vend + threadIdx.x; i < c.end; i += kOptThreads) {
    float pv = p[i];
    if (!skip) {
      float mv = m[i], vv = v[i];
      adamw_one(pv, g[i], mv, vv, a, mult, step_size, bc2_sqrt);
      p[i] = pv; m[i] = mv; v[i] = vv;
    }
    if (EMA) s[i] = __fadd_rn(__fmul_rn(a.ema_omd, pv), __fmul_rn(a.ema_d, s[i]));
  }
}

int check_table(const char* who, int T, int64_t total_chunks) {
  if (T <= 0 || total_chunks <= 0 || total_chunks > 0x7fffffffLL) {
    set_last_error("%s: bad table T=%d chunks=%lld", who, T, static_cast<long long>(total_chunks));
    return CNX_ERR_BAD_ARG;
  }
  return CNX_OK;
}

}  // namespace

int optim_grad_norm(const void* const* grad_ptrs, const int64_t* numels,
                    const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                    const float* scale, float max_norm, float* partials, float* state,
                    cudaStream_t stream) {
  int rc = check_table("optim_grad_norm", num_tensors, total_chunks);
  if (rc != CNX_OK) return rc;
  if (partials == nullptr || state == nullptr) {
    set_last_error("optim_grad_norm: partials (float[total_chunks]) and state (float[8]) required");
    return CNX_ERR_BAD_ARG;
  }
  grad_norm_kernel<<<static_cast<unsigned>(total_chunks), kOptThreads, 0, stream>>>(
      grad_ptrs, numels, chunk_prefix, num_tensors, scale, max_norm, partials, state);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("optim_grad_norm launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

// ---------------------------------------------------------------- allocation-free variant
// The per-chunk partial sums of grad_norm_kernel in library-owned memory instead of a caller
// buffer.  kStaticNormChunks floats (256 KiB) are held by every device that loads the library, from
// module load until exit.  grad_norm_kernel writes partials[i] for every i < total_chunks before
// the last CTA reads any of them, so nothing in the array has to start at zero and a call does not
// depend on what the previous call left there; two calls that overlap on the GPU (different
// streams) race on it.
static __device__ __align__(256) float g_static_norm_partials[kStaticNormChunks];

namespace {
std::mutex g_static_mu;
float* g_static_partials_addr[64] = {};      // per device slot: DEVICE address of the array above
}  // namespace

int optim_grad_norm_static(const void* const* grad_ptrs, const int64_t* numels,
                           const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                           const float* scale, float max_norm, float* state, cudaStream_t stream) {
  int rc = check_table("optim_grad_norm_static", num_tensors, total_chunks);
  if (rc != CNX_OK) return rc;
  if (total_chunks > kStaticNormChunks) {
    set_last_error("optim_grad_norm_static: %lld chunks of %d elements exceed the static scratch "
                   "(max %d chunks = %lld elements incl. per-tensor round-up); use "
                   "cnx_optim_grad_norm with a caller buffer",
                   static_cast<long long>(total_chunks), kEmaChunkElems, kStaticNormChunks,
                   static_cast<long long>(kStaticNormChunks) * kEmaChunkElems);
    return CNX_ERR_BAD_ARG;
  }
  if (state == nullptr) {
    set_last_error("optim_grad_norm_static: state (float[8]) required");
    return CNX_ERR_BAD_ARG;
  }
  // In host code `g_static_norm_partials` is a host-side stand-in, not a device address: resolve it
  // once per device, with that device current (it is: the caller's pointers live on it).
  const int slot = current_device_slot();
  float* partials = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mu);
    partials = g_static_partials_addr[slot];
    if (partials == nullptr) {
      void* addr = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&addr, g_static_norm_partials);
      if (e != cudaSuccess || addr == nullptr) {
        set_last_error("optim_grad_norm_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        return CNX_ERR_LAUNCH;
      }
      partials = g_static_partials_addr[slot] = static_cast<float*>(addr);
    }
  }
  grad_norm_kernel<<<static_cast<unsigned>(total_chunks), kOptThreads, 0, stream>>>(
      grad_ptrs, numels, chunk_prefix, num_tensors, scale, max_norm, partials, state);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("optim_grad_norm_static launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int optim_adamw_ema(void* const* param_ptrs, const void* const* grad_ptrs, void* const* m_ptrs,
                    void* const* v_ptrs, void* const* shadow_ptrs, const int64_t* numels,
                    const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks, float lr,
                    float beta1, float beta2, float eps, float weight_decay, float ema_one_minus_decay,
                    float ema_decay, const float* state, cudaStream_t stream) {
  int rc = check_table("optim_adamw_ema", num_tensors, total_chunks);
  if (rc != CNX_OK) return rc;
  if (state == nullptr || param_ptrs == nullptr || grad_ptrs == nullptr || m_ptrs == nullptr ||
      v_ptrs == nullptr) {
    set_last_error("optim_adamw_ema: null table or state");
    return CNX_ERR_BAD_ARG;
  }
  AdamWArgs a{lr, beta1, beta2, eps, weight_decay, ema_one_minus_decay, ema_decay};
  const unsigned grid = static_cast<unsigned>(total_chunks);
  if (shadow_ptrs != nullptr)
    adamw_ema_kernel<true><<<grid, kOptThreads, 0, stream>>>(param_ptrs, grad_ptrs, m_ptrs, v_ptrs,
                                                             shadow_ptrs, numels, chunk_prefix,
                                                             num_tensors, a, state);
  else
    adamw_ema_kernel<false><<<grid, kOptThreads, 0, stream>>>(param_ptrs, grad_ptrs, m_ptrs, v_ptrs,
                                                              nullptr, numels, chunk_prefix,
                                                              num_tensors, a, state);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("optim_adamw_ema launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace cnx
