// elementwise.cu — the three single-kernel glue ops of the training hot path:
//   * mix_batch       batch-level MixUp/CutMix driven by a host plan
//                     (restates MixUpCutMixDataset.mixup/cutmix, /root/reference/dataset.py:81-155)
//   * softce_fwd_bwd  label-smoothed soft-target cross-entropy, loss + dlogits in one pass
//                     (nn.CrossEntropyLoss(label_smoothing) on prob targets,
//                      /root/reference/train_advanced_v4.py:626,221)
//   * ema_lerp_multi  multi-tensor EMA update  s <- (1-d)*p + d*s
//                     (EMA.update, /root/reference/train_advanced_v4.py:75-80)
// All three are HBM-bound: 16-byte vector accesses, no shared-memory staging (no reuse).
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

// ------------------------------------------------------------------ mix_batch
// grid: (ceil(CH*H*W/4 / 256), B).  One float4 (4 consecutive pixels of a row) per thread.
// mixup pixels use separate rounded multiplies and a rounded add so the result is bit-identical
// to torch's  data1 * lam + data2 * (1 - lam)  on float32 tensors.
__global__ void __launch_bounds__(256)
mix_batch_kernel(const float* __restrict__ x, const int64_t* __restrict__ labels,
                 const MixPlanEntry* __restrict__ plan, float* __restrict__ out,
                 float* __restrict__ targets, int CH, int H, int W, int NC, float on_value,
                 float off_value) {
  const int b = blockIdx.y;
  const MixPlanEntry e = plan[b];
  const size_t per = static_cast<size_t>(CH) * H * W;
  const float* xa = x + b * per;
  const float* xb = x + static_cast<size_t>(e.partner) * per;
  float* o = out + b * per;

  // soft target row (first NC threads of block x == 0)
  if (blockIdx.x == 0) {
    const int64_t la = labels[b];
    const int64_t lb = labels[e.partner];
    for (int c = threadIdx.x; c < NC; c += blockDim.x) {
      // (smoothed) one-hot rows blended like the pixels: y1 * lam + y2 * (1 - lam); with
      // on = 1, off = 0 the products are exact, i.e. the reference's  oh1*lam + oh2*(1-lam)
      const float ya = (c == la) ? on_value : off_value;
      const float yb = (c == lb) ? on_value : off_value;
      float t;
      if (e.mode == 0) {
        t = ya;
      } else {
        t = __fadd_rn(__fmul_rn(ya, e.lam), __fmul_rn(yb, e.one_minus_lam));
      }
      targets[static_cast<size_t>(b) * NC + c] = t;
    }
  }

  const size_t nvec = per >> 2;  // W % 4 == 0 is checked by the launcher
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= nvec) return;
  const size_t el = i << 2;
  const float4 va = __ldg(reinterpret_cast<const float4*>(xa) + i);
  float4 r = va;
  if (e.mode == 1) {
    const float4 vb = __ldg(reinterpret_cast<const float4*>(xb) + i);
    r.x = __fadd_rn(__fmul_rn(va.x, e.lam), __fmul_rn(vb.x, e.one_minus_lam));
    r.y = __fadd_rn(__fmul_rn(va.y, e.lam), __fmul_rn(vb.y, e.one_minus_lam));
    r.z = __fadd_rn(__fmul_rn(va.z, e.lam), __fmul_rn(vb.z, e.one_minus_lam));
    r.w = __fadd_rn(__fmul_rn(va.w, e.lam), __fmul_rn(vb.w, e.one_minus_lam));
  } else if (e.mode == 2) {
    const int xw = static_cast<int>(el % W);
    const int yh = static_cast<int>((el / W) % H);
    if (yh >= e.bby1 && yh < e.bby2 && xw + 3 >= e.bbx1 && xw < e.bbx2) {
      const float4 vb = __ldg(reinterpret_cast<const float4*>(xb) + i);
      if (xw + 0 >= e.bbx1 && xw + 0 < e.bbx2) r.x = vb.x;
      if (xw + 1 >= e.bbx1 && xw + 1 < e.bbx2) r.y = vb.y;
      if (xw + 2 >= e.bbx1 && xw + 2 < e.bbx2) r.z = vb.z;
      if (xw + 3 >= e.bbx1 && xw + 3 < e.bbx2) r.w = vb.w;
    }
  }
  reinterpret_cast<float4*>(o)[i] = r;
}

// ------------------------------------------------------------- softce_fwd_bwd
// One CTA, one warp per row (rows strided over the 32 warps); per-row losses are reduced in a
// fixed order so the result is deterministic.  B*NC is a few thousand elements: launch-latency
// bound by construction, so the only goal is ONE launch and no host sync.
constexpr int kCeWarps = 32;
__global__ void __launch_bounds__(kCeWarps * 32)
softce_kernel(const float* __restrict__ z, const float* __restrict__ t, float* __restrict__ loss,
              float* __restrict__ dz, int B, int NC, float eps) {
  __shared__ float s_loss[kCeWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float unif = eps / static_cast<float>(NC);
  const float invB = 1.0f / static_cast<float>(B);
  float wloss = 0.f;
  for (int row = warp; row < B; row += kCeWarps) {
    const float* zr = z + static_cast<size_t>(row) * NC;
    const float* tr = t + static_cast<size_t>(row) * NC;
    float mx = -INFINITY;
    for (int c = lane; c < NC; c += 32) mx = fmaxf(mx, zr[c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float se = 0.f, st = 0.f, stz = 0.f;
    for (int c = lane; c < NC; c += 32) {
      const float zc = zr[c] - mx;
      const float tc = (1.0f - eps) * tr[c] + unif;
      se += expf(zc);
      st += tc;
      stz += tc * zc;
    }
    se = warp_sum(se);
    st = warp_sum(st);
    stz = warp_sum(stz);
    const float lse = logf(se);
    // -sum_c tc * (zc - lse) = lse * sum(tc) - sum(tc * zc)
    wloss += lse * st - stz;
    if (dz != nullptr) {
      const float inv_se = 1.0f / se;
      float* dr = dz + static_cast<size_t>(row) * NC;
      for (int c = lane; c < NC; c += 32) {
        const float tc = (1.0f - eps) * tr[c] + unif;
        dr[c] = (expf(zr[c] - mx) * inv_se * st - tc) * invB;
      }
    }
  }
  if (lane == 0) s_loss[warp] = wloss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float acc = 0.f;
    for (int w = 0; w < kCeWarps; ++w) acc += s_loss[w];
    *loss = acc * invB;
  }
}

// ------------------------------------------------------------- ema_lerp_multi
// One CTA per 8192-element chunk; chunk_prefix[t] is the first chunk of tensor t, so a CTA finds
// its tensor with a binary search over <= a few hundred entries (L2-resident table).
__global__ void __launch_bounds__(256)
ema_lerp_kernel(const void* const* __restrict__ pptr, void* const* __restrict__ sptr,
                const int64_t* __restrict__ numels, const int64_t* __restrict__ prefix, int T,
                float omd, float d) {
  const int64_t chunk = blockIdx.x;
  int lo = 0, hi = T;  // find t with prefix[t] <= chunk < prefix[t+1]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (prefix[mid] <= chunk) lo = mid; else hi = mid;
  }
  const int t = lo;
  const float* __restrict__ p = static_cast<const float*>(pptr[t]);
  float* __restrict__ s = static_cast<float*>(sptr[t]);
  const int64_t n = numels[t];
  const int64_t start = (chunk - prefix[t]) * kEmaChunkElems;
  const int64_t end = min(start + static_cast<int64_t>(kEmaChunkElems), n);
  const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(s)) & 15) == 0;
  if (aligned) {
    const int64_t vend = start + ((end - start) & ~static_cast<int64_t>(3));
    for (int64_t i = start + threadIdx.x * 4; i < vend; i += 256 * 4) {
      const float4 pv = __ldg(reinterpret_cast<const float4*>(p + i));
      float4 sv = *reinterpret_cast<const float4*>(s + i);
      sv.x = __fadd_rn(__fmul_rn(omd, pv.x), __fmul_rn(d, sv.x));
      sv.y = __fadd_rn(__fmul_rn(omd, pv.y), __fmul_rn(d, sv.y));
      sv.z = __fadd_rn(__fmul_rn(omd, pv.z), __fmul_rn(d, sv.z));
      sv.w = __fadd_rn(__fmul_rn(omd, pv.w), __fmul_rn(d, sv.w));
      *reinterpret_cast<float4*>(s + i) = sv;
    }
    for (int64_t i = vend + threadIdx.x; i < end; i += 256)
      s[i] = __fadd_rn(__fmul_rn(omd, p[i]), __fmul_rn(d, s[i]));
  } else {
    for (int64_t i = start + threadIdx.x; i < end; i += 256)
      s[i] = __fadd_rn(__fmul_rn(omd, p[i]), __fmul_rn(d, s[i]));
  }
}

}  // namespace

int mix_batch(const float* x, const int64_t* labels, const MixPlanEntry* plan, float* out,
              float* targets, int B, int CH, int H, int W, int num_classes, float on_value,
              float off_value, cudaStream_t stream) {
  if (B <= 0 || CH <= 0 || H <= 0 || W <= 0 || num_classes <= 0 || (W % 4) != 0) {
    set_last_error("mix_batch: bad shape B=%d CH=%d H=%d W=%d NC=%d (W must be a multiple of 4)", B,
                   CH, H, W, num_classes);
    return CNX_ERR_BAD_SHAPE;
  }
  if (x == out) {
    set_last_error("mix_batch: out must not alias x (partners read un-mixed pixels)");
    return CNX_ERR_BAD_ARG;
  }
  const size_t nvec = static_cast<size_t>(CH) * H * W / 4;
  dim3 grid(static_cast<unsigned>((nvec + 255) / 256), B);
  mix_batch_kernel<<<grid, 256, 0, stream>>>(x, labels, plan, out, targets, CH, H, W, num_classes,
                                             on_value, off_value);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("mix_batch launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int softce_fwd_bwd(const float* logits, const float* targets, float* loss, float* dlogits, int B,
                   int NC, float label_smoothing, cudaStream_t stream) {
  if (B <= 0 || NC <= 0) {
    set_last_error("softce_fwd_bwd: bad shape B=%d NC=%d", B, NC);
    return CNX_ERR_BAD_SHAPE;
  }
  softce_kernel<<<1, kCeWarps * 32, 0, stream>>>(logits, targets, loss, dlogits, B, NC,
                                                 label_smoothing);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("softce_fwd_bwd launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

int ema_lerp_multi(const void* const* param_ptrs, void* const* shadow_ptrs, const int64_t* numels,
                   const int64_t* chunk_prefix, int num_tensors, int64_t total_chunks,
                   float one_minus_decay, float decay, cudaStream_t stream) {
  if (num_tensors <= 0 || total_chunks <= 0 || total_chunks > 0x7fffffffLL) {
    set_last_error("ema_lerp_multi: bad table T=%d chunks=%lld", num_tensors,
                   static_cast<long long>(total_chunks));
    return CNX_ERR_BAD_ARG;
  }
  ema_lerp_kernel<<<static_cast<unsigned>(total_chunks), 256, 0, stream>>>(
      param_ptrs, shadow_ptrs, numels, chunk_prefix, num_tensors, one_minus_decay, decay);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("ema_lerp_multi launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace cnx
