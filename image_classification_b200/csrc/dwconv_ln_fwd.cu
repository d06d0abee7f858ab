// dwconv_ln_fwd.cu — fused depthwise 7x7 conv (+bias) + channels-last LayerNorm, forward.
//
// Replaces, for one ConvNeXt block, the sequence  conv_dw -> permute -> norm  of timm's
// ConvNeXtBlock.forward (un-vendored; reached from /root/reference/train_advanced_v4.py:134-135,
// restated in SURVEY.md §3.3).  The LN affine (norm.weight / norm.bias) is NOT applied here:
// it is folded into fc1's weight/bias by cnx_prep_block_weights, so this kernel emits the
// normalised activations xhat (bf16) that are both fc1's A operand and the tensor LN-backward
// needs, plus per-pixel mean / rstd (fp32).
//
// Layout / mapping (B200-first, not a translation of any library kernel):
//   * x is NHWC bf16.  A CTA owns a TH x TW pixel tile of one image for a slab of CS <= 128
//     channels; the (TH+6) x (TW+6) x CS halo tile is fetched by ONE 4-D TMA box whose
//     out-of-bounds rows/cols are zero-filled by the hardware (that IS the conv padding).
//   * channels wider than 128 are split over the CTAs of a thread-block cluster (cluster dim z =
//     C / CS <= 8).  LayerNorm's per-pixel statistics are combined across the cluster through
//     distributed shared memory with Chan's parallel mean/M2 formula (no E[x^2]-E[x]^2).
//   * a thread owns 2 adjacent channels (one bf16x2 word: a warp reads 128 contiguous bytes of
//     shared memory, conflict-free) and a 2x4 output micro-tile; the 49 taps for its two channels
//     live in registers as fp32 pairs and the inner loop is packed fma.rn.f32x2.
//
// Algorithmic bytes per launch: 2 * M * C * 2 B (read x, write xhat) + 8*M B stats.
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int kThreads = 256;
constexpr int kLanesPerPG = 64;             // threads per pixel group (2 warps = 128 channels)
constexpr int kPG = kThreads / kLanesPerPG; // pixel groups per CTA
constexpr int MH = 2, MW = 4;               // output micro-tile per thread
constexpr int kPixPerChunk = kPG * MH * MW; // 32

struct FwdParams {
  int* timeout_flag;
  __nv_bfloat16* xhat;   // [B,H,W,C]
  float* mean;           // [B*H*W]
  float* rstd;           // [B*H*W]
  const float* wt;       // [49][C] tap-major depthwise weights
  const float* bias;     // [C]
  int B, H, W, C;
  int CS;                // channels per slab (C / gridDim.z), even, <= 128
  int TH, TW;            // tile (multiples of MH, MW)
  int tiles_w;
  float eps;
};

template <bool kCluster>
__global__ void __launch_bounds__(kThreads, 1)
dwconv_ln_fwd_kernel(const __grid_constant__ CUtensorMap x_map, const FwdParams p) {
  extern __shared__ __align__(128) unsigned char smem_dyn[];
  // TMA destinations must be 128-byte aligned; do not rely on the toolchain's placement
  unsigned char* smem_raw = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);
  __shared__ __align__(8) uint64_t tma_bar;
  __shared__ float2 s_part[2][kPixPerChunk][2];  // (mean_w, M2_w) per chunk pixel per warp-in-group
  __shared__ float2 s_final[kPixPerChunk];       // (mean, rstd)

  const uint32_t* tile = reinterpret_cast<const uint32_t*>(smem_raw);

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int pg = tid / kLanesPerPG;
  const int lpg = tid % kLanesPerPG;   // lane within pixel group -> channel pair
  const int wig = lpg >> 5;            // warp within group (0/1)
  const int CS = p.CS, CS2 = CS >> 1;
  const bool ch_active = (2 * lpg) < CS;
  const int slab = blockIdx.z;
  const int c0 = slab * CS;            // first channel of this slab
  const int n = blockIdx.y;
  const int th_i = blockIdx.x / p.tiles_w, tw_i = blockIdx.x % p.tiles_w;
  const int h0 = th_i * p.TH, w0 = tw_i * p.TW;
  const int HT = p.TH + 6, WT = p.TW + 6;

  if (tid == 0) {
    mbar_init(&tma_bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    const uint32_t bytes = static_cast<uint32_t>(HT) * WT * CS * 2u;
    mbar_arrive_expect_tx(&tma_bar, bytes);
    tma_load_4d(smem_raw, &x_map, &tma_bar, c0, w0 - 3, h0 - 3, n);
  }

  // taps + bias for this thread's two channels (overlaps the TMA flight)
  float2 wreg[49];
  float2 breg = make_float2(0.f, 0.f);
  if (ch_active) {
    const float2* wp = reinterpret_cast<const float2*>(p.wt + c0 + 2 * lpg);
    const int cstride2 = p.C >> 1;
#pragma unroll
    for (int t = 0; t < 49; ++t) wreg[t] = __ldg(wp + static_cast<size_t>(t) * cstride2);
    breg = __ldg(reinterpret_cast<const float2*>(p.bias + c0 + 2 * lpg));
  } else {
#pragma unroll
    for (int t = 0; t < 49; ++t) wreg[t] = make_float2(0.f, 0.f);
  }
  // channels this warp contributes to the statistics
  const int n_w = max(0, min(64, CS - 64 * wig));
  const float inv_nw = n_w > 0 ? 1.0f / static_cast<float>(n_w) : 0.f;

  mbar_wait(&tma_bar, 0, p.timeout_flag);

  const int mt_w = p.TW / MW;
  const int nmt = (p.TH / MH) * mt_w;
  const int nchunk = (nmt + kPG - 1) / kPG;
  const uint32_t nranks = kCluster ? cluster_nctarank() : 1u;

  for (int chunk = 0; chunk < nchunk; ++chunk) {
    const int mt = chunk * kPG + pg;
    const bool mt_valid = mt < nmt;
    const int oy = mt_valid ? (mt / mt_w) * MH : 0;
    const int ox = mt_valid ? (mt % mt_w) * MW : 0;
    const int buf = chunk & 1;

    float2 acc[MH][MW];
#pragma unroll
    for (int o = 0; o < MH; ++o)
#pragma unroll
      for (int j = 0; j < MW; ++j) acc[o][j] = breg;

    if (mt_valid && ch_active) {
      const uint32_t* base = tile + (static_cast<size_t>(oy) * WT + ox) * CS2 + lpg;
#pragma unroll
      for (int r = 0; r < MH + 6; ++r) {
        float2 in[MW + 6];
#pragma unroll
        for (int j = 0; j < MW + 6; ++j)
          in[j] = bf16x2_to_f32x2(base[(static_cast<size_t>(r) * WT + j) * CS2]);
#pragma unroll
        for (int o = 0; o < MH; ++o) {
          const int kh = r - o;
          if (kh >= 0 && kh < 7) {
#pragma unroll
            for (int kw = 0; kw < 7; ++kw)
#pragma unroll
              for (int j = 0; j < MW; ++j)
                acc[o][j] = ffma2(in[j + kw], wreg[kh * 7 + kw], acc[o][j]);
          }
        }
      }
    }

    // ---- per-warp statistics (two-pass inside the warp: mean, then centred M2)
#pragma unroll
    for (int o = 0; o < MH; ++o)
#pragma unroll
      for (int j = 0; j < MW; ++j) {
        const float s = ch_active ? (acc[o][j].x + acc[o][j].y) : 0.f;
        const float mean_w = warp_sum(s) * inv_nw;
        const float dx = acc[o][j].x - mean_w, dy = acc[o][j].y - mean_w;
        const float m2 = warp_sum(ch_active ? (dx * dx + dy * dy) : 0.f);
        if (lane == 0) s_part[buf][pg * (MH * MW) + o * MW + j][wig] = make_float2(mean_w, m2);
      }

    if (kCluster) {
      cluster_arrive();
      cluster_wait();
    } else {
      __syncthreads();
    }

    // ---- combine the per-warp partials of all slabs (Chan): threads 0..31, one chunk pixel each
    if (tid < kPixPerChunk) {
      float tot_mean = 0.f;
      const float inv_c = 1.0f / static_cast<float>(p.C);
      float m2sum = 0.f;
      auto load_part = [&](uint32_t r, int w) -> float2 {
        if (kCluster) return dsmem_ld_f32x2(dsmem_map(smem_u32(&s_part[buf][tid][w]), r));
        return s_part[buf][tid][w];
      };
      for (uint32_t r = 0; r < nranks; ++r) {
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          const int ng = max(0, min(64, CS - 64 * w));
          const float2 v = load_part(r, w);
          tot_mean += v.x * static_cast<float>(ng);
          m2sum += (ng > 0) ? v.y : 0.f;
        }
      }
      tot_mean *= inv_c;
      for (uint32_t r = 0; r < nranks; ++r) {
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          const int ng = max(0, min(64, CS - 64 * w));
          const float d = load_part(r, w).x - tot_mean;
          m2sum += static_cast<float>(ng) * d * d;
        }
      }
      const float var = m2sum * inv_c;
      const float rstd = rsqrtf(var + p.eps);
      s_final[tid] = make_float2(tot_mean, rstd);
      if (slab == 0) {
        const int pgi = tid / (MH * MW), pi = tid % (MH * MW);
        const int mt2 = chunk * kPG + pgi;
        if (mt2 < nmt) {
          const int hh = h0 + (mt2 / mt_w) * MH + pi / MW;
          const int ww = w0 + (mt2 % mt_w) * MW + pi % MW;
          if (hh < p.H && ww < p.W) {
            const size_t m = (static_cast<size_t>(n) * p.H + hh) * p.W + ww;
            p.mean[m] = tot_mean;
            p.rstd[m] = rstd;
          }
        }
      }
    }
    __syncthreads();

    // ---- normalise + store (bf16x2 per thread; a warp writes 128 contiguous bytes per pixel)
    if (mt_valid && ch_active) {
#pragma unroll
      for (int o = 0; o < MH; ++o)
#pragma unroll
        for (int j = 0; j < MW; ++j) {
          const int hh = h0 + oy + o, ww = w0 + ox + j;
          if (hh < p.H && ww < p.W) {
            const float2 st = s_final[pg * (MH * MW) + o * MW + j];
            const float a = (acc[o][j].x - st.x) * st.y;
            const float b = (acc[o][j].y - st.x) * st.y;
            const size_t m = (static_cast<size_t>(n) * p.H + hh) * p.W + ww;
            uint32_t* dst = reinterpret_cast<uint32_t*>(p.xhat + m * p.C + c0) + lpg;
            *dst = f32x2_to_bf16x2(a, b);
          }
        }
    }
  }

  if (kCluster) {  // keep this CTA's shared memory alive until every peer has read it
    cluster_arrive();
    cluster_wait();
  }
}

// tile heuristic: multiples of the micro-tile, halo tile <= ~100 KB so two CTAs fit per SM
void pick_tile(int H, int W, int CS, int* TH, int* TW) {
  auto round_up = [](int v, int m) { return (v + m - 1) / m * m; };
  int tw = round_up(W, MW);
  if (tw > 16) {  // split wide rows into near-equal multiples of MW, <= 16 wide
    const int parts = (W + 15) / 16;
    tw = round_up((W + parts - 1) / parts, MW);
  }
  int th = round_up(H, MH);
  const int budget = 100 * 1024;
  while (th > MH && (th + 6) * (tw + 6) * CS * 2 > budget) {
    const int parts = (H + th - 1) / th + 1;
    th = round_up((H + parts - 1) / parts, MH);
  }
  *TH = th;
  *TW = tw;
}

}  // namespace

int dwconv_ln_fwd(const void* x, const float* wt, const float* bias, void* xhat, float* mean,
                  float* rstd, int B, int H, int W, int C, float eps, cudaStream_t stream) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (C % 8) != 0) {
    set_last_error("dwconv_ln_fwd: bad shape B=%d H=%d W=%d C=%d (C must be a multiple of 8)", B,
                   H, W, C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int nslab = (C + 127) / 128;
  if (nslab > 8 || (C % nslab) != 0 || ((C / nslab) % 8) != 0) {
    set_last_error("dwconv_ln_fwd: C=%d cannot be split into <=8 equal slabs of <=128 channels", C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int CS = C / nslab;
  int TH, TW;
  pick_tile(H, W, CS, &TH, &TW);

  CUtensorMap map;
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  const uint32_t box[4] = {(uint32_t)CS, (uint32_t)(TW + 6), (uint32_t)(TH + 6), 1u};
  int rc = make_tensor_map(&map, x, 2, 4, dims, strides, box, 0);
  if (rc != CNX_OK) return rc;

  FwdParams p;
  p.timeout_flag = timeout_flag_ptr();
  if (p.timeout_flag == nullptr) {
    set_last_error("could not map the timeout flag (cudaHostAlloc failed)");
    return CNX_ERR_DRIVER;
  }
  p.xhat = static_cast<__nv_bfloat16*>(xhat);
  p.mean = mean;
  p.rstd = rstd;
  p.wt = wt;
  p.bias = bias;
  p.B = B; p.H = H; p.W = W; p.C = C;
  p.CS = CS; p.TH = TH; p.TW = TW;
  p.tiles_w = (W + TW - 1) / TW;
  p.eps = eps;
  const int tiles_h = (H + TH - 1) / TH;
  const size_t smem = static_cast<size_t>(TH + 6) * (TW + 6) * CS * 2 + 128;

  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(p.tiles_w * tiles_h, B, nslab);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = nslab;
  cfg.attrs = attr;
  cfg.numAttrs = 1;

  cudaError_t e;
  if (nslab > 1) {
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(dwconv_ln_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           200 * 1024);
      attr_set = true;
    }
    e = cudaLaunchKernelEx(&cfg, dwconv_ln_fwd_kernel<true>, map, p);
  } else {
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(dwconv_ln_fwd_kernel<false>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      attr_set = true;
    }
    e = cudaLaunchKernelEx(&cfg, dwconv_ln_fwd_kernel<false>, map, p);
  }
  if (e != cudaSuccess) {
    set_last_error("dwconv_ln_fwd launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace cnx
