// dwconv_ln_bwd.cu — fused LayerNorm-backward + depthwise 7x7 conv backward (data + filter + bias)
// + shortcut-gradient add, for one ConvNeXt block.
//
// Autograd backward of  conv_dw -> permute -> norm  plus the residual edge of timm's
// ConvNeXtBlock (driven by scaler.scale(loss).backward(),
// /root/reference/train_advanced_v4.py:223).  With the LN affine folded into fc1 (prep.cu) the
// incoming gradient is dxhat = dL/d(xhat) straight out of the fc1 data-gradient GEMM, so
//   dconv = rstd * (dxhat - mean_c(dxhat) - xhat * mean_c(dxhat * xhat))            (LN backward)
//   dx    = dy + sum_taps dconv[p - off] * w[tap]                                   (bwd-data + shortcut)
//   dw[tap] = sum_p x[p] * dconv[p - off],  dbias = sum_p dconv[p],  dy_colsum = sum_p dy[p]
//
// One CTA owns a TH x TW pixel tile x a slab of <=128 channels; slabs of a wider C form a
// thread-block cluster and exchange the two per-pixel LN-backward sums through DSMEM.  dxhat and
// xhat halo tiles arrive by TMA (zero-filled outside the image, which makes dconv vanish there);
// dconv is materialised in place (bf16) over the dxhat tile, then a thread walks 2x4 output
// micro-tiles for its two channels with packed fp32 FMAs, keeping the 49 filter-gradient
// accumulators in registers and reading the taps from shared memory.
//
// Algorithmic bytes per launch: read dxhat, xhat, x, dy + write dx = 5 * M * C * 2 B
// (SURVEY.md §8d counts 4*E*s for the save-and-re-read variant; the fifth pass is the shortcut
// gradient dy that eager PyTorch adds in a separate kernel).
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int kThreads = 256;
constexpr int kLanesPerPG = 64;
constexpr int kPG = kThreads / kLanesPerPG;
constexpr int MH = 2, MW = 4;
constexpr int kAccRows = 51;  // 49 taps + dbias + dy_colsum

struct BwdParams {
  int* timeout_flag;
  const float* rstd;           // [B*H*W]
  const __nv_bfloat16* x;      // [B,H,W,C] block input
  const __nv_bfloat16* dy;     // [B,H,W,C] gradient of block output
  const float* wt;             // [49][C]
  __nv_bfloat16* dx;           // [B,H,W,C]
  float* dwt;                  // [49][C]  (+=)
  float* dbias;                // [C]      (+=)
  float* dy_colsum;            // [C]      (+=)
  int B, H, W, C, CS, TH, TW, tiles_w;
  uint32_t tile_bytes;         // bytes of one halo tile (multiple of 128)
};

template <bool kCluster>
__global__ void __launch_bounds__(kThreads, 1)
dwconv_ln_bwd_kernel(const __grid_constant__ CUtensorMap dxhat_map,
                     const __grid_constant__ CUtensorMap xhat_map, const BwdParams p) {
  extern __shared__ __align__(128) unsigned char smem_dyn[];
  unsigned char* smem_raw = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);
  __shared__ __align__(8) uint64_t tma_bar;
  __shared__ float2 s_part[448];   // per halo pixel (sum dxhat, sum dxhat*xhat) of this slab
  __shared__ float4 s_coef[448];   // per halo pixel (rstd, rstd*m1, rstd*m2, -)

  uint32_t* tile_d = reinterpret_cast<uint32_t*>(smem_raw);                 // dxhat -> dconv
  uint32_t* tile_x = reinterpret_cast<uint32_t*>(smem_raw + p.tile_bytes);  // xhat -> taps

  const int tid = threadIdx.x;
  const int pg = tid / kLanesPerPG;
  const int lpg = tid % kLanesPerPG;
  const int CS = p.CS, CS2 = CS >> 1;
  const bool ch_active = (2 * lpg) < CS;
  const int slab = blockIdx.z;
  const int c0 = slab * CS;
  const int n = blockIdx.y;
  const int th_i = blockIdx.x / p.tiles_w, tw_i = blockIdx.x % p.tiles_w;
  const int h0 = th_i * p.TH, w0 = tw_i * p.TW;
  const int HT = p.TH + 6, WT = p.TW + 6;
  const int HP = HT * WT;

  if (tid == 0) {
    mbar_init(&tma_bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    const uint32_t bytes = static_cast<uint32_t>(HP) * CS * 2u;
    mbar_arrive_expect_tx(&tma_bar, 2 * bytes);
    tma_load_4d(tile_d, &dxhat_map, &tma_bar, c0, w0 - 3, h0 - 3, n);
    tma_load_4d(tile_x, &xhat_map, &tma_bar, c0, w0 - 3, h0 - 3, n);
  }
  mbar_wait(&tma_bar, 0, p.timeout_flag);

  // ---- phase 2: per-halo-pixel partial sums over this slab (thread per pixel, rotated channel
  //      order so the 32 lanes of a warp hit 32 different banks)
  for (int hp = tid; hp < HP; hp += kThreads) {
    float s1 = 0.f, s2 = 0.f;
    const uint32_t* pd = tile_d + static_cast<size_t>(hp) * CS2;
    const uint32_t* px = tile_x + static_cast<size_t>(hp) * CS2;
    int c = tid % CS2;
    for (int i = 0; i < CS2; ++i) {
      const float2 d = bf16x2_to_f32x2(pd[c]);
      const float2 xh = bf16x2_to_f32x2(px[c]);
      s1 += d.x + d.y;
      s2 = fmaf(d.x, xh.x, s2);
      s2 = fmaf(d.y, xh.y, s2);
      if (++c == CS2) c = 0;
    }
    s_part[hp] = make_float2(s1, s2);
  }
  if (kCluster) {
    cluster_arrive();
    cluster_wait();
  } else {
    __syncthreads();
  }

  // ---- phase 3: combine slabs, fold rstd in
  {
    const uint32_t nranks = kCluster ? cluster_nctarank() : 1u;
    const float inv_c = 1.0f / static_cast<float>(p.C);
    for (int hp = tid; hp < HP; hp += kThreads) {
      float s1 = 0.f, s2 = 0.f;
      if (kCluster) {
        for (uint32_t r = 0; r < nranks; ++r) {
          const float2 v = dsmem_ld_f32x2(dsmem_map(smem_u32(&s_part[hp]), r));
          s1 += v.x;
          s2 += v.y;
        }
      } else {
        const float2 v = s_part[hp];
        s1 = v.x;
        s2 = v.y;
      }
      const int hh = h0 - 3 + hp / WT, ww = w0 - 3 + hp % WT;
      float rs = 0.f;
      if (hh >= 0 && hh < p.H && ww >= 0 && ww < p.W)
        rs = __ldg(p.rstd + (static_cast<size_t>(n) * p.H + hh) * p.W + ww);
      s_coef[hp] = make_float4(rs, rs * s1 * inv_c, rs * s2 * inv_c, 0.f);
    }
  }
  __syncthreads();

  // ---- phase 4: dconv = a*dxhat - b - c*xhat, in place over the dxhat tile (bf16)
  float2 acc_dbias = make_float2(0.f, 0.f);
  if (ch_active) {
    for (int hp = pg; hp < HP; hp += kPG) {
      const float4 cf = s_coef[hp];
      const size_t idx = static_cast<size_t>(hp) * CS2 + lpg;
      const float2 d = bf16x2_to_f32x2(tile_d[idx]);
      const float2 xh = bf16x2_to_f32x2(tile_x[idx]);
      const float dcx = fmaf(cf.x, d.x, -cf.y) - cf.z * xh.x;
      const float dcy = fmaf(cf.x, d.y, -cf.y) - cf.z * xh.y;
      tile_d[idx] = f32x2_to_bf16x2(dcx, dcy);
      const int hy = hp / WT, hx = hp % WT;
      if (hy >= 3 && hy < p.TH + 3 && hx >= 3 && hx < p.TW + 3) {  // interior: counted once
        acc_dbias.x += dcx;
        acc_dbias.y += dcy;
      }
    }
  }
  __syncthreads();   // xhat tile is dead from here on: reuse it for the taps

  // ---- taps of this slab -> shared memory [49][CS2] float2
  float2* s_w = reinterpret_cast<float2*>(tile_x);
  for (int i = tid; i < 49 * CS2; i += kThreads) {
    const int t = i / CS2, l = i % CS2;
    s_w[i] = __ldg(reinterpret_cast<const float2*>(p.wt + static_cast<size_t>(t) * p.C + c0) + l);
  }
  __syncthreads();

  // ---- phase 5: bwd-data + bwd-filter over 2x4 micro-tiles
  float2 dw_acc[49];
#pragma unroll
  for (int t = 0; t < 49; ++t) dw_acc[t] = make_float2(0.f, 0.f);
  float2 acc_dy = make_float2(0.f, 0.f);

  const int mt_w = p.TW / MW;
  const int nmt = (p.TH / MH) * mt_w;
  if (ch_active) {
    for (int mt = pg; mt < nmt; mt += kPG) {
      const int oy = (mt / mt_w) * MH, ox = (mt % mt_w) * MW;
      float2 xq[MH][MW], dxa[MH][MW];
      bool ok[MH][MW];
#pragma unroll
      for (int o = 0; o < MH; ++o)
#pragma unroll
        for (int j = 0; j < MW; ++j) {
          const int hh = h0 + oy + o, ww = w0 + ox + j;
          ok[o][j] = hh < p.H && ww < p.W;
          xq[o][j] = make_float2(0.f, 0.f);
          dxa[o][j] = make_float2(0.f, 0.f);
          if (ok[o][j]) {
            const size_t m = (static_cast<size_t>(n) * p.H + hh) * p.W + ww;
            xq[o][j] = bf16x2_to_f32x2(
                __ldg(reinterpret_cast<const uint32_t*>(p.x + m * p.C + c0) + lpg));
            dxa[o][j] = bf16x2_to_f32x2(
                __ldg(reinterpret_cast<const uint32_t*>(p.dy + m * p.C + c0) + lpg));
            acc_dy.x += dxa[o][j].x;
            acc_dy.y += dxa[o][j].y;
          }
        }
      const uint32_t* base = tile_d + (static_cast<size_t>(oy) * WT + ox) * CS2 + lpg;
#pragma unroll
      for (int r = 0; r < MH + 6; ++r) {
        float2 in[MW + 6];
#pragma unroll
        for (int j = 0; j < MW + 6; ++j)
          in[j] = bf16x2_to_f32x2(base[(static_cast<size_t>(r) * WT + j) * CS2]);
#pragma unroll
        for (int o = 0; o < MH; ++o) {
          const int rr = r - o;            // dconv row offset relative to the output row
          if (rr >= 0 && rr < 7) {
            const int kh = 6 - rr;
#pragma unroll
            for (int cc = 0; cc < 7; ++cc) {
              const int kw = 6 - cc;
              const float2 wv = s_w[(kh * 7 + kw) * CS2 + lpg];
#pragma unroll
              for (int j = 0; j < MW; ++j) {
                dxa[o][j] = ffma2(in[j + cc], wv, dxa[o][j]);
                dw_acc[kh * 7 + kw] = ffma2(in[j + cc], xq[o][j], dw_acc[kh * 7 + kw]);
              }
            }
          }
        }
      }
#pragma unroll
      for (int o = 0; o < MH; ++o)
#pragma unroll
        for (int j = 0; j < MW; ++j)
          if (ok[o][j]) {
            const int hh = h0 + oy + o, ww = w0 + ox + j;
            const size_t m = (static_cast<size_t>(n) * p.H + hh) * p.W + ww;
            *(reinterpret_cast<uint32_t*>(p.dx + m * p.C + c0) + lpg) =
                f32x2_to_bf16x2(dxa[o][j].x, dxa[o][j].y);
          }
    }
  }
  __syncthreads();   // tiles and taps are dead: reuse the whole buffer for the cross-group reduce

  // ---- reduce the 51 per-channel accumulators over the 4 pixel groups, then one atomic each
  float2* s_red = reinterpret_cast<float2*>(smem_raw);   // [kPG][51][CS2]
  if (ch_active) {
#pragma unroll
    for (int t = 0; t < 49; ++t) s_red[(pg * kAccRows + t) * CS2 + lpg] = dw_acc[t];
    s_red[(pg * kAccRows + 49) * CS2 + lpg] = acc_dbias;
    s_red[(pg * kAccRows + 50) * CS2 + lpg] = acc_dy;
  }
  __syncthreads();
  for (int i = tid; i < kAccRows * CS2; i += kThreads) {
    const int t = i / CS2, l = i % CS2;
    float2 s = s_red[i];
#pragma unroll
    for (int g = 1; g < kPG; ++g) {
      const float2 v = s_red[g * kAccRows * CS2 + i];
      s.x += v.x;
      s.y += v.y;
    }
    float* dst;
    if (t < 49) dst = p.dwt + static_cast<size_t>(t) * p.C + c0 + 2 * l;
    else if (t == 49) dst = p.dbias + c0 + 2 * l;
    else dst = p.dy_colsum + c0 + 2 * l;
    atomicAdd(dst, s.x);
    atomicAdd(dst + 1, s.y);
  }

  if (kCluster) {   // peers may still be reading s_part through DSMEM
    cluster_arrive();
    cluster_wait();
  }
}

void pick_tile_bwd(int H, int W, int CS, int* TH, int* TW) {
  auto round_up = [](int v, int m) { return (v + m - 1) / m * m; };
  int tw = round_up(W, MW);
  if (tw > 16) {
    const int parts = (W + 15) / 16;
    tw = round_up((W + parts - 1) / parts, MW);
  }
  int th = round_up(H, MH);
  // two halo tiles must fit, and the pixel tables hold 448 halo pixels
  while (th > MH && ((th + 6) * (tw + 6) * CS * 2 > 100 * 1024 || (th + 6) * (tw + 6) > 448)) {
    const int parts = (H + th - 1) / th + 1;
    th = round_up((H + parts - 1) / parts, MH);
  }
  *TH = th;
  *TW = tw;
}

}  // namespace

size_t dwconv_ln_bwd_workspace_bytes(int, int, int, int) { return 0; }

int dwconv_ln_bwd(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                  const void* dy, const float* wt, void* dx, float* dwt, float* dbias,
                  float* dy_colsum, int B, int H, int W, int C, void* /*workspace*/,
                  size_t /*ws_bytes*/, cudaStream_t stream) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (C % 8) != 0) {
    set_last_error("dwconv_ln_bwd: bad shape B=%d H=%d W=%d C=%d", B, H, W, C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int nslab = (C + 127) / 128;
  if (nslab > 8 || (C % nslab) != 0 || ((C / nslab) % 8) != 0) {
    set_last_error("dwconv_ln_bwd: C=%d cannot be split into <=8 equal slabs of <=128 channels", C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int CS = C / nslab;
  int TH, TW;
  pick_tile_bwd(H, W, CS, &TH, &TW);
  if ((TH + 6) * (TW + 6) > 448) {
    set_last_error("dwconv_ln_bwd: tile %dx%d exceeds the halo-pixel table", TH, TW);
    return CNX_ERR_BAD_SHAPE;
  }

  CUtensorMap map_d, map_x;
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  const uint32_t box[4] = {(uint32_t)CS, (uint32_t)(TW + 6), (uint32_t)(TH + 6), 1u};
  int rc = make_tensor_map(&map_d, dxhat, 2, 4, dims, strides, box, 0);
  if (rc != CNX_OK) return rc;
  rc = make_tensor_map(&map_x, xhat, 2, 4, dims, strides, box, 0);
  if (rc != CNX_OK) return rc;

  BwdParams p;
  p.timeout_flag = timeout_flag_ptr();
  if (p.timeout_flag == nullptr) {
    set_last_error("could not map the timeout flag (cudaHostAlloc failed)");
    return CNX_ERR_DRIVER;
  }
  p.rstd = rstd;
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.dy = static_cast<const __nv_bfloat16*>(dy);
  p.wt = wt;
  p.dx = static_cast<__nv_bfloat16*>(dx);
  p.dwt = dwt; p.dbias = dbias; p.dy_colsum = dy_colsum;
  p.B = B; p.H = H; p.W = W; p.C = C; p.CS = CS; p.TH = TH; p.TW = TW;
  p.tiles_w = (W + TW - 1) / TW;
  const int tiles_h = (H + TH - 1) / TH;
  const size_t tile = (static_cast<size_t>(TH + 6) * (TW + 6) * CS * 2 + 127) & ~size_t(127);
  p.tile_bytes = static_cast<uint32_t>(tile);
  size_t smem = 2 * tile;
  const size_t red = static_cast<size_t>(kPG) * kAccRows * (CS / 2) * sizeof(float2);
  const size_t taps = static_cast<size_t>(49) * (CS / 2) * sizeof(float2);
  if (smem < red) smem = red;
  if (tile < taps) smem = tile + taps > smem ? tile + taps : smem;
  smem += 128;   // manual 128-byte alignment slack
  if (smem > 212 * 1024) {   // + ~11 KB of static tables must stay under the 227 KB CTA limit
    set_last_error("dwconv_ln_bwd: shared-memory request %zu exceeds the SM", smem);
    return CNX_ERR_BAD_SHAPE;
  }

  cudaMemsetAsync(dwt, 0, sizeof(float) * 49 * C, stream);
  cudaMemsetAsync(dbias, 0, sizeof(float) * C, stream);
  cudaMemsetAsync(dy_colsum, 0, sizeof(float) * C, stream);

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
      cudaFuncSetAttribute(dwconv_ln_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           212 * 1024);
      attr_set = true;
    }
    e = cudaLaunchKernelEx(&cfg, dwconv_ln_bwd_kernel<true>, map_d, map_x, p);
  } else {
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(dwconv_ln_bwd_kernel<false>,
                           cudaFuncAttributeMaxDynamicSharedMemorySize, 212 * 1024);
      attr_set = true;
    }
    e = cudaLaunchKernelEx(&cfg, dwconv_ln_bwd_kernel<false>, map_d, map_x, p);
  }
  if (e != cudaSuccess) {
    set_last_error("dwconv_ln_bwd launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace cnx
