// dwconv7.cu — depthwise 7x7 convolution engine (forward and backward) for the ConvNeXt block,
// plus the host launchers behind cnx_dwconv_ln_fwd / cnx_dwconv_ln_bwd.
//
// Replaces  conv_dw -> permute -> norm  of timm's ConvNeXtBlock.forward and its autograd backward
// (un-vendored; reached from /root/reference/train_advanced_v4.py:134-135 and :223, restated in
// SURVEY.md §3.3).  v2 design: the depthwise conv is FMA-bound on B200
// (49 fp32 FMAs per element forward, 98 backward, against 4 bytes of HBM traffic), so the kernels
// are built to keep the FMA pipe busy rather than around the LayerNorm reduction:
//
//   * persistent CTAs (one per SM), each bound to ONE 128-channel slab for its whole life, walking
//     8 x TW pixel tiles; the (8+6) x (TW+6) x 128 halo tile of the NEXT work item is fetched by a
//     4-D TMA box (out-of-image rows/cols zero-filled by hardware = the conv padding) into the
//     other half of a double buffer while the current one is being computed;
//   * a thread owns 2 adjacent channels (one bf16x2 word; a warp reads 128 contiguous bytes of
//     shared memory, conflict free) and a 2 x 7 output micro-tile: 104 shared loads feed
//     686 packed fma.rn.f32x2 (SASS FFMA2) forward, 1372 backward;
//   * forward keeps the 49 taps in registers for the CTA's life; backward keeps the 49
//     filter-gradient accumulators in registers across ALL its tiles (one atomic per
//     (tap, channel) per CTA at the very end) and reads the taps from shared memory;
//   * no cross-channel dependency remains in these kernels, hence no cluster / DSMEM exchange:
//     the channels-last LayerNorm (and its backward) run as row kernels (ln_rows.cu) on data that
//     is L2-resident for every stage but the first.  Under the reference's autocast the conv
//     output is rounded to half precision before LayerNorm as well.
//
// Algorithmic bytes: forward pair 2*M*C*2 B, backward pair 4*M*C*2 B
// (dxhat, xhat... see DESIGN.md) — the intermediate u / dconv round trips are extra traffic that
// is reported, not counted.
#include "cnx_common.cuh"
#include "cnx_kernels.h"
#include <stdlib.h>

namespace cnx {

bool dwconv_ln_fwd_is_fused(int C);

// ln_rows.cu
int ln_rows_fwd(void* u_inout, float* mean, float* rstd, int64_t M, int C, float eps,
                cudaStream_t stream);
int ln_rows_bwd(const void* dxhat, const void* xhat, const float* rstd, void* dconv, int64_t M,
                int C, cudaStream_t stream);

namespace {

constexpr int kThreads = 256;
constexpr int kGroups = 4;          // pixel groups per CTA (64 lanes = 128 channels each)
constexpr int kWords = 64;          // bf16x2 words per pixel of a slab in shared memory
constexpr int kSlab = 128;
constexpr int MH = 2, MW = 7;       // output micro-tile per thread
constexpr int kTH = 8;              // tile height (4 micro-tile rows)
constexpr int kAccRows = 51;        // 49 taps + dbias + dy_colsum

struct ConvParams {
  int* timeout_flag;
  const float* wt;             // [49][C] tap-major
  const float* bias;           // [C] (forward)
  __nv_bfloat16* out;          // forward: u [B,H,W,C]; backward: dx
  const __nv_bfloat16* dy;     // backward: gradient of the block output (shortcut edge)
  float* dwt;                  // [49][C]  (+=)
  float* dbias;                // [C]      (+=)
  float* dy_colsum;            // [C]      (+=)
  float* mean;                 // fused LayerNorm (forward, C <= 128): [B*H*W]
  float* rstd;
  float eps;
  int B, H, W, C;
  int nslab, tiles_w, tiles_h, nitems;
};

__device__ __forceinline__ unsigned char* align128(unsigned char* p) {
  return p + ((128u - (smem_u32(p) & 127u)) & 127u);
}

struct Item { int n, h0, w0; };
__device__ __forceinline__ Item decode_item(const ConvParams& p, int item, int TW) {
  const int per_img = p.tiles_w * p.tiles_h;
  Item it;
  it.n = item / per_img;
  const int r = item - it.n * per_img;
  const int th_i = r / p.tiles_w;
  it.h0 = th_i * kTH;
  it.w0 = (r - th_i * p.tiles_w) * TW;
  return it;
}

__device__ __forceinline__ void group_sync(int pg) {
  asm volatile("bar.sync %0, 64;" ::"r"(1 + pg) : "memory");
}

// Sum of 16 per-lane values over the 32 lanes of a warp by halving: 16 shuffles instead of 80.
// On return v[0] holds, on lanes 2k and 2k+1, the warp total of value index k.
__device__ __forceinline__ void warp_reduce16(float (&v)[16], int lane) {
#pragma unroll
  for (int step = 0; step < 4; ++step) {
    const int n = 8 >> step;                 // values kept after this step
    const bool hi = (lane >> (4 - step)) & 1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (i < n) {
        const float keep = hi ? v[n + i] : v[i];
        const float send = hi ? v[i] : v[n + i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16 >> step);
      }
    }
  }
  v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// ------------------------------------------------------------------------------ forward
template <int TW, bool FUSE_LN>
__global__ void __launch_bounds__(kThreads, 1)
dwconv7_fwd_kernel(const __grid_constant__ CUtensorMap x_map, const ConvParams p) {
  constexpr int WT = TW + 6, HT = kTH + 6;
  constexpr int NCG = TW / MW;
  constexpr uint32_t kBufBytes = HT * WT * kWords * 4;
  extern __shared__ __align__(128) unsigned char smem_dyn[];
  unsigned char* smem_raw = align128(smem_dyn);
  __shared__ __align__(8) uint64_t full_bar[2];
  __shared__ float2 s_sum[kGroups][16];      // FUSE_LN: per pixel (warp 0, warp 1) partial sums
  __shared__ float2 s_sq[kGroups][16];

  const int tid = threadIdx.x;
  const int pg = tid >> 6;
  const int cp = tid & 63;
  const int lane = tid & 31;
  const int wg = (tid >> 5) & 1;             // warp inside the pixel group
  const int slab = blockIdx.x % p.nslab;
  const int worker = blockIdx.x / p.nslab;
  const int nworkers = gridDim.x / p.nslab;
  const int c0 = slab * kSlab;
  const bool ch_active = (c0 + 2 * cp) < p.C;
  const float inv_c = 1.0f / static_cast<float>(p.C);
  const size_t pstride = static_cast<size_t>(p.C >> 1);

  if (tid == 0) {
    tma_prefetch_desc(&x_map);
    mbar_init(&full_bar[0], 1);
    mbar_init(&full_bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  pdl_wait();                 // first global access (taps, TMA) comes after the previous kernel
  pdl_launch_dependents();    // persistent single-wave grid: the next kernel may stage its prologue

  auto issue = [&](int item, int b) {
    const Item it = decode_item(p, item, TW);
    mbar_arrive_expect_tx(&full_bar[b], kBufBytes);
    tma_load_4d(smem_raw + b * kBufBytes, &x_map, &full_bar[b], c0, it.w0 - 3, it.h0 - 3, it.n);
  };
  if (tid == 0 && worker < p.nitems) issue(worker, 0);

  // taps + bias of this thread's two channels: registers for the CTA's whole life
  float2 wreg[49];
  float2 breg = make_float2(0.f, 0.f);
  if (ch_active) {
    const float2* wp = reinterpret_cast<const float2*>(p.wt + c0 + 2 * cp);
    const int cstride2 = p.C >> 1;
#pragma unroll
    for (int t = 0; t < 49; ++t) wreg[t] = __ldg(wp + static_cast<size_t>(t) * cstride2);
    breg = __ldg(reinterpret_cast<const float2*>(p.bias + c0 + 2 * cp));
  } else {
#pragma unroll
    for (int t = 0; t < 49; ++t) wreg[t] = make_float2(0.f, 0.f);
  }

  int it = 0;
  for (int item = worker; item < p.nitems; item += nworkers, ++it) {
    const int b = it & 1;
    const int next = item + nworkers;
    // buffer b^1 was last read in iteration it-1, which ended with __syncthreads
    if (tid == 0 && next < p.nitems) issue(next, b ^ 1);
    const Item ti = decode_item(p, item, TW);
    mbar_wait(&full_bar[b], (it >> 1) & 1, p.timeout_flag);

    const uint32_t* tile = reinterpret_cast<const uint32_t*>(smem_raw + b * kBufBytes);
    const int rows_valid = min(kTH, p.H - ti.h0);
    const int cols_valid = min(TW, p.W - ti.w0);
    const int nmt = ((rows_valid + MH - 1) / MH) * NCG;
    for (int mt = pg; mt < nmt; mt += kGroups) {
      const int rp = mt / NCG, cg = mt - rp * NCG;
      const int oy = rp * MH, ox = cg * MW;
      if (ox >= cols_valid) continue;             // uniform per pixel group
      float2 acc[MH][MW];
#pragma unroll
      for (int o = 0; o < MH; ++o)
#pragma unroll
        for (int j = 0; j < MW; ++j) acc[o][j] = breg;
      const uint32_t* base = tile + (oy * WT + ox) * kWords + cp;
#pragma unroll
      for (int r = 0; r < MH + 6; ++r) {
        float2 in[MW + 6];
#pragma unroll
        for (int j = 0; j < MW + 6; ++j) in[j] = bf16x2_to_f32x2(base[(r * WT + j) * kWords]);
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
      if (FUSE_LN) {
        // ---- channels-last LayerNorm of the 14 pixels of this micro-tile, all C (<= 128) channels
        // of which live in the 64 lanes of this pixel group.  Statistics of the bf16-ROUNDED conv
        // output (the reference's autocast hands LayerNorm a half-precision tensor), two-pass.
        float2 u[MH][MW];
        float red[16];
#pragma unroll
        for (int o = 0; o < MH; ++o)
#pragma unroll
          for (int j = 0; j < MW; ++j) {
            u[o][j] = ch_active ? bf16x2_to_f32x2(f32x2_to_bf16x2(acc[o][j].x, acc[o][j].y))
                                : make_float2(0.f, 0.f);
            red[o * MW + j] = u[o][j].x + u[o][j].y;
          }
        red[14] = 0.f; red[15] = 0.f;
        warp_reduce16(red, lane);
        if ((lane & 1) == 0) reinterpret_cast<float*>(&s_sum[pg][lane >> 1])[wg] = red[0];
        group_sync(pg);
        float mean[MH * MW];
#pragma unroll
        for (int i = 0; i < MH * MW; ++i) {
          const float2 t = s_sum[pg][i];
          mean[i] = (t.x + t.y) * inv_c;
        }
#pragma unroll
        for (int o = 0; o < MH; ++o)
#pragma unroll
          for (int j = 0; j < MW; ++j) {
            const float m = mean[o * MW + j];
            u[o][j].x -= m;
            u[o][j].y -= m;
            red[o * MW + j] = ch_active ? fmaf(u[o][j].x, u[o][j].x, u[o][j].y * u[o][j].y) : 0.f;
          }
        red[14] = 0.f; red[15] = 0.f;
        warp_reduce16(red, lane);
        if ((lane & 1) == 0) reinterpret_cast<float*>(&s_sq[pg][lane >> 1])[wg] = red[0];
        group_sync(pg);
#pragma unroll
        for (int o = 0; o < MH; ++o) {
          const int hh = ti.h0 + oy + o;
          const size_t prow = (static_cast<size_t>(ti.n) * p.H + hh) * p.W + ti.w0 + ox;
          uint32_t* dst = reinterpret_cast<uint32_t*>(p.out + prow * p.C + c0) + cp;
#pragma unroll
          for (int j = 0; j < MW; ++j) {
            const float2 t = s_sq[pg][o * MW + j];
            const float rs = rsqrtf((t.x + t.y) * inv_c + p.eps);
            if (hh < p.H && ox + j < cols_valid) {
              if (ch_active) dst[j * pstride] = f32x2_to_bf16x2(u[o][j].x * rs, u[o][j].y * rs);
              if (cp == o * MW + j) {          // one lane per pixel records the statistics
                p.mean[prow + j] = mean[o * MW + j];
                p.rstd[prow + j] = rs;
              }
            }
          }
        }
      } else if (ch_active) {
#pragma unroll
        for (int o = 0; o < MH; ++o) {
          const int hh = ti.h0 + oy + o;
          if (hh < p.H) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(
                                p.out + ((static_cast<size_t>(ti.n) * p.H + hh) * p.W + ti.w0 + ox) *
                                            p.C + c0) + cp;
#pragma unroll
            for (int j = 0; j < MW; ++j)
              if (ox + j < cols_valid) dst[j * pstride] = f32x2_to_bf16x2(acc[o][j].x, acc[o][j].y);
          }
        }
      }
    }
    __syncthreads();   // everyone is done with buffer b before it is refilled
  }
}

// ----------------------------------------------------------------------------- backward
// inputs: dconv = dL/d(conv output) (halo tile), x (interior tile), dy (global, shortcut edge)
//   dx[p]   = dy[p] + sum_off dconv[p - off] * w[off]
//   dw[off] = sum_p x[p] * dconv[p - off],  dbias = sum_p dconv[p],  dy_colsum = sum_p dy[p]
template <int TW>
__global__ void __launch_bounds__(kThreads, 1)
dwconv7_bwd_kernel(const __grid_constant__ CUtensorMap dconv_map,
                   const __grid_constant__ CUtensorMap x_map, const ConvParams p) {
  constexpr int WT = TW + 6, HT = kTH + 6;
  constexpr int NCG = TW / MW;
  constexpr uint32_t kHaloBytes = HT * WT * kWords * 4;
  constexpr uint32_t kXBytes = kTH * TW * kWords * 4;
  constexpr uint32_t kBufBytes = kHaloBytes + kXBytes;
  constexpr uint32_t kTapBytes = 49 * kWords * 8;
  extern __shared__ __align__(128) unsigned char smem_dyn[];
  unsigned char* smem_raw = align128(smem_dyn);
  __shared__ __align__(8) uint64_t full_bar[2];
  float2* s_w = reinterpret_cast<float2*>(smem_raw + 2 * kBufBytes);   // [49][64] float2

  const int tid = threadIdx.x;
  const int pg = tid >> 6;
  const int cp = tid & 63;
  const int slab = blockIdx.x % p.nslab;
  const int worker = blockIdx.x / p.nslab;
  const int nworkers = gridDim.x / p.nslab;
  const int c0 = slab * kSlab;
  const bool ch_active = (c0 + 2 * cp) < p.C;

  if (tid == 0) {
    tma_prefetch_desc(&dconv_map);
    tma_prefetch_desc(&x_map);
    mbar_init(&full_bar[0], 1);
    mbar_init(&full_bar[1], 1);
    fence_mbar_init();
  }
  pdl_wait();
  pdl_launch_dependents();    // persistent single-wave grid: the next kernel may stage its prologue
  // taps of this slab -> shared memory (zero for channels past C)
  for (int i = tid; i < 49 * kWords; i += kThreads) {
    const int t = i >> 6, l = i & 63;
    float2 v = make_float2(0.f, 0.f);
    if (c0 + 2 * l < p.C)
      v = __ldg(reinterpret_cast<const float2*>(p.wt + static_cast<size_t>(t) * p.C + c0) + l);
    s_w[i] = v;
  }
  __syncthreads();

  auto issue = [&](int item, int b) {
    const Item it = decode_item(p, item, TW);
    unsigned char* dst = smem_raw + b * kBufBytes;
    mbar_arrive_expect_tx(&full_bar[b], kBufBytes);
    tma_load_4d(dst, &dconv_map, &full_bar[b], c0, it.w0 - 3, it.h0 - 3, it.n);
    tma_load_4d(dst + kHaloBytes, &x_map, &full_bar[b], c0, it.w0, it.h0, it.n);
  };
  if (tid == 0 && worker < p.nitems) issue(worker, 0);

  float2 dw_acc[49];
#pragma unroll
  for (int t = 0; t < 49; ++t) dw_acc[t] = make_float2(0.f, 0.f);
  float2 acc_db = make_float2(0.f, 0.f), acc_dy = make_float2(0.f, 0.f);

  int it = 0;
  for (int item = worker; item < p.nitems; item += nworkers, ++it) {
    const int b = it & 1;
    const int next = item + nworkers;
    if (tid == 0 && next < p.nitems) issue(next, b ^ 1);
    const Item ti = decode_item(p, item, TW);
    mbar_wait(&full_bar[b], (it >> 1) & 1, p.timeout_flag);

    const uint32_t* tile_d = reinterpret_cast<const uint32_t*>(smem_raw + b * kBufBytes);
    const uint32_t* tile_x = reinterpret_cast<const uint32_t*>(smem_raw + b * kBufBytes + kHaloBytes);
    const int rows_valid = min(kTH, p.H - ti.h0);
    const int cols_valid = min(TW, p.W - ti.w0);
    const int nmt = ((rows_valid + MH - 1) / MH) * NCG;
    for (int mt = pg; mt < nmt; mt += kGroups) {
      const int rp = mt / NCG, cg = mt - rp * NCG;
      const int oy = rp * MH, ox = cg * MW;
      if (ox >= cols_valid) continue;
      // shortcut gradient: issue the global loads now, consume after the FMA loop
      uint32_t dyv[MH][MW];
      const size_t pstride = static_cast<size_t>(p.C >> 1);
#pragma unroll
      for (int o = 0; o < MH; ++o) {
        const int hh = ti.h0 + oy + o;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(
                                  p.dy + ((static_cast<size_t>(ti.n) * p.H + min(hh, p.H - 1)) * p.W +
                                          ti.w0 + ox) * p.C + c0) + cp;
#pragma unroll
        for (int j = 0; j < MW; ++j)
          dyv[o][j] = (ch_active && hh < p.H && ox + j < cols_valid) ? __ldg(src + j * pstride) : 0u;
      }
      float2 xq[MH][MW], dxa[MH][MW];
#pragma unroll
      for (int o = 0; o < MH; ++o)
#pragma unroll
        for (int j = 0; j < MW; ++j) {
          xq[o][j] = bf16x2_to_f32x2(tile_x[((oy + o) * TW + ox + j) * kWords + cp]);
          dxa[o][j] = make_float2(0.f, 0.f);
        }
      const uint32_t* base = tile_d + (oy * WT + ox) * kWords + cp;
#pragma unroll
      for (int r = 0; r < MH + 6; ++r) {
        float2 in[MW + 6];
#pragma unroll
        for (int j = 0; j < MW + 6; ++j) in[j] = bf16x2_to_f32x2(base[(r * WT + j) * kWords]);
#pragma unroll
        for (int o = 0; o < MH; ++o) {
          const int rr = r - o;            // dconv row minus output row, + 3
          if (rr >= 0 && rr < 7) {
            const int kh = 6 - rr;
            if (rr == 3) {                 // dconv at the output pixels themselves
#pragma unroll
              for (int j = 0; j < MW; ++j) {
                acc_db.x += in[j + 3].x;
                acc_db.y += in[j + 3].y;
              }
            }
#pragma unroll
            for (int cc = 0; cc < 7; ++cc) {
              const int kw = 6 - cc;
              const float2 wv = s_w[(kh * 7 + kw) * kWords + cp];
#pragma unroll
              for (int j = 0; j < MW; ++j) {
                dxa[o][j] = ffma2(in[j + cc], wv, dxa[o][j]);
                dw_acc[kh * 7 + kw] = ffma2(in[j + cc], xq[o][j], dw_acc[kh * 7 + kw]);
              }
            }
          }
        }
      }
      if (ch_active) {
#pragma unroll
        for (int o = 0; o < MH; ++o) {
          const int hh = ti.h0 + oy + o;
          if (hh < p.H) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(
                                p.out + ((static_cast<size_t>(ti.n) * p.H + hh) * p.W + ti.w0 + ox) *
                                            p.C + c0) + cp;
#pragma unroll
            for (int j = 0; j < MW; ++j)
              if (ox + j < cols_valid) {
                const float2 g = bf16x2_to_f32x2(dyv[o][j]);
                acc_dy.x += g.x;
                acc_dy.y += g.y;
                dst[j * pstride] = f32x2_to_bf16x2(dxa[o][j].x + g.x, dxa[o][j].y + g.y);
              }
          }
        }
      }
    }
    __syncthreads();
  }

  // ---- reduce the 51 per-channel accumulators over the 4 pixel groups, one atomic each per CTA
  float2* s_red = reinterpret_cast<float2*>(smem_raw);   // [kGroups][51][64] float2 = 104 KB
#pragma unroll
  for (int t = 0; t < 49; ++t) s_red[(pg * kAccRows + t) * kWords + cp] = dw_acc[t];
  s_red[(pg * kAccRows + 49) * kWords + cp] = acc_db;
  s_red[(pg * kAccRows + 50) * kWords + cp] = acc_dy;
  __syncthreads();
  for (int i = tid; i < kAccRows * kWords; i += kThreads) {
    const int t = i >> 6, l = i & 63;
    if (c0 + 2 * l >= p.C) continue;
    float2 s = s_red[i];
#pragma unroll
    for (int g = 1; g < kGroups; ++g) {
      const float2 v = s_red[g * kAccRows * kWords + i];
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
}

// ------------------------------------------------------------------------------ host side
int check_shape(const char* who, int B, int H, int W, int C) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (C % 8) != 0) {
    set_last_error("%s: bad shape B=%d H=%d W=%d C=%d (C must be a multiple of 8)", who, B, H, W, C);
    return CNX_ERR_BAD_SHAPE;
  }
  if ((C + kSlab - 1) / kSlab > 8) {
    set_last_error("%s: C=%d needs more than 8 slabs of 128 channels", who, C);
    return CNX_ERR_BAD_SHAPE;
  }
  return CNX_OK;
}

void fill_geometry(ConvParams& p, int B, int H, int W, int C, int TW) {
  p.B = B; p.H = H; p.W = W; p.C = C;
  p.nslab = (C + kSlab - 1) / kSlab;
  p.tiles_w = (W + TW - 1) / TW;
  p.tiles_h = (H + kTH - 1) / kTH;
  p.nitems = B * p.tiles_w * p.tiles_h;
}

int grid_size(const ConvParams& p) {
  int workers = sm_count() / p.nslab;
  if (workers < 1) workers = 1;
  if (workers > p.nitems) workers = p.nitems;
  return workers * p.nslab;
}

int make_nhwc_map(CUtensorMap* map, const void* base, int B, int H, int W, int C, int box_w,
                  int box_h) {
  const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  const uint32_t box[4] = {(uint32_t)kSlab, (uint32_t)box_w, (uint32_t)box_h, 1u};
  return make_tensor_map(map, base, 2, 4, dims, strides, box, 0);
}

template <int TW, bool FUSE_LN>
int launch_fwd(const void* x, ConvParams& p, int B, int H, int W, int C, cudaStream_t stream) {
  fill_geometry(p, B, H, W, C, TW);
  CUtensorMap map;
  int rc = make_nhwc_map(&map, x, B, H, W, C, TW + 6, kTH + 6);
  if (rc != CNX_OK) return rc;
  const size_t smem = 2ull * (kTH + 6) * (TW + 6) * kWords * 4 + 128;
  rc = ensure_dyn_smem(dwconv7_fwd_kernel<TW, FUSE_LN>, smem, "dwconv7_fwd");
  if (rc != CNX_OK) return rc;
  cudaError_t e = launch_pdl(dwconv7_fwd_kernel<TW, FUSE_LN>, grid_size(p), kThreads, smem, stream,
                             1u, map, p);
  if (e != cudaSuccess) {
    set_last_error("dwconv7_fwd launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

template <int TW>
int launch_bwd(const void* dconv, const void* x, ConvParams& p, int B, int H, int W, int C,
               cudaStream_t stream) {
  fill_geometry(p, B, H, W, C, TW);
  CUtensorMap map_d, map_x;
  int rc = make_nhwc_map(&map_d, dconv, B, H, W, C, TW + 6, kTH + 6);
  if (rc != CNX_OK) return rc;
  rc = make_nhwc_map(&map_x, x, B, H, W, C, TW, kTH);
  if (rc != CNX_OK) return rc;
  const size_t smem = 2ull * ((kTH + 6) * (TW + 6) + kTH * TW) * kWords * 4 + 49 * kWords * 8 + 128;
  rc = ensure_dyn_smem(dwconv7_bwd_kernel<TW>, smem, "dwconv7_bwd");
  if (rc != CNX_OK) return rc;
  cudaError_t e = launch_pdl(dwconv7_bwd_kernel<TW>, grid_size(p), kThreads, smem, stream, 1u, map_d,
                             map_x, p);
  if (e != cudaSuccess) {
    set_last_error("dwconv7_bwd launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace

int dwconv_ln_fwd(const void* x, const float* wt, const float* bias, void* xhat, float* mean,
                  float* rstd, int B, int H, int W, int C, float eps, cudaStream_t stream) {
  int rc = check_shape("dwconv_ln_fwd", B, H, W, C);
  if (rc != CNX_OK) return rc;
  ConvParams p{};
  p.timeout_flag = timeout_flag_ptr();
  if (p.timeout_flag == nullptr) {
    set_last_error("could not map the timeout flag (cudaHostAlloc failed)");
    return CNX_ERR_DRIVER;
  }
  p.wt = wt;
  p.bias = bias;
  p.out = static_cast<__nv_bfloat16*>(xhat);
  p.mean = mean; p.rstd = rstd; p.eps = eps;
  if (dwconv_ln_fwd_is_fused(C)) {
    // one slab holds every channel of a pixel: LayerNorm happens in the conv kernel's epilogue
    return (W > 7) ? launch_fwd<14, true>(x, p, B, H, W, C, stream)
                   : launch_fwd<7, true>(x, p, B, H, W, C, stream);
  }
  // wider maps: the conv output u lands in the xhat buffer ...
  rc = (W > 7) ? launch_fwd<14, false>(x, p, B, H, W, C, stream)
               : launch_fwd<7, false>(x, p, B, H, W, C, stream);
  if (rc != CNX_OK) return rc;
  // ... and is normalised in place, row by row (L2-resident for every such stage)
  return ln_rows_fwd(xhat, mean, rstd, static_cast<int64_t>(B) * H * W, C, eps, stream);
}

// The fused kernel leaves DRAM traffic at the algorithmic 2*E*s but measured no faster than conv +
// row kernel (stage 1, B=128: see DESIGN.md 3.1); CNX_DWCONV_NO_FUSE=1 selects the two-kernel path.
bool dwconv_ln_fwd_is_fused(int C) {
  static const bool no_fuse = getenv("CNX_DWCONV_NO_FUSE") != nullptr;
  return C <= kSlab && !no_fuse;
}

size_t dwconv_ln_bwd_workspace_bytes(int B, int H, int W, int C) {
  return static_cast<size_t>(B) * H * W * C * 2;   // dconv, bf16
}

int dwconv_ln_bwd(const void* dxhat, const void* xhat, const float* rstd, const void* x,
                  const void* dy, const float* wt, void* dx, float* dwt, float* dbias,
                  float* dy_colsum, int B, int H, int W, int C, void* workspace, size_t ws_bytes,
                  int accumulate, cudaStream_t stream) {
  int rc = check_shape("dwconv_ln_bwd", B, H, W, C);
  if (rc != CNX_OK) return rc;
  const size_t need = dwconv_ln_bwd_workspace_bytes(B, H, W, C);
  if (workspace == nullptr || ws_bytes < need ||
      (reinterpret_cast<uintptr_t>(workspace) & 15) != 0) {
    set_last_error("dwconv_ln_bwd: needs a 16-byte aligned workspace of %zu bytes (got %zu)", need,
                   workspace ? ws_bytes : size_t(0));
    return CNX_ERR_WORKSPACE;
  }
  rc = ln_rows_bwd(dxhat, xhat, rstd, workspace, static_cast<int64_t>(B) * H * W, C, stream);
  if (rc != CNX_OK) return rc;
  return dwconv_bwd(workspace, x, dy, wt, dx, dwt, dbias, dy_colsum, B, H, W, C, accumulate,
                    stream);
}

int dwconv_bwd(const void* dconv, const void* x, const void* dy, const float* wt, void* dx,
               float* dwt, float* dbias, float* dy_colsum, int B, int H, int W, int C,
               int accumulate, cudaStream_t stream) {
  int rc = check_shape("dwconv_bwd", B, H, W, C);
  if (rc != CNX_OK) return rc;
  ConvParams p{};
  p.timeout_flag = timeout_flag_ptr();
  if (p.timeout_flag == nullptr) {
    set_last_error("could not map the timeout flag (cudaHostAlloc failed)");
    return CNX_ERR_DRIVER;
  }
  p.wt = wt;
  p.out = static_cast<__nv_bfloat16*>(dx);
  p.dy = static_cast<const __nv_bfloat16*>(dy);
  p.dwt = dwt; p.dbias = dbias; p.dy_colsum = dy_colsum;
  if (!accumulate) {
    if ((rc = zero_async(dwt, sizeof(float) * 49 * C, stream, "dwconv_bwd")) != CNX_OK) return rc;
    if ((rc = zero_async(dbias, sizeof(float) * C, stream, "dwconv_bwd")) != CNX_OK) return rc;
    if ((rc = zero_async(dy_colsum, sizeof(float) * C, stream, "dwconv_bwd")) != CNX_OK) return rc;
  }
  return (W > 7) ? launch_bwd<14>(dconv, x, p, B, H, W, C, stream)
                 : launch_bwd<7>(dconv, x, p, B, H, W, C, stream);
}

}  // namespace cnx
