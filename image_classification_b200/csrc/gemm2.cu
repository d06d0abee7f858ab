// gemm2.cu — persistent, warp-specialised bf16 GEMM on 5th-gen tensor cores for the ConvNeXt MLP
// (fc1 / fc2 and their data- and weight-gradients), with the block's elementwise work fused into
// the epilogue.  Supersedes mlp_gemm.cu's one-tile-per-CTA kernel.
//
// Replaces mlp.fc1 -> act(GELU) -> mlp.fc2 -> *gamma -> +shortcut of timm's ConvNeXtBlock.forward
// and their autograd backward (reached from
// /root/reference/train_advanced_v4.py:134-135, backward from :223).
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )        fp32 accumulate in tensor memory
//
// One CTA per SM, 10 warps, looping over 128 x BN output tiles (BN = 256 or 128):
//   warp 0 / lane 0 : TMA producer — cp.async.bulk.tensor into a ring of 128B-swizzled stages,
//                     running ahead across tile boundaries
//   warp 1          : TMEM allocator; lane 0 = MMA issuer — tcgen05.mma.cta_group::1.kind::f16
//                     (128 x BN x 16), TWO accumulators of BN columns in tensor memory so tile
//                     i+1's main loop overlaps tile i's epilogue; tcgen05.commit frees ring slots
//                     and publishes the accumulator
//   warps 2..9      : epilogue — tcgen05.ld (thread = accumulator row, two warps per TMEM lane
//                     quarter, one per column half), fused math in registers, results staged in
//                     128B-swizzled shared memory (conflict-free 16-byte stores) and written by TMA
//                     (cp.async.bulk.tensor store: fully coalesced, clips ragged M / N edges);
//                     residual / pre-activation operands arrive by TMA into the same staging
//                     layout one chunk ahead of use.  Staging chunks are 32 rows x 64 columns
//                     (128B swizzle), or 32 x 32 (64B swizzle) for the epilogues without a second
//                     operand: the 32 KB that frees go to one more operand stage.
//                     Variants: 16 epilogue warps (two per staging pair) for the short-K fc1;
//                     CTA pairs (cta_group::2, 256-row tiles, each CTA stages half of B).
// A CTA keeps the same column block for its whole life whenever the tile grid allows it, so the
// per-column sums of the GELU-gradient epilogue stay in registers (one atomic per column per CTA).
// GELU and its derivative are degree-6/7 odd minimax polynomials evaluated with packed
// fma.rn.f32x2 (|err| < 3e-4, below the bf16 output resolution).  Measured: the
// epilogue arithmetic is free; the loop is paced by operand bytes in flight per SM (ring depth),
// and every byte the epilogue moves competes with the operand stream.
// Operands may be K-major (row-major [rows,K]) or MN-major (stored [K,rows]); the latter is what
// the weight-gradient GEMMs need (K = tokens), so no activation transpose ever touches HBM.
#include "cnx_common.cuh"
#include "cnx_kernels.h"
#include <stdlib.h>

namespace cnx {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                 // 64 bf16 = 128 B = one swizzle row
constexpr int kDuos = 8;               // (TMEM lane quarter, column half) pairs = staging buffer pairs
constexpr int EPI_ATOMIC_F32 = 5;      // internal: split-K fp32 reduction

// PAIR: two CTAs (a cluster of 2 on one TPC) share a 256 x BN tile through cta_group::2 MMAs; each
// stages its own 128 rows of A and HALF of B, which halves the B bytes written into and read out
// of every SM's shared memory (ncu on the single-CTA kernel: tensor pipe capped near 63 % because
// TMA writes + MMA operand reads exceed the 128 B/clk shared-memory port).
// CW = columns of one staging chunk (a 32-row x CW bf16 box of the output, per epilogue warp):
// 64 (128B swizzle) or 32 (64B swizzle).  Every GEMM here is paced by operand bytes in flight
// towards L2 (ring depth sweep, DESIGN.md 3.2: 3 / 4 / 5 / 6 stages = 62 / 56 / 52 / 50 us for the
// stage-3 fc1 data-gradient), so the epilogues that need no residual operand stage through half-size
// chunks and give the freed 32 KB to a sixth operand stage.
template <int BN, bool PAIR, bool STAGED, int CW = 64>
struct Cfg {
  static constexpr int kTileM = PAIR ? 2 * BM : BM;
  static constexpr int kBRows = PAIR ? BN / 2 : BN;      // rows of B staged by one CTA
  static constexpr uint32_t kABytes = BM * BK * 2;
  static constexpr uint32_t kBBytes = kBRows * BK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr uint32_t kChunkBytes = 32 * CW * 2;
  // bf16 epilogues stage through shared memory; the fp32 reduction epilogue does not
  static constexpr uint32_t kEpiBytes = STAGED ? kDuos * 2 * kChunkBytes : 0;
  static constexpr uint32_t kBarBytes = 512;
  static constexpr uint32_t kSmemMax = 232448;           // 227 KB opt-in limit per CTA
  // as many operand stages as fit: the main loop is bound by bytes in flight towards L2
  static constexpr int kFit = (kSmemMax - 1024 - kBarBytes - kEpiBytes) / kStageBytes;
  static constexpr int kStages = kFit > 8 ? 8 : kFit;
  static constexpr uint32_t kTotal = kStages * kStageBytes + kEpiBytes + kBarBytes + 1024;
  static constexpr uint32_t kTmemCols = 2 * BN;
  static_assert(kStages >= 3, "operand ring too shallow");
  static_assert(CW == 64 || CW == 32, "staging chunks are 64 or 32 columns wide");
};

struct GemmTuning { int stages = -1, pair = -1, bn = -1, splits = -1, epi_warps = 0, chunk_cols = 0; };
GemmTuning g_tune;                 // development overrides (cnx_debug_gemm_tuning)

struct Gemm2Params {
  int* timeout_flag;
  int M, N, K;
  int a_mn, b_mn;
  int tiles_m, tiles_n, splits, kb_per_split, nkb_total, items;
  int stages;                      // operand-ring depth in use (<= Cfg::kStages)
  float* D0f; int64_t ldd0;        // fp32 output (EPI_STORE_F32 / EPI_ATOMIC_F32)
  const float* bias;
  float* colsum;
  const float2* rowvec2;           // EPI_GELU_GRAD: [N] (r[n], -b[n]) (nullable)
  float* rowstats;                 // EPI_GELU_GRAD: [M][2] += ; EPI_LN_BWD: [M][2] read
  const float* rstd;               // EPI_LN_BWD: [M]
  float inv_n;                     // EPI_LN_BWD: 1 / N
};

// ---- packed-fp32 helpers (one issue slot, two lanes)
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  float2 d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<uint64_t&>(d))
      : "l"(reinterpret_cast<uint64_t&>(a)), "l"(reinterpret_cast<uint64_t&>(b)));
  return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 d;
  asm("add.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<uint64_t&>(d))
      : "l"(reinterpret_cast<uint64_t&>(a)), "l"(reinterpret_cast<uint64_t&>(b)));
  return d;
}
__device__ __forceinline__ float2 splat(float v) { return make_float2(v, v); }
__device__ __forceinline__ float2 clamp4(float2 z) {
  return make_float2(fminf(fmaxf(z.x, -4.0f), 4.0f), fminf(fmaxf(z.y, -4.0f), 4.0f));
}
// Phi(z) = 0.5 + zc * P(zc^2), zc = clamp(z, +-4): odd minimax fit of the normal CDF, degree 6 in
// zc^2: |err| < 1.05e-4 on [-4, 4] evaluated in fp32 (Phi(4) = 1 - 3.2e-5 beyond).  The FMA pipe,
// not the tensor pipe, paces the GELU-carrying epilogues (a 128 x 256 tile at K = 512 is 2048 clk
// of MMA; the polynomial alone was 12 FFMA2 = 24 pipe clk per two elements: 3072 clk per tile per
// scheduler), so the fit is as short as the bf16 result allows: the degree-8 fit of round 1 (5e-6)
// bought nothing below the output's 2^-9 rounding.  |gelu error| <= |z| * 1.05e-4 <= 4.2e-4.
__device__ __forceinline__ float2 normcdf2(float2 z) {
  const float2 zc = clamp4(z);
  const float2 t = fmul2(zc, zc);
  float2 p = splat(2.816107978e-08f);
  p = ffma2(p, t, splat(-1.891893589e-06f));
  p = ffma2(p, t, splat(5.419047763e-05f));
  p = ffma2(p, t, splat(-8.789816718e-04f));
  p = ffma2(p, t, splat(9.112954784e-03f));
  p = ffma2(p, t, splat(-6.538836144e-02f));
  p = ffma2(p, t, splat(3.985269153e-01f));
  return ffma2(zc, p, splat(0.5f));
}
__device__ __forceinline__ float2 gelu2(float2 z) { return fmul2(z, normcdf2(z)); }
// d/dz [z Phi(z)] = Phi(z) + z phi(z) = 0.5 + zc * Q(zc^2), degree 7: |err| < 2.8e-4 on [-4, 4]
// (values span [-0.13, 1.13]; the product dh * gelu' is rounded to bf16, 2^-9)
__device__ __forceinline__ float2 gelu_grad2(float2 z) {
  const float2 zc = clamp4(z);
  const float2 t = fmul2(zc, zc);
  float2 p = splat(-1.641978307e-08f);
  p = ffma2(p, t, splat(1.213805920e-06f));
  p = ffma2(p, t, splat(-3.845956586e-05f));
  p = ffma2(p, t, splat(6.876453804e-04f));
  p = ffma2(p, t, splat(-7.687441414e-03f));
  p = ffma2(p, t, splat(5.591482525e-02f));
  p = ffma2(p, t, splat(-2.620298446e-01f));
  p = ffma2(p, t, splat(7.967216433e-01f));
  return ffma2(zc, p, splat(0.5f));
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c,
                                             uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c),
               "r"(d)
               : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr)
               : "memory");
  return r;
}
__device__ __forceinline__ uint32_t ld_shared_b32(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(addr) : "memory");
  return r;
}
__device__ __forceinline__ void st_global_v4f(float* p, float a, float b, float c, float d) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_add_v4f(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}

struct ItemCoord { int m_blk, n_blk, kb_begin, kb_end; };
__device__ __forceinline__ ItemCoord decode(const Gemm2Params& p, int w) {
  ItemCoord c;
  c.n_blk = w % p.tiles_n;
  const int rest = w / p.tiles_n;
  c.m_blk = rest % p.tiles_m;
  const int split = rest / p.tiles_m;
  c.kb_begin = split * p.kb_per_split;
  c.kb_end = min(p.nkb_total, c.kb_begin + p.kb_per_split);
  return c;
}

// EW = epilogue warps: 8 (one warp per (lane quarter, column half): both 32-column halves of every
// 64-column chunk) or 16 (two warps per such pair, one 32-column half each, sharing the pair's
// staging buffers through a 64-thread named barrier).  The GELU-carrying epilogues are latency-
// bound with 2 warps per scheduler (ncu r01: issue slots 37 % busy, tensor pipe 40 %); 4 warps
// per scheduler hide the tcgen05.ld / shared-memory / polynomial latencies.
template <int BN, int EPI, bool PAIR, int EW, int CW = 64>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
             const __grid_constant__ CUtensorMap map_d0, const __grid_constant__ CUtensorMap map_d1,
             const __grid_constant__ CUtensorMap map_r, const Gemm2Params p) {
  using L = Cfg<BN, PAIR, !(EPI == EPI_STORE_F32 || EPI == EPI_ATOMIC_F32), CW>;
  constexpr int kStages = L::kStages;
  constexpr int kTileM = L::kTileM;
  constexpr uint32_t kChunkBytes = L::kChunkBytes;
  static_assert(CW == 64 || EW == 8, "32-column staging chunks: 8-warp epilogue only");
  constexpr bool kHasR = (EPI == EPI_BIAS_RESIDUAL || EPI == EPI_GELU_GRAD || EPI == EPI_LN_BWD);
  constexpr bool kF32Out = (EPI == EPI_STORE_F32 || EPI == EPI_ATOMIC_F32);
  constexpr int kChunks = BN / 2 / CW;         // staging chunks per epilogue warp per tile

  extern __shared__ unsigned char smem_dyn[];
  // 1024-byte alignment: the 128B swizzle pattern is a function of address bits [7,10)
  const uint32_t raw = smem_u32(smem_dyn);
  unsigned char* smem = smem_dyn + ((1024u - (raw & 1023u)) & 1023u);
  unsigned char* epi_base = smem + kStages * L::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_base + L::kEpiBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;     // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;      // [2]
  uint64_t* r_bar = tmem_empty_bar + 2;              // [kDuos]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(r_bar + kDuos);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;       // 0 = leader of the CTA pair
  const int w_first = PAIR ? (blockIdx.x >> 1) : blockIdx.x; // both CTAs of a pair walk the same items
  const int w_stride = PAIR ? (gridDim.x >> 1) : gridDim.x;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    if (!kF32Out) tma_prefetch_desc(&map_d0);
    if (EPI == EPI_BIAS_GELU) tma_prefetch_desc(&map_d1);
    if (kHasR) tma_prefetch_desc(&map_r);
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], PAIR ? 2 * EW : EW);
    }
#pragma unroll
    for (int w = 0; w < kDuos; ++w) mbar_init(&r_bar[w], 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    if (PAIR) tmem_alloc_pair(tmem_slot, L::kTmemCols);
    else tmem_alloc(tmem_slot, L::kTmemCols);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();    // the peer's barriers must exist before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // the prologue above overlapped the previous kernel's tail; from here on global memory is touched
  pdl_wait();
  pdl_launch_dependents();    // persistent single-wave grid: the next kernel may stage its prologue

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer
      int stage = 0;
      uint32_t phase = 0;
      for (int w = w_first; w < p.items; w += w_stride) {
        const ItemCoord c = decode(p, w);
        const int m0 = c.m_blk * kTileM + static_cast<int>(rank) * BM;
        const int n0 = c.n_blk * BN + static_cast<int>(rank) * (BN - L::kBRows);   // pair: own half of B
        for (int kb = c.kb_begin; kb < c.kb_end; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1, p.timeout_flag);
          unsigned char* sa = smem + stage * L::kStageBytes;
          unsigned char* sb = sa + L::kABytes;
          const int k0 = kb * BK;
          if (PAIR) {
            // both CTAs' bytes are credited to the LEADER's full barrier, which it arms for both
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * L::kStageBytes);
            const uint32_t fb = dsmem_map(smem_u32(&full_bar[stage]), 0);
            if (!p.a_mn) {
              tma_load_2d_pair(sa, &map_a, fb, k0, m0);
            } else {
#pragma unroll
              for (int a = 0; a < BM / 64; ++a)
                tma_load_2d_pair(sa + a * (BK * 128), &map_a, fb, m0 + a * 64, k0);
            }
            if (!p.b_mn) {
              tma_load_2d_pair(sb, &map_b, fb, k0, n0);                   // box [BN/2 rows][64 k]
            } else {
#pragma unroll
              for (int a = 0; a < L::kBRows / 64; ++a)
                tma_load_2d_pair(sb + a * (BK * 128), &map_b, fb, n0 + a * 64, k0);
            }
          } else {
            mbar_arrive_expect_tx(&full_bar[stage], L::kStageBytes);
            if (!p.a_mn) {
              tma_load_2d(sa, &map_a, &full_bar[stage], k0, m0);          // box [128 rows][64 k]
            } else {
#pragma unroll
              for (int a = 0; a < BM / 64; ++a)                           // box [64 k][64 rows]
                tma_load_2d(sa + a * (BK * 128), &map_a, &full_bar[stage], m0 + a * 64, k0);
            }
            if (!p.b_mn) {
              tma_load_2d(sb, &map_b, &full_bar[stage], k0, n0);          // box [BN rows][64 k]
            } else {
#pragma unroll
              for (int a = 0; a < BN / 64; ++a)
                tma_load_2d(sb + a * (BK * 128), &map_b, &full_bar[stage], n0 + a * 64, k0);
            }
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      // -------------------------------------------------------------- MMA issuer
      const uint32_t idesc = umma_idesc_bf16(kTileM, BN, p.a_mn, p.b_mn);
      int stage = 0;
      uint32_t phase = 0;
      int i = 0;
      for (int w = w_first; w < p.items; w += w_stride, ++i) {
        const ItemCoord c = decode(p, w);
        const int ab = i & 1;
        mbar_wait(&tmem_empty_bar[ab], ((i >> 1) & 1) ^ 1, p.timeout_flag);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + ab * BN;
        for (int kb = c.kb_begin; kb < c.kb_end; ++kb) {
          mbar_wait(&full_bar[stage], phase, p.timeout_flag);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * L::kStageBytes);
          const uint32_t sb = sa + L::kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: 16 k-elements = 32 B inside the swizzled 128 B row; 8-row groups 1024 B
            // apart.  MN-major: 16 k-rows = 2048 B; 8-k groups 1024 B apart (SBO), 64-wide MN
            // atoms BK*128 B apart (LBO).
            const uint64_t da = p.a_mn ? umma_desc_sw128(sa + k * 2048, BK * 128, 1024)
                                       : umma_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = p.b_mn ? umma_desc_sw128(sb + k * 2048, BK * 128, 1024)
                                       : umma_desc_sw128(sb + k * 32, 16, 1024);
            if (PAIR) umma_bf16_pair(tmem_d, da, db, idesc, (kb > c.kb_begin || k > 0) ? 1u : 0u);
            else umma_bf16(tmem_d, da, db, idesc, (kb > c.kb_begin || k > 0) ? 1u : 0u);
          }
          // slot reusable (in both CTAs of a pair) once these MMAs have read it
          if (PAIR) umma_commit_pair(&empty_bar[stage], 3);
          else umma_commit(&empty_bar[stage]);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        if (PAIR) umma_commit_pair(&tmem_full_bar[ab], 3);   // accumulator complete (both CTAs)
        else umma_commit(&tmem_full_bar[ab]);
      }
    }
  } else if constexpr (EW == 16) {
    // ------------------------------------------------- epilogue warps, two per staging buffer pair
    static_assert(!kF32Out, "the fp32 reduction epilogue runs with 8 epilogue warps");
    const int ew = warp - 2;                   // 0..15
    const int q = warp & 3;                    // TMEM lane quarter this warp may read
    const int part = ew >> 2;                  // 0..3
    const int half = part >> 1;                // column half of the tile
    const int mysub = part & 1;                // this warp's 32-column half of every 64-column chunk
    const int duo = half * 4 + q;              // the warps (mysub 0, 1) of a duo share buffers
    const bool lead = (mysub == 0) && (lane == 0);     // issues the duo's TMA loads / stores
    unsigned char* buf0 = epi_base + duo * 2 * kChunkBytes;  // output staging (D0)
    unsigned char* buf1 = buf0 + kChunkBytes;                // D1 staging or R landing zone
    const uint32_t buf0_u = smem_u32(buf0), buf1_u = smem_u32(buf1);
    const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;
    const uint32_t sw = static_cast<uint32_t>(lane & 7);
    uint64_t* rbar = &r_bar[duo];
    uint32_t r_jobs = 0;
    auto duo_sync = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + duo) : "memory"); };

    auto chunk_n = [&](const ItemCoord& c, int ch) { return c.n_blk * BN + half * (BN / 2) + ch * 64; };
    auto next_valid = [&](int& w, int& ch) {
      while (w < p.items) {
        const ItemCoord c = decode(p, w);
        for (; ch < kChunks; ++ch)
          if (chunk_n(c, ch) < p.N) return;
        ch = 0;
        w += w_stride;
      }
    };
    auto issue_r = [&](int w, int ch) {
      const ItemCoord c = decode(p, w);
      mbar_arrive_expect_tx(rbar, kChunkBytes);
      tma_load_2d(buf1, &map_r, rbar, chunk_n(c, ch),
                  c.m_blk * kTileM + static_cast<int>(rank) * BM + q * 32);
    };
    if (kHasR && lead) {
      int w = w_first, ch = 0;
      next_valid(w, ch);
      if (w < p.items) issue_r(w, ch);
    }

    float2 cs_acc[kChunks];                    // column sums over this warp's 16 rows
#pragma unroll
    for (int ch = 0; ch < kChunks; ++ch) cs_acc[ch] = make_float2(0.f, 0.f);
    int cs_nblk = -1;
    auto flush_colsum = [&]() {
      if (EPI == EPI_GELU_GRAD && p.colsum != nullptr && cs_nblk >= 0) {
#pragma unroll
        for (int ch = 0; ch < kChunks; ++ch) {
          const int n = cs_nblk * BN + half * (BN / 2) + ch * 64 + 2 * lane;
          if (n < p.N) {
            atomicAdd(p.colsum + n, cs_acc[ch].x);
            atomicAdd(p.colsum + n + 1, cs_acc[ch].y);
          }
          cs_acc[ch] = make_float2(0.f, 0.f);
        }
      }
    };

    int i = 0;
    for (int w = w_first; w < p.items; w += w_stride, ++i) {
      const ItemCoord c = decode(p, w);
      const int ab = i & 1;
      const int row0 = c.m_blk * kTileM + static_cast<int>(rank) * BM + q * 32;
      if (EPI == EPI_GELU_GRAD && c.n_blk != cs_nblk) {
        flush_colsum();
        cs_nblk = c.n_blk;
      }
      mbar_wait(&tmem_full_bar[ab], (i >> 1) & 1, p.timeout_flag);
      tc_fence_after();
      const uint32_t tmem_acc = tmem_base + ab * BN + (static_cast<uint32_t>(q * 32) << 16) +
                                half * (BN / 2);
#pragma unroll
      for (int ch = 0; ch < kChunks; ++ch) {
        const int n_g = chunk_n(c, ch);
        if (n_g >= p.N) continue;                        // uniform over the duo

        if (kHasR) {
          mbar_wait(rbar, r_jobs & 1, p.timeout_flag);
          ++r_jobs;
        }
        uint32_t pk[16];
        uint32_t hk[EPI == EPI_BIAS_GELU ? 16 : 1];
#pragma unroll
        for (int piece = 0; piece < 2; ++piece) {
          const int col = mysub * 32 + piece * 16;       // column offset inside the chunk
          uint32_t v[16];
          tmem_ld_32x16(tmem_acc + ch * 64 + col, v);
          uint32_t rr[8];
          if (kHasR) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const uint32_t unit = static_cast<uint32_t>(mysub * 4 + piece * 2 + k);
              const uint4 u = ld_shared_v4(buf1_u + row_off + ((unit ^ sw) << 4));
              rr[4 * k] = u.x; rr[4 * k + 1] = u.y; rr[4 * k + 2] = u.z; rr[4 * k + 3] = u.w;
            }
          }
          tmem_ld_wait();
          const int n = n_g + col;
          float2 f[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            f[j] = make_float2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
          if (EPI == EPI_BIAS_GELU || EPI == EPI_BIAS_RESIDUAL || EPI == EPI_BIAS) {
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              if (n + j < p.N) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n + j));
                f[j / 2] = fadd2(f[j / 2], make_float2(b.x, b.y));
                f[j / 2 + 1] = fadd2(f[j / 2 + 1], make_float2(b.z, b.w));
              }
            }
          }
          if (EPI == EPI_BIAS_RESIDUAL) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              f[j] = fadd2(f[j], bf16x2_to_f32x2(rr[j]));
            }
          }
          if (EPI == EPI_GELU_GRAD) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmul2(f[j], gelu_grad2(bf16x2_to_f32x2(rr[j])));
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) pk[piece * 8 + j] = f32x2_to_bf16x2(f[j].x, f[j].y);
          if (EPI == EPI_BIAS_GELU) {
            // activation of the ROUNDED pre-activation (what backward will see), second output
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float2 h2 = gelu2(bf16x2_to_f32x2(pk[piece * 8 + j]));
              hk[piece * 8 + j] = f32x2_to_bf16x2(h2.x, h2.y);
            }
          }
        }

        if (ch == kChunks - 1) {
          // last TMEM read of this tile is done: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (PAIR) mbar_arrive_cluster(dsmem_map(smem_u32(&tmem_empty_bar[ab]), 0));
            else mbar_arrive(&tmem_empty_bar[ab]);
          }
        }

        // A: the previous chunk's TMA stores have drained the staging buffers, and both warps are
        // done reading this chunk's R (and the previous chunk's column sums)
        if (lead) tma_store_wait_read();
        duo_sync();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint32_t unit = static_cast<uint32_t>(mysub * 4 + k);
          st_shared_v4(buf0_u + row_off + ((unit ^ sw) << 4), pk[4 * k], pk[4 * k + 1],
                       pk[4 * k + 2], pk[4 * k + 3]);
          if (EPI == EPI_BIAS_GELU)
            st_shared_v4(buf1_u + row_off + ((unit ^ sw) << 4), hk[4 * k], hk[4 * k + 1],
                         hk[4 * k + 2], hk[4 * k + 3]);
        }
        fence_proxy_async();          // make this lane's st.shared visible to the TMA engine
        duo_sync();                   // B: the whole 32 x 64 chunk is staged
        if (lead) {
          if (kHasR) {                // R of this chunk is consumed: fetch the next one
            int w2 = w, ch2 = ch + 1;
            next_valid(w2, ch2);
            if (w2 < p.items) issue_r(w2, ch2);
          }
          tma_store_2d(&map_d0, buf0, n_g, row0);
          if (EPI == EPI_BIAS_GELU) tma_store_2d(&map_d1, buf1, n_g, row0);
          tma_store_commit();
        }
        if (EPI == EPI_GELU_GRAD) {
          // column sums of the ROUNDED dz: lane l owns columns 2l, 2l+1 of the chunk, this warp
          // the rows [16 * mysub, +16) (rows past M hold acc = 0 -> contribute nothing)
          const uint32_t cchunk = static_cast<uint32_t>(lane >> 2), cword = static_cast<uint32_t>(lane & 3);
          float2 s = cs_acc[ch];
#pragma unroll
          for (int r0 = 0; r0 < 16; ++r0) {
            const int r = mysub * 16 + r0;
            const float2 t2 = bf16x2_to_f32x2(
                ld_shared_b32(buf0_u + r * 128 + ((cchunk ^ (r & 7)) << 4) + (cword << 2)));
            s = fadd2(s, t2);
          }
          cs_acc[ch] = s;
        }
      }

      if (chunk_n(c, kChunks - 1) >= p.N) {
        // the last chunk was skipped (ragged N): release the accumulator here
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (PAIR) mbar_arrive_cluster(dsmem_map(smem_u32(&tmem_empty_bar[ab]), 0));
          else mbar_arrive(&tmem_empty_bar[ab]);
        }
      }
    }
    flush_colsum();
    if (lead) tma_store_wait_all();
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int ew = warp - 2;
    const int q = warp & 3;                    // TMEM lane quarter this warp may read
    const int half = ew >> 2;                  // column half of the tile
    unsigned char* buf0 = epi_base + ew * 2 * kChunkBytes;   // output staging (D0)
    unsigned char* buf1 = buf0 + kChunkBytes;                // D1 staging or R landing zone
    const uint32_t buf0_u = smem_u32(buf0), buf1_u = smem_u32(buf1);
    // staging rows are CW bf16 wide; 16-byte units are XOR-swizzled like the TMA box (128B swizzle:
    // unit ^ (row & 7), 64B swizzle: unit ^ ((row >> 1) & 3)) so the row-per-lane stores and loads
    // are bank-conflict-free
    constexpr int kSubs = CW / 32;             // 32-column TMEM reads per chunk
    constexpr uint32_t kRowBytes = CW * 2;
    const uint32_t row_off = static_cast<uint32_t>(lane) * kRowBytes;
    const uint32_t sw = (CW == 64) ? static_cast<uint32_t>(lane & 7) : static_cast<uint32_t>((lane >> 1) & 3);
    uint64_t* rbar = &r_bar[ew];
    uint32_t r_jobs = 0;                       // R chunks consumed so far (phase of rbar)

    auto chunk_n = [&](const ItemCoord& c, int ch) { return c.n_blk * BN + half * (BN / 2) + ch * CW; };
    // first valid (n < N) chunk at or after (w, ch); w >= items when none is left
    auto next_valid = [&](int& w, int& ch) {
      while (w < p.items) {
        const ItemCoord c = decode(p, w);
        for (; ch < kChunks; ++ch)
          if (chunk_n(c, ch) < p.N) return;
        ch = 0;
        w += w_stride;
      }
    };
    auto issue_r = [&](int w, int ch) {
      const ItemCoord c = decode(p, w);
      mbar_arrive_expect_tx(rbar, kChunkBytes);
      tma_load_2d(buf1, &map_r, rbar, chunk_n(c, ch),
                  c.m_blk * kTileM + static_cast<int>(rank) * BM + q * 32);
    };
    if (kHasR && lane == 0) {
      int w = w_first, ch = 0;
      next_valid(w, ch);
      if (w < p.items) issue_r(w, ch);
    }

    float2 cs_acc[kChunks];                    // column sums (EPI_GELU_GRAD): lane -> cols 2l, 2l+1
#pragma unroll
    for (int ch = 0; ch < kChunks; ++ch) cs_acc[ch] = make_float2(0.f, 0.f);
    int cs_nblk = -1;
    auto flush_colsum = [&]() {
      if (EPI == EPI_GELU_GRAD && p.colsum != nullptr && cs_nblk >= 0) {
#pragma unroll
        for (int ch = 0; ch < kChunks; ++ch) {
          // CW == 32: lanes l and l ^ 16 hold partial sums (even / odd rows) of the same columns
          const int n = cs_nblk * BN + half * (BN / 2) + ch * CW + 2 * (lane & (CW / 2 - 1));
          if (n < p.N) {
            atomicAdd(p.colsum + n, cs_acc[ch].x);
            atomicAdd(p.colsum + n + 1, cs_acc[ch].y);
          }
          cs_acc[ch] = make_float2(0.f, 0.f);
        }
      }
    };

    int i = 0;
    for (int w = w_first; w < p.items; w += w_stride, ++i) {
      const ItemCoord c = decode(p, w);
      const int ab = i & 1;
      const int row0 = c.m_blk * kTileM + static_cast<int>(rank) * BM + q * 32;
      if (EPI == EPI_GELU_GRAD && c.n_blk != cs_nblk) {
        flush_colsum();
        cs_nblk = c.n_blk;
      }
      // LayerNorm backward by linearity (EPI_LN_BWD): out = rs * acc + c0 + xhat * c1 with the
      // per-row constants rs = rstd, c0 = -rstd * s1 / N, c1 = -rstd * s2 / N, where
      // s1 = sum_n dxhat[m,n] and s2 = sum_n dxhat[m,n] * xhat[m,n] came out of the dz GEMM's
      // epilogue (below): dxhat = dz W1f, so s1 = dz . rowsum(W1f) and, because
      // z = xhat W1f^T + b1f,  s2 = dz . (z - b1f) — no pass over dxhat is needed.
      float ln_rs = 0.f, ln_c0 = 0.f, ln_c1 = 0.f;
      if (EPI == EPI_LN_BWD && row0 + lane < p.M) {
        const float2 st2 = __ldg(reinterpret_cast<const float2*>(p.rowstats) + row0 + lane);
        ln_rs = __ldg(p.rstd + row0 + lane);
        ln_c0 = -ln_rs * st2.x * p.inv_n;
        ln_c1 = -ln_rs * st2.y * p.inv_n;
      }
      float2 st_s1 = make_float2(0.f, 0.f), st_s2 = make_float2(0.f, 0.f);   // EPI_GELU_GRAD row sums
      mbar_wait(&tmem_full_bar[ab], (i >> 1) & 1, p.timeout_flag);
      tc_fence_after();
      const uint32_t tmem_acc = tmem_base + ab * BN + (static_cast<uint32_t>(q * 32) << 16) +
                                half * (BN / 2);

#pragma unroll
      for (int ch = 0; ch < kChunks; ++ch) {
        const int n_g = chunk_n(c, ch);
        if (n_g >= p.N) continue;                        // warp-uniform

        if (kF32Out) {
          const int row = row0 + lane;
#pragma unroll
          for (int sub = 0; sub < kSubs; ++sub) {
            uint32_t v[32];
            tmem_ld_32x32(tmem_acc + ch * CW + sub * 32, v);
            tmem_ld_wait();
            const int n = n_g + sub * 32;
            if (row < p.M) {
              float* d = p.D0f + static_cast<int64_t>(row) * p.ldd0 + n;
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                if (n + j < p.N) {
                  if (EPI == EPI_ATOMIC_F32)
                    red_add_v4f(d + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                  else
                    st_global_v4f(d + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                  __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                }
              }
            }
          }
          continue;
        }

        if (kHasR) {
          mbar_wait(rbar, r_jobs & 1, p.timeout_flag);
          ++r_jobs;
        }

#pragma unroll
        for (int sub = 0; sub < kSubs; ++sub) {
          uint32_t v[32];
          tmem_ld_32x32(tmem_acc + ch * CW + sub * 32, v);
          uint32_t rr[16];
          if (kHasR) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint4 u = ld_shared_v4(buf1_u + row_off + (((sub * 4 + k) ^ sw) << 4));
              rr[4 * k] = u.x; rr[4 * k + 1] = u.y; rr[4 * k + 2] = u.z; rr[4 * k + 3] = u.w;
            }
          }
          tmem_ld_wait();
          const int n = n_g + sub * 32;

          float2 f[16];
#pragma unroll
          for (int j = 0; j < 16; ++j)
            f[j] = make_float2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));

          if (EPI == EPI_BIAS_GELU || EPI == EPI_BIAS_RESIDUAL || EPI == EPI_BIAS) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (n + j < p.N) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n + j));
                f[j / 2] = fadd2(f[j / 2], make_float2(b.x, b.y));
                f[j / 2 + 1] = fadd2(f[j / 2 + 1], make_float2(b.z, b.w));
              }
            }
          }
          if (EPI == EPI_BIAS_RESIDUAL) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              f[j] = fadd2(f[j], bf16x2_to_f32x2(rr[j]));
            }
          }
          if (EPI == EPI_GELU_GRAD) {
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = fmul2(f[j], gelu_grad2(bf16x2_to_f32x2(rr[j])));
          }
          if (EPI == EPI_LN_BWD) {
            const float2 rs2 = splat(ln_rs), c02 = splat(ln_c0), c12 = splat(ln_c1);
#pragma unroll
            for (int j = 0; j < 16; ++j)
              f[j] = ffma2(bf16x2_to_f32x2(rr[j]), c12, ffma2(f[j], rs2, c02));
          }

          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) pk[j] = f32x2_to_bf16x2(f[j].x, f[j].y);
          if (EPI == EPI_GELU_GRAD && p.rowstats != nullptr && n + 31 < p.N) {
            // row sums for the LayerNorm backward, taken over the ROUNDED dz (what the fc1
            // data-gradient GEMM will consume) and the rounded z it was computed from
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float4 rb = __ldg(reinterpret_cast<const float4*>(p.rowvec2 + n + 2 * j));
              const float2 dzr = bf16x2_to_f32x2(pk[j]);
              st_s1 = ffma2(dzr, make_float2(rb.x, rb.z), st_s1);
              st_s2 = ffma2(dzr, fadd2(bf16x2_to_f32x2(rr[j]), make_float2(rb.y, rb.w)), st_s2);
            }
          }
          uint32_t hk[EPI == EPI_BIAS_GELU ? 16 : 1];
          if (EPI == EPI_BIAS_GELU) {
            // activation of the ROUNDED pre-activation (what backward will see), second output
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 h2 = gelu2(bf16x2_to_f32x2(pk[j]));
              hk[j] = f32x2_to_bf16x2(h2.x, h2.y);
            }
          }
          if (sub == 0) {
            // the staging buffers must have been drained by the previous chunk's TMA stores; waiting
            // here (not at the top of the chunk) lets the drain overlap this chunk's first TMEM
            // read and arithmetic
            if (lane == 0) tma_store_wait_read();
            __syncwarp();
          }
#pragma unroll
          for (int k = 0; k < 4; ++k)
            st_shared_v4(buf0_u + row_off + (((sub * 4 + k) ^ sw) << 4), pk[4 * k], pk[4 * k + 1],
                         pk[4 * k + 2], pk[4 * k + 3]);
          if (EPI == EPI_BIAS_GELU) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              st_shared_v4(buf1_u + row_off + (((sub * 4 + k) ^ sw) << 4), hk[4 * k],
                           hk[4 * k + 1], hk[4 * k + 2], hk[4 * k + 3]);
          }
        }

        if (ch == kChunks - 1) {
          // last TMEM read of this tile is done: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (PAIR) mbar_arrive_cluster(dsmem_map(smem_u32(&tmem_empty_bar[ab]), 0));
            else mbar_arrive(&tmem_empty_bar[ab]);
          }
        }

        fence_proxy_async();          // make this lane's st.shared visible to the TMA engine
        __syncwarp();                 // all lanes stored (and finished reading R from buf1)
        if (lane == 0) {
          if (kHasR) {                // R of this chunk is consumed: fetch the next one
            int w2 = w, ch2 = ch + 1;
            next_valid(w2, ch2);
            if (w2 < p.items) issue_r(w2, ch2);
          }
          tma_store_2d(&map_d0, buf0, n_g, row0);
          if (EPI == EPI_BIAS_GELU) tma_store_2d(&map_d1, buf1, n_g, row0);
          tma_store_commit();
        }
        if (EPI == EPI_GELU_GRAD) {
          // column sums of the ROUNDED dz (what the weight-gradient GEMM will consume): lane l
          // owns columns 2l, 2l+1 of this chunk; 32 conflict-free word reads down the rows
          // (rows past M hold acc = 0 -> contribute nothing)
          float2 s = cs_acc[ch];
          if (CW == 64) {
            const uint32_t cchunk = static_cast<uint32_t>(lane >> 2), cword = static_cast<uint32_t>(lane & 3);
#pragma unroll
            for (int r = 0; r < 32; ++r) {
              const float2 t2 = bf16x2_to_f32x2(
                  ld_shared_b32(buf0_u + r * 128 + ((cchunk ^ (r & 7)) << 4) + (cword << 2)));
              s = fadd2(s, t2);
            }
          } else {
            // 64-byte rows: lanes 0-15 walk the even rows, lanes 16-31 the odd rows (one 128-byte
            // line per step, conflict-free); lane & 15 -> columns 2 (lane & 15), +1
            const uint32_t cchunk = static_cast<uint32_t>((lane & 15) >> 2), cword = static_cast<uint32_t>(lane & 3);
            const int rpar = lane >> 4;
#pragma unroll
            for (int i2 = 0; i2 < 16; ++i2) {
              const int r = 2 * i2 + rpar;
              const float2 t2 = bf16x2_to_f32x2(
                  ld_shared_b32(buf0_u + r * 64 + ((cchunk ^ (i2 & 3)) << 4) + (cword << 2)));
              s = fadd2(s, t2);
            }
          }
          cs_acc[ch] = s;
        }
      }

      if (EPI == EPI_GELU_GRAD && p.rowstats != nullptr && row0 + lane < p.M) {
        atomicAdd(p.rowstats + 2 * (row0 + lane), st_s1.x + st_s1.y);
        atomicAdd(p.rowstats + 2 * (row0 + lane) + 1, st_s2.x + st_s2.y);
      }
      if (kF32Out || chunk_n(c, kChunks - 1) >= p.N) {
        // the last chunk was skipped (ragged N) or the fp32 path ran: release the accumulator here
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (PAIR) mbar_arrive_cluster(dsmem_map(smem_u32(&tmem_empty_bar[ab]), 0));
          else mbar_arrive(&tmem_empty_bar[ab]);
        }
      }
    }
    flush_colsum();
    if (lane == 0) tma_store_wait_all();
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();    // the leader's MMAs read the peer's shared memory until the end
  else __syncthreads();
  if (warp == 1) {
    if (PAIR) tmem_dealloc_pair(tmem_base, L::kTmemCols);
    else tmem_dealloc(tmem_base, L::kTmemCols);
  }
}

template <int BN, int EPI, bool PAIR, int EW, int CW = 64>
int launch2(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& md0,
            const CUtensorMap& md1, const CUtensorMap& mr, const Gemm2Params& p_in, int grid,
            cudaStream_t stream) {
  using L = Cfg<BN, PAIR, !(EPI == EPI_STORE_F32 || EPI == EPI_ATOMIC_F32), CW>;
  Gemm2Params p = p_in;
  p.stages = (g_tune.stages >= 2 && g_tune.stages < L::kStages) ? g_tune.stages : L::kStages;
  const int rc_attr = ensure_dyn_smem(gemm2_kernel<BN, EPI, PAIR, EW, CW>, L::kTotal, "gemm_bf16");
  if (rc_attr != CNX_OK) return rc_attr;
  cudaError_t e = launch_pdl(gemm2_kernel<BN, EPI, PAIR, EW, CW>, grid, 64 + 32 * EW, L::kTotal,
                             stream, PAIR ? 2u : 1u, ma, mb, md0, md1, mr, p);
  if (e != cudaSuccess) {
    set_last_error("gemm_bf16 launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

template <int BN, bool PAIR>
int dispatch(int epi, int ew, bool narrow, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& md0,
             const CUtensorMap& md1, const CUtensorMap& mr, const Gemm2Params& p, int grid,
             cudaStream_t stream) {
#define CNX_GEMM_CASE(E)                                                                   \
  case E:                                                                                  \
    return ew == 16 ? launch2<BN, E, PAIR, 16>(ma, mb, md0, md1, mr, p, grid, stream)      \
                    : launch2<BN, E, PAIR, 8>(ma, mb, md0, md1, mr, p, grid, stream);
  // epilogues without a residual operand: 32-column staging chunks with 8 warps (one more stage)
#define CNX_GEMM_CASE_NARROW(E)                                                            \
  case E:                                                                                  \
    return ew == 16 ? launch2<BN, E, PAIR, 16>(ma, mb, md0, md1, mr, p, grid, stream)      \
           : narrow ? launch2<BN, E, PAIR, 8, 32>(ma, mb, md0, md1, mr, p, grid, stream)   \
                    : launch2<BN, E, PAIR, 8>(ma, mb, md0, md1, mr, p, grid, stream);
  switch (epi) {
    CNX_GEMM_CASE_NARROW(EPI_STORE_BF16)
    CNX_GEMM_CASE_NARROW(EPI_BIAS)
    CNX_GEMM_CASE_NARROW(EPI_BIAS_GELU)
    CNX_GEMM_CASE(EPI_BIAS_RESIDUAL)
    CNX_GEMM_CASE(EPI_GELU_GRAD)
    case EPI_LN_BWD:     return launch2<BN, EPI_LN_BWD, PAIR, 8>(ma, mb, md0, md1, mr, p, grid, stream);
    case EPI_STORE_F32:  return launch2<BN, EPI_STORE_F32, PAIR, 8>(ma, mb, md0, md1, mr, p, grid, stream);
    case EPI_ATOMIC_F32: return launch2<BN, EPI_ATOMIC_F32, PAIR, 8>(ma, mb, md0, md1, mr, p, grid, stream);
    default: break;
  }
#undef CNX_GEMM_CASE
#undef CNX_GEMM_CASE_NARROW
  set_last_error("gemm_bf16: unsupported epilogue %d", epi);
  return CNX_ERR_BAD_ARG;
}

// Persistent-loop cost of a tiling: (rounds) x (tile width), for `units` workers (SMs, or CTA
// pairs) over tiles of tile_m rows.  `workers_out` = workers to launch.
double tile_cost(int M, int N, int BN, int tile_m, int units, int* workers_out) {
  const int tiles_m = (M + tile_m - 1) / tile_m, tiles_n = (N + BN - 1) / BN;
  const int items = tiles_m * tiles_n;
  // option 1: every worker keeps one column block for its whole life (column sums stay in
  // registers); option 2: all workers, column block changes from item to item
  int w1 = (tiles_n <= units) ? (units / tiles_n) * tiles_n : units;
  if (w1 > items) w1 = items;
  int w2 = units < items ? units : items;
  const int r1 = (items + w1 - 1) / w1, r2 = (items + w2 - 1) / w2;
  int workers = w1, rounds = r1;
  if (r2 < r1) { workers = w2; rounds = r2; }
  *workers_out = workers;
  // the main loop is bound by operand bytes pulled from L2 (tools/gemm_sweep.py): an item costs
  // the rows of A plus the rows of B one CTA stages per k-block
  const int b_rows = (tile_m > BM) ? BN / 2 : BN;
  return static_cast<double>(rounds) * (BM + b_rows);
}

}  // namespace

int gemm_debug_max_clusters(int cluster_size) {
  // how many clusters of `cluster_size` pair-kernel CTAs (full shared-memory footprint) the device
  // can hold at once: GPC fragmentation shows up here (development probe)
  using L = Cfg<256, true, true>;
  auto kern = gemm2_kernel<256, EPI_STORE_BF16, true, 8>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
  if (cluster_size > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(sm_count() / cluster_size * cluster_size);
  cfg.blockDim = dim3(64 + 32 * 8);
  cfg.dynamicSmemBytes = L::kTotal;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_size;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = -1;
  cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
  if (e != cudaSuccess) { cudaGetLastError(); return -1; }
  return n;
}

void gemm_debug_tuning(int stages, int pair, int bn, int splits) {
  g_tune.stages = stages; g_tune.pair = pair; g_tune.bn = bn; g_tune.splits = splits;
}

void gemm_debug_epilogue_warps(int n) { g_tune.epi_warps = (n == 8 || n == 16) ? n : 0; }

void gemm_debug_staging_chunk(int cols) { g_tune.chunk_cols = (cols == 64) ? 64 : 0; }

int gemm_bf16(const GemmArgs& a, cudaStream_t stream) {
  if (a.M <= 0 || a.N <= 0 || a.K <= 0 || (a.N % 8) != 0 || (a.lda % 8) != 0 || (a.ldb % 8) != 0) {
    set_last_error("gemm_bf16: bad shape M=%d N=%d K=%d lda=%lld ldb=%lld (N, lda, ldb must be "
                   "multiples of 8)", a.M, a.N, a.K, (long long)a.lda, (long long)a.ldb);
    return CNX_ERR_BAD_SHAPE;
  }
  if (!a.a_mn_major && (a.K % 8) != 0) {
    set_last_error("gemm_bf16: K=%d must be a multiple of 8 for K-major operands", a.K);
    return CNX_ERR_BAD_SHAPE;
  }
  auto misaligned = [](const void* q) { return q && (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
  if (misaligned(a.A) || misaligned(a.B) || misaligned(a.D0) || misaligned(a.D1) ||
      misaligned(a.bias) || misaligned(a.R) || misaligned(a.colsum)) {
    set_last_error("gemm_bf16: every operand pointer must be 16-byte aligned");
    return CNX_ERR_BAD_ARG;
  }
  if ((a.ldd0 % 8) != 0 || (a.D1 && (a.ldd1 % 8) != 0) || (a.R && (a.ldr % 8) != 0)) {
    set_last_error("gemm_bf16: output/residual leading dimensions must be multiples of 8");
    return CNX_ERR_BAD_SHAPE;
  }
  int epi = a.epilogue;
  const bool need_bias = (epi == EPI_BIAS_GELU || epi == EPI_BIAS_RESIDUAL || epi == EPI_BIAS);
  const bool need_r = (epi == EPI_BIAS_RESIDUAL || epi == EPI_GELU_GRAD || epi == EPI_LN_BWD);
  if (epi < 0 || (epi > EPI_STORE_F32 && epi != EPI_BIAS && epi != EPI_LN_BWD) || (need_bias && !a.bias) ||
      (need_r && !a.R) || (epi == EPI_BIAS_GELU && !a.D1) ||
      (epi == EPI_LN_BWD && (!a.rowstats || !a.rstd)) ||
      (epi == EPI_GELU_GRAD && ((a.rowvec2 != nullptr) != (a.rowstats != nullptr)))) {
    set_last_error("gemm_bf16: unsupported epilogue %d or missing operand", a.epilogue);
    return CNX_ERR_BAD_ARG;
  }

  const int nsm = sm_count();
  const int nkb = (a.K + BK - 1) / BK;
  Gemm2Params p{};
  p.timeout_flag = timeout_flag_ptr();
  if (p.timeout_flag == nullptr) {
    set_last_error("could not map the timeout flag (cudaHostAlloc failed)");
    return CNX_ERR_DRIVER;
  }
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.a_mn = a.a_mn_major ? 1 : 0;
  p.b_mn = a.b_mn_major ? 1 : 0;
  p.bias = a.bias;
  p.colsum = a.colsum;
  p.rowvec2 = reinterpret_cast<const float2*>(a.rowvec2);
  p.rowstats = a.rowstats;
  p.rstd = a.rstd;
  p.inv_n = 1.0f / static_cast<float>(a.N);
  p.D0f = static_cast<float*>(a.D0);
  p.ldd0 = a.ldd0;
  p.nkb_total = nkb;

  // CTA pairs (256-row tiles through cta_group::2) whenever there are at least two row tiles
  // (CNX_GEMM_NO_PAIR=1 forces the single-CTA kernel: A/B debugging switch)
  static const bool no_pair = getenv("CNX_GEMM_NO_PAIR") != nullptr;
  // measured with tools/gemm_sweep.py (profiles/r01_gemm_sweep_v2.txt), same process A/B: every one of
  // these GEMMs is bound by operand bytes pulled from L2 (~8 TB/s chip-wide), so the pair tile
  // (2/3 of the operand bytes per MAC) wins wherever the main loop dominates; with short K the
  // two CTAs' heavy epilogues gate each other's next tile and the single-CTA kernel stays ahead
  bool pair = false;
  switch (epi) {
    case EPI_STORE_F32:     pair = true; break;
    case EPI_STORE_BF16:
    case EPI_BIAS:
    case EPI_LN_BWD:
    case EPI_BIAS_RESIDUAL: pair = a.K >= 1024 || (a.K >= 512 && a.N >= 1024); break;
    case EPI_BIAS_GELU:     pair = a.K >= 256; break;
    case EPI_GELU_GRAD:     pair = a.K >= 512; break;
    default: break;
  }
  pair = pair && !no_pair && a.M > BM && nsm >= 2;
  if (g_tune.pair >= 0) pair = g_tune.pair != 0 && a.M > BM && nsm >= 2;
  const int tile_m = pair ? 2 * BM : BM;
  const int units = pair ? nsm / 2 : nsm;

  // column-tile width: fewer persistent rounds wins; 256 preferred (less operand traffic per MAC)
  int BN = 256, workers = 0;
  {
    int w256, w128;
    const double c256 = tile_cost(a.M, a.N, 256, tile_m, units, &w256);
    const double c128 = tile_cost(a.M, a.N, 128, tile_m, units, &w128);
    // split-K weight gradients are one long main loop per item: the wide tile always wins there
    const bool wide = (epi == EPI_STORE_F32) ? (a.N > 128) : !(a.N <= 128 || c128 < 0.95 * c256);
    bool use256 = wide;
    // fc1 of the first stage (K = 128: two k-blocks per tile): the narrow tile's shorter epilogue
    // turns around faster (measured 172 vs 188 us)
    if (epi == EPI_BIAS_GELU && a.K <= 128 && a.N > 128) use256 = false;
    if (g_tune.bn == 128 || g_tune.bn == 256) use256 = g_tune.bn == 256;
    if (use256) { BN = 256; workers = w256; } else { BN = 128; workers = w128; }
  }
  p.tiles_m = (a.M + tile_m - 1) / tile_m;
  p.tiles_n = (a.N + BN - 1) / BN;
  int splits = 1;
  if (epi == EPI_STORE_F32) {
    // split-K so that the weight-gradient GEMMs (small M x N, K = tokens) fill the machine; the
    // caller zero-fills D0 (fp32 vector reductions accumulate into it).  Pick the split count
    // that minimises (persistent rounds) x (k-blocks per item), keeping >= 4 k-blocks per item.
    const int tiles = p.tiles_m * p.tiles_n;
    int max_splits = nkb / 4;
    if (max_splits > 2 * units) max_splits = 2 * units;
    long best = -1;
    for (int s = 1; s <= max_splits; ++s) {
      const int kb = (nkb + s - 1) / s;
      const int eff = (nkb + kb - 1) / kb;
      const long rounds = (static_cast<long>(tiles) * eff + units - 1) / units;
      const long cost = rounds * (kb + 12);    // + ~12 k-blocks of prologue/epilogue/atomics per item
      if (best < 0 || cost < best) { best = cost; splits = eff; }
    }
  }
  if (epi == EPI_STORE_F32 && g_tune.splits >= 1) splits = g_tune.splits < nkb ? g_tune.splits : nkb;
  int kbps = (nkb + splits - 1) / splits;
  splits = (nkb + kbps - 1) / kbps;             // no empty split
  // D0 += acc in every case (the header's contract): callers may accumulate several launches
  // (gradient-accumulation windows) into one buffer
  if (epi == EPI_STORE_F32) epi = EPI_ATOMIC_F32;
  p.splits = splits;
  p.kb_per_split = kbps;
  p.items = p.tiles_m * p.tiles_n * splits;
  if (splits > 1) workers = p.items < units ? p.items : units;
  const int grid = pair ? 2 * workers : workers;

  // 16 epilogue warps measured (profiles/r02_microbench_epi_warps_*.json): fc1+GELU 176 -> 170 us
  // and 96 -> 94 us at K = 128 / 256 (stages 1-2, HBM-bound: more stores in flight), no change at
  // K >= 512, GELU' (dz) 2 us SLOWER everywhere: so only the short-K fc1 uses them.
  // cnx_debug_gemm_epilogue_warps() forces 8 or 16 for every bf16 epilogue (A/B measurements)
  int ew = (epi == EPI_BIAS_GELU && a.K <= 256) ? 16 : 8;
  if (g_tune.epi_warps != 0) ew = g_tune.epi_warps;
  if (epi == EPI_LN_BWD || (epi == EPI_GELU_GRAD && a.rowstats != nullptr)) ew = 8;   // 8-warp path only
  // Narrow staging chunks (one more operand stage) for the 8-warp epilogues that read no residual
  // operand: measured (tools/gpu_ring.sh) fc1+GELU at K = 512 67.6 -> 62.5 us, fc1 data-gradient
  // 51.9 -> 50.0 us; with a residual operand the 2 KB loads one chunk ahead no longer cover their
  // latency (GELU-gradient 72 -> 84 us), so those keep 64-column chunks.  CNX_GEMM_WIDE_CHUNKS=1
  // or cnx_debug_gemm_staging_chunk(64) force 64 everywhere (A/B switch, parity tests).
  static const bool wide_env = getenv("CNX_GEMM_WIDE_CHUNKS") != nullptr;
  const bool wide_chunks = wide_env || g_tune.chunk_cols == 64;
  const bool narrow = !wide_chunks && ew == 8 &&
                      (epi == EPI_STORE_BF16 || epi == EPI_BIAS || epi == EPI_BIAS_GELU);

  CUtensorMap ma, mb, md0, md1, mr;
  int rc;
  {
    // K-major: stored [rows][K] -> dims (K, rows), box (64, tile rows).
    // MN-major: stored [K][rows] -> dims (rows, K), box (64 rows, 64 k) per 64-wide atom.
    const uint64_t dims_k[2] = {(uint64_t)a.K, (uint64_t)a.M};
    const uint64_t dims_mn[2] = {(uint64_t)a.M, (uint64_t)a.K};
    const uint64_t st[1] = {(uint64_t)a.lda * 2};
    const uint32_t box_k[2] = {BK, BM};
    const uint32_t box_mn[2] = {64, BK};
    rc = make_tensor_map(&ma, a.A, 2, 2, a.a_mn_major ? dims_mn : dims_k, st,
                         a.a_mn_major ? box_mn : box_k, 3);
    if (rc != CNX_OK) return rc;
  }
  {
    const uint64_t dims_k[2] = {(uint64_t)a.K, (uint64_t)a.N};
    const uint64_t dims_mn[2] = {(uint64_t)a.N, (uint64_t)a.K};
    const uint64_t st[1] = {(uint64_t)a.ldb * 2};
    const uint32_t box_k[2] = {BK, (uint32_t)(pair ? BN / 2 : BN)};   // a pair member stages half of B
    const uint32_t box_mn[2] = {64, BK};
    rc = make_tensor_map(&mb, a.B, 2, 2, a.b_mn_major ? dims_mn : dims_k, st,
                         a.b_mn_major ? box_mn : box_k, 3);
    if (rc != CNX_OK) return rc;
  }
  // epilogue maps: [M][N] bf16 row-major, box 64 cols x 32 rows, 128B swizzle (staging layout)
  // (32 cols x 32 rows, 64B swizzle for the narrow-chunk epilogues)
  auto mn_map = [&](CUtensorMap* m, const void* base, int64_t ld) {
    const uint64_t dims[2] = {(uint64_t)a.N, (uint64_t)a.M};
    const uint64_t st[1] = {(uint64_t)ld * 2};
    const uint32_t box[2] = {narrow ? 32u : 64u, 32};
    return make_tensor_map(m, base, 2, 2, dims, st, box, narrow ? 2 : 3);
  };
  const bool f32_out = (epi == EPI_STORE_F32 || epi == EPI_ATOMIC_F32);
  md0 = ma; md1 = ma; mr = ma;    // placeholders for unused maps (never dereferenced)
  if (!f32_out) {
    rc = mn_map(&md0, a.D0, a.ldd0);
    if (rc != CNX_OK) return rc;
  }
  if (epi == EPI_BIAS_GELU) {
    rc = mn_map(&md1, a.D1, a.ldd1);
    if (rc != CNX_OK) return rc;
  }
  if (need_r) {
    rc = mn_map(&mr, a.R, a.ldr);
    if (rc != CNX_OK) return rc;
  }

  if (pair)
    return BN == 256 ? dispatch<256, true>(epi, ew, narrow, ma, mb, md0, md1, mr, p, grid, stream)
                     : dispatch<128, true>(epi, ew, narrow, ma, mb, md0, md1, mr, p, grid, stream);
  return BN == 256 ? dispatch<256, false>(epi, ew, narrow, ma, mb, md0, md1, mr, p, grid, stream)
                   : dispatch<128, false>(epi, ew, narrow, ma, mb, md0, md1, mr, p, grid, stream);
}

}  // namespace cnx
