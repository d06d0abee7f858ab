// mlp_gemm.cu — bf16 GEMM on 5th-gen tensor cores for the ConvNeXt MLP (fc1 / fc2 and their
// data- and weight-gradients), with the block's elementwise work fused into the epilogue.
//
// Replaces mlp.fc1 -> act(GELU) -> mlp.fc2 -> *gamma -> +shortcut of timm's ConvNeXtBlock.forward
// and their autograd backward (reached from
// /root/reference/train_advanced_v4.py:134-135, backward from :223).
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )        fp32 accumulate in tensor memory
//
// Structure (one CTA per 128 x BN output tile, 2 CTAs co-resident per SM so one tile's epilogue
// overlaps the other's main loop):
//   warp 0 / lane 0 : TMA producer  — cp.async.bulk.tensor into a kStages-deep smem ring
//                      (128B-swizzled tiles), completion on "full" mbarriers
//   warp 1 / lane 0 : MMA issuer    — tcgen05.mma.cta_group::1.kind::f16, 128 x BN x 16 per
//                      instruction, accumulator in TMEM; tcgen05.commit frees ring slots
//   warp 2          : TMEM allocator
//   all 8 warps     : epilogue      — tcgen05.ld (thread = accumulator row), fused math,
//                      16-byte global stores / vector reductions
// Operands may be K-major (row-major [rows,K]) or MN-major (stored [K,rows]); the latter is what
// the weight-gradient GEMMs need (K = tokens), so no activation transpose ever touches HBM.
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;             // 64 bf16 = 128 B = one swizzle row
constexpr int kStages = 3;
constexpr int kGemmThreads = 256;
constexpr int EPI_ATOMIC_F32 = 5;  // internal: split-K fp32 reduction (EPI_STORE_F32 with splits>1)

template <int BN>
struct SmemLayout {
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTotal = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GemmParams {
  int* timeout_flag;
  int M, N, K;
  int a_mn, b_mn;
  int kb_per_split;
  void* D0; int64_t ldd0;
  void* D1; int64_t ldd1;
  const float* bias;
  const __nv_bfloat16* R; int64_t ldr;
  float* colsum;
};

// erf via Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7): 1 MUFU.RCP + 1 MUFU.EX2 + ~10 FMA-pipe ops,
// about a third of libdevice erff's issue slots — the epilogue is issue-bound, not accuracy-bound
// (outputs are rounded to bf16, eps 3.9e-3).
__device__ __forceinline__ float fast_erf(float x) {
  const float ax = fabsf(x);
  const float t = __frcp_rn(fmaf(0.3275911f, ax, 1.0f));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  p *= t;
  const float e = __expf(-ax * ax);
  const float r = fmaf(-p, e, 1.0f);
  return copysignf(r, x);
}
__device__ __forceinline__ float fast_gelu(float x) {
  return 0.5f * x * (1.0f + fast_erf(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float fast_gelu_grad(float x) {
  const float cdf = 0.5f * (1.0f + fast_erf(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
  return fmaf(x, pdf, cdf);
}

// lane l ends with sum over the 32 lanes of v[l] (butterfly transpose-reduce: 31 shuffles)
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int step = 16; step >= 1; step >>= 1) {
    const bool up = (lane & step) != 0;
#pragma unroll
    for (int i = 0; i < step; ++i) {
      const float send = up ? v[i] : v[i + step];
      const float keep = up ? v[i + step] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, step);
    }
  }
  return v[0];
}

__device__ __forceinline__ void st_global_v4(void* p, uint32_t a, uint32_t b, uint32_t c,
                                             uint32_t d) {
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}
__device__ __forceinline__ uint4 ld_global_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void red_add_v4_f32(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}

template <int BN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 2)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const GemmParams p) {
  using L = SmemLayout<BN>;
  extern __shared__ unsigned char smem_dyn[];
  // 1024-byte alignment: the 128B swizzle pattern is a function of address bits [7,10)
  const uint32_t raw = smem_u32(smem_dyn);
  const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
  unsigned char* smem = smem_dyn + pad;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * L::kStageBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full_bar = empty_bar + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * BM;
  const int nkb_total = (p.K + BK - 1) / BK;
  const int kb_begin = blockIdx.z * p.kb_per_split;
  const int kb_end = min(nkb_total, kb_begin + p.kb_per_split);
  const int nkb = kb_end - kb_begin;   // launcher guarantees >= 1

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    for (int i = 0; i < nkb; ++i) {
      const int s = i % kStages;
      const uint32_t ph = (i / kStages) & 1;
      mbar_wait(&empty_bar[s], ph ^ 1, p.timeout_flag);
      unsigned char* sa = smem + s * L::kStageBytes;
      unsigned char* sb = sa + L::kABytes;
      mbar_arrive_expect_tx(&full_bar[s], L::kStageBytes);
      const int k0 = (kb_begin + i) * BK;
      if (!p.a_mn) {
        tma_load_2d(sa, &map_a, &full_bar[s], k0, m0);                  // box [128 rows][64 k]
      } else {
#pragma unroll
        for (int a = 0; a < BM / 64; ++a)                               // box [64 k][64 rows]
          tma_load_2d(sa + a * (BK * 128), &map_a, &full_bar[s], m0 + a * 64, k0);
      }
      if (!p.b_mn) {
        tma_load_2d(sb, &map_b, &full_bar[s], k0, n0);
      } else {
#pragma unroll
        for (int a = 0; a < BN / 64; ++a)
          tma_load_2d(sb + a * (BK * 128), &map_b, &full_bar[s], n0 + a * 64, k0);
      }
    }
  } else if (warp == 1 && lane == 0) {
    // -------------------------------------------------------------- MMA issuer
    const uint32_t idesc = umma_idesc_bf16(BM, BN, p.a_mn, p.b_mn);
    for (int i = 0; i < nkb; ++i) {
      const int s = i % kStages;
      const uint32_t ph = (i / kStages) & 1;
      mbar_wait(&full_bar[s], ph, p.timeout_flag);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + s * L::kStageBytes);
      const uint32_t sb = sa + L::kABytes;
#pragma unroll
      for (int k = 0; k < BK / 16; ++k) {
        // K-major: 16 k-elements = 32 B inside the swizzled 128 B row; 8-row groups 1024 B apart.
        // MN-major: 16 k-rows = 2048 B; 8-k groups 1024 B apart (SBO), 64-wide MN atoms BK*128 B
        //           apart (LBO).
        const uint64_t da = p.a_mn ? umma_desc_sw128(sa + k * 2048, BK * 128, 1024)
                                   : umma_desc_sw128(sa + k * 32, 16, 1024);
        const uint64_t db = p.b_mn ? umma_desc_sw128(sb + k * 2048, BK * 128, 1024)
                                   : umma_desc_sw128(sb + k * 32, 16, 1024);
        umma_bf16(tmem_base, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
      }
      umma_commit(&empty_bar[s]);          // slot reusable once these MMAs have read it
    }
    umma_commit(tmem_full_bar);            // accumulator complete
  }
  __syncwarp();

  // ------------------------------------------------------------------ epilogue
  mbar_wait(tmem_full_bar, 0, p.timeout_flag);
  tc_fence_after();

  const int q = warp & 3;                  // TMEM lane quarter this warp may read
  const int half = warp >> 2;              // column half
  const int row = m0 + q * 32 + lane;
  const bool row_ok = row < p.M;
  constexpr int kColsPerWarp = BN / 2;

#pragma unroll 1
  for (int cc = 0; cc < kColsPerWarp; cc += 32) {
    const int col0 = half * kColsPerWarp + cc;      // column within the tile
    const int n = n0 + col0;                        // global column of v[0]
    uint32_t v[32];
    tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + col0, v);
    tmem_ld_wait();
    if (n >= p.N) continue;                         // warp-uniform
    const bool full_chunk = (n + 32) <= p.N;        // N % 8 == 0 is guaranteed by the launcher

    if (EPI == EPI_STORE_F32 || EPI == EPI_ATOMIC_F32) {
      if (row_ok) {
        float* d = static_cast<float*>(p.D0) + static_cast<int64_t>(row) * p.ldd0 + n;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          if (full_chunk || (n + j) < p.N) {
            if (EPI == EPI_ATOMIC_F32) {
              red_add_v4_f32(d + j, __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                             __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
            } else {
              st_global_v4(d + j, v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
          }
        }
      }
      continue;
    }

    float f[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);

    if (EPI == EPI_BIAS_GELU || EPI == EPI_BIAS_RESIDUAL) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        if (full_chunk || (n + j) < p.N) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + n + j));
          f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
        }
      }
    }

    if (EPI == EPI_BIAS_RESIDUAL || EPI == EPI_GELU_GRAD) {
      // R (shortcut) or Z (pre-activation): this thread's row, 32 bf16 = 4 x 16 B
      if (row_ok) {
        const __nv_bfloat16* r = p.R + static_cast<int64_t>(row) * p.ldr + n;
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          if (full_chunk || (n + j) < p.N) {
            const uint4 u = ld_global_nc_v4(r + j);
            const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 rv = bf16x2_to_f32x2(w4[t]);
              if (EPI == EPI_BIAS_RESIDUAL) {
                f[j + 2 * t] += rv.x;
                f[j + 2 * t + 1] += rv.y;
              } else {
                f[j + 2 * t] *= fast_gelu_grad(rv.x);
                f[j + 2 * t + 1] *= fast_gelu_grad(rv.y);
              }
            }
          } else {
#pragma unroll
            for (int t = 0; t < 8; ++t) f[j + t] = 0.f;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = 0.f;
      }
    }

    // pack to bf16 (round once), store D0
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = f32x2_to_bf16x2(f[2 * j], f[2 * j + 1]);
    if (row_ok) {
      __nv_bfloat16* d = static_cast<__nv_bfloat16*>(p.D0) + static_cast<int64_t>(row) * p.ldd0 + n;
#pragma unroll
      for (int j = 0; j < 32; j += 8)
        if (full_chunk || (n + j) < p.N)
          st_global_v4(d + j, pk[j / 2], pk[j / 2 + 1], pk[j / 2 + 2], pk[j / 2 + 3]);
    }

    if (EPI == EPI_BIAS_GELU) {
      // activation of the ROUNDED pre-activation (what backward will see), second output
      uint32_t hk[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float2 zr = bf16x2_to_f32x2(pk[j]);
        hk[j] = f32x2_to_bf16x2(fast_gelu(zr.x), fast_gelu(zr.y));
      }
      if (row_ok) {
        __nv_bfloat16* d = static_cast<__nv_bfloat16*>(p.D1) + static_cast<int64_t>(row) * p.ldd1 + n;
#pragma unroll
        for (int j = 0; j < 32; j += 8)
          if (full_chunk || (n + j) < p.N)
            st_global_v4(d + j, hk[j / 2], hk[j / 2 + 1], hk[j / 2 + 2], hk[j / 2 + 3]);
      }
    }

    if (EPI == EPI_GELU_GRAD) {
      if (p.colsum != nullptr) {
        // column sums of the ROUNDED dz (what the weight-gradient GEMM will consume)
        float g[32];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float2 t2 = bf16x2_to_f32x2(pk[j]);
          g[2 * j] = t2.x;
          g[2 * j + 1] = t2.y;
        }
        const float cs = warp_transpose_sum(g, lane);
        if ((n + lane) < p.N) atomicAdd(p.colsum + n + lane, cs);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, BN);
}

template <int BN, int EPI>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, int splits,
           cudaStream_t stream) {
  using L = SmemLayout<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_kernel<BN, EPI>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
    if (e != cudaSuccess) {
      set_last_error("gemm_bf16: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return CNX_ERR_LAUNCH;
    }
    attr_set = true;
  }
  dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, splits);
  gemm_bf16_kernel<BN, EPI><<<grid, kGemmThreads, L::kTotal, stream>>>(ma, mb, p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("gemm_bf16 launch: %s", cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace

int gemm_bf16_v1(const GemmArgs& a, cudaStream_t stream) {
  if (a.M <= 0 || a.N <= 0 || a.K <= 0 || (a.N % 8) != 0 || (a.lda % 8) != 0 || (a.ldb % 8) != 0) {
    set_last_error("gemm_bf16: bad shape M=%d N=%d K=%d lda=%lld ldb=%lld (N, lda, ldb must be "
                   "multiples of 8)", a.M, a.N, a.K, (long long)a.lda, (long long)a.ldb);
    return CNX_ERR_BAD_SHAPE;
  }
  if (!a.a_mn_major && (a.K % 8) != 0) {
    set_last_error("gemm_bf16: K=%d must be a multiple of 8 for K-major operands", a.K);
    return CNX_ERR_BAD_SHAPE;
  }
  constexpr int BN = 128;
  CUtensorMap ma, mb;
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
    const uint32_t box_k[2] = {BK, BN};
    const uint32_t box_mn[2] = {64, BK};
    rc = make_tensor_map(&mb, a.B, 2, 2, a.b_mn_major ? dims_mn : dims_k, st,
                         a.b_mn_major ? box_mn : box_k, 3);
    if (rc != CNX_OK) return rc;
  }

  GemmParams p;
  p.timeout_flag = timeout_flag_ptr();
  if (p.timeout_flag == nullptr) {
    set_last_error("could not map the timeout flag (cudaHostAlloc failed)");
    return CNX_ERR_DRIVER;
  }
  p.M = a.M; p.N = a.N; p.K = a.K;
  p.a_mn = a.a_mn_major ? 1 : 0;
  p.b_mn = a.b_mn_major ? 1 : 0;
  p.D0 = a.D0; p.ldd0 = a.ldd0;
  p.D1 = a.D1; p.ldd1 = a.ldd1;
  p.bias = a.bias;
  p.R = static_cast<const __nv_bfloat16*>(a.R); p.ldr = a.ldr;
  p.colsum = a.colsum;

  const int nkb = (a.K + BK - 1) / BK;
  int splits = 1;
  int epi = a.epilogue;
  if (epi == EPI_STORE_F32) {
    // split-K so that the weight-gradient GEMMs (tiny M x N, K = tokens) fill the machine;
    // the caller zero-fills D0 (fp32 vector reductions accumulate into it).
    const int tiles = ((a.M + BM - 1) / BM) * ((a.N + BN - 1) / BN);
    const int target = 2 * sm_count();
    splits = (target + tiles - 1) / tiles;
    if (splits > nkb / 4) splits = nkb / 4;     // keep >= 4 k-blocks per CTA
    if (splits < 1) splits = 1;
    if (splits > 1) epi = EPI_ATOMIC_F32;
  }
  int kbps = (nkb + splits - 1) / splits;
  splits = (nkb + kbps - 1) / kbps;             // no empty split
  p.kb_per_split = kbps;

  auto misaligned = [](const void* q) { return q && (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
  if (misaligned(a.D0) || misaligned(a.D1) || misaligned(a.bias) || misaligned(a.R) ||
      misaligned(a.colsum)) {
    set_last_error("gemm_bf16: every operand pointer must be 16-byte aligned");
    return CNX_ERR_BAD_ARG;
  }
  if ((a.ldd0 % 8) != 0 || (a.D1 && (a.ldd1 % 8) != 0) || (a.R && (a.ldr % 8) != 0)) {
    set_last_error("gemm_bf16: output/residual leading dimensions must be multiples of 8");
    return CNX_ERR_BAD_SHAPE;
  }
  switch (epi) {
    case EPI_STORE_BF16:    return launch<BN, EPI_STORE_BF16>(ma, mb, p, splits, stream);
    case EPI_BIAS_GELU:
      if (!a.bias || !a.D1) break;
      return launch<BN, EPI_BIAS_GELU>(ma, mb, p, splits, stream);
    case EPI_BIAS_RESIDUAL:
      if (!a.bias || !a.R) break;
      return launch<BN, EPI_BIAS_RESIDUAL>(ma, mb, p, splits, stream);
    case EPI_GELU_GRAD:
      if (!a.R) break;
      return launch<BN, EPI_GELU_GRAD>(ma, mb, p, splits, stream);
    case EPI_STORE_F32:     return launch<BN, EPI_STORE_F32>(ma, mb, p, splits, stream);
    case EPI_ATOMIC_F32:    return launch<BN, EPI_ATOMIC_F32>(ma, mb, p, splits, stream);
    default: break;
  }
  set_last_error("gemm_bf16: unsupported epilogue %d or missing operand", a.epilogue);
  return CNX_ERR_BAD_ARG;
}

}  // namespace cnx
