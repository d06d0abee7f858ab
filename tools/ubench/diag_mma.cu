// diag_mma.cu — feasibility probe for a tensor-core depthwise 7x7 convolution.
//
// Depthwise conv has no operand shared across channels, so on tcgen05 it can only run as
//   D[128 pixels, 16 ch] += A_t[128 pixels, 16 ch] * diag(w[t, 16 ch])     (one MMA per tap and
// 16-channel group; 1/16 of the MMA is useful).  Whether that beats the FMA pipe depends on three
// things this program measures on one SM:
//   U1  the A operand of tap (kh, kw) is the SAME 128B-swizzled pixel tile read at a start address
//       shifted by (kh*WT + kw) rows of 128 B — not 1024-byte aligned: does the descriptor's
//       base-offset field make that work, and with which value?
//   U2  a 16 x 16 diagonal B tile in the no-swizzle K-major core-matrix layout (512 B per tile);
//   U3  issue rate of M=128, N=16, K=16 MMAs from one thread (cycles per MMA).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I image_classification_b200/csrc \
//             -o tools/ubench/diag_mma tools/ubench/diag_mma.cu
#include "cnx_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

using namespace cnx;

constexpr int kRows = 512;          // pixel rows in the shared-memory tile (128 B each, 64 channels)
constexpr int kTaps = 49;
constexpr int kWT = 20;             // padded tile width of the 14-wide stage-3 images
constexpr int kGroups = 4;          // 16-channel groups of the 64-channel slab

__device__ __forceinline__ uint64_t desc_sw128_bo(uint32_t addr, uint32_t sbo, uint32_t base_off) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr & 0x3ffffu) >> 4);
  d |= static_cast<uint64_t>((16u >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;
  d |= static_cast<uint64_t>(base_off & 7u) << 49;
  d |= 2ull << 61;
  return d;
}
__device__ __forceinline__ uint64_t desc_noswz(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr & 0x3ffffu) >> 4);
  d |= static_cast<uint64_t>((lbo >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;
  return d;                         // layout type 0 = no swizzle
}

// mode: 0 = base offset 0, 1 = base offset (addr >> 7) & 7;  swap_lbo: exchange LBO / SBO of B
__global__ void __launch_bounds__(128)
diag_mma_kernel(const __nv_bfloat16* __restrict__ A, const float* __restrict__ w, float* __restrict__ D,
                long long* __restrict__ cycles, int r0, int mode, int swap_lbo, int reps, int nb) {
  extern __shared__ unsigned char smem_dyn[];
  const uint32_t raw = smem_u32(smem_dyn);
  unsigned char* smem = smem_dyn + ((1024u - (raw & 1023u)) & 1023u);
  unsigned char* sA = smem;                                  // kRows x 128 B, 128B swizzle
  unsigned char* sB = sA + kRows * 128;                      // kTaps x kGroups x 512 B
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x;

  for (int i = tid; i < kRows * 64; i += 128) {              // element (row r, channel c)
    const int r = i >> 6, c = i & 63;
    const uint32_t off = r * 128 + (((c >> 3) ^ (r & 7)) << 4) + (c & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = A[i];
  }
  for (int i = tid; i < kTaps * kGroups * 256; i += 128) {   // B[t][g][n][k] = (n == k) ? w : 0
    const int k = i & 15, n = (i >> 4) & 15, tg = i >> 8;
    const int g = tg % kGroups, t = tg / kGroups;
    const float v = (n == k) ? w[t * 64 + g * 16 + n] : 0.f;
    // core matrix (ni, ki) = 8 rows x 16 B; inside: row-major 16 B per row
    const uint32_t off = tg * 512 + ((n >> 3) * 2 + (k >> 3)) * 128 + (n & 7) * 16 + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = __float2bfloat16_rn(v);
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (tid < 32) tmem_alloc(&tmem_slot, 512);
  fence_proxy_async();                                       // generic-proxy smem writes -> MMA
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t idesc = umma_idesc_bf16(128, 16, 0, 0);

  if (tid == 0) {
    const long long t0 = clock64();
    for (int rep = 0; rep < reps; ++rep) {
      for (int t = 0; t < kTaps; ++t) {
        const int shift = (t / 7) * kWT + (t % 7);
        const uint32_t a_row = smem_u32(sA) + (r0 + shift) * 128;
        const uint32_t bo = mode ? ((a_row >> 7) & 7u) : 0u;
        for (int mb = 0; mb < nb; ++mb) {        // nb independent 128-pixel blocks (U3 only: the
                                                 // blocks overlap in the tile, results unused)
#pragma unroll
          for (int g = 0; g < kGroups; ++g) {
            const uint64_t da = desc_sw128_bo(a_row + mb * 16 * 128 + g * 32, 1024, bo);
            const uint32_t b_addr = smem_u32(sB) + (t * kGroups + g) * 512;
            const uint64_t db = swap_lbo ? desc_noswz(b_addr, 256, 128) : desc_noswz(b_addr, 128, 256);
            umma_bf16(tmem + mb * 64 + g * 16, da, db, idesc, (t > 0 || rep > 0) ? 1u : 0u);
          }
        }
      }
    }
    umma_commit(&bar);
    const long long t1 = clock64();
    int dummy = 0;
    mbar_wait(&bar, 0, &dummy);
    const long long t2 = clock64();
    cycles[0] = t1 - t0;
    cycles[1] = t2 - t0;
  }
  __syncthreads();
  tc_fence_after();
  {
    const int warp = tid >> 5;
    uint32_t v[32];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      tmem_ld_32x32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + h * 32, v);
      tmem_ld_wait();
      for (int j = 0; j < 32; ++j) D[tid * 64 + h * 32 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tmem, 512);
}

int main() {
  std::vector<__nv_bfloat16> hA(kRows * 64);
  std::vector<float> hAf(kRows * 64), hw(kTaps * 64);
  srand(1);
  for (int i = 0; i < kRows * 64; ++i) {
    hA[i] = __float2bfloat16_rn((rand() % 2001 - 1000) / 500.0f);
    hAf[i] = __bfloat162float(hA[i]);
  }
  for (int i = 0; i < kTaps * 64; ++i)
    hw[i] = __bfloat162float(__float2bfloat16_rn((rand() % 2001 - 1000) / 4000.0f));
  __nv_bfloat16* dA; float *dw, *dD; long long* dc;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dw, hw.size() * 4); cudaMalloc(&dD, 128 * 64 * 4);
  cudaMalloc(&dc, 16);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dw, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice);
  const size_t smem = kRows * 128 + kTaps * kGroups * 512 + 1024;
  cudaFuncSetAttribute(diag_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  std::vector<float> hD(128 * 64);
  for (int r0 : {0, 3, 21}) {
    std::vector<double> ref(128 * 64, 0.0);
    for (int m = 0; m < 128; ++m)
      for (int c = 0; c < 64; ++c)
        for (int t = 0; t < kTaps; ++t)
          ref[m * 64 + c] += (double)hAf[(r0 + (t / 7) * kWT + (t % 7) + m) * 64 + c] * hw[t * 64 + c];
    for (int mode = 0; mode < 2; ++mode)
      for (int swap = 0; swap < 2; ++swap) {
        cudaMemset(dD, 0, 128 * 64 * 4);
        diag_mma_kernel<<<1, 128, smem>>>(dA, dw, dD, dc, r0, mode, swap, 1, 1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("r0=%d mode=%d swap=%d: %s\n", r0, mode, swap, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
        double err = 0, mx = 0;
        for (int i = 0; i < 128 * 64; ++i) { err = fmax(err, fabs(hD[i] - ref[i])); mx = fmax(mx, fabs(ref[i])); }
        printf("{\"r0\": %d, \"base_offset_mode\": %d, \"swap_lbo_sbo\": %d, \"max_abs_err\": %.3e, \"ref_max\": %.3f}\n",
               r0, mode, swap, err, mx);
      }
  }
  // U3: issue rate (results are garbage sums, only the clock matters)
  for (int nb : {1, 2, 4, 8}) {
    long long hc[2];
    const int reps = 4;
    diag_mma_kernel<<<1, 128, smem>>>(dA, dw, dD, dc, 0, 0, 0, reps, nb);
    cudaDeviceSynchronize();
    cudaMemcpy(hc, dc, 16, cudaMemcpyDeviceToHost);
    const int n = reps * kTaps * kGroups * nb;
    printf("{\"independent_accumulators\": %d, \"mmas\": %d, \"issue_cycles_per_mma\": %.1f, "
           "\"complete_cycles_per_mma\": %.1f}\n", 4 * nb, n, (double)hc[0] / n, (double)hc[1] / n);
  }
  return 0;
}
