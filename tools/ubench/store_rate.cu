// store_rate.cu — how fast can the epilogue warps of the GEMM push a [M, N] bf16 tile stream out of
// shared memory?  Same work split as gemm2.cu (148 persistent CTAs, 8 epilogue warps, each warp owns
// [32 rows x 64 columns] chunks of a 128 x 256 tile), no MMA, no arithmetic: only the stores.
//   mode 0  TMA 2D store, box [32 rows][128 B] (128B swizzle), one output          (plain epilogue)
//   mode 1  same, two outputs per chunk                                             (fc1: z and h)
//   mode 2  TMA 2D store of the same 4 KB but into a tensor whose rows are 128 B long, so every box
//           is 4 KB of CONTIGUOUS global memory (is it the 32 scattered rows that cost?)
//   mode 3  cp.async.bulk shared -> global, 4 KB contiguous
//   mode 4  st.global.v4 from registers, lane = row, 64 B per lane                  (fp32-style path)
//   mode 5  st.global.v4 from registers, fully coalesced (512 B per warp instruction)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I image_classification_b200/csrc \
//             -o tools/ubench/store_rate tools/ubench/store_rate.cu
#include "cnx_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

using namespace cnx;

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                             const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                             CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                             CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeFn enc, void* base, uint64_t rows, uint64_t cols) {
  CUtensorMap m;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t str[1] = {cols * 2};
  cuuint32_t box[2] = {64, 32};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, str, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

__global__ void __launch_bounds__(256)
store_kernel(const __grid_constant__ CUtensorMap map0, const __grid_constant__ CUtensorMap map1,
             const __grid_constant__ CUtensorMap map_lin, __nv_bfloat16* d0, __nv_bfloat16* d1,
             int M, int N, int mode) {
  extern __shared__ unsigned char smem_dyn[];
  const uint32_t raw = smem_u32(smem_dyn);
  unsigned char* smem = smem_dyn + ((1024u - (raw & 1023u)) & 1023u);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* buf0 = smem + warp * 8192;
  unsigned char* buf1 = buf0 + 4096;
  for (int i = lane; i < 2048; i += 32) reinterpret_cast<uint32_t*>(buf0)[i] = 0x3f803f80u;
  fence_proxy_async();
  __syncwarp();
  const int q = warp & 3, half = warp >> 2;
  const int nblk = N / 256, tiles = (M / 128) * nblk;
  const uint4 val = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int m_blk = t / nblk, n_blk = t % nblk;
    const int row0 = m_blk * 128 + q * 32;
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      const int n = n_blk * 256 + half * 128 + ch * 64;
      if (mode <= 3) {
        if (lane == 0) tma_store_wait_read();
        __syncwarp();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (mode == 0 || mode == 1) {
            tma_store_2d(&map0, buf0, n, row0);
            if (mode == 1) tma_store_2d(&map1, buf1, n, row0);
          } else if (mode == 2) {
            // chunk index in a [M*N/64, 64] view: 32 consecutive 128-byte rows
            const long long chunk = (static_cast<long long>(row0 / 32) * (N / 64) + n / 64);
            tma_store_2d(&map_lin, buf0, 0, static_cast<int>(chunk * 32));
          } else {
            const long long chunk = (static_cast<long long>(row0 / 32) * (N / 64) + n / 64);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 4096;" ::"l"(
                             reinterpret_cast<uint64_t>(d0 + chunk * 2048)),
                         "r"(smem_u32(buf0))
                         : "memory");
          }
          tma_store_commit();
        }
      } else if (mode == 4) {
        __nv_bfloat16* p = d0 + static_cast<long long>(row0 + lane) * N + n;
#pragma unroll
        for (int k = 0; k < 8; ++k) *reinterpret_cast<uint4*>(p + k * 8) = val;
      } else {
        const long long chunk = (static_cast<long long>(row0 / 32) * (N / 64) + n / 64);
        __nv_bfloat16* p = d0 + chunk * 2048;
#pragma unroll
        for (int k = 0; k < 8; ++k) *reinterpret_cast<uint4*>(p + (k * 32 + lane) * 8) = val;
      }
    }
  }
  if (lane == 0) tma_store_wait_all();
}

int main() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  EncodeFn enc = reinterpret_cast<EncodeFn>(fn);
  const size_t flush_bytes = 256u << 20;
  void* flush;
  cudaMalloc(&flush, flush_bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int smem = 8 * 8192 + 1024;
  cudaFuncSetAttribute(store_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int shapes[3][2] = {{25088, 2048}, {100352, 1024}, {6272, 4096}};
  for (auto& sh : shapes) {
    const int M = sh[0], N = sh[1];
    __nv_bfloat16 *d0, *d1;
    cudaMalloc(&d0, (size_t)M * N * 2); cudaMalloc(&d1, (size_t)M * N * 2);
    CUtensorMap m0 = make_map(enc, d0, M, N), m1 = make_map(enc, d1, M, N);
    CUtensorMap ml = make_map(enc, d0, (uint64_t)M * N / 64, 64);
    for (int mode = 0; mode <= 5; ++mode) {
      for (int grid : {148, 296}) {
        std::vector<float> ts;
        for (int it = 0; it < 9; ++it) {
          cudaMemsetAsync(flush, 0, flush_bytes);
          cudaEventRecord(e0);
          store_kernel<<<grid, 256, smem>>>(m0, m1, ml, d0, d1, M, N, mode);
          cudaEventRecord(e1);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("mode %d: %s\n", mode, cudaGetErrorString(e)); return 1; }
          float ms; cudaEventElapsedTime(&ms, e0, e1);
          if (it >= 2) ts.push_back(ms);
        }
        std::sort(ts.begin(), ts.end());
        const double us = ts[ts.size() / 2] * 1e3;
        const double bytes = (double)M * N * 2 * (mode == 1 ? 2 : 1);
        printf("{\"M\": %d, \"N\": %d, \"mode\": %d, \"ctas\": %d, \"us\": %.1f, \"TB/s\": %.2f}\n", M, N,
               mode, grid, us, bytes / us * 1e-6);
      }
    }
    cudaFree(d0); cudaFree(d1);
  }
  return 0;
}
