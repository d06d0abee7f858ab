// Microbenchmark: issue rate of scalar FFMA vs packed FFMA2 (fma.rn.f32x2) vs a mix, per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_rate fma_rate.cu && ./fma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(d))
      : "l"(reinterpret_cast<uint64_t&>(a)), "l"(reinterpret_cast<uint64_t&>(b)), "l"(reinterpret_cast<uint64_t&>(c)));
  return d;
}
__device__ __forceinline__ float ffma1(float a, float b, float c) {
  float d;
  asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
template <int MODE>
__global__ void k(float* out, int iters, float s, long long* cyc) {
  float2 a[8]; float b[16];
  for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  for (int i = 0; i < 16; ++i) b[i] = threadIdx.x * 0.002f + i;
  float2 m = make_float2(s, s * 0.5f), c = make_float2(0.25f, 0.125f);
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {            // 16 scalar FFMA
#pragma unroll
      for (int i = 0; i < 16; ++i) b[i] = ffma1(b[i], s, 0.25f + 0.f * i);
    } else if (MODE == 1) {     // 8 FFMA2 (= 16 FMAs)
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = ffma2(a[i], m, c);
    } else if (MODE == 2) {     // 8 FFMA2 + 16 FFMA interleaved (= 32 FMAs)
#pragma unroll
      for (int i = 0; i < 8; ++i) { a[i] = ffma2(a[i], m, c); b[2*i] = ffma1(b[2*i], s, 0.25f); b[2*i+1] = ffma1(b[2*i+1], s, 0.25f); }
    } else {                    // scalar FFMA with register (non-immediate) addend
#pragma unroll
      for (int i = 0; i < 16; ++i) b[i] = ffma1(b[i], s, m.y);
    }
  }
  long long t1 = clock64();
  float r = 0;
  for (int i = 0; i < 8; ++i) r += a[i].x + a[i].y;
  for (int i = 0; i < 16; ++i) r += b[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  const int iters = 20000;
  const char* names[4] = {"16xFFMA(imm c)", "8xFFMA2", "8xFFMA2+16xFFMA", "16xFFMA(reg c)"};
  for (int threads : {128, 256, 512, 1024}) {
    for (int mode = 0; mode < 4; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, threads>>>(out, iters, 1.0001f, cyc);
        if (mode == 1) k<1><<<148, threads>>>(out, iters, 1.0001f, cyc);
        if (mode == 2) k<2><<<148, threads>>>(out, iters, 1.0001f, cyc);
        if (mode == 3) k<3><<<148, threads>>>(out, iters, 1.0001f, cyc);
        cudaDeviceSynchronize();
      }
      const double fmas = (mode == 2 ? 32.0 : 16.0) * iters * threads;
      printf("threads/SM=%4d %-18s cycles=%lld  FMA/clk/SM=%.1f\n", threads, names[mode], *cyc, fmas / *cyc);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
