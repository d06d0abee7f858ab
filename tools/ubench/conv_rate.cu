// Microbenchmark: register-resident 1x7 conv inner loop (acc[j] += in[j+kw] * w[kw]) with packed
// FFMA2 vs scalar FFMA, to see what the register file lets each sustain.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(d))
      : "l"(reinterpret_cast<uint64_t&>(a)), "l"(reinterpret_cast<uint64_t&>(b)), "l"(reinterpret_cast<uint64_t&>(c)));
  return d;
}
template <int MODE>
__global__ void k(float* out, int iters, float s, long long* cyc) {
  float2 in[13], w[7], acc[2][7];
  for (int i = 0; i < 13; ++i) in[i] = make_float2(threadIdx.x * 0.001f + i * s, i * 0.5f * s);
  for (int i = 0; i < 7; ++i) w[i] = make_float2(0.01f * i * s, 0.02f * i * s);
  for (int o = 0; o < 2; ++o) for (int i = 0; i < 7; ++i) acc[o][i] = make_float2(0.f, 0.f);
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {          // packed, taps as operand B (as in dwconv7.cu)
#pragma unroll
      for (int o = 0; o < 2; ++o)
#pragma unroll
      for (int kw = 0; kw < 7; ++kw)
#pragma unroll
        for (int j = 0; j < 7; ++j) acc[o][j] = ffma2(in[j + kw], w[kw], acc[o][j]);
    } else if (MODE == 1) {   // scalar
#pragma unroll
      for (int o = 0; o < 2; ++o)
#pragma unroll
      for (int kw = 0; kw < 7; ++kw)
#pragma unroll
        for (int j = 0; j < 7; ++j) {
          acc[o][j].x = fmaf(in[j + kw].x, w[kw].x, acc[o][j].x);
          acc[o][j].y = fmaf(in[j + kw].y, w[kw].y, acc[o][j].y);
        }
    } else if (MODE == 2) {   // packed, j outer / kw inner (accumulator chain dependent)
#pragma unroll
      for (int o = 0; o < 2; ++o)
#pragma unroll
      for (int j = 0; j < 7; ++j)
#pragma unroll
        for (int kw = 0; kw < 7; ++kw) acc[o][j] = ffma2(in[j + kw], w[kw], acc[o][j]);
    } else {                  // packed, taps as operand A
#pragma unroll
      for (int o = 0; o < 2; ++o)
#pragma unroll
      for (int kw = 0; kw < 7; ++kw)
#pragma unroll
        for (int j = 0; j < 7; ++j) acc[o][j] = ffma2(w[kw], in[j + kw], acc[o][j]);
    }
    in[it % 13].x += 1e-9f;   // keep the loop body from being hoisted
  }
  long long t1 = clock64();
  float r = 0;
  for (int o = 0; o < 2; ++o) for (int i = 0; i < 7; ++i) r += acc[o][i].x + acc[o][i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  const int iters = 4000;
  const char* names[4] = {"FFMA2 taps=B", "scalar FFMA", "FFMA2 j-outer", "FFMA2 taps=A"};
  for (int threads : {256, 512}) {
    for (int mode = 0; mode < 4; ++mode) {
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, threads>>>(out, iters, 1.0001f, cyc);
        if (mode == 1) k<1><<<148, threads>>>(out, iters, 1.0001f, cyc);
        if (mode == 2) k<2><<<148, threads>>>(out, iters, 1.0001f, cyc);
        if (mode == 3) k<3><<<148, threads>>>(out, iters, 1.0001f, cyc);
        cudaDeviceSynchronize();
      }
      const double fmas = 2.0 * 98 * iters * threads;
      printf("threads/SM=%4d %-16s cycles=%lld  FMA/clk/SM=%.1f\n", threads, names[mode], *cyc, fmas / *cyc);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
