// prep.cu — weight folding before the block's GEMMs and its adjoint after them.
//
// The block computes   y = x + gamma * (fc2(gelu(fc1(ln_w * xhat + ln_b))))
// Both diagonal scalings are folded into the GEMM weights once per weight version so that no
// activation-sized elementwise pass exists for them:
//     W1f = fc1_w * diag(ln_w)      b1f = fc1_b + fc1_w @ ln_b
//     W2f = diag(gamma) * fc2_w     b2f = gamma * fc2_b
// (bf16 copies in both orientations: the data-gradient GEMMs want the transposed weight K-major).
// unfold_block_grads applies the exact adjoint of that map to the gradients of the folded
// parameters, producing gradients for the nine timm-named parameters of the block.
#include "cnx_common.cuh"
#include "cnx_kernels.h"

namespace cnx {

namespace {

// in [R][Cc] fp32 (row-major) -> out [R][Cc] bf16 and outT [Cc][R] bf16, scaled by
// row_scale[r] (nullable) and col_scale[c] (nullable).  32x32 tile (bx, by) through padded smem.
__device__ __forceinline__ void scale_cast_transpose_tile(
    float (*tile)[33], int bx, int by, const float* __restrict__ in,
    const float* __restrict__ row_scale, const float* __restrict__ col_scale,
    __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ outT, int R, int Cc) {
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int c = bx * 32 + tx;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int r = by * 32 + ty + i;
    float v = 0.f;
    if (r < R && c < Cc) {
      v = in[static_cast<size_t>(r) * Cc + c];
      if (row_scale) v *= row_scale[r];
      if (col_scale) v *= col_scale[c];
      out[static_cast<size_t>(r) * Cc + c] = __float2bfloat16_rn(v);
    }
    tile[ty + i][tx] = v;
  }
  __syncthreads();
  const int r2 = by * 32 + tx;  // transposed: threads run along original rows
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    const int c2 = bx * 32 + ty + i;
    if (r2 < R && c2 < Cc)
      outT[static_cast<size_t>(c2) * R + r2] = __float2bfloat16_rn(tile[tx][ty + i]);
  }
}

struct PrepArgs {
  const float *dw_w, *ln_w, *ln_b, *fc1_w, *fc1_b, *fc2_w, *fc2_b, *gamma;
  float* wt_dw;
  __nv_bfloat16 *w1f, *w1fT, *w2f, *w2fT;
  float *b1f, *b2f;
  float2* rb;                // [4C] (rowsum_c bf16(W1f[j,c]), -b1f[j]): LayerNorm-backward by linearity
  int C;
  int n_t1, n_t2, n_b1;      // CTA counts of the first three sections
};

// The whole fold of one block in ONE launch (four launches until round 2: ~1 ms per optimizer
// step over 36 blocks, most of it launch overhead).  CTA sections, in order:
//   [0, n_t1)          fc1_w [4C][C] tiles, columns scaled by ln_w  -> w1f, w1fT
//   [n_t1, +n_t2)      fc2_w [C][4C] tiles, rows scaled by gamma    -> w2f, w2fT
//   [.., +n_b1)        b1f[j] = fc1_b[j] + sum_c fc1_w[j,c] * ln_b[c]   (one warp per row j)
//   the rest           wt_dw[t][c] = dw_w[c][t];  b2f[c] = gamma[c] * fc2_b[c]
__global__ void __launch_bounds__(256) prep_block_kernel(const PrepArgs a) {
  __shared__ float tile[32][33];
  const int C = a.C, R4 = 4 * a.C;
  int b = blockIdx.x;
  if (b < a.n_t1) {
    const int nbx = (C + 31) / 32;
    scale_cast_transpose_tile(tile, b % nbx, b / nbx, a.fc1_w, nullptr, a.ln_w, a.w1f, a.w1fT, R4, C);
    return;
  }
  b -= a.n_t1;
  if (b < a.n_t2) {
    const int nbx = (R4 + 31) / 32;
    scale_cast_transpose_tile(tile, b % nbx, b / nbx, a.fc2_w, a.gamma, nullptr, a.w2f, a.w2fT, C, R4);
    return;
  }
  b -= a.n_t2;
  if (b < a.n_b1) {
    const int row = b * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= R4) return;
    float s = 0.f, r = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float w = a.fc1_w[static_cast<size_t>(row) * C + c];
      s = fmaf(w, a.ln_b[c], s);
      r += __bfloat162float(__float2bfloat16_rn(w * a.ln_w[c]));   // the value stored in w1f
    }
    s = warp_sum(s);
    r = warp_sum(r);
    if (lane == 0) {
      const float b = a.fc1_b[row] + s;
      a.b1f[row] = b;
      if (a.rb != nullptr) a.rb[row] = make_float2(r, -b);
    }
    return;
  }
  b -= a.n_b1;
  const int i = b * 256 + threadIdx.x;
  if (i < 49 * C) {
    const int t = i / C, c = i % C;
    a.wt_dw[i] = a.dw_w[c * 49 + t];
  }
  if (i < C) a.b2f[i] = a.gamma[i] * a.fc2_b[i];
}

// fc1 side of the adjoint.  grid (ceil(C/32), row_splits), block 32x8.
//   d_fc1_w[j,c] = dW1f[j,c]*ln_w[c] + db1f[j]*ln_b[c]
//   d_ln_w[c]   += sum_j dW1f[j,c]*fc1_w[j,c]        d_ln_b[c] += sum_j db1f[j]*fc1_w[j,c]
__global__ void __launch_bounds__(256)
unfold_fc1_kernel(const float* __restrict__ dw1f, const float* __restrict__ db1f,
                  const float* __restrict__ fc1_w, const float* __restrict__ ln_w,
                  const float* __restrict__ ln_b, float* __restrict__ d_fc1_w,
                  float* __restrict__ d_ln_w, float* __restrict__ d_ln_b, int R, int Cc,
                  int rows_per_block) {
  __shared__ float s1[8][33], s2[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  const int r_begin = blockIdx.y * rows_per_block;
  const int r_end = min(R, r_begin + rows_per_block);
  float a1 = 0.f, a2 = 0.f;
  if (c < Cc) {
    const float g = ln_w[c], bt = ln_b[c];
    for (int r = r_begin + ty; r < r_end; r += 8) {
      const size_t idx = static_cast<size_t>(r) * Cc + c;
      const float dw = dw1f[idx], w = fc1_w[idx], db = db1f[r];
      d_fc1_w[idx] = fmaf(dw, g, db * bt);
      a1 = fmaf(dw, w, a1);
      a2 = fmaf(db, w, a2);
    }
  }
  s1[ty][tx] = a1;
  s2[ty][tx] = a2;
  __syncthreads();
  if (ty == 0 && c < Cc) {
#pragma unroll
    for (int i = 1; i < 8; ++i) { a1 += s1[i][tx]; a2 += s2[i][tx]; }
    atomicAdd(d_ln_w + c, a1);
    atomicAdd(d_ln_b + c, a2);
  }
}

// fc2 side of the adjoint: one CTA per output channel c (row of fc2_w [C][4C]).
//   d_fc2_w[c,j] = gamma[c]*dW2f[c,j]   d_fc2_b[c] = gamma[c]*db2f[c]
//   d_gamma[c]   = sum_j dW2f[c,j]*fc2_w[c,j] + db2f[c]*fc2_b[c]
// One launch for everything but the fc1 side: CTAs [0, C) = fc2 rows (as above), the CTAs after
// them transpose the depthwise taps (d_dw_w[c][t] = dwt_dw[t][c]) and zero d_ln_w / d_ln_b, which
// the fc1 kernel launched next accumulates into.
__global__ void __launch_bounds__(256)
unfold_fc2_dw_kernel(const float* __restrict__ dw2f, const float* __restrict__ db2f,
                     const float* __restrict__ fc2_w, const float* __restrict__ fc2_b,
                     const float* __restrict__ gamma, float* __restrict__ d_fc2_w,
                     float* __restrict__ d_fc2_b, float* __restrict__ d_gamma,
                     const float* __restrict__ dwt, float* __restrict__ d_dw_w,
                     float* __restrict__ d_ln_w, float* __restrict__ d_ln_b, int Cc, int R4) {
  __shared__ float sred[8];
  if (static_cast<int>(blockIdx.x) >= Cc) {
    const int i = (blockIdx.x - Cc) * 256 + threadIdx.x;
    if (i < 49 * Cc) {
      const int c = i / 49, t = i % 49;
      d_dw_w[i] = dwt[t * Cc + c];
    }
    if (i < Cc) {
      d_ln_w[i] = 0.f;
      d_ln_b[i] = 0.f;
    }
    return;
  }
  const int c = blockIdx.x;
  const float g = gamma[c];
  float a = 0.f;
  for (int j = threadIdx.x; j < R4; j += 256) {
    const size_t idx = static_cast<size_t>(c) * R4 + j;
    const float dw = dw2f[idx];
    d_fc2_w[idx] = g * dw;
    a = fmaf(dw, fc2_w[idx], a);
  }
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sred[i];
    d_gamma[c] = fmaf(db2f[c], fc2_b[c], t);
    d_fc2_b[c] = g * db2f[c];
  }
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("%s launch: %s", what, cudaGetErrorString(e));
    return CNX_ERR_LAUNCH;
  }
  return CNX_OK;
}

}  // namespace

int prep_block_weights(const float* dw_w, const float* ln_w, const float* ln_b, const float* fc1_w,
                       const float* fc1_b, const float* fc2_w, const float* fc2_b,
                       const float* ls_gamma, float* wt_dw, void* w1f, void* w1fT, float* b1f,
                       void* w2f, void* w2fT, float* b2f, float* rb, int C, cudaStream_t stream) {
  if (C <= 0 || (C % 8) != 0) {
    set_last_error("prep_block_weights: C=%d must be a positive multiple of 8", C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int R4 = 4 * C;
  PrepArgs a;
  a.dw_w = dw_w; a.ln_w = ln_w; a.ln_b = ln_b; a.fc1_w = fc1_w; a.fc1_b = fc1_b; a.fc2_w = fc2_w;
  a.fc2_b = fc2_b; a.gamma = ls_gamma; a.wt_dw = wt_dw;
  a.w1f = static_cast<__nv_bfloat16*>(w1f); a.w1fT = static_cast<__nv_bfloat16*>(w1fT);
  a.w2f = static_cast<__nv_bfloat16*>(w2f); a.w2fT = static_cast<__nv_bfloat16*>(w2fT);
  a.b1f = b1f; a.b2f = b2f; a.rb = reinterpret_cast<float2*>(rb); a.C = C;
  a.n_t1 = ((C + 31) / 32) * ((R4 + 31) / 32);
  a.n_t2 = a.n_t1;
  a.n_b1 = (R4 + 7) / 8;
  const int n_small = (49 * C + 255) / 256;
  prep_block_kernel<<<a.n_t1 + a.n_t2 + a.n_b1 + n_small, 256, 0, stream>>>(a);
  return check_launch("prep_block_weights");
}

int unfold_block_grads(const float* dw1f, const float* db1f, const float* dw2f, const float* db2f,
                       const float* dwt_dw, const float* ln_w, const float* ln_b,
                       const float* fc1_w, const float* fc2_w, const float* fc2_b,
                       const float* ls_gamma, float* d_dw_w, float* d_ln_w, float* d_ln_b,
                       float* d_fc1_w, float* d_fc2_w, float* d_fc2_b, float* d_ls_gamma, int C,
                       cudaStream_t stream) {
  if (C <= 0 || (C % 8) != 0) {
    set_last_error("unfold_block_grads: C=%d must be a positive multiple of 8", C);
    return CNX_ERR_BAD_SHAPE;
  }
  const int R4 = 4 * C;
  unfold_fc2_dw_kernel<<<C + (49 * C + 255) / 256, 256, 0, stream>>>(
      dw2f, db2f, fc2_w, fc2_b, ls_gamma, d_fc2_w, d_fc2_b, d_ls_gamma, dwt_dw, d_dw_w, d_ln_w,
      d_ln_b, C, R4);
  {
    const int rows_per_block = 128;
    dim3 grid((C + 31) / 32, (R4 + rows_per_block - 1) / rows_per_block);
    unfold_fc1_kernel<<<grid, 256, 0, stream>>>(dw1f, db1f, fc1_w, ln_w, ln_b, d_fc1_w, d_ln_w,
                                                d_ln_b, R4, C, rows_per_block);
  }
  return check_launch("unfold_block_grads");
}

}  // namespace cnx
