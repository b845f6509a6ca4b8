// Small layout / reduction helpers that keep the tabular model families (heart-disease MLP, VFL bottom / top models, the
// VAEs: layer widths 2 .. 256, most of them not multiples of 32) on the tcgen05 GEMM:
//   pad2d   : dst[r, c] = src[r, c] for c < cols, `fill` in column `fill_col` (if >= 0), 0 elsewhere - one launch builds the
//             zero-padded (and bias-augmented: a column of ones in X / the bias vector in W) operand the GEMM needs.
//   col_sum : out[c] = sum_r x[r, c]  (bias gradients), deterministic two-stage tree per column block.
// Replaces (reference): nothing explicit - ATen pads / reduces implicitly inside nn.Linear (lab/tutorial_2a/centralized.py:24-28,
// lab/tutorial_2b/vfl.py:21-22,37-39, lab/tutorial_2a/generative-modeling.py:50-75).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

__global__ void __launch_bounds__(256) pad2d_kernel(const float* __restrict__ src, int64_t ld_src, float* __restrict__ dst, int64_t ld_dst,
                                                    int rows, int cols, int rows_dst, int cols_dst, int fill_col,
                                                    const float* __restrict__ fill_vec, float fill_val) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t total = (int64_t)rows_dst * cols_dst;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols_dst), c = (int)(i % cols_dst);
    float v = 0.f;
    if (r < rows) {
      if (c < cols) v = src[(int64_t)r * ld_src + c];
      else if (c == fill_col) v = fill_vec ? fill_vec[r] : fill_val;
    }
    dst[(int64_t)r * ld_dst + c] = v;
  }
}

// one block per 32 columns; 8 warps stride over the rows, then a shared-memory tree (fixed order: deterministic)
__global__ void __launch_bounds__(256) col_sum_kernel(const float* __restrict__ x, int64_t ld, int rows, int cols, float* __restrict__ out,
                                                      float alpha, float beta) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float part[8][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < cols)
    for (int r = w; r < rows; r += 8) s += x[(int64_t)r * ld + c];
  part[w][lane] = s;
  __syncthreads();
  if (w == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][lane];
    out[c] = alpha * t + (beta != 0.f ? beta * out[c] : 0.f);
  }
}

}  // namespace

extern "C" {

int ddl_pad2d(const float* src, int64_t ld_src, int rows, int cols, float* dst, int64_t ld_dst, int rows_dst, int cols_dst,
              int fill_col, const float* fill_vec, float fill_val, void* stream) {
  DDL_REQUIRE(rows >= 0 && cols >= 0 && rows_dst >= rows && cols_dst >= cols && ld_dst >= cols_dst, "bad pad2d shapes");
  DDL_REQUIRE(fill_col < cols_dst, "fill column outside the destination");
  if (rows_dst == 0 || cols_dst == 0) return DDL_OK;
  int64_t blocks = ((int64_t)rows_dst * cols_dst + 255) / 256;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  DDL_CUDA_CHECK(ddl_launch_pdl(pad2d_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, src, ld_src, dst, ld_dst, rows,
                                cols, rows_dst, cols_dst, fill_col, fill_vec, fill_val));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_col_sum(const float* x, int64_t ld, int rows, int cols, float* out, float alpha, float beta, void* stream) {
  DDL_REQUIRE(rows >= 0 && cols >= 1 && ld >= cols, "bad col_sum shapes");
  DDL_CUDA_CHECK(ddl_launch_pdl(col_sum_kernel, dim3((cols + 31) / 32), dim3(256), 0, (cudaStream_t)stream, x, ld, rows, cols, out, alpha,
                                beta));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
