// K16 (part) / K11: RMSNorm forward + backward, one warp per row.
//   y = x * rsqrt(mean(x^2) + eps) * w
//   dx = r * (w*dy) - x * r^3 * mean(x * w*dy) ,  dw += sum_rows dy * x * r      (r = rsqrt(mean(x^2)+eps))
// Replaces (reference): the RMSNorm inside simplellm's LLaMA blocks (lab/tutorial_1b/primer/intro.py:17,28).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

__global__ void __launch_bounds__(256) rmsnorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          float* __restrict__ y, float* __restrict__ rstd, int64_t rows, int d,
                                                          float eps) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const float* xr = x + row * d;
  float ss = 0.f;
  for (int i = lane; i < d; i += 32) { const float v = xr[i]; ss = fmaf(v, v, ss); }
  ss = ddl::warp_sum(ss);
  const float r = rsqrtf(ss / (float)d + eps);
  if (lane == 0 && rstd) rstd[row] = r;
  float* yr = y + row * d;
  for (int i = lane; i < d; i += 32) yr[i] = xr[i] * r * w[i];
}

__global__ void __launch_bounds__(256) rmsnorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          const float* __restrict__ dy, const float* __restrict__ rstd,
                                                          float* __restrict__ dx, float* __restrict__ dw, int64_t rows, int d) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const float* xr = x + row * d;
  const float* gr = dy + row * d;
  const float r = rstd[row];
  float dot = 0.f;
  for (int i = lane; i < d; i += 32) dot = fmaf(xr[i], w[i] * gr[i], dot);
  dot = ddl::warp_sum(dot);
  const float c = r * r * r * dot / (float)d;
  float* dxr = dx + row * d;
  for (int i = lane; i < d; i += 32) {
    dxr[i] = r * w[i] * gr[i] - xr[i] * c;
    atomicAdd(dw + i, gr[i] * xr[i] * r);
  }
}

}  // namespace

extern "C" {

int ddl_rmsnorm_fwd(const float* x, const float* w, float* y, float* rstd, int64_t rows, int d, float eps, void* stream) {
  if (rows == 0) return DDL_OK;
  const int64_t blocks = (rows + 7) / 8;
  DDL_CUDA_CHECK(ddl_launch_pdl(rmsnorm_fwd_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, x, w, y, rstd, rows, d, eps));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_rmsnorm_bwd(const float* x, const float* w, const float* dy, const float* rstd, float* dx, float* dw, int64_t rows, int d,
                    void* stream) {
  if (rows == 0) return DDL_OK;
  const int64_t blocks = (rows + 7) / 8;
  DDL_CUDA_CHECK(ddl_launch_pdl(rmsnorm_bwd_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, x, w, dy, rstd, dx, dw, rows, d));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
