// K15: optimizer steps over flat fp32 arenas (one launch updates a whole model, or the whole
// [G, size] stack of the clients packed on this GPU).
// Replaces (reference): torch.optim.SGD at lab/tutorial_1a/hfl_complete.py:196,265,306,319;
// Adam at lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:33,67; AdamW at lab/tutorial_2b/vfl.py:50,82.
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                  float* __restrict__ buf, float lr, float wd, float mom,
                                                  float gscale, int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    float4 w = reinterpret_cast<float4*>(p)[v];
    float4 d = ddl::ld_stream(reinterpret_cast<const float4*>(g) + v);
    d.x = fmaf(wd, w.x, gscale * d.x); d.y = fmaf(wd, w.y, gscale * d.y);
    d.z = fmaf(wd, w.z, gscale * d.z); d.w = fmaf(wd, w.w, gscale * d.w);
    if (buf != nullptr) {
      float4 b = reinterpret_cast<float4*>(buf)[v];
      b.x = fmaf(mom, b.x, d.x); b.y = fmaf(mom, b.y, d.y); b.z = fmaf(mom, b.z, d.z); b.w = fmaf(mom, b.w, d.w);
      reinterpret_cast<float4*>(buf)[v] = b;
      d = b;
    }
    w.x = fmaf(-lr, d.x, w.x); w.y = fmaf(-lr, d.y, w.y); w.z = fmaf(-lr, d.z, w.z); w.w = fmaf(-lr, d.w, w.w);
    reinterpret_cast<float4*>(p)[v] = w;
  }
}

struct AdamArgs {
  float lr, b1, b2, eps, wd, bc1, bc2_rsqrt, gscale;
  int decoupled;
};

__device__ __forceinline__ void adam1(float& w, float g, float& m, float& v, const AdamArgs& a) {
  g *= a.gscale;
  if (a.decoupled) w *= (1.f - a.lr * a.wd);
  else g = fmaf(a.wd, w, g);
  m = fmaf(a.b1, m, (1.f - a.b1) * g);
  v = fmaf(a.b2, v, (1.f - a.b2) * g * g);
  const float denom = sqrtf(v) * a.bc2_rsqrt + a.eps;
  w -= (a.lr / a.bc1) * (m / denom);
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, AdamArgs a,
                                                   int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 w = reinterpret_cast<float4*>(p)[i];
    const float4 d = ddl::ld_stream(reinterpret_cast<const float4*>(g) + i);
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    adam1(w.x, d.x, mm.x, vv.x, a); adam1(w.y, d.y, mm.y, vv.y, a);
    adam1(w.z, d.z, mm.z, vv.z, a); adam1(w.w, d.w, mm.w, vv.w, a);
    reinterpret_cast<float4*>(p)[i] = w;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}

// Same update with the step counter read from device memory: inside a CUDA graph the counter cannot be a by-value argument.
__global__ void __launch_bounds__(256) adam_devstep_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                           float* __restrict__ v, AdamArgs a, const int* __restrict__ step_ptr,
                                                           int64_t nvec) {
  pdl_launch_dependents();
  pdl_wait();
  const float step = (float)max(1, *step_ptr);
  a.bc1 = 1.f - powf(a.b1, step);
  a.bc2_rsqrt = 1.f / sqrtf(1.f - powf(a.b2, step));
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 w = reinterpret_cast<float4*>(p)[i];
    const float4 d = ddl::ld_stream(reinterpret_cast<const float4*>(g) + i);
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    adam1(w.x, d.x, mm.x, vv.x, a); adam1(w.y, d.y, mm.y, vv.y, a);
    adam1(w.z, d.z, mm.z, vv.z, a); adam1(w.w, d.w, mm.w, vv.w, a);
    reinterpret_cast<float4*>(p)[i] = w;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}

int grid_for(int64_t nvec) {
  int64_t g = (nvec + 255) / 256;
  if (g > (int64_t)kNumSMs * 8) g = (int64_t)kNumSMs * 8;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace

extern "C" {

int ddl_sgd_step(float* params, const float* grads, float* momentum_buf, int64_t n, float lr, float weight_decay,
                 float momentum, float grad_scale, void* stream) {
  DDL_REQUIRE(n % 4 == 0, "n must be a multiple of 4 floats (arenas are padded)");
  if (n == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(sgd_kernel, dim3(grid_for(n / 4)), dim3(256), 0, (cudaStream_t)stream, params, grads, momentum != 0.f ? momentum_buf : nullptr,
                                                                 lr, weight_decay, momentum, grad_scale, n / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, int step,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled,
                  float grad_scale, void* stream) {
  DDL_REQUIRE(n % 4 == 0, "n must be a multiple of 4 floats (arenas are padded)");
  DDL_REQUIRE(step >= 1, "Adam step counter starts at 1");
  if (n == 0) return DDL_OK;
  AdamArgs a;
  a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps; a.wd = weight_decay; a.decoupled = decoupled;
  a.gscale = grad_scale;
  a.bc1 = 1.f - powf(beta1, (float)step);
  a.bc2_rsqrt = 1.f / sqrtf(1.f - powf(beta2, (float)step));
  DDL_CUDA_CHECK(ddl_launch_pdl(adam_kernel, dim3(grid_for(n / 4)), dim3(256), 0, (cudaStream_t)stream, params, grads, exp_avg, exp_avg_sq, a, n / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const int* step_dev,
                      float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled, float grad_scale,
                      void* stream) {
  DDL_REQUIRE(n % 4 == 0, "n must be a multiple of 4 floats (arenas are padded)");
  DDL_REQUIRE(step_dev != nullptr, "device step counter is NULL");
  if (n == 0) return DDL_OK;
  AdamArgs a;
  a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps; a.wd = weight_decay; a.decoupled = decoupled;
  a.gscale = grad_scale;
  a.bc1 = 1.f; a.bc2_rsqrt = 1.f;
  DDL_CUDA_CHECK(ddl_launch_pdl(adam_devstep_kernel, dim3(grid_for(n / 4)), dim3(256), 0, (cudaStream_t)stream, params, grads, exp_avg, exp_avg_sq, a, step_dev, n / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
