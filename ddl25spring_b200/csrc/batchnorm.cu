// K11: BatchNorm 1d / 2d forward (train + eval) and backward, one CTA per channel.
//
// x is [N, C, S] (S = H*W, or 1 for BatchNorm1d): channel c of sample n is the contiguous run x[n][c][0..S).
//   train fwd : mean_c, var_c (biased) over N*S; y = (x - mean) * rstd * gamma + beta; running stats updated with
//               `momentum` (running_var gets the unbiased variance, like torch); mean / rstd saved for backward.
//   eval fwd  : y = (x - running_mean) * rsqrt(running_var + eps) * gamma + beta.
//   backward  : dgamma = sum dy * xhat, dbeta = sum dy,
//               dx = gamma * rstd * (dy - dbeta / M - xhat * dgamma / M),  M = N * S.
// Two statistics passes + one normalisation pass; the per-channel working set of the lab models (<= 256 x 1024
// floats) stays in L2 between passes.
// Replaces (reference): nn.BatchNorm1d in the tabular VAE / VFL-VAE (lab/tutorial_2a/generative-modeling.py:21-45,
// lab/tutorial_2b/exercise_3.py:17-23,40-46,63-81) and BatchNorm2d of ResNet-18 / DCGAN [north star].
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void channel_index(int64_t i, int S, int C, int c, int64_t& off) {
  const int64_t n = i / S, s = i % S;
  off = (n * C + c) * (int64_t)S + s;
}

__global__ void __launch_bounds__(kThreads) bn_fwd_train_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, float* __restrict__ y,
                                                                float* __restrict__ save_mean, float* __restrict__ save_rstd,
                                                                float* __restrict__ run_mean, float* __restrict__ run_var,
                                                                int N, int C, int S, float eps, float momentum) {
  __shared__ float red[32];
  const int c = blockIdx.x;
  const int64_t M = (int64_t)N * S;
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < M; i += kThreads) { int64_t o; channel_index(i, S, C, c, o); s += x[o]; }
  const float mean = ddl::block_sum(s, red) / (float)M;
  float v = 0.f;
  for (int64_t i = threadIdx.x; i < M; i += kThreads) { int64_t o; channel_index(i, S, C, c, o); const float d = x[o] - mean; v = fmaf(d, d, v); }
  const float var = ddl::block_sum(v, red) / (float)M;
  const float rstd = rsqrtf(var + eps);
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  for (int64_t i = threadIdx.x; i < M; i += kThreads) { int64_t o; channel_index(i, S, C, c, o); y[o] = (x[o] - mean) * rstd * g + b; }
  if (threadIdx.x == 0) {
    save_mean[c] = mean;
    save_rstd[c] = rstd;
    if (run_mean) {
      const float unbiased = M > 1 ? var * (float)M / (float)(M - 1) : var;
      run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
      run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
    }
  }
}

__global__ void __launch_bounds__(kThreads) bn_fwd_eval_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, const float* __restrict__ run_mean,
                                                               const float* __restrict__ run_var, float* __restrict__ y, int64_t total,
                                                               int C, int S, float eps) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)((i / S) % C);
    const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    y[i] = (x[i] - run_mean[c]) * rsqrtf(run_var[c] + eps) * g + b;
  }
}

__global__ void __launch_bounds__(kThreads) bn_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                          const float* __restrict__ gamma, const float* __restrict__ save_mean,
                                                          const float* __restrict__ save_rstd, float* __restrict__ dx,
                                                          float* __restrict__ dgamma, float* __restrict__ dbeta, int N, int C, int S) {
  __shared__ float red[32];
  const int c = blockIdx.x;
  const int64_t M = (int64_t)N * S;
  const float mean = save_mean[c], rstd = save_rstd[c];
  float s1 = 0.f, s2 = 0.f;
  for (int64_t i = threadIdx.x; i < M; i += kThreads) {
    int64_t o; channel_index(i, S, C, c, o);
    const float g = dy[o];
    s1 += g;
    s2 = fmaf(g, (x[o] - mean) * rstd, s2);
  }
  const float db = ddl::block_sum(s1, red);
  const float dg = ddl::block_sum(s2, red);
  const float gam = gamma ? gamma[c] : 1.f;
  const float k = gam * rstd, mb = db / (float)M, mg = dg / (float)M;
  for (int64_t i = threadIdx.x; i < M; i += kThreads) {
    int64_t o; channel_index(i, S, C, c, o);
    dx[o] = k * (dy[o] - mb - (x[o] - mean) * rstd * mg);
  }
  if (threadIdx.x == 0) {
    if (dgamma) dgamma[c] = dg;
    if (dbeta) dbeta[c] = db;
  }
}

}  // namespace

extern "C" {

int ddl_batchnorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* save_mean, float* save_rstd,
                      float* running_mean, float* running_var, int N, int C, int S, float eps, float momentum, int training,
                      void* stream) {
  DDL_REQUIRE(N > 0 && C > 0 && S > 0, "empty BatchNorm input");
  cudaStream_t st = (cudaStream_t)stream;
  if (training) {
    DDL_REQUIRE(save_mean && save_rstd, "training mode needs save_mean / save_rstd buffers");
    bn_fwd_train_kernel<<<C, kThreads, 0, st>>>(x, gamma, beta, y, save_mean, save_rstd, running_mean, running_var, N, C, S, eps, momentum);
  } else {
    DDL_REQUIRE(running_mean && running_var, "eval mode needs running statistics");
    const int64_t total = (int64_t)N * C * S;
    int64_t blocks = (total + kThreads - 1) / kThreads;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    bn_fwd_eval_kernel<<<(int)blocks, kThreads, 0, st>>>(x, gamma, beta, running_mean, running_var, y, total, C, S, eps);
  }
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_batchnorm_bwd(const float* x, const float* dy, const float* gamma, const float* save_mean, const float* save_rstd, float* dx,
                      float* dgamma, float* dbeta, int N, int C, int S, void* stream) {
  DDL_REQUIRE(N > 0 && C > 0 && S > 0, "empty BatchNorm input");
  bn_bwd_kernel<<<C, kThreads, 0, (cudaStream_t)stream>>>(x, dy, gamma, save_mean, save_rstd, dx, dgamma, dbeta, N, C, S);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
