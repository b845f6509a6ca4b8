// K13 / K14 / K17: fused loss kernels (forward value + input gradient in one pass) and top-1 eval.
// Replaces (reference): F.log_softmax + F.nll_loss (lab/tutorial_1a/hfl_complete.py:62,78,251);
// CrossEntropyLoss with float one-hot targets (lab/tutorial_2b/vfl.py:51,79); causalLLMLoss
// (lab/tutorial_1b/primer/intro.py:29); MSE(sum)+KL (lab/tutorial_2a/generative-modeling.py:121-130,
// lab/tutorial_2b/exercise_3.py:140-147); argmax/eq/sum accuracy (hfl_complete.py:172-183).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

// One warp per row; C arbitrary (strided over lanes). target_idx (int64) or target_prob (float[N,C]).
// loss_sum += sum_rows( -sum_c t_c * logp_c ) * loss_scale  ;  grad = (softmax * sum_t - t) * grad_scale
// rows with target index == ignore_index contribute nothing.
__global__ void __launch_bounds__(256) softmax_xent_kernel(const float* __restrict__ logits, int64_t ld,
                                                           const int64_t* __restrict__ tidx, int64_t tstride,
                                                           const float* __restrict__ tprob, float* __restrict__ grad,
                                                           float* __restrict__ loss_sum, float* __restrict__ n_valid,
                                                           int64_t N, int C, float loss_scale, float grad_scale,
                                                           int64_t ignore_index) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  float local_loss = 0.f, local_cnt = 0.f;
  for (int64_t r = warp; r < N; r += nwarps) {
    const float* x = logits + r * ld;
    float* gx = grad ? grad + r * ld : nullptr;
    int64_t t = -1;
    if (tidx) {
      t = tidx[r * tstride];
      if (t == ignore_index) {
        if (gx) for (int c = lane; c < C; c += 32) gx[c] = 0.f;
        continue;
      }
    }
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, x[c]);
    mx = ddl::warp_max(mx);
    float se = 0.f, st = 0.f, stx = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float v = x[c];
      se += __expf(v - mx);
      if (tprob) { const float tp = tprob[r * C + c]; st += tp; stx += tp * v; }
    }
    se = ddl::warp_sum(se);
    const float lse = mx + __logf(se);
    if (tprob) { st = ddl::warp_sum(st); stx = ddl::warp_sum(stx); }
    else { st = 1.f; stx = x[t]; }
    if (lane == 0) { local_loss += st * lse - stx; local_cnt += 1.f; }
    if (gx) {
      for (int c = lane; c < C; c += 32) {
        const float p = __expf(x[c] - lse);
        const float tc = tprob ? tprob[r * C + c] : (c == t ? 1.f : 0.f);
        gx[c] = (p * st - tc) * grad_scale;
      }
    }
  }
  if (lane == 0 && (local_loss != 0.f || local_cnt != 0.f)) {
    atomicAdd(loss_sum, local_loss * loss_scale);
    if (n_valid) atomicAdd(n_valid, local_cnt);
  }
}

// Wide rows (the LLaMA vocabulary: C = 32 000): one BLOCK per row with 128-bit accesses - with one warp per row the 767 rows of
// the lab batch keep only 96 CTAs busy, each lane walking 1000 scalar loads three times (259 us measured).  Index targets only.
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = ddl::warp_max(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (lane < nw) ? red[lane] : -INFINITY;
  return ddl::warp_max(v);
}

__global__ void __launch_bounds__(256) softmax_xent_wide_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ tidx,
                                                                int64_t tstride, float* __restrict__ grad, float* __restrict__ loss_sum,
                                                                float* __restrict__ n_valid, int C4, float loss_scale, float grad_scale,
                                                                int64_t ignore_index) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  const int64_t r = blockIdx.x;
  const float4* x = reinterpret_cast<const float4*>(logits + r * ld);
  float4* gx = grad ? reinterpret_cast<float4*>(grad + r * ld) : nullptr;
  const int64_t t = tidx[r * tstride];
  if (t == ignore_index) {
    if (gx) for (int c = threadIdx.x; c < C4; c += blockDim.x) gx[c] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  float mx = -INFINITY;
  for (int c = threadIdx.x; c < C4; c += blockDim.x) {
    const float4 v = x[c];
    mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
  }
  mx = block_max(mx, red);
  float se = 0.f;
  for (int c = threadIdx.x; c < C4; c += blockDim.x) {
    const float4 v = x[c];
    se += __expf(v.x - mx) + __expf(v.y - mx) + __expf(v.z - mx) + __expf(v.w - mx);
  }
  se = ddl::block_sum(se, red);
  const float lse = mx + __logf(se);
  if (threadIdx.x == 0) {
    atomicAdd(loss_sum, (lse - logits[r * ld + t]) * loss_scale);
    if (n_valid) atomicAdd(n_valid, 1.f);
  }
  if (gx) {
    const int tc = (int)(t >> 2), tl = (int)(t & 3);
    for (int c = threadIdx.x; c < C4; c += blockDim.x) {
      const float4 v = x[c];
      float4 g = make_float4(__expf(v.x - lse) * grad_scale, __expf(v.y - lse) * grad_scale, __expf(v.z - lse) * grad_scale,
                             __expf(v.w - lse) * grad_scale);
      if (c == tc) (&g.x)[tl] -= grad_scale;
      gx[c] = g;
    }
  }
}

__global__ void __launch_bounds__(256) eval_top1_kernel(const float* __restrict__ logits, int64_t ld,
                                                        const int64_t* __restrict__ target, int64_t N, int C,
                                                        int* __restrict__ correct) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int cnt = 0;
  for (int64_t r = warp; r < N; r += nwarps) {
    const float* x = logits + r * ld;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = x[c];
      if (v > best) { best = v; bi = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }   // first max wins (torch.argmax)
    }
    if (lane == 0 && (int64_t)bi == target[r]) ++cnt;
  }
  if (lane == 0 && cnt) atomicAdd(correct, cnt);
}

// Binary cross-entropy on logits against a CONSTANT target t (the DCGAN losses: real = 1, fake = 0): one block,
// *loss += scale * sum_i [max(x,0) - x t + log(1 + exp(-|x|))],  grad_i = scale * (sigmoid(x_i) - t).
// Replaces (reference / north star): F.binary_cross_entropy_with_logits in the federated DCGAN local step.
__global__ void __launch_bounds__(256) bce_logits_const_kernel(const float* __restrict__ x, int64_t n, float target, float scale,
                                                               float* __restrict__ loss, float* __restrict__ grad) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = x[i];
    acc += fmaxf(v, 0.f) - v * target + log1pf(__expf(-fabsf(v)));
    if (grad) grad[i] = scale * (1.f / (1.f + __expf(-v)) - target);
  }
  acc = ddl::block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(loss, acc * scale);
}

// VAE loss: MSE(sum)(xr, x) - 1/2 sum(1 + lv - mu^2 - exp(lv)); gradients written when non-null.
__global__ void __launch_bounds__(256) vae_loss_kernel(const float* __restrict__ xr, const float* __restrict__ x,
                                                       int64_t nx, const float* __restrict__ mu,
                                                       const float* __restrict__ lv, int64_t nz,
                                                       float* __restrict__ loss, float* __restrict__ g_xr,
                                                       float* __restrict__ g_mu, float* __restrict__ g_lv,
                                                       float gscale) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  float acc = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t i = i0; i < nx; i += stride) {
    const float d = xr[i] - x[i];
    acc = fmaf(d, d, acc);
    if (g_xr) g_xr[i] = 2.f * d * gscale;
  }
  for (int64_t i = i0; i < nz; i += stride) {
    const float m = mu[i], l = lv[i], e = __expf(l);
    acc += -0.5f * (1.f + l - m * m - e);
    if (g_mu) g_mu[i] = m * gscale;
    if (g_lv) g_lv[i] = 0.5f * (e - 1.f) * gscale;
  }
  acc = ddl::block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(loss, acc);
}

int warp_grid(int64_t rows) {
  int64_t g = (rows + 7) / 8;  // 8 warps per 256-thread block
  if (g > (int64_t)kNumSMs * 8) g = (int64_t)kNumSMs * 8;
  return (int)(g < 1 ? 1 : g);
}

}  // namespace

extern "C" {

int ddl_softmax_xent(const float* logits, int64_t ld, const int64_t* target_idx, int64_t target_stride,
                     const float* target_prob, float* grad, float* loss_sum, float* n_valid, int64_t N, int C,
                     float loss_scale, float grad_scale, int64_t ignore_index, void* stream) {
  DDL_REQUIRE((target_idx != nullptr) != (target_prob != nullptr), "pass exactly one of target_idx / target_prob");
  DDL_REQUIRE(loss_sum != nullptr, "loss_sum must be a device float (pre-zeroed by the caller)");
  if (N == 0) return DDL_OK;
  if (target_idx != nullptr && C >= 2048 && C % 4 == 0 && ld % 4 == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0 &&
      (grad == nullptr || (reinterpret_cast<uintptr_t>(grad) & 15) == 0) && N <= 0x7fffffff) {
    DDL_CUDA_CHECK(ddl_launch_pdl(softmax_xent_wide_kernel, dim3((unsigned)N), dim3(256), 0, (cudaStream_t)stream, logits, ld, target_idx,
                                  target_stride, grad, loss_sum, n_valid, C / 4, loss_scale, grad_scale, ignore_index));
    DDL_LAUNCH_CHECK();
    return DDL_OK;
  }
  DDL_CUDA_CHECK(ddl_launch_pdl(softmax_xent_kernel, dim3(warp_grid(N)), dim3(256), 0, (cudaStream_t)stream, logits, ld, target_idx, target_stride, target_prob,
                                                                       grad, loss_sum, n_valid, N, C, loss_scale,
                                                                       grad_scale, ignore_index));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_bce_logits_const(const float* logits, int64_t n, float target, float scale, float* loss_accum, float* grad, void* stream) {
  DDL_REQUIRE(n >= 0 && loss_accum != nullptr, "bad bce arguments");
  if (n == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(bce_logits_const_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, logits, n, target, scale, loss_accum, grad));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_eval_top1(const float* logits, int64_t ld, const int64_t* target, int64_t N, int C, int* correct,
                  void* stream) {
  if (N == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(eval_top1_kernel, dim3(warp_grid(N)), dim3(256), 0, (cudaStream_t)stream, logits, ld, target, N, C, correct));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_vae_loss(const float* x_recon, const float* x, int64_t nx, const float* mu, const float* logvar, int64_t nz,
                 float* loss, float* grad_x_recon, float* grad_mu, float* grad_logvar, float grad_scale,
                 void* stream) {
  const int64_t n = nx > nz ? nx : nz;
  int64_t g = (n + 255) / 256;
  if (g > kNumSMs * 4) g = kNumSMs * 4;
  if (g < 1) g = 1;
  DDL_CUDA_CHECK(ddl_launch_pdl(vae_loss_kernel, dim3((int)g), dim3(256), 0, (cudaStream_t)stream, x_recon, x, nx, mu, logvar, nz, loss, grad_x_recon, grad_mu,
                                                            grad_logvar, grad_scale));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
