// K16 (element-wise part of the LLaMA block): fused QKV split + head transpose + rotary embedding, and SwiGLU.
//
//   rope_qkv_fwd : qkv [B, T, 3, H, hd] (the output of the fused QKV GEMM) -> q, k, v [B, H, T, hd], with q and k rotated pair-wise
//                  by (cos, sin)[t, p]:  y_even = x_even c - x_odd s,  y_odd = x_even s + x_odd c       (one pass instead of
//                  view/permute/contiguous + two 6-kernel apply_rope chains)
//   rope_qkv_bwd : the transpose: d(qkv) from dq, dk, dv (inverse rotation for q and k)
//   swiglu_fwd/bwd: h = silu(g) * u;  dg = dh u silu'(g),  du = dh silu(g)
//
// Replaces (reference): simplellm's LLaMA block internals (lab/tutorial_1b/primer/intro.py:17,28 via `LLama`), which run as a
// dozen ATen element-wise kernels per layer.
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

template <bool BWD>
__global__ void __launch_bounds__(256) rope_qkv_kernel(float* __restrict__ qkv, const float* __restrict__ cs, const float* __restrict__ sn,
                                                       float* __restrict__ q, float* __restrict__ k, float* __restrict__ v, int B, int T,
                                                       int H, int hd, int hp) {
  // hp >= hd: row pitch of q / k / v.  Forward writes zeros into the pad columns [hd, hp) so the tensors are directly usable as
  // MN-major operands of the attention GEMMs (attention_tc.cu); backward ignores them.
  pdl_launch_dependents();
  pdl_wait();
  const int P = hd / 2, PP = hp / 2;
  const int64_t total = (int64_t)B * T * 3 * H * PP;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int p = (int)(r % PP); r /= PP;
    const int h = (int)(r % H); r /= H;
    const int w = (int)(r % 3); r /= 3;
    const int t = (int)(r % T);
    const int b = (int)(r / T);
    float* dstbase = w == 0 ? q : (w == 1 ? k : v);
    float2* dst = reinterpret_cast<float2*>(dstbase + (((int64_t)b * H + h) * T + t) * hp) + p;
    if (p >= P) {
      if (!BWD) *dst = make_float2(0.f, 0.f);
      continue;
    }
    float2* src = reinterpret_cast<float2*>(qkv) + ((((int64_t)b * T + t) * 3 + w) * H + h) * P + p;   // [b, t, w, h, p] pairs
    const float c = w < 2 ? cs[t * P + p] : 1.f, s = w < 2 ? sn[t * P + p] : 0.f;
    if (!BWD) {
      const float2 x = *src;
      *dst = make_float2(x.x * c - x.y * s, x.x * s + x.y * c);
    } else {
      const float2 g = *dst;
      *src = make_float2(g.x * c + g.y * s, -g.x * s + g.y * c);
    }
  }
}

__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const float4* __restrict__ g, const float4* __restrict__ u, float4* __restrict__ h,
                                                         int64_t n4) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = g[i], b = u[i];
    float4 o;
    o.x = a.x / (1.f + __expf(-a.x)) * b.x; o.y = a.y / (1.f + __expf(-a.y)) * b.y;
    o.z = a.z / (1.f + __expf(-a.z)) * b.z; o.w = a.w / (1.f + __expf(-a.w)) * b.w;
    h[i] = o;
  }
}

__device__ __forceinline__ void swiglu_grad(float g, float u, float dh, float& dg, float& du) {
  const float sig = 1.f / (1.f + __expf(-g));
  const float silu = g * sig;
  du = dh * silu;
  dg = dh * u * (sig + silu * (1.f - sig));
}

__global__ void __launch_bounds__(256) swiglu_bwd_kernel(const float4* __restrict__ g, const float4* __restrict__ u,
                                                         const float4* __restrict__ dh, float4* __restrict__ dg, float4* __restrict__ du,
                                                         int64_t n4) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = g[i], b = u[i], d = dh[i];
    float4 x, y;
    swiglu_grad(a.x, b.x, d.x, x.x, y.x); swiglu_grad(a.y, b.y, d.y, x.y, y.y);
    swiglu_grad(a.z, b.z, d.z, x.z, y.z); swiglu_grad(a.w, b.w, d.w, x.w, y.w);
    dg[i] = x;
    du[i] = y;
  }
}

// Gate and up projections computed by ONE GEMM against the stacked weight [Wg; Wu]: gu is [M, 2F] with the gate in columns
// [0, F) and the up projection in [F, 2F) of every row.  h = silu(gate) * up is [M, F]; the backward writes d(gu) in the same
// layout so that ONE dgrad GEMM and ONE wgrad GEMM serve both projections.
__global__ void __launch_bounds__(256) swiglu_gu_fwd_kernel(const float4* __restrict__ gu, float4* __restrict__ h, int64_t M, int F4) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t n4 = M * F4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = i / F4;
    const int f = (int)(i - m * F4);
    const float4 a = gu[m * 2 * F4 + f], b = gu[m * 2 * F4 + F4 + f];
    float4 o;
    o.x = a.x / (1.f + __expf(-a.x)) * b.x; o.y = a.y / (1.f + __expf(-a.y)) * b.y;
    o.z = a.z / (1.f + __expf(-a.z)) * b.z; o.w = a.w / (1.f + __expf(-a.w)) * b.w;
    h[i] = o;
  }
}

__global__ void __launch_bounds__(256) swiglu_gu_bwd_kernel(const float4* __restrict__ gu, const float4* __restrict__ dh,
                                                            float4* __restrict__ dgu, int64_t M, int F4) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t n4 = M * F4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = i / F4;
    const int f = (int)(i - m * F4);
    const float4 a = gu[m * 2 * F4 + f], b = gu[m * 2 * F4 + F4 + f], d = dh[i];
    float4 x, y;
    swiglu_grad(a.x, b.x, d.x, x.x, y.x); swiglu_grad(a.y, b.y, d.y, x.y, y.y);
    swiglu_grad(a.z, b.z, d.z, x.z, y.z); swiglu_grad(a.w, b.w, d.w, x.w, y.w);
    dgu[m * 2 * F4 + f] = x;
    dgu[m * 2 * F4 + F4 + f] = y;
  }
}

int blocks_for(int64_t work) {
  int64_t b = (work + 255) / 256;
  if (b > kNumSMs * 16) b = kNumSMs * 16;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace

extern "C" {

int ddl_rope_qkv(float* qkv, const float* cos_tab, const float* sin_tab, float* q, float* k, float* v, int B, int T, int H, int hd,
                 int hd_pitch, int backward, void* stream) {
  DDL_REQUIRE(hd % 2 == 0 && B > 0 && T > 0 && H > 0, "rope_qkv: head dim must be even");
  if (hd_pitch <= 0) hd_pitch = hd;
  DDL_REQUIRE(hd_pitch >= hd && hd_pitch % 2 == 0, "rope_qkv: row pitch %d must be even and >= head dim %d", hd_pitch, hd);
  const int64_t total = (int64_t)B * T * 3 * H * (hd_pitch / 2);
  if (backward)
    DDL_CUDA_CHECK(ddl_launch_pdl(rope_qkv_kernel<true>, dim3(blocks_for(total)), dim3(256), 0, (cudaStream_t)stream, qkv, cos_tab, sin_tab,
                                  q, k, v, B, T, H, hd, hd_pitch));
  else
    DDL_CUDA_CHECK(ddl_launch_pdl(rope_qkv_kernel<false>, dim3(blocks_for(total)), dim3(256), 0, (cudaStream_t)stream, qkv, cos_tab, sin_tab,
                                  q, k, v, B, T, H, hd, hd_pitch));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_swiglu_fwd(const float* g, const float* u, float* h, int64_t n, void* stream) {
  DDL_REQUIRE(n % 4 == 0, "swiglu: n %% 4 != 0");
  DDL_CUDA_CHECK(ddl_launch_pdl(swiglu_fwd_kernel, dim3(blocks_for(n / 4)), dim3(256), 0, (cudaStream_t)stream,
                                reinterpret_cast<const float4*>(g), reinterpret_cast<const float4*>(u), reinterpret_cast<float4*>(h), n / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_swiglu_bwd(const float* g, const float* u, const float* dh, float* dg, float* du, int64_t n, void* stream) {
  DDL_REQUIRE(n % 4 == 0, "swiglu: n %% 4 != 0");
  DDL_CUDA_CHECK(ddl_launch_pdl(swiglu_bwd_kernel, dim3(blocks_for(n / 4)), dim3(256), 0, (cudaStream_t)stream,
                                reinterpret_cast<const float4*>(g), reinterpret_cast<const float4*>(u), reinterpret_cast<const float4*>(dh),
                                reinterpret_cast<float4*>(dg), reinterpret_cast<float4*>(du), n / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_swiglu_gu_fwd(const float* gu, float* h, int64_t M, int F, void* stream) {
  DDL_REQUIRE(M > 0 && F > 0 && F % 4 == 0, "swiglu_gu: F=%d must be a positive multiple of 4", F);
  DDL_CUDA_CHECK(ddl_launch_pdl(swiglu_gu_fwd_kernel, dim3(blocks_for(M * (F / 4))), dim3(256), 0, (cudaStream_t)stream,
                                reinterpret_cast<const float4*>(gu), reinterpret_cast<float4*>(h), M, F / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_swiglu_gu_bwd(const float* gu, const float* dh, float* dgu, int64_t M, int F, void* stream) {
  DDL_REQUIRE(M > 0 && F > 0 && F % 4 == 0, "swiglu_gu: F=%d must be a positive multiple of 4", F);
  DDL_CUDA_CHECK(ddl_launch_pdl(swiglu_gu_bwd_kernel, dim3(blocks_for(M * (F / 4))), dim3(256), 0, (cudaStream_t)stream,
                                reinterpret_cast<const float4*>(gu), reinterpret_cast<const float4*>(dh), reinterpret_cast<float4*>(dgu), M, F / 4));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"

// ---- embedding: out[n, :] = W[tok[n], :] and its scatter-add backward dW[tok[n], :] += g[n, :] (rows equal to padding_idx get
// no gradient).  One warp per token row, 128-bit accesses; the backward uses red.global.add (token rows repeat within a batch).
// Replaces (reference): nn.Embedding inside simplellm's LLama / LLamaFirstStage.embed (lab/tutorial_1b/PP/1F1B/intro_PP_1F1B.py:53).
namespace {

__global__ void __launch_bounds__(256) embedding_fwd_kernel(const float* __restrict__ W, const int64_t* __restrict__ tok, float* __restrict__ out,
                                                            int64_t n, int d4, int64_t vocab) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= n) return;
  int64_t t = tok[r];
  t = t < 0 ? 0 : (t >= vocab ? vocab - 1 : t);
  const float4* src = reinterpret_cast<const float4*>(W) + t * d4;
  float4* dst = reinterpret_cast<float4*>(out) + r * d4;
  for (int c = lane; c < d4; c += 32) dst[c] = src[c];
}

__global__ void __launch_bounds__(256) embedding_bwd_kernel(const float* __restrict__ g, const int64_t* __restrict__ tok, float* __restrict__ dW,
                                                            int64_t n, int d, int64_t vocab, int64_t padding_idx) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= n) return;
  const int64_t t = tok[r];
  if (t == padding_idx || t < 0 || t >= vocab) return;
  const float* src = g + r * d;
  float* dst = dW + t * d;
  for (int c = lane; c < d; c += 32) atomicAdd(dst + c, src[c]);
}

}  // namespace

extern "C" {

int ddl_embedding_fwd(const float* W, const int64_t* tokens, float* out, int64_t n_tokens, int d, int64_t vocab, void* stream) {
  DDL_REQUIRE(d % 4 == 0 && d > 0 && vocab > 0, "embedding dim must be a multiple of 4");
  if (n_tokens == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(embedding_fwd_kernel, dim3((unsigned)((n_tokens + 7) / 8)), dim3(256), 0, (cudaStream_t)stream, W, tokens, out,
                                n_tokens, d / 4, vocab));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_embedding_bwd(const float* grad_out, const int64_t* tokens, float* dW_accum, int64_t n_tokens, int d, int64_t vocab,
                      int64_t padding_idx, void* stream) {
  DDL_REQUIRE(d > 0 && vocab > 0, "bad embedding shape");
  if (n_tokens == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(embedding_bwd_kernel, dim3((unsigned)((n_tokens + 7) / 8)), dim3(256), 0, (cudaStream_t)stream, grad_out, tokens,
                                dW_accum, n_tokens, d, vocab, padding_idx));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
