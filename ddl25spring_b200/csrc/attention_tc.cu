// K16: causal self-attention with every matrix product on the tcgen05 tensor cores.
//
// At the labs' shape (T = 256, head dim 48, B*H = 18) the whole T x T score matrix of a layer is 4.7 MB, so the products
// are issued as GROUPED GEMMs (one group per (batch, head)) on gemm_tcgen05.cu - TMA-fed, TMEM accumulators, 3xTF32 so the
// result is fp32-grade - and this file supplies the three row-wise kernels between them:
//
//   forward :  S = scale Q K^T  (GEMM)  ->  causal_softmax (in place: P, zeros above the diagonal)  ->  O = P V  (GEMM)
//   backward:  attn_bwd_prep (dO -> 32-column-padded copy, delta = rowsum(dO * O))  ->  dP = dO V^T  (GEMM)
//              ->  attn_ds (dS = scale P (dP - delta), in place)  ->  dV = P^T dO,  dQ = dS K,  dK = dS^T Q  (GEMMs; the
//              transposed operands are consumed MN-major straight from P / dS, no transpose pass)
//
// Q, K, V, O, dO, dQ, dK, dV are [G, T, HP] with the head dim zero-padded to HP = a multiple of 32 (48 -> 64) by the fused
// QKV split + RoPE kernel (llama_ops.cu), which is what makes them legal MN-major GEMM operands.
// Replaces (reference): the attention inside simplellm's LLaMA blocks (lab/tutorial_1b/primer/intro.py:17,28), and the CUDA-core
// flash kernels of attention.cu (kept for shapes these GEMMs do not take: T not a multiple of 32).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

// one warp per row of S [rows = G*T, T]: row r attends to columns 0 .. (r % T)
__global__ void __launch_bounds__(256) causal_softmax_kernel(float* __restrict__ S, int T, int64_t rows) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int t = (int)(r % T);
  float* row = S + r * T;
  float m = -INFINITY;
  for (int c = lane; c <= t; c += 32) m = fmaxf(m, row[c]);
  m = ddl::warp_max(m);
  float s = 0.f;
  for (int c = lane; c <= t; c += 32) {
    const float e = __expf(row[c] - m);
    row[c] = e;
    s += e;
  }
  s = ddl::warp_sum(s);
  const float inv = 1.f / s;
  for (int c = lane; c < T; c += 32) row[c] = c <= t ? row[c] * inv : 0.f;
}

// one warp per (g, t) row: dOp[row, 0:HP] = dO[b, h, t, 0:hd] zero-padded (dO given by strides), delta[row] = sum_d dO * O
__global__ void __launch_bounds__(256) attn_bwd_prep_kernel(const float* __restrict__ dO, int64_t sb, int64_t sh, int64_t st,
                                                            const float* __restrict__ O, float* __restrict__ dOp,
                                                            float* __restrict__ delta, int H, int T, int hd, int HP, int64_t rows) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int t = (int)(r % T);
  const int64_t g = r / T;
  const int h = (int)(g % H);
  const int64_t b = g / H;
  const float* src = dO + b * sb + h * sh + (int64_t)t * st;
  float acc = 0.f;
  for (int c = lane; c < HP; c += 32) {
    const float d = c < hd ? src[c] : 0.f;
    dOp[r * HP + c] = d;
    acc = fmaf(d, O[r * HP + c], acc);
  }
  acc = ddl::warp_sum(acc);
  if (lane == 0) delta[r] = acc;
}

// dS = scale * P * (dP - delta[row]), in place in dP; float4 per thread, T % 4 == 0
__global__ void __launch_bounds__(256) attn_ds_kernel(const float4* __restrict__ P, float4* __restrict__ dP, const float* __restrict__ delta,
                                                      int T4, int64_t n4, float scale) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = delta[i / T4];
    const float4 p = P[i];
    float4 x = dP[i];
    x.x = scale * p.x * (x.x - d); x.y = scale * p.y * (x.y - d); x.z = scale * p.z * (x.z - d); x.w = scale * p.w * (x.w - d);
    dP[i] = x;
  }
}

}  // namespace

extern "C" {

int ddl_causal_softmax(float* S, int64_t rows, int T, void* stream) {
  DDL_REQUIRE(rows >= 0 && T >= 1, "bad softmax shape");
  if (rows == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(causal_softmax_kernel, dim3((unsigned)((rows + 7) / 8)), dim3(256), 0, (cudaStream_t)stream, S, T, rows));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_attn_bwd_prep(const float* dO, int64_t stride_b, int64_t stride_h, int64_t stride_t, const float* O_padded, float* dO_padded,
                      float* delta, int B, int H, int T, int hd, int HP, void* stream) {
  DDL_REQUIRE(hd >= 1 && HP >= hd, "bad head dims");
  const int64_t rows = (int64_t)B * H * T;
  if (rows == 0) return DDL_OK;
  DDL_CUDA_CHECK(ddl_launch_pdl(attn_bwd_prep_kernel, dim3((unsigned)((rows + 7) / 8)), dim3(256), 0, (cudaStream_t)stream, dO, stride_b,
                                stride_h, stride_t, O_padded, dO_padded, delta, H, T, hd, HP, rows));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_attn_ds(const float* P, float* dP_inout, const float* delta, int64_t rows, int T, float scale, void* stream) {
  DDL_REQUIRE(T % 4 == 0, "T must be a multiple of 4");
  const int64_t n4 = rows * (T / 4);
  if (n4 == 0) return DDL_OK;
  int64_t blocks = (n4 + 255) / 256;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  DDL_CUDA_CHECK(ddl_launch_pdl(attn_ds_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, reinterpret_cast<const float4*>(P),
                                reinterpret_cast<float4*>(dP_inout), delta, T / 4, n4, scale));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
