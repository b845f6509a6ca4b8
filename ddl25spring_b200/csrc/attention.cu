// K16: causal self-attention forward + backward for the labs' LLaMA shape (T = 256, head dim 48, 6 heads, B <= 3).
//
// q, k, v, o: [B*H, T, D] fp32.  The problem is tiny (18 heads x 256 x 256 scores), far below one tensor-core tile
// per SM, so this is a CUDA-core flash-style kernel: one THREAD per query row (forward, dQ) or per key row (dK, dV);
// K/V (resp. Q/dO) tiles are staged in shared memory and every thread of a warp reads the same row -> broadcast,
// no bank conflicts; online softmax keeps the running max / sum in registers; nothing of size T x T is stored.
// Backward recomputes the probabilities from the saved log-sum-exp.
// Replaces (reference): the attention inside simplellm's LLaMA blocks (lab/tutorial_1b/primer/intro.py:17,28).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

constexpr int kRows = 128;   // query (or key) rows per CTA = threads per CTA
constexpr int kTile = 32;    // rows of the other operand staged per step

template <int D>
__global__ void __launch_bounds__(kRows) attn_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                         const float* __restrict__ v, float* __restrict__ o, float* __restrict__ lse,
                                                         int T, float scale) {
  __shared__ float Ks[kTile][D], Vs[kTile][D];
  const int bh = blockIdx.y;
  const int i = blockIdx.x * kRows + threadIdx.x;
  const float* qb = q + (int64_t)bh * T * D;
  const float* kb = k + (int64_t)bh * T * D;
  const float* vb = v + (int64_t)bh * T * D;
  float qr[D], acc[D];
#pragma unroll
  for (int d = 0; d < D; ++d) { qr[d] = i < T ? qb[(int64_t)i * D + d] * scale : 0.f; acc[d] = 0.f; }
  float m = -INFINITY, l = 0.f;
  const int last = min(T, (blockIdx.x + 1) * kRows);          // causal: keys beyond this CTA's last query are never needed
  for (int j0 = 0; j0 < last; j0 += kTile) {
    __syncthreads();
    for (int e = threadIdx.x; e < kTile * D; e += kRows) {
      const int r = e / D, d = e % D;
      const bool ok = j0 + r < T;
      Ks[r][d] = ok ? kb[(int64_t)(j0 + r) * D + d] : 0.f;
      Vs[r][d] = ok ? vb[(int64_t)(j0 + r) * D + d] : 0.f;
    }
    __syncthreads();
    const int jn = min(kTile, i - j0 + 1);                    // keys j0 .. min(j0+kTile-1, i)
    for (int r = 0; r < jn; ++r) {
      float s = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) s = fmaf(qr[d], Ks[r][d], s);
      const float mn = fmaxf(m, s);
      const float corr = __expf(m - mn), p = __expf(s - mn);
      l = l * corr + p;
#pragma unroll
      for (int d = 0; d < D; ++d) acc[d] = fmaf(acc[d], corr, p * Vs[r][d]);
      m = mn;
    }
  }
  if (i < T) {
    const float inv = 1.f / l;
    float* ob = o + ((int64_t)bh * T + i) * D;
#pragma unroll
    for (int d = 0; d < D; ++d) ob[d] = acc[d] * inv;
    lse[(int64_t)bh * T + i] = m + __logf(l);
  }
}

// dQ: one thread per query row.  delta_i = sum_d dO_i,d * O_i,d is computed here and stored for the dK/dV kernels.
template <int D>
__global__ void __launch_bounds__(kRows) attn_bwd_dq_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                            const float* __restrict__ v, const float* __restrict__ o,
                                                            const float* __restrict__ dO, const float* __restrict__ lse,
                                                            float* __restrict__ dq, float* __restrict__ delta, int T, float scale) {
  __shared__ float Ks[kTile][D], Vs[kTile][D];
  const int bh = blockIdx.y;
  const int i = blockIdx.x * kRows + threadIdx.x;
  const int64_t base = (int64_t)bh * T * D;
  float qr[D], go[D], acc[D];
  float dl = 0.f;
#pragma unroll
  for (int d = 0; d < D; ++d) {
    qr[d] = i < T ? q[base + (int64_t)i * D + d] * scale : 0.f;
    go[d] = i < T ? dO[base + (int64_t)i * D + d] : 0.f;
    dl = fmaf(go[d], i < T ? o[base + (int64_t)i * D + d] : 0.f, dl);
    acc[d] = 0.f;
  }
  const float li = i < T ? lse[(int64_t)bh * T + i] : 0.f;
  if (i < T) delta[(int64_t)bh * T + i] = dl;
  const int last = min(T, (blockIdx.x + 1) * kRows);
  for (int j0 = 0; j0 < last; j0 += kTile) {
    __syncthreads();
    for (int e = threadIdx.x; e < kTile * D; e += kRows) {
      const int r = e / D, d = e % D;
      const bool ok = j0 + r < T;
      Ks[r][d] = ok ? k[base + (int64_t)(j0 + r) * D + d] : 0.f;
      Vs[r][d] = ok ? v[base + (int64_t)(j0 + r) * D + d] : 0.f;
    }
    __syncthreads();
    const int jn = min(kTile, i - j0 + 1);
    for (int r = 0; r < jn; ++r) {
      float s = 0.f, dp = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) { s = fmaf(qr[d], Ks[r][d], s); dp = fmaf(go[d], Vs[r][d], dp); }
      const float ds = __expf(s - li) * (dp - dl) * scale;
#pragma unroll
      for (int d = 0; d < D; ++d) acc[d] = fmaf(ds, Ks[r][d], acc[d]);
    }
  }
  if (i < T) {
#pragma unroll
    for (int d = 0; d < D; ++d) dq[base + (int64_t)i * D + d] = acc[d];
  }
}

// dK (WHICH = 0) or dV (WHICH = 1): one thread per key row, looping over the queries at or after it.
template <int D, int WHICH>
__global__ void __launch_bounds__(kRows) attn_bwd_dkv_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                             const float* __restrict__ v, const float* __restrict__ dO,
                                                             const float* __restrict__ lse, const float* __restrict__ delta,
                                                             float* __restrict__ out, int T, float scale) {
  __shared__ float Qs[kTile][D], Gs[kTile][D];
  __shared__ float Ls[kTile], Ds[kTile];
  const int bh = blockIdx.y;
  const int j = blockIdx.x * kRows + threadIdx.x;
  const int64_t base = (int64_t)bh * T * D;
  float kr[D], vr[D], acc[D];
#pragma unroll
  for (int d = 0; d < D; ++d) {
    kr[d] = j < T ? k[base + (int64_t)j * D + d] : 0.f;
    vr[d] = (WHICH == 0 && j < T) ? v[base + (int64_t)j * D + d] : 0.f;
    acc[d] = 0.f;
  }
  const int first = blockIdx.x * kRows;                       // queries before this CTA's first key never attend to it
  for (int i0 = first; i0 < T; i0 += kTile) {
    __syncthreads();
    for (int e = threadIdx.x; e < kTile * D; e += kRows) {
      const int r = e / D, d = e % D;
      const bool ok = i0 + r < T;
      Qs[r][d] = ok ? q[base + (int64_t)(i0 + r) * D + d] * scale : 0.f;
      Gs[r][d] = ok ? dO[base + (int64_t)(i0 + r) * D + d] : 0.f;
    }
    if (threadIdx.x < kTile) {
      const bool ok = i0 + threadIdx.x < T;
      Ls[threadIdx.x] = ok ? lse[(int64_t)bh * T + i0 + threadIdx.x] : INFINITY;   // p = exp(s - inf) = 0 for padding rows
      Ds[threadIdx.x] = ok ? delta[(int64_t)bh * T + i0 + threadIdx.x] : 0.f;
    }
    __syncthreads();
    for (int r = max(0, j - i0); r < kTile; ++r) {            // only queries i >= j
      float s = 0.f;
#pragma unroll
      for (int d = 0; d < D; ++d) s = fmaf(Qs[r][d], kr[d], s);
      const float p = __expf(s - Ls[r]);
      if (WHICH == 1) {
#pragma unroll
        for (int d = 0; d < D; ++d) acc[d] = fmaf(p, Gs[r][d], acc[d]);
      } else {
        float dp = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) dp = fmaf(Gs[r][d], vr[d], dp);
        const float ds = p * (dp - Ds[r]);                    // Qs already carries `scale`
#pragma unroll
        for (int d = 0; d < D; ++d) acc[d] = fmaf(ds, Qs[r][d], acc[d]);
      }
    }
  }
  if (j < T) {
#pragma unroll
    for (int d = 0; d < D; ++d) out[base + (int64_t)j * D + d] = acc[d];
  }
}

template <int D>
int launch_fwd(const float* q, const float* k, const float* v, float* o, float* lse, int BH, int T, float scale, cudaStream_t st) {
  dim3 grid((T + kRows - 1) / kRows, BH);
  attn_fwd_kernel<D><<<grid, kRows, 0, st>>>(q, k, v, o, lse, T, scale);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

template <int D>
int launch_bwd(const float* q, const float* k, const float* v, const float* o, const float* dO, const float* lse, float* dq, float* dk,
               float* dv, float* delta, int BH, int T, float scale, cudaStream_t st) {
  dim3 grid((T + kRows - 1) / kRows, BH);
  attn_bwd_dq_kernel<D><<<grid, kRows, 0, st>>>(q, k, v, o, dO, lse, dq, delta, T, scale);
  DDL_LAUNCH_CHECK();
  attn_bwd_dkv_kernel<D, 0><<<grid, kRows, 0, st>>>(q, k, v, dO, lse, delta, dk, T, scale);
  DDL_LAUNCH_CHECK();
  attn_bwd_dkv_kernel<D, 1><<<grid, kRows, 0, st>>>(q, k, v, dO, lse, delta, dv, T, scale);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // namespace

extern "C" {

int ddl_causal_attention_fwd(const float* q, const float* k, const float* v, float* o, float* lse, int BH, int T, int D, float scale,
                             void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  switch (D) {
    case 16: return launch_fwd<16>(q, k, v, o, lse, BH, T, scale, st);
    case 32: return launch_fwd<32>(q, k, v, o, lse, BH, T, scale, st);
    case 48: return launch_fwd<48>(q, k, v, o, lse, BH, T, scale, st);
    case 64: return launch_fwd<64>(q, k, v, o, lse, BH, T, scale, st);
  }
  DDL_REQUIRE(false, "head dim %d not supported (16, 32, 48, 64)", D);
}

int ddl_causal_attention_bwd(const float* q, const float* k, const float* v, const float* o, const float* dO, const float* lse,
                             float* dq, float* dk, float* dv, float* delta_scratch, int BH, int T, int D, float scale, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  switch (D) {
    case 16: return launch_bwd<16>(q, k, v, o, dO, lse, dq, dk, dv, delta_scratch, BH, T, scale, st);
    case 32: return launch_bwd<32>(q, k, v, o, dO, lse, dq, dk, dv, delta_scratch, BH, T, scale, st);
    case 48: return launch_bwd<48>(q, k, v, o, dO, lse, dq, dk, dv, delta_scratch, BH, T, scale, st);
    case 64: return launch_bwd<64>(q, k, v, o, dO, lse, dq, dk, dv, delta_scratch, BH, T, scale, st);
  }
  DDL_REQUIRE(false, "head dim %d not supported (16, 32, 48, 64)", D);
}

}  // extern "C"
