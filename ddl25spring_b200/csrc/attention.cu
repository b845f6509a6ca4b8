// K16: causal self-attention forward + backward for the labs' LLaMA shape (T = 256, head dim 48, 6 heads, B <= 3).
//
// q, k, v, o: [B*H, T, D] fp32.  The problem is tiny (18 heads x 256 x 256 scores), far below one tensor-core tile
// per SM, so this is a CUDA-core flash-style kernel: SIXTEEN LANES per query row (forward, dQ) or per key row (dK + dV in
// one pass) = 4 key-splits x 4 lanes holding a quarter of the head dimension each; K/V (resp. Q/dO) tiles are staged in
// shared memory; online softmax keeps the running max / sum in registers; nothing of size T x T is stored.
// Backward recomputes the probabilities from the saved log-sum-exp.
// Replaces (reference): the attention inside simplellm's LLaMA blocks (lab/tutorial_1b/primer/intro.py:17,28).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

// Work split: a QUERY (forward, dQ) or KEY (dK/dV) row is owned by 16 adjacent lanes = 4 key-splits x 4 dim-lanes.  A dim-lane
// holds D/4 of the head dimension (dot products are finished with two xor-shuffles inside the quad); the four splits walk
// interleaved quarters of the other operand's rows and are merged at the end (online-softmax merge in the forward, plain sums
// in the backward).  A CTA of 128 threads covers 8 rows, so the lab shape (18 heads x 256 tokens) launches 576 CTAs = 16 warps
// per SM.  History: one thread per row, 36 CTAs: 90 us per kernel; 4 lanes per row, 144 CTAs = 4 warps per SM:
// 72 us, still latency-bound on one warp per scheduler; this layout hides the FMA / shuffle / exp latencies behind 16 warps.
constexpr int kThreads = 128;
constexpr int kRows = kThreads / 16;  // rows (queries or keys) per CTA
constexpr int kTile = 32;             // rows of the other operand staged in shared memory per step

__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}
__device__ __forceinline__ float split_sum(float v) {       // over the 4 key-splits of a row
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  return v;
}

template <int D>
__device__ __forceinline__ void load_tile(float (*dst)[D], const float* __restrict__ src, int row0, int T, float mul) {
  for (int e = threadIdx.x; e < kTile * D / 4; e += kThreads) {
    const int r = e / (D / 4), d = (e % (D / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < T) v = *reinterpret_cast<const float4*>(src + (int64_t)(row0 + r) * D + d);
    v.x *= mul; v.y *= mul; v.z *= mul; v.w *= mul;
    *reinterpret_cast<float4*>(&dst[r][d]) = v;
  }
}

// dot / axpy of a lane's D/4 slice against a shared-memory row, four floats per LDS.128
template <int D4>
__device__ __forceinline__ float dot_slice(const float (&a)[D4], const float* __restrict__ row) {
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < D4; c += 4) {
    const float4 v = *reinterpret_cast<const float4*>(row + c);
    s = fmaf(a[c], v.x, s); s = fmaf(a[c + 1], v.y, s); s = fmaf(a[c + 2], v.z, s); s = fmaf(a[c + 3], v.w, s);
  }
  return s;
}
template <int D4>
__device__ __forceinline__ void axpy_slice(float (&acc)[D4], float p, const float* __restrict__ row) {
#pragma unroll
  for (int c = 0; c < D4; c += 4) {
    const float4 v = *reinterpret_cast<const float4*>(row + c);
    acc[c] = fmaf(p, v.x, acc[c]); acc[c + 1] = fmaf(p, v.y, acc[c + 1]);
    acc[c + 2] = fmaf(p, v.z, acc[c + 2]); acc[c + 3] = fmaf(p, v.w, acc[c + 3]);
  }
}

struct Lane {
  int row, warp_first, warp_last, split, g;
};
__device__ __forceinline__ Lane lane_of() {
  Lane l;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  l.g = lane & 3;
  l.split = (lane >> 2) & 3;
  l.warp_first = blockIdx.x * kRows + w * 2;
  l.warp_last = l.warp_first + 1;
  l.row = l.warp_first + (lane >> 4);
  return l;
}

template <int D>
__global__ void __launch_bounds__(kThreads) attn_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                            const float* __restrict__ v, float* __restrict__ o, float* __restrict__ lse,
                                                            int T, float scale) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int D4 = D / 4;
  __shared__ __align__(16) float Ks[kTile][D], Vs[kTile][D];
  const Lane L = lane_of();
  const int bh = blockIdx.y, i = L.row, g = L.g;
  const int64_t base = (int64_t)bh * T * D;
  float qr[D4], acc[D4];
#pragma unroll
  for (int d = 0; d < D4; ++d) { qr[d] = i < T ? q[base + (int64_t)i * D + g * D4 + d] * scale : 0.f; acc[d] = 0.f; }
  float m = -INFINITY, l = 0.f;
  const int last = min(T, (blockIdx.x + 1) * kRows);          // causal: keys beyond this CTA's last query are never needed
  for (int j0 = 0; j0 < last; j0 += kTile) {
    __syncthreads();
    load_tile<D>(Ks, k + base, j0, T, 1.f);
    load_tile<D>(Vs, v + base, j0, T, 1.f);
    __syncthreads();
    const int jn = min(kTile, min(L.warp_last, T - 1) - j0 + 1);   // uniform per warp (shuffles inside the loop)
    for (int rr = 0; rr < jn; rr += 4) {
      const int r = rr + L.split;                                  // <= 31
      const float s = quad_sum(dot_slice<D4>(qr, &Ks[r][g * D4]));
      if (r < jn && j0 + r <= i) {                                 // the quad of a (row, split) agrees
        const float mn = fmaxf(m, s);
        const float corr = __expf(m - mn), p = __expf(s - mn);
        l = l * corr + p;
#pragma unroll
        for (int d = 0; d < D4; ++d) acc[d] *= corr;
        axpy_slice<D4>(acc, p, &Vs[r][g * D4]);
        m = mn;
      }
    }
  }
  // merge the four key-splits of each row (online-softmax merge); all lanes take part
#pragma unroll
  for (int off = 4; off <= 8; off <<= 1) {
    const float mo = __shfl_xor_sync(0xffffffffu, m, off), lo = __shfl_xor_sync(0xffffffffu, l, off);
    const float mn = fmaxf(m, mo);
    const float ca = m == -INFINITY ? 0.f : __expf(m - mn), cb = mo == -INFINITY ? 0.f : __expf(mo - mn);
    l = l * ca + lo * cb;
#pragma unroll
    for (int d = 0; d < D4; ++d) {
      const float ao = __shfl_xor_sync(0xffffffffu, acc[d], off);
      acc[d] = acc[d] * ca + ao * cb;
    }
    m = mn;
  }
  if (i < T && L.split == 0) {
    const float inv = 1.f / l;
#pragma unroll
    for (int d = 0; d < D4; ++d) o[base + (int64_t)i * D + g * D4 + d] = acc[d] * inv;
    if (g == 0) lse[(int64_t)bh * T + i] = m + __logf(l);
  }
}

// dQ.  delta_i = sum_d dO_i,d * O_i,d is computed here and stored for the dK/dV kernel.
template <int D>
__global__ void __launch_bounds__(kThreads) attn_bwd_dq_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                               const float* __restrict__ v, const float* __restrict__ o,
                                                               const float* __restrict__ dO, const float* __restrict__ lse,
                                                               float* __restrict__ dq, float* __restrict__ delta, int T, float scale) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int D4 = D / 4;
  __shared__ __align__(16) float Ks[kTile][D], Vs[kTile][D];
  const Lane L = lane_of();
  const int bh = blockIdx.y, i = L.row, g = L.g;
  const int64_t base = (int64_t)bh * T * D;
  float qr[D4], go[D4], acc[D4];
  float dl = 0.f;
#pragma unroll
  for (int d = 0; d < D4; ++d) {
    const int64_t off = base + (int64_t)i * D + g * D4 + d;
    qr[d] = i < T ? q[off] * scale : 0.f;
    go[d] = i < T ? dO[off] : 0.f;
    dl = fmaf(go[d], i < T ? o[off] : 0.f, dl);
    acc[d] = 0.f;
  }
  dl = quad_sum(dl);
  const float li = i < T ? lse[(int64_t)bh * T + i] : 0.f;
  if (i < T && g == 0 && L.split == 0) delta[(int64_t)bh * T + i] = dl;
  const int last = min(T, (blockIdx.x + 1) * kRows);
  for (int j0 = 0; j0 < last; j0 += kTile) {
    __syncthreads();
    load_tile<D>(Ks, k + base, j0, T, 1.f);
    load_tile<D>(Vs, v + base, j0, T, 1.f);
    __syncthreads();
    const int jn = min(kTile, min(L.warp_last, T - 1) - j0 + 1);
    for (int rr = 0; rr < jn; rr += 4) {
      const int r = rr + L.split;
      const float s = quad_sum(dot_slice<D4>(qr, &Ks[r][g * D4]));
      const float dp = quad_sum(dot_slice<D4>(go, &Vs[r][g * D4]));
      if (r < jn && j0 + r <= i) {
        const float ds = __expf(s - li) * (dp - dl) * scale;
        axpy_slice<D4>(acc, ds, &Ks[r][g * D4]);
      }
    }
  }
#pragma unroll
  for (int d = 0; d < D4; ++d) acc[d] = split_sum(acc[d]);
  if (i < T && L.split == 0) {
#pragma unroll
    for (int d = 0; d < D4; ++d) dq[base + (int64_t)i * D + g * D4 + d] = acc[d];
  }
}

// dK and dV in one pass: the lanes of a key row loop over the queries at or after it.
template <int D>
__global__ void __launch_bounds__(kThreads) attn_bwd_dkv_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                                const float* __restrict__ v, const float* __restrict__ dO,
                                                                const float* __restrict__ lse, const float* __restrict__ delta,
                                                                float* __restrict__ dk, float* __restrict__ dv, int T, float scale) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int D4 = D / 4;
  __shared__ __align__(16) float Qs[kTile][D], Gs[kTile][D];
  __shared__ float Ls[kTile], Ds[kTile];
  const Lane L = lane_of();
  const int bh = blockIdx.y, j = L.row, g = L.g;
  const int64_t base = (int64_t)bh * T * D;
  float kr[D4], vr[D4], acck[D4], accv[D4];
#pragma unroll
  for (int d = 0; d < D4; ++d) {
    const int64_t off = base + (int64_t)j * D + g * D4 + d;
    kr[d] = j < T ? k[off] : 0.f;
    vr[d] = j < T ? v[off] : 0.f;
    acck[d] = 0.f; accv[d] = 0.f;
  }
  const int first = blockIdx.x * kRows;                       // queries before this CTA's first key never attend to it
  for (int i0 = first; i0 < T; i0 += kTile) {
    __syncthreads();
    load_tile<D>(Qs, q + base, i0, T, scale);
    load_tile<D>(Gs, dO + base, i0, T, 1.f);
    if (threadIdx.x < kTile) {
      const bool ok = i0 + threadIdx.x < T;
      Ls[threadIdx.x] = ok ? lse[(int64_t)bh * T + i0 + threadIdx.x] : INFINITY;   // p = exp(s - inf) = 0 for padding rows
      Ds[threadIdx.x] = ok ? delta[(int64_t)bh * T + i0 + threadIdx.x] : 0.f;
    }
    __syncthreads();
    for (int rr = max(0, L.warp_first - i0); rr < kTile; rr += 4) {   // uniform per warp; only queries i >= j contribute
      const int r = min(rr + L.split, kTile - 1);
      const float s = quad_sum(dot_slice<D4>(kr, &Qs[r][g * D4]));
      const float dp = quad_sum(dot_slice<D4>(vr, &Gs[r][g * D4]));
      if (rr + L.split < kTile && i0 + r >= j) {
        const float p = __expf(s - Ls[r]);
        const float ds = p * (dp - Ds[r]);                      // Qs already carries `scale`
        axpy_slice<D4>(accv, p, &Gs[r][g * D4]);
        axpy_slice<D4>(acck, ds, &Qs[r][g * D4]);
      }
    }
  }
#pragma unroll
  for (int d = 0; d < D4; ++d) { acck[d] = split_sum(acck[d]); accv[d] = split_sum(accv[d]); }
  if (j < T && L.split == 0) {
#pragma unroll
    for (int d = 0; d < D4; ++d) {
      dk[base + (int64_t)j * D + g * D4 + d] = acck[d];
      dv[base + (int64_t)j * D + g * D4 + d] = accv[d];
    }
  }
}

template <int D>
int launch_fwd(const float* q, const float* k, const float* v, float* o, float* lse, int BH, int T, float scale, cudaStream_t st) {
  dim3 grid((T + kRows - 1) / kRows, BH);
  DDL_CUDA_CHECK(ddl_launch_pdl(attn_fwd_kernel<D>, dim3(grid), dim3(kThreads), 0, st, q, k, v, o, lse, T, scale));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

template <int D>
int launch_bwd(const float* q, const float* k, const float* v, const float* o, const float* dO, const float* lse, float* dq, float* dk,
               float* dv, float* delta, int BH, int T, float scale, cudaStream_t st) {
  dim3 grid((T + kRows - 1) / kRows, BH);
  DDL_CUDA_CHECK(ddl_launch_pdl(attn_bwd_dq_kernel<D>, dim3(grid), dim3(kThreads), 0, st, q, k, v, o, dO, lse, dq, delta, T, scale));
  DDL_LAUNCH_CHECK();
  DDL_CUDA_CHECK(ddl_launch_pdl(attn_bwd_dkv_kernel<D>, dim3(grid), dim3(kThreads), 0, st, q, k, v, dO, lse, delta, dk, dv, T, scale));
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
