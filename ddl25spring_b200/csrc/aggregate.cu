// Aggregation kernels over flat fp32 update rows (K1, K4, K5, K6 and the Krum score/select
// epilogue of K3 in SURVEY §8).
//
// Every kernel addresses its inputs through a ROW-POINTER TABLE passed by value in the kernel
// parameters: row k is `rows[k] + d`.  A row may live in this GPU's HBM or be a peer-mapped
// pointer into another GPU's HBM (parallel/symm.py hands those out), so the same code is the
// single-GPU aggregator and the per-GPU half of the coordinate-sharded NVLink exchange
// (GPU g reduces coordinates [d0, d1) of every client on every GPU and stores the result to
// all replicas through the `outs` table).
//
// Replaces (reference): the CPU list-of-tensors aggregation in
// lab/tutorial_1a/hfl_complete.py:291-306 (FedSGD) and :370-383 (FedAvg).
#include "common.cuh"
#include "ddl25_kernels.h"
#include "select.cuh"

namespace {

constexpr int kMaxRows = DDL_MAX_ROWS;   // rows per launch
constexpr int kMaxOuts = DDL_MAX_OUTS;   // broadcast destinations per launch

struct RowTable {
  const float* rows[kMaxRows];
  float w[kMaxRows];
};
struct OutTable {
  float* outs[kMaxOuts];
};

// ------------------------------------------------------------------------------------------
// K1: out[d] = alpha * base[d] + beta * sum_k w[k] * rows[k][d]   for d in [d0, d1)
//   FedAvg : alpha = 0, beta = 1               (global model overwrite, hfl_complete.py:380-383)
//   FedSGD : base = params, alpha = 1, beta = -lr  (server SGD step fused, hfl_complete.py:301-306)
// and the result is stored to every pointer in `outs` (global replicas on all GPUs and/or the
// packed clients' weight rows = next round's broadcast, hfl_complete.py:356).
// Grid-stride over float4 groups; K loads in flight per thread.
// ------------------------------------------------------------------------------------------
template <int UNROLL>
__global__ void __launch_bounds__(256) weighted_sum_kernel(RowTable tab, int K, const float* __restrict__ base,
                                                           float alpha, float beta, OutTable ot, int n_out,
                                                           int64_t v0, int64_t v1, int peer_rows,
                                                           const float* __restrict__ dev_w) {
  // weights either ride in the parameter block (host-known n_k / sum n) or are read from device
  // memory (Krum's selection weights are produced by krum_select on the GPU: no host round trip)
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < v1; v += stride) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int k = 0;
    // rows with weight 0 are neither loaded nor accumulated: Krum's selection weights are mostly zero (only the chosen
    // rows cross HBM / NVLink), and a rejected row may hold NaN / Inf, which 0 * x would turn into NaN
    for (; k + UNROLL <= K; k += UNROLL) {
      float4 x[UNROLL];
      float w[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        w[u] = dev_w ? __ldg(dev_w + k + u) : tab.w[k + u];
        const float4* p = reinterpret_cast<const float4*>(tab.rows[k + u]) + v;
        x[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (w[u] != 0.f) x[u] = peer_rows ? ddl::ld_coherent(p) : ddl::ld_stream(p);
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        acc.x = fmaf(w[u], x[u].x, acc.x); acc.y = fmaf(w[u], x[u].y, acc.y);
        acc.z = fmaf(w[u], x[u].z, acc.z); acc.w = fmaf(w[u], x[u].w, acc.w);
      }
    }
    for (; k < K; ++k) {
      const float w = dev_w ? __ldg(dev_w + k) : tab.w[k];
      if (w == 0.f) continue;
      const float4* p = reinterpret_cast<const float4*>(tab.rows[k]) + v;
      const float4 x = peer_rows ? ddl::ld_coherent(p) : ddl::ld_stream(p);
      acc.x = fmaf(w, x.x, acc.x); acc.y = fmaf(w, x.y, acc.y);
      acc.z = fmaf(w, x.z, acc.z); acc.w = fmaf(w, x.w, acc.w);
    }
    float4 r = make_float4(beta * acc.x, beta * acc.y, beta * acc.z, beta * acc.w);
    if (base != nullptr) {
      const float4 b = reinterpret_cast<const float4*>(base)[v];
      r.x = fmaf(alpha, b.x, r.x); r.y = fmaf(alpha, b.y, r.y);
      r.z = fmaf(alpha, b.z, r.z); r.w = fmaf(alpha, b.w, r.w);
    }
    for (int o = 0; o < n_out; ++o) reinterpret_cast<float4*>(ot.outs[o])[v] = r;
  }
}

// ------------------------------------------------------------------------------------------
// K4 / K5: coordinate-wise median and trimmed mean.
// One thread per coordinate: the K values of that coordinate are loaded (one coalesced 128 B
// line per warp per row) into registers and run through a straight-line Batcher odd-even merge
// network (select.cuh; KP = K padded to a power of two), then reduced over ranks [lo, hi).  A
// window that is exactly the median takes the network pruned to the two centre ranks.
//   median (odd K)  : lo = hi-1 = (K-1)/2
//   median (even K) : lo = K/2-1, hi = K/2+1 (mean of the middle pair) or lower median
//   trimmed mean    : lo = trim, hi = K - trim
// ------------------------------------------------------------------------------------------
template <int KP, bool MEDIAN>
__global__ void __launch_bounds__(128) rank_select_mean_kernel(RowTable tab, int K, int lo, int hi, OutTable ot,
                                                               int n_out, int64_t d0, int64_t d1, int one) {
  const int64_t d = d0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= d1) return;
  float a[KP];
#pragma unroll
  for (int k = 0; k < KP; ++k)
    a[k] = (k < K) ? __ldcg(tab.rows[k] + d) : (MEDIAN ? ddl::median_pad(k, K, KP) : __int_as_float(0x7f800000));
  const float r = MEDIAN ? ddl::median_of<KP>(a, hi - lo == 2, one) : ddl::window_mean<KP>(a, lo, hi, one);
  for (int o = 0; o < n_out; ++o) ot.outs[o][d] = r;
}

// ------------------------------------------------------------------------------------------
// K6: Byzantine perturbation fused into the client's update write-out.
//   sign-flip : out = g - s * (u - g)    (weights message; for a gradient message g == nullptr -> -s*u)
//   optional additive uniform noise of amplitude `noise` (counter-based RNG)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) attack_perturb_kernel(const float* __restrict__ u, const float* __restrict__ g,
                                                             float* __restrict__ out, float scale, float noise,
                                                             uint32_t seed, int64_t nvec) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const float4 x = reinterpret_cast<const float4*>(u)[v];
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g != nullptr) b = reinterpret_cast<const float4*>(g)[v];
    float4 r;
    r.x = b.x - scale * (x.x - b.x); r.y = b.y - scale * (x.y - b.y);
    r.z = b.z - scale * (x.z - b.z); r.w = b.w - scale * (x.w - b.w);
    if (noise != 0.f) {
      const uint32_t i = (uint32_t)(v * 4);
      r.x += noise * (2.f * ddl::rng_uniform(seed, i) - 1.f);
      r.y += noise * (2.f * ddl::rng_uniform(seed, i + 1) - 1.f);
      r.z += noise * (2.f * ddl::rng_uniform(seed, i + 2) - 1.f);
      r.w += noise * (2.f * ddl::rng_uniform(seed, i + 3) - 1.f);
    }
    reinterpret_cast<float4*>(out)[v] = r;
  }
}

// ------------------------------------------------------------------------------------------
// Krum epilogue (K3 `krum_select`): from the K x K Gram matrix G = U U^T
//   dist_ij = G_ii + G_jj - 2 G_ij ; score_i = sum of the (K - f - 2) smallest dist_ij, j != i
//   sel[0..multi) = indices of the `multi` smallest scores (stable), and the FedAvg-style
//   selection weights w[i] = 1/multi for selected rows, 0 otherwise.
// One block of 32 warps; a warp owns a row at a time (K <= 128).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) krum_select_kernel(const float* __restrict__ G, int ldg, int K, int f, int multi,
                                                           float* __restrict__ scores, int* __restrict__ sel,
                                                           float* __restrict__ selw) {
  // One warp per row i (32 warps take rows w, w + 32, ...): the K - 1 distances of the row go to shared memory, every lane
  // ranks its distances against the whole row (ties by index) and the m = K - f - 2 smallest are summed with a fixed-order
  // warp reduction.  (The first version kept an insertion-sorted list per THREAD in local memory: 118 us at K = 64.)
  __shared__ float s_score[128];
  __shared__ float s_diag[128];
  __shared__ float s_d[32][128];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int t = threadIdx.x; t < 128; t += blockDim.x) {
    s_diag[t] = t < K ? G[(int64_t)t * ldg + t] : 0.f;
    s_score[t] = __int_as_float(0x7f800000);
  }
  __syncthreads();
  const int m = K - f - 2;
  const float inf = __int_as_float(0x7f800000);
  for (int i = w; i < K; i += 32) {
    for (int j = lane; j < K; j += 32) {
      // a non-finite distance (NaN / Inf rows, or a Gram entry that overflowed fp32) must count as "infinitely far":
      // fmaxf alone would turn NaN into 0 and make the poisoned row the best candidate
      float dj = s_diag[i] + s_diag[j] - 2.f * G[(int64_t)i * ldg + j];
      dj = (dj == dj && fabsf(dj) <= 3.0e38f) ? fmaxf(dj, 0.f) : inf;
      s_d[w][j] = (j == i) ? inf : dj;
    }
    __syncwarp();
    float acc = 0.f;
    for (int j = lane; j < K; j += 32) {
      if (j == i) continue;
      const float dj = s_d[w][j];
      int rank = 0;
      for (int l = 0; l < K; ++l) {
        const float dl = s_d[w][l];
        rank += (l != i) && ((dl < dj) || (dl == dj && l < j));
      }
      if (rank < m) acc += dj;
    }
    acc = ddl::warp_sum(acc);
    if (!(s_diag[i] == s_diag[i]) || fabsf(s_diag[i]) > 3.0e38f) acc = inf;   // a non-finite row is never selected
    if (lane == 0) { s_score[i] = acc; scores[i] = acc; }
    __syncwarp();
  }
  __syncthreads();
  const int i = threadIdx.x;
  if (i < K) {
    // rank of row i among the scores (stable: ties broken by index)
    const float sc = s_score[i];
    int rank = 0;
    for (int j = 0; j < K; ++j) rank += (s_score[j] < sc) || (s_score[j] == sc && j < i);
    if (rank < multi) sel[rank] = i;
    selw[i] = rank < multi ? 1.f / (float)multi : 0.f;
  }
}

int grid_for(int64_t work_items, int block, int per_sm) {
  int64_t g = (work_items + block - 1) / block;
  const int64_t cap = (int64_t)kNumSMs * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

template <int KP>
int launch_rank_select(const RowTable& tab, int K, int lo, int hi, const OutTable& ot, int n_out, int64_t d0,
                       int64_t d1, cudaStream_t st) {
  const int64_t n = d1 - d0;
  const int block = 128;
  const int64_t grid = (n + block - 1) / block;
  const bool med = (lo == (K - 1) / 2 && hi == lo + 1) || (K % 2 == 0 && lo == K / 2 - 1 && hi == K / 2 + 1);
  if (med) rank_select_mean_kernel<KP, true><<<(unsigned)grid, block, 0, st>>>(tab, K, lo, hi, ot, n_out, d0, d1, 1);
  else rank_select_mean_kernel<KP, false><<<(unsigned)grid, block, 0, st>>>(tab, K, lo, hi, ot, n_out, d0, d1, 1);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int fill_tables(RowTable& tab, OutTable& ot, const float* const* rows, const float* weights, int K,
                float* const* outs, int n_out, uintptr_t align_mask = 15) {
  DDL_REQUIRE(K >= 1 && K <= kMaxRows, "K=%d outside [1, %d]", K, kMaxRows);
  DDL_REQUIRE(n_out >= 1 && n_out <= kMaxOuts, "n_out=%d outside [1, %d]", n_out, kMaxOuts);
  for (int k = 0; k < K; ++k) {
    DDL_REQUIRE(rows[k] != nullptr && (reinterpret_cast<uintptr_t>(rows[k]) & align_mask) == 0, "row %d null or misaligned", k);
    tab.rows[k] = rows[k];
    tab.w[k] = weights ? weights[k] : 1.f;
  }
  for (int o = 0; o < n_out; ++o) {
    DDL_REQUIRE(outs[o] != nullptr && (reinterpret_cast<uintptr_t>(outs[o]) & align_mask) == 0, "out %d null or misaligned", o);
    ot.outs[o] = outs[o];
  }
  return DDL_OK;
}

}  // namespace

extern "C" {

int ddl_weighted_sum_rows(const float* const* rows, const float* weights, const float* dev_weights, int K,
                          const float* base, float alpha, float beta, float* const* outs, int n_out, int64_t d0,
                          int64_t d1, int peer_rows, int max_ctas, void* stream) {
  DDL_REQUIRE(d0 % 4 == 0 && d1 % 4 == 0 && d1 >= d0, "coordinate range [%lld, %lld) must be float4 aligned",
              (long long)d0, (long long)d1);
  if (d1 == d0) return DDL_OK;
  RowTable tab;
  OutTable ot;
  if (int e = fill_tables(tab, ot, rows, weights, K, outs, n_out)) return e;
  const int64_t v0 = d0 / 4, v1 = d1 / 4;
  int grid = grid_for(v1 - v0, 256, 8);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  weighted_sum_kernel<8><<<grid, 256, 0, (cudaStream_t)stream>>>(tab, K, base, alpha, beta, ot, n_out, v0, v1,
                                                                  peer_rows, dev_weights);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_weighted_sum(const float* updates, int64_t ld, int K, const float* weights, const float* dev_weights,
                     const float* base, float alpha, float beta, float* out, int64_t D, void* stream) {
  DDL_REQUIRE(K >= 1, "K must be >= 1");
  DDL_REQUIRE(ld % 4 == 0 && D % 4 == 0, "ld and D must be multiples of 4 floats");
  // more rows than one launch's table holds: chain launches, accumulating into `out`
  int done = 0;
  while (done < K) {
    const int kk = (K - done) < kMaxRows ? (K - done) : kMaxRows;
    const float* rows[kMaxRows];
    for (int k = 0; k < kk; ++k) rows[k] = updates + (int64_t)(done + k) * ld;
    float* outs[1] = {out};
    const float* b = done == 0 ? base : out;
    const float a = done == 0 ? alpha : 1.f;
    if (int e = ddl_weighted_sum_rows(rows, weights ? weights + done : nullptr,
                                      dev_weights ? dev_weights + done : nullptr, kk, b, a, beta, outs, 1, 0, D, 0, 0,
                                      stream))
      return e;
    done += kk;
  }
  return DDL_OK;
}

int ddl_rank_select_mean_rows(const float* const* rows, int K, int lo, int hi, float* const* outs, int n_out,
                              int64_t d0, int64_t d1, void* stream) {
  DDL_REQUIRE(K >= 1 && K <= 128, "rank-select supports 1 <= K <= 128 rows, got %d", K);
  DDL_REQUIRE(0 <= lo && lo < hi && hi <= K, "bad rank window [%d, %d) for K=%d", lo, hi, K);
  if (d1 <= d0) return DDL_OK;
  RowTable tab;
  OutTable ot;
  if (int e = fill_tables(tab, ot, rows, nullptr, K, outs, n_out, 3)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  if (K <= 4) return launch_rank_select<4>(tab, K, lo, hi, ot, n_out, d0, d1, st);
  if (K <= 8) return launch_rank_select<8>(tab, K, lo, hi, ot, n_out, d0, d1, st);
  if (K <= 16) return launch_rank_select<16>(tab, K, lo, hi, ot, n_out, d0, d1, st);
  if (K <= 32) return launch_rank_select<32>(tab, K, lo, hi, ot, n_out, d0, d1, st);
  if (K <= 64) return launch_rank_select<64>(tab, K, lo, hi, ot, n_out, d0, d1, st);
  return launch_rank_select<128>(tab, K, lo, hi, ot, n_out, d0, d1, st);
}

static int rank_select_contig(const float* updates, int64_t ld, int K, int lo, int hi, float* out, int64_t D,
                              void* stream) {
  DDL_REQUIRE(K >= 1 && K <= 128, "rank-select supports 1 <= K <= 128 rows, got %d", K);
  const float* rows[128];
  for (int k = 0; k < K; ++k) rows[k] = updates + (int64_t)k * ld;
  float* outs[1] = {out};
  return ddl_rank_select_mean_rows(rows, K, lo, hi, outs, 1, 0, D, stream);
}

int ddl_coordinate_median(const float* updates, int64_t ld, int K, int lower, float* out, int64_t D, void* stream) {
  int lo, hi;
  if ((K & 1) || lower) { lo = (K - 1) / 2; hi = lo + 1; }
  else { lo = K / 2 - 1; hi = K / 2 + 1; }
  return rank_select_contig(updates, ld, K, lo, hi, out, D, stream);
}

int ddl_trimmed_mean(const float* updates, int64_t ld, int K, int trim, float* out, int64_t D, void* stream) {
  DDL_REQUIRE(trim >= 0 && 2 * trim < K, "trim=%d invalid for K=%d", trim, K);
  return rank_select_contig(updates, ld, K, trim, K - trim, out, D, stream);
}

int ddl_attack_perturb(const float* update, const float* global_or_null, float* out, float scale, float noise,
                       uint32_t seed, int64_t D, void* stream) {
  DDL_REQUIRE(D % 4 == 0, "D must be a multiple of 4 floats");
  if (D == 0) return DDL_OK;
  attack_perturb_kernel<<<grid_for(D / 4, 256, 8), 256, 0, (cudaStream_t)stream>>>(update, global_or_null, out, scale,
                                                                                    noise, seed, D / 4);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_krum_select(const float* gram, int ldg, int K, int n_byzantine, int multi, float* scores, int* sel,
                    float* selw, void* stream) {
  DDL_REQUIRE(K >= 1 && K <= 128, "Krum select supports K <= 128, got %d", K);
  DDL_REQUIRE(K - n_byzantine - 2 >= 1, "Krum needs K >= f + 3 (K=%d, f=%d)", K, n_byzantine);
  DDL_REQUIRE(multi >= 1 && multi <= K, "multi=%d outside [1, K]", multi);
  krum_select_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(gram, ldg, K, n_byzantine, multi, scores, sel, selw);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
