// Packed MnistCnn engine — classifier head (K13 + K15 fused).
//
//  head_train : for every client g (one CTA each): dropout2 -> fc2 -> log_softmax -> NLL (mean over
//               the client's valid rows) -> d logits -> dW_fc2 / db_fc2 / d h1 -> ReLU' and dropout2
//               masks -> dz1 (input of the fc1 backward GEMMs) and db_fc1; then applies the SGD step
//               to fc2.weight, fc2.bias and fc1.bias in place.  fc2 is 128 -> 10: far too small for
//               tensor cores, so one warp walks a row with shuffle reductions.
//  head_eval  : logits + top-1 count for Server.test().
//
// Replaces (reference): dropout2 / fc2 / log_softmax / F.nll_loss / backward / optimizer.step of
// MnistCnn (lab/tutorial_1a/hfl_complete.py:59-62,78-80) and the accuracy count of Server.test (:172-183).
#include "common.cuh"
#include "ddl25_kernels.h"
#include "mnist_common.cuh"

namespace mnist {
StepInfo make_step(int G, int B, const int* bsz, uint32_t seed, float p1, float p2, float lr);
}

namespace {
using namespace mnist;

constexpr int kHeadWarps = 8;
constexpr int kHeadParts = 4;                         // CTAs per client (rows interleaved): a single CTA per client was latency-bound
constexpr int kPartial = NCLS * FC1 + FC1 + NCLS;   // dW2 | db1 | db2 partial sums of one warp

__global__ void __launch_bounds__(kHeadWarps * 32, 1)
head_train_kernel(float* __restrict__ W, const float* __restrict__ h1, const int64_t* __restrict__ Y, int64_t y_client_stride,
                  const int* __restrict__ perm, int64_t perm_stride, int sample_off, float* __restrict__ dz1,
                  float* __restrict__ part_out, const StepInfo si) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float sm[];
  float* s_w2 = sm;                       // [10][128]
  float* s_b2 = s_w2 + NCLS * FC1;        // [10] (+6 pad)
  float* s_part = s_b2 + 16;              // [kHeadWarps][kPartial]
  __shared__ float s_loss[kHeadWarps];
  const int g = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* Wg = W + (int64_t)g * P;
  for (int i = threadIdx.x; i < NCLS * FC1; i += blockDim.x) s_w2[i] = Wg[OFF_F2W + i];
  if (threadIdx.x < NCLS) s_b2[threadIdx.x] = Wg[OFF_F2B + threadIdx.x];
  __syncthreads();

  const int n = si.bsz[g];
  const float inv_n = n > 0 ? 1.f / (float)n : 0.f;
  const float keep_scale = 1.f / (1.f - si.p2);
  float dw2[NCLS][4];
  float db1[4] = {0.f, 0.f, 0.f, 0.f};
  float db2 = 0.f, loss = 0.f;
#pragma unroll
  for (int c = 0; c < NCLS; ++c)
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) dw2[c][jj] = 0.f;

  for (int b = blockIdx.y * kHeadWarps + warp; b < si.B; b += kHeadWarps * kHeadParts) {
    const int64_t row = (int64_t)g * si.B + b;
    float4* dz = reinterpret_cast<float4*>(dz1 + row * FC1) + lane;
    if (b >= n) {                      // rows beyond this client's minibatch must contribute nothing
      *dz = make_float4(0.f, 0.f, 0.f, 0.f);
      continue;
    }
    const float4 hr = reinterpret_cast<const float4*>(h1 + row * FC1)[lane];
    const float raw[4] = {hr.x, hr.y, hr.z, hr.w};
    float h[4];
    bool keep[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      keep[jj] = si.p2 > 0.f ? ddl::rng_uniform(si.seed ^ 0x5bd1e995u, (uint32_t)(row * FC1 + lane * 4 + jj)) >= si.p2 : true;
      h[jj] = keep[jj] ? raw[jj] * keep_scale : 0.f;
    }
    float logit[NCLS];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) {
      const float4 w = reinterpret_cast<const float4*>(s_w2 + c * FC1)[lane];
      float p = h[0] * w.x + h[1] * w.y + h[2] * w.z + h[3] * w.w;
      p = ddl::warp_sum(p) + s_b2[c];
      logit[c] = p;
      mx = fmaxf(mx, p);
    }
    float se = 0.f;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) se += __expf(logit[c] - mx);
    const float lse = mx + __logf(se);
    const int idx = perm ? perm[(int64_t)g * perm_stride + sample_off + b] : sample_off + b;
    const int t = (int)Y[(int64_t)g * y_client_stride + idx];
    float dh[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < NCLS; ++c) {
      const float dl = (__expf(logit[c] - lse) - (c == t ? 1.f : 0.f)) * inv_n;
      if (c == t) loss += (lse - logit[c]) * inv_n;
      if (lane == c) db2 += dl;
      const float4 w = reinterpret_cast<const float4*>(s_w2 + c * FC1)[lane];
      dh[0] = fmaf(dl, w.x, dh[0]); dh[1] = fmaf(dl, w.y, dh[1]);
      dh[2] = fmaf(dl, w.z, dh[2]); dh[3] = fmaf(dl, w.w, dh[3]);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) dw2[c][jj] = fmaf(dl, h[jj], dw2[c][jj]);
    }
    float o[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      o[jj] = (keep[jj] && raw[jj] > 0.f) ? dh[jj] * keep_scale : 0.f;
      db1[jj] += o[jj];
    }
    *dz = make_float4(o[0], o[1], o[2], o[3]);
  }
  // ---- deterministic cross-warp reduction of the parameter gradients, then the SGD step
  float* part = s_part + warp * kPartial;
#pragma unroll
  for (int c = 0; c < NCLS; ++c)
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) part[c * FC1 + lane * 4 + jj] = dw2[c][jj];
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) part[NCLS * FC1 + lane * 4 + jj] = db1[jj];
  if (lane < NCLS) part[NCLS * FC1 + FC1 + lane] = db2;
  if (lane == 0) s_loss[warp] = loss;
  __syncthreads();
  // block partial -> global scratch [g][part][kPartial + 1]; head_apply_kernel sums the parts in a fixed order
  float* gp = part_out + ((int64_t)g * kHeadParts + blockIdx.y) * (kPartial + 1);
  for (int i = threadIdx.x; i < kPartial; i += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kHeadWarps; ++w) s += s_part[w * kPartial + i];
    gp[i] = s;
  }
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < kHeadWarps; ++w) s += s_loss[w];
    gp[kPartial] = s;
  }
}

__global__ void __launch_bounds__(256) head_apply_kernel(float* __restrict__ W, const float* __restrict__ part, float* __restrict__ loss_out,
                                                         float lr) {
  pdl_launch_dependents();
  pdl_wait();
  const int g = blockIdx.x;
  float* Wg = W + (int64_t)g * P;
  const float* gp = part + (int64_t)g * kHeadParts * (kPartial + 1);
  for (int i = threadIdx.x; i <= kPartial; i += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < kHeadParts; ++q) s += gp[q * (kPartial + 1) + i];
    if (i == kPartial) {
      if (loss_out) loss_out[g] = s;
      continue;
    }
    float* dst = i < NCLS * FC1 ? Wg + OFF_F2W + i
                                : (i < NCLS * FC1 + FC1 ? Wg + OFF_F1B + (i - NCLS * FC1) : Wg + OFF_F2B + (i - NCLS * FC1 - FC1));
    *dst -= lr * s;
  }
}

// eval: one warp per row, weights of "client" 0 (the global model); counts top-1 hits
__global__ void __launch_bounds__(256) head_eval_kernel(const float* __restrict__ W, const float* __restrict__ h1,
                                                        const int64_t* __restrict__ Y, int sample_off, int nrows,
                                                        float* __restrict__ logits_out, int* __restrict__ correct) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_w2[NCLS * FC1];
  __shared__ float s_b2[16];
  for (int i = threadIdx.x; i < NCLS * FC1; i += blockDim.x) s_w2[i] = W[OFF_F2W + i];
  if (threadIdx.x < NCLS) s_b2[threadIdx.x] = W[OFF_F2B + threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nw = (gridDim.x * blockDim.x) >> 5;
  int hits = 0;
  for (int row = wid; row < nrows; row += nw) {
    const float4 hr = reinterpret_cast<const float4*>(h1 + (int64_t)row * FC1)[lane];
    float best = -INFINITY;
    int arg = 0;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) {
      const float4 w = reinterpret_cast<const float4*>(s_w2 + c * FC1)[lane];
      float p = hr.x * w.x + hr.y * w.y + hr.z * w.z + hr.w * w.w;
      p = ddl::warp_sum(p) + s_b2[c];
      if (logits_out && lane == 0) logits_out[(int64_t)row * NCLS + c] = p;
      if (p > best) { best = p; arg = c; }
    }
    if (Y && lane == 0 && arg == (int)Y[sample_off + row]) ++hits;
  }
  if (correct && lane == 0 && hits) atomicAdd(correct, hits);
}

}  // namespace

extern "C" {

int ddl_mnist_head_scratch_floats(int G) { return G * kHeadParts * (kPartial + 1); }

int ddl_mnist_head_train(float* W, const float* h1, const int64_t* Y, int64_t y_client_stride, const int* perm,
                         int64_t perm_stride, int sample_off, float* dz1, float* loss_out, float* scratch, int G, int B,
                         const int* bsz, float p_drop, uint32_t seed, float lr, void* stream) {
  DDL_REQUIRE(scratch != nullptr, "head_train needs ddl_mnist_head_scratch_floats(G) floats of scratch");
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  DDL_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "dropout probability must be in [0, 1)");
  StepInfo si = make_step(G, B, bsz, seed, 0.f, p_drop, lr);
  const int smem = (NCLS * FC1 + 16 + kHeadWarps * kPartial) * (int)sizeof(float);
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(head_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured.done();
  }
  dim3 grid(G, kHeadParts);
  DDL_CUDA_CHECK(ddl_launch_pdl(head_train_kernel, dim3(grid), dim3(kHeadWarps * 32), smem, (cudaStream_t)stream, W, h1, Y, y_client_stride, perm, perm_stride,
                                                                           sample_off, dz1, scratch, si));
  DDL_LAUNCH_CHECK();
  DDL_CUDA_CHECK(ddl_launch_pdl(head_apply_kernel, dim3(G), dim3(256), 0, (cudaStream_t)stream, W, scratch, loss_out, lr));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_mnist_head_eval(const float* W, const float* h1, const int64_t* Y, int sample_off, int nrows, float* logits_out,
                        int* correct, void* stream) {
  if (nrows <= 0) return DDL_OK;
  int blocks = (nrows + 7) / 8;
  if (blocks > kNumSMs * 4) blocks = kNumSMs * 4;
  DDL_CUDA_CHECK(ddl_launch_pdl(head_eval_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, W, h1, Y, sample_off, nrows, logits_out, correct));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
