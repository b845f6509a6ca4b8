// Packed MnistCnn engine — CUDA-core backward kernels and the small-tensor SGD step.
//
//  pool_relu_bwd : d pooled -> d conv2 pre-activation on the 26-pitch NHWC grid, through the saved
//                  arg-max / dropout / ReLU mask; also per-sample conv2.bias gradient partials.
//  conv1_wgrad   : conv1.weight / conv1.bias gradient partials per sample (Cin = 1: 9 taps).
//  small_update  : ordered reduction of all partials (deterministic), SGD on conv1 / conv2 weights and
//                  biases, and re-pack of the new conv2 weights into the tensor-core tile layouts.
//
// Replaces (reference): autograd backward of max_pool2d / relu / dropout1 / conv1 and the SGD step
// on those tensors (lab/tutorial_1a/hfl_complete.py:51-56,79-80).
#include "common.cuh"
#include "ddl25_kernels.h"
#include "mnist_common.cuh"

namespace mnist {
StepInfo make_step(int G, int B, const int* bsz, uint32_t seed, float p1, float p2, float lr);
}

namespace {
using namespace mnist;

constexpr int CH_PITCH = 145;   // padded channel stride of the staged pooled-gradient (bank spread)

__global__ void __launch_bounds__(256) pool_relu_bwd_kernel(const float* __restrict__ dpool, const uint8_t* __restrict__ pmask,
                                                            float* __restrict__ dy2, float* __restrict__ c2b_part,
                                                            const StepInfo si) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_dp[C2 * CH_PITCH];
  __shared__ uint8_t s_mk[C2 * CH_PITCH];
  __shared__ float s_red[8 * C2];
  const int row = blockIdx.x;
  const int g = row / si.B, b = row % si.B;
  const int n = si.bsz[g];
  if (b > n) return;                               // rows past the guard sample are never read
  float4* out = reinterpret_cast<float4*>(dy2 + (int64_t)row * PIX1 * C2);
  if (b == n) {                                    // guard sample: the wgrad k-loop may overrun into it
    for (int i = threadIdx.x; i < PIX1 * C2 / 4; i += 256) out[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  const float* dp = dpool + (int64_t)row * FLAT;
  const uint8_t* mk = pmask + (int64_t)row * FLAT;
  for (int i = threadIdx.x; i < FLAT; i += 256) {
    const int c = i / (HP * HP), r = i % (HP * HP);
    s_dp[c * CH_PITCH + r] = dp[i];
    s_mk[c * CH_PITCH + r] = mk[i];
  }
  __syncthreads();
  const float scale = 1.f / (1.f - si.p1);
  const int q = threadIdx.x & 15;                  // this thread always serves channels 4q..4q+3
  float bsum[4] = {0.f, 0.f, 0.f, 0.f};
  for (int item = threadIdx.x; item < PIX1 * 16; item += 256) {
    const int p = item >> 4;
    const int h = p / H1, w = p % H1;
    float o[4] = {0.f, 0.f, 0.f, 0.f};
    if (h < H2 && w < H2) {
      const int pr = (h >> 1) * HP + (w >> 1);
      const int me = ((h & 1) << 1) | (w & 1);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int idx = (4 * q + jj) * CH_PITCH + pr;
        const int m = s_mk[idx];
        if ((m & 4) && (m & 3) == me) o[jj] = ddl::to_tf32(s_dp[idx] * scale);
        bsum[jj] += o[jj];
      }
    }
    out[p * 16 + q] = make_float4(o[0], o[1], o[2], o[3]);
  }
  // lanes l and l+16 serve the same channel quad: fold them, then sum the 8 warps in a fixed order
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    bsum[jj] += __shfl_xor_sync(0xffffffffu, bsum[jj], 16);
    if ((threadIdx.x & 31) < 16) s_red[(threadIdx.x >> 5) * C2 + q * 4 + jj] = bsum[jj];
  }
  __syncthreads();
  if (threadIdx.x < C2) {
    float s = 0.f;
    for (int k = 0; k < 8; ++k) s += s_red[k * C2 + threadIdx.x];
    c2b_part[(int64_t)row * C2 + threadIdx.x] = s;
  }
}

// per-sample partial of conv1 gradients: part[row][0..287] = dW[c][tap], part[row][288..319] = db[c]
__global__ void __launch_bounds__(256) conv1_wgrad_kernel(const float* __restrict__ X, int64_t x_client_stride,
                                                          const int* __restrict__ perm, int64_t perm_stride, int sample_off,
                                                          const float* __restrict__ da1, float* __restrict__ c1_part,
                                                          const StepInfo si) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float img[IMG * IMG];
  __shared__ float s_red[256 * 41];
  const int row = blockIdx.x;
  const int g = row / si.B, b = row % si.B;
  if (b >= si.bsz[g]) return;
  const int idx = perm ? perm[(int64_t)g * perm_stride + sample_off + b] : sample_off + b;
  const float4* src = reinterpret_cast<const float4*>(X + (int64_t)g * x_client_stride + (int64_t)idx * (IMG * IMG));
  for (int i = threadIdx.x; i < IMG * IMG / 4; i += 256) reinterpret_cast<float4*>(img)[i] = src[i];
  __syncthreads();
  const int q = threadIdx.x & 7;
  float acc[4][9], accb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int jj = 0; jj < 4; ++jj)
#pragma unroll
    for (int t = 0; t < 9; ++t) acc[jj][t] = 0.f;
  const float4* d = reinterpret_cast<const float4*>(da1 + (int64_t)row * PIX1 * C1);
  for (int item = threadIdx.x; item < PIX1 * 8; item += 256) {
    const int p = item >> 3;
    const int h = p / H1, w = p % H1;
    const float4 gv = ddl::ld_stream(d + p * 8 + q);
    const float gg[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const float xv = img[(h + r) * IMG + w + s];
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) acc[jj][r * 3 + s] = fmaf(gg[jj], xv, acc[jj][r * 3 + s]);
      }
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) accb[jj] += gg[jj];
  }
  float* mine = s_red + threadIdx.x * 41;
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
#pragma unroll
    for (int t = 0; t < 9; ++t) mine[jj * 9 + t] = acc[jj][t];
    mine[36 + jj] = accb[jj];
  }
  __syncthreads();
  float* out = c1_part + (int64_t)row * 320;
  for (int o = threadIdx.x; o < 320; o += 256) {
    int qq, slot;
    if (o < 288) { const int c = o / 9, t = o % 9; qq = c >> 2; slot = (c & 3) * 9 + t; }
    else { const int c = o - 288; qq = c >> 2; slot = 36 + (c & 3); }
    float s = 0.f;
    for (int k = 0; k < 32; ++k) s += s_red[(qq + 8 * k) * 41 + slot];
    out[o] = s;
  }
}

// One launch: reduce partials in a fixed order, SGD, re-pack conv2 weights for the next step.
//  wg_part [G][n_cta][3 (r)][128 (s*32+ci)][64 (co)]   conv2 weight-gradient partials
//  c2b_part[G*B][64], c1_part[G*B][320]
__global__ void __launch_bounds__(256) small_update_kernel(float* __restrict__ W, const float* __restrict__ wg_part, int n_cta,
                                                           const float* __restrict__ c2b_part, const float* __restrict__ c1_part,
                                                           float* __restrict__ wt_fwd, float* __restrict__ wt_dgrad,
                                                           const StepInfo si) {
  pdl_launch_dependents();
  pdl_wait();
  const int g = blockIdx.y;
  const int n = si.bsz[g];
  float* Wg = W + (int64_t)g * P;
  const int nthreads = gridDim.x * blockDim.x;
  const int total = 9 * C1 * C2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += nthreads) {
    {
      const int co = i & 63, ci = (i >> 6) & 31, tap = i >> 11;
      const int r = tap / 3, s = tap % 3;
      float* wp = Wg + OFF_C2W + (co * C1 + ci) * 9 + tap;
      float wv = *wp;
      if (n > 0) {
        const float* pp = wg_part + (((int64_t)g * n_cta * 3 + r) * 128 + s * 32 + ci) * 64 + co;
        float sum = 0.f;
#pragma unroll 6
        for (int jc = 0; jc < n_cta; ++jc) sum += pp[(int64_t)jc * 3 * 128 * 64];
        wv -= si.lr * sum;
        *wp = wv;
      }
      const float tv = ddl::to_tf32(wv);
      wt_fwd[(int64_t)g * 9 * C2 * C1 + (tap * C2 + co) * C1 + ci] = tv;
      wt_dgrad[(int64_t)g * 9 * C2 * C1 + ((tap * 2 + (co >> 5)) * C1 + ci) * 32 + (co & 31)] = tv;
    }
  }
  // small reductions over the client's rows: one warp per output, lanes stride over the rows
  if (n > 0) {
    const int lane = threadIdx.x & 31;
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = nthreads >> 5;
    for (int k = wid; k < C2 + 320; k += nwarps) {
      float sum = 0.f;
      if (k < C2) {
        for (int b = lane; b < n; b += 32) sum += c2b_part[((int64_t)g * si.B + b) * C2 + k];
      } else {
        for (int b = lane; b < n; b += 32) sum += c1_part[((int64_t)g * si.B + b) * 320 + (k - C2)];
      }
      sum = ddl::warp_sum(sum);
      if (lane == 0) {
        const int o = k - C2;
        float* dst = k < C2 ? Wg + OFF_C2B + k : Wg + (o < 288 ? OFF_C1W + o : OFF_C1B + (o - 288));
        *dst -= si.lr * sum;
      }
    }
  }
}

}  // namespace

extern "C" {

int ddl_mnist_pool_bwd(const float* dpool, const uint8_t* pmask, float* dy2, float* c2b_part, int G, int B,
                       const int* bsz, float p_drop, void* stream) {
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  StepInfo si = make_step(G, B, bsz, 0, p_drop, 0.f, 0.f);
  DDL_CUDA_CHECK(ddl_launch_pdl(pool_relu_bwd_kernel, dim3(G * B), dim3(256), 0, (cudaStream_t)stream, dpool, pmask, dy2, c2b_part, si));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_mnist_conv1_wgrad(const float* X, int64_t x_client_stride, const int* perm, int64_t perm_stride, int sample_off,
                          const float* da1, float* c1_part, int G, int B, const int* bsz, void* stream) {
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  StepInfo si = make_step(G, B, bsz, 0, 0.f, 0.f, 0.f);
  DDL_CUDA_CHECK(ddl_launch_pdl(conv1_wgrad_kernel, dim3(G * B), dim3(256), 0, (cudaStream_t)stream, X, x_client_stride, perm, perm_stride, sample_off, da1, c1_part, si));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_mnist_small_update(float* W, const float* wg_part, int n_cta, const float* c2b_part, const float* c1_part,
                           float* wt_fwd, float* wt_dgrad, int G, int B, const int* bsz, float lr, void* stream) {
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  StepInfo si = make_step(G, B, bsz, 0, 0.f, 0.f, lr);
  dim3 grid(72, G);          // one conv2 weight per thread: the 18-partial reduction is latency-bound, so go wide
  DDL_CUDA_CHECK(ddl_launch_pdl(small_update_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, W, wg_part, n_cta, c2b_part, c1_part, wt_fwd, wt_dgrad, si));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
