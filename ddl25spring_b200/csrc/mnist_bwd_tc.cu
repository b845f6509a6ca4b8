// Packed MnistCnn engine — conv2 backward on the tensor cores (K9 dgrad / wgrad).
//
//  conv2_dgrad : da1[q][ci] = relu'(a1[q][ci]) * sum_{tap,co} dy2[q - off(tap)][co] * W2[co][ci][tap]
//                implicit GEMM, M = 128 flat positions, N = 32, K = 9 taps x 64 couts.  One staged
//                window of dy2 (two 32-column halves) serves all 9 taps through shifted start addresses.
//  conv2_wgrad : dW2[co][ci][r][s] = sum_q a1[q + r*26 + s][ci] * dy2[q][co]
//                implicit GEMM with the reduction over positions.  A = a1 window read M-major with
//                FOUR OVERLAPPING atoms (atom s starts s rows later: LBO = one 128 B row), so a single
//                MMA covers the three horizontal taps of a filter row (+1 unused atom); B = dy2 N-major.
//                Both operands use the 128B swizzle with 32B atoms (the only legal layout for 32-bit
//                MN-major data).  Each CTA writes its partial sums; small_update reduces them in order.
//
// Both rely on the swizzle being a pure function of the shared-memory address (csrc/dbg_probe.cu).
// Replaces (reference): autograd backward of conv2 (lab/tutorial_1a/hfl_complete.py:53,79).
#include "common.cuh"
#include "ddl25_kernels.h"
#include "mnist_common.cuh"
#include "tc05.cuh"
#include "tmap.h"

namespace mnist {
StepInfo make_step(int G, int B, const int* bsz, uint32_t seed, float p1, float p2, float lr);
}

namespace {
using namespace mnist;
using namespace tc05;

constexpr int MAX_OFF = 2 * H1 + 2;                    // 54: largest tap offset on the flat grid
constexpr int WIN_BYTES = WIN_ROWS * 128;              // 23 552

// ------------------------------------------------------------------------------------------ dgrad
constexpr int DG_STAGES = 3;
constexpr int DG_SUB = 1;                              // accumulators per tile (interleaving chains did not help: DESIGN.md finding 7)
constexpr int DG_EPI_GROUPS = 2;                       // epilogue warpgroups on alternating tiles / TMEM buffers
constexpr int DG_THREADS = 64 + DG_EPI_GROUPS * 128;
constexpr int DG_STAGE_BYTES = 2 * WIN_BYTES;          // two 32-column halves of the dy2 window
constexpr int DG_W_BYTES = 9 * 2 * C1 * 128;           // [tap][half][ci 32][co_l 32] = 72 KB
constexpr int DG_SMEM = DG_W_BYTES + DG_STAGES * DG_STAGE_BYTES + 256 + 1024;

struct DgradParams {
  const float* a1;
  float* da1;
  int ctas_per_client;
  StepInfo si;
};

__global__ void __launch_bounds__(DG_THREADS, 1)
conv2_dgrad_kernel(const __grid_constant__ CUtensorMap tm_dy2, const __grid_constant__ CUtensorMap tm_w, const DgradParams prm) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_w = smem;
  uint8_t* s_win = smem + DG_W_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_win + DG_STAGES * DG_STAGE_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + DG_STAGES;
  uint64_t* acc_full = bars + 2 * DG_STAGES;
  uint64_t* acc_empty = bars + 2 * DG_STAGES + 2;
  uint64_t* w_bar = bars + 2 * DG_STAGES + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * DG_STAGES + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const StepInfo& si = prm.si;
  const int g = blockIdx.x / prm.ctas_per_client;
  const int j = blockIdx.x % prm.ctas_per_client;
  const int npos = si.bsz[g] * PIX1;                   // valid flat positions of this client
  const int ntiles = (npos + TILE_M - 1) / TILE_M;
  const int64_t pos0 = (int64_t)g * si.B * PIX1;       // first flat position of this client

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tm_dy2); tma_prefetch_desc(&tm_w); }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < DG_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
    mbar_init(w_bar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<2 * DG_SUB * C1>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one() && ntiles > 0) {
      mbar_arrive_expect_tx(w_bar, DG_W_BYTES);
      for (int i = 0; i < 3; ++i) tma_load_2d(s_w + i * (DG_W_BYTES / 3), &tm_w, w_bar, 0, g * 576 + i * 192);
      uint32_t phase = 0;
      int s = 0;
      for (int t = j; t < ntiles; t += prm.ctas_per_client) {
        mbar_wait(&empty_bar[s], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], DG_STAGE_BYTES);
        const int r0 = (int)(pos0 + (int64_t)t * TILE_M) - MAX_OFF;      // may be negative: TMA zero-fills
        tma_load_2d(s_win + s * DG_STAGE_BYTES, &tm_dy2, &full_bar[s], 0, r0);
        tma_load_2d(s_win + s * DG_STAGE_BYTES + WIN_BYTES, &tm_dy2, &full_bar[s], 32, r0);
        if (++s == DG_STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (ntiles > 0) {
      constexpr uint32_t idesc = make_idesc(FMT_TF32, TILE_M, C1, false, false);
      mbar_wait(w_bar, 0);
      uint32_t phase = 0, acc_phase = 0;
      int s = 0, buf = 0;
      for (int t = j; t < ntiles; t += prm.ctas_per_client) {
        mbar_wait(&acc_empty[buf], acc_phase ^ 1);
        mbar_wait(&full_bar[s], phase);
        fence_after_thread_sync();
        if (elect_one()) {
          const uint64_t a_base = smem_desc_sw128(smem_u32(s_win + s * DG_STAGE_BYTES), 16, 1024);
          const uint64_t b_base = smem_desc_sw128(smem_u32(s_w), 16, 1024);
          const uint32_t acc = tmem_base + buf * (DG_SUB * C1);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
              for (int kk = 0; kk < 4; ++kk)    // window row (q - off): shift = 54 - off
                mma_tf32(acc + (((tap * 2 + hh) * 4 + kk) % DG_SUB) * C1, desc_advance(a_base, hh * WIN_BYTES + (MAX_OFF - ((tap / 3) * H1 + (tap % 3))) * 128 + kk * 32),
                         desc_advance(b_base, (tap * 2 + hh) * (C1 * 128) + kk * 32), idesc, ((tap * 2 + hh) * 4 + kk) >= DG_SUB);
            }
          }
          mma_commit(&empty_bar[s]);
          mma_commit(&acc_full[buf]);
        }
        __syncwarp();
        if (++s == DG_STAGES) { s = 0; phase ^= 1; }
        if (++buf == 2) { buf = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // epilogue: ReLU mask from a1, coalesced 128 B rows straight to global
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    uint32_t acc_phase = 0;
    const int buf = (warp - 2) >> 2;                  // warpgroup w drains TMEM buffer w = every other tile
    for (int t = j + buf * prm.ctas_per_client; t < ntiles; t += DG_EPI_GROUPS * prm.ctas_per_client) {
      mbar_wait(&acc_full[buf], acc_phase);
      fence_after_thread_sync();
      float v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * (DG_SUB * C1), v);
#pragma unroll
      for (int sub = 1; sub < DG_SUB; ++sub) {
        float u[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * (DG_SUB * C1) + sub * C1, u);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += u[i];
      }
      fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      const int lp = t * TILE_M + m;
      if (lp < npos) {
        const float4* act = reinterpret_cast<const float4*>(prm.a1 + (pos0 + lp) * C1);
        float4* dst = reinterpret_cast<float4*>(prm.da1 + (pos0 + lp) * C1);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 a = act[i];
          float4 o;
          o.x = a.x > 0.f ? v[4 * i] : 0.f; o.y = a.y > 0.f ? v[4 * i + 1] : 0.f;
          o.z = a.z > 0.f ? v[4 * i + 2] : 0.f; o.w = a.w > 0.f ? v[4 * i + 3] : 0.f;
          dst[i] = o;
        }
      }
      acc_phase ^= 1;
    }
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<2 * DG_SUB * C1>(tmem_base);
  }
}

}  // namespace

extern "C" int ddl_mnist_conv2_dgrad(const float* wt_dgrad, const float* dy2, const float* a1, float* da1, int G, int B,
                                     const int* bsz, void* stream) {
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  StepInfo si = make_step(G, B, bsz, 0, 0.f, 0.f, 0.f);
  CUtensorMap tm_d, tm_w;
  {
    uint64_t dims[2] = {(uint64_t)C2, (uint64_t)G * B * PIX1};
    uint64_t strides[1] = {256};
    uint32_t box[2] = {32, (uint32_t)WIN_ROWS};
    if (int e = ddl_encode_tmap(&tm_d, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dy2, dims, strides, box, 1)) return e;
  }
  {
    uint64_t dims[2] = {32, (uint64_t)G * 576};
    uint64_t strides[1] = {128};
    uint32_t box[2] = {32, 192};
    if (int e = ddl_encode_tmap(&tm_w, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, wt_dgrad, dims, strides, box, 1)) return e;
  }
  DgradParams prm;
  prm.a1 = a1; prm.da1 = da1; prm.si = si; prm.ctas_per_client = kNumSMs / G;
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(conv2_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DG_SMEM));
    configured.done();
  }
  DDL_CUDA_CHECK(ddl_launch_pdl(conv2_dgrad_kernel, dim3(prm.ctas_per_client * G), dim3(DG_THREADS), DG_SMEM, (cudaStream_t)stream, tm_d, tm_w, prm));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}
