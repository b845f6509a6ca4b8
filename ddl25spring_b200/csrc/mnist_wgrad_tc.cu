// Packed MnistCnn engine — conv2 weight gradient on the tensor cores (see mnist_bwd_tc.cu header).
//
//   D_r[(s, ci)][co] = sum_q a1[q + r*26 + s][ci] * dy2[q][co]        r = 0..2 (one accumulator each)
//
// The reduction index q (flat position) is the MMA K dimension, so both operands are MN-major:
//   A: rows of the staged a1 window are the K rows; the M = 128 "columns" are 4 atoms of 32 channels
//      where atom s is the same data shifted by s rows (LBO = 128 B) -> taps s = 0,1,2 (+1 unused);
//   B: dy2 window, N = 64 couts = two 32-column halves staged as separate tiles (LBO = tile size).
// Each CTA owns a slice of one client's positions and writes its 3 x 96 x 64 partial sums.
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

constexpr int WG_KB = 64;                              // positions per k-block
constexpr int WG_AROWS = 120;                          // 64 + 2*26 + 3 = 119 -> multiple of 8
constexpr int WG_A_BYTES = WG_AROWS * 128;             // 15 360
constexpr int WG_BH_BYTES = WG_KB * 128;               // one 32-column half of dy2: 8 192
constexpr int WG_STAGE_BYTES = WG_A_BYTES + 2 * WG_BH_BYTES;   // 31 744
constexpr int WG_STAGES = 6;
constexpr int WG_SMEM = WG_STAGES * WG_STAGE_BYTES + 256 + 1024;

struct WgradParams {
  float* part;           // [G][ctas_per_client][3][128][64]
  int ctas_per_client;
  StepInfo si;
};

__global__ void __launch_bounds__(192, 1)
conv2_wgrad_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_dy2, const WgradParams prm) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + WG_STAGES * WG_STAGE_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + WG_STAGES;
  uint64_t* acc_bar = bars + 2 * WG_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * WG_STAGES + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const StepInfo& si = prm.si;
  const int g = blockIdx.x / prm.ctas_per_client;
  const int j = blockIdx.x % prm.ctas_per_client;
  const int npos = si.bsz[g] * PIX1;
  const int nkb = (npos + WG_KB - 1) / WG_KB;
  // contiguous slice of k-blocks for this CTA
  const int per = (nkb + prm.ctas_per_client - 1) / prm.ctas_per_client;
  const int kb0 = j * per;
  int kb1 = kb0 + per;
  if (kb1 > nkb) kb1 = nkb;
  const int my_kb = kb1 > kb0 ? kb1 - kb0 : 0;
  const int cpos0 = g * si.B * PIX1;                   // first flat position of this client (a1 map is 2-D)

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tm_a1); tma_prefetch_desc(&tm_dy2); }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(acc_bar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<256>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      uint32_t phase = 0;
      int s = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[s], phase ^ 1);
        uint8_t* st = smem + s * WG_STAGE_BYTES;
        mbar_arrive_expect_tx(&full_bar[s], WG_STAGE_BYTES);
        const int q0 = kb * WG_KB;                     // client-local position
        tma_load_2d(st, &tm_a1, &full_bar[s], 0, cpos0 + q0);
        tma_load_3d(st + WG_A_BYTES, &tm_dy2, &full_bar[s], 0, q0, g);
        tma_load_3d(st + WG_A_BYTES + WG_BH_BYTES, &tm_dy2, &full_bar[s], 32, q0, g);
        if (++s == WG_STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, C2, true, true);
    uint32_t phase = 0;
    int s = 0;
    for (int kb = 0; kb < my_kb; ++kb) {
      mbar_wait(&full_bar[s], phase);
      fence_after_thread_sync();
      if (elect_one()) {
        const uint32_t a_win = smem_u32(smem + s * WG_STAGE_BYTES);
        const uint64_t a_base = smem_desc_sw128(a_win, 128, 512, LAYOUT_SW128_BASE32B);
        const uint64_t b_base = smem_desc_sw128(a_win + WG_A_BYTES, WG_BH_BYTES, 512, LAYOUT_SW128_BASE32B);
        // kk outer / r inner: consecutive MMAs go to the three different accumulators, so the dependent
        // chains (latency-bound at N = 64) overlap
#pragma unroll
        for (int kk = 0; kk < WG_KB / 8; ++kk) {
#pragma unroll
          for (int r = 0; r < 3; ++r)
            mma_tf32(tmem_base + r * C2, desc_advance(a_base, (kk * 8 + r * H1) * 128), desc_advance(b_base, kk * 8 * 128),
                     idesc, (kb | kk) != 0);
        }
        mma_commit(&empty_bar[s]);
        if (kb == my_kb - 1) mma_commit(acc_bar);
      }
      __syncwarp();
      if (++s == WG_STAGES) { s = 0; phase ^= 1; }
    }
  } else {
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;                 // (s, ci) = (m / 32, m % 32); s == 3 is the unused atom
    if (my_kb > 0) {
      mbar_wait(acc_bar, 0);
      fence_after_thread_sync();
    }
    float* out = prm.part + ((int64_t)(g * prm.ctas_per_client + j) * 3) * 128 * 64;
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c0 = 0; c0 < C2; c0 += 32) {
        float v[32];
        if (my_kb > 0) {
          tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + r * C2 + c0, v);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0.f;
        }
        if (m < 96) {
          float4* dst = reinterpret_cast<float4*>(out + ((int64_t)r * 128 + m) * 64 + c0);
#pragma unroll
          for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      }
    }
    fence_before_thread_sync();
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<256>(tmem_base);
  }
}

}  // namespace

extern "C" int ddl_mnist_conv2_wgrad(const float* a1, const float* dy2, float* wg_part, int G, int B, const int* bsz,
                                     void* stream) {
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  StepInfo si = make_step(G, B, bsz, 0, 0.f, 0.f, 0.f);
  CUtensorMap tm_a, tm_d;
  {
    uint64_t dims[2] = {(uint64_t)C1, (uint64_t)G * B * PIX1};
    uint64_t strides[1] = {128};
    uint32_t box[2] = {32, (uint32_t)WG_AROWS};
    if (int e = ddl_encode_tmap(&tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, a1, dims, strides, box, 2)) return e;
  }
  {
    // 3-D so that positions past the client's B*676 rows are zero-filled instead of read from the next client
    uint64_t dims[3] = {(uint64_t)C2, (uint64_t)B * PIX1, (uint64_t)G};
    uint64_t strides[2] = {256, (uint64_t)B * PIX1 * 256};
    uint32_t box[3] = {32, (uint32_t)WG_KB, 1};
    if (int e = ddl_encode_tmap(&tm_d, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, dy2, dims, strides, box, 2)) return e;
  }
  WgradParams prm;
  prm.part = wg_part; prm.si = si; prm.ctas_per_client = kNumSMs / G;
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
    configured.done();
  }
  DDL_CUDA_CHECK(ddl_launch_pdl(conv2_wgrad_kernel, dim3(prm.ctas_per_client * G), dim3(192), WG_SMEM, (cudaStream_t)stream, tm_a, tm_d, prm));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

extern "C" int ddl_mnist_wgrad_ctas(int G) { return G >= 1 && G <= mnist::MAX_G ? kNumSMs / G : 0; }
