// Packed MnistCnn engine — forward kernels (K9 conv fprop, K12 fused ReLU/pool/dropout epilogue).
//
//  conv1_relu      : direct 3x3 conv (Cin = 1) + bias + ReLU on CUDA cores, gathers each client's
//                    minibatch through its shuffle permutation, writes NHWC tf32-rounded a1.
//  prep_conv2      : re-packs conv2.weight of every client into the two K-major tile layouts the
//                    tensor-core kernels consume (fprop: [tap][cout][cin]; dgrad: [tap][half][cin][cout/2]).
//  conv2_pool_fwd  : implicit-GEMM 3x3 conv (32 -> 64) on tcgen05 + bias + ReLU + 2x2 max-pool +
//                    dropout in the epilogue.  One TMA load stages a 184-row window of the flat
//                    26-pitch activation grid; the 9 filter taps are 9 start-address offsets into
//                    that same window (the 128B-swizzle is a pure function of the shared-memory
//                    address, verified on B200 by csrc/dbg_probe.cu), so each input byte crosses
//                    L2 -> SMEM once instead of nine times.  TF32 operands, fp32 accumulation in
//                    TMEM (the reference's cuDNN convs run TF32 by default: torch.backends.cudnn.allow_tf32).
//
// Replaces (reference): MnistCnn.forward conv1/relu/conv2/relu/max_pool2d/dropout1
// (lab/tutorial_1a/hfl_complete.py:50-56) for all clients packed on the GPU at once.
#include "common.cuh"
#include "ddl25_kernels.h"
#include "mnist_common.cuh"
#include "tc05.cuh"
#include "tmap.h"

namespace {
using namespace mnist;
using namespace tc05;

// ------------------------------------------------------------------------------------------ conv1
__global__ void __launch_bounds__(256) conv1_relu_kernel(const float* __restrict__ W, const float* __restrict__ X,
                                                         int64_t x_client_stride, const int* __restrict__ perm,
                                                         int64_t perm_stride, int sample_off, float* __restrict__ a1,
                                                         const StepInfo si) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float img[IMG * IMG];
  const int row = blockIdx.x;
  const int g = row / si.B, b = row % si.B;
  if (b >= si.bsz[g]) return;
  const int idx = perm ? perm[(int64_t)g * perm_stride + sample_off + b] : sample_off + b;
  const float4* src = reinterpret_cast<const float4*>(X + (int64_t)g * x_client_stride + (int64_t)idx * (IMG * IMG));
  for (int i = threadIdx.x; i < IMG * IMG / 4; i += 256) reinterpret_cast<float4*>(img)[i] = src[i];
  // this thread always serves channels 4q..4q+3: keep their 36 weights + 4 biases in registers
  const int q = threadIdx.x & 7;
  const float* Wg = W + (int64_t)g * P;
  float w[4][9], bias[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    bias[j] = Wg[OFF_C1B + 4 * q + j];
#pragma unroll
    for (int t = 0; t < 9; ++t) w[j][t] = Wg[OFF_C1W + (4 * q + j) * 9 + t];
  }
  __syncthreads();
  float* dst = a1 + (int64_t)row * PIX1 * C1;
  for (int item = threadIdx.x; item < PIX1 * 8; item += 256) {
    const int p = item >> 3;
    const int h = p / H1, x0 = p % H1;
    float acc[4] = {bias[0], bias[1], bias[2], bias[3]};
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const float xv = img[(h + r) * IMG + x0 + s];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = fmaf(xv, w[j][r * 3 + s], acc[j]);
      }
    float4 o;
    o.x = ddl::to_tf32(fmaxf(acc[0], 0.f)); o.y = ddl::to_tf32(fmaxf(acc[1], 0.f));
    o.z = ddl::to_tf32(fmaxf(acc[2], 0.f)); o.w = ddl::to_tf32(fmaxf(acc[3], 0.f));
    reinterpret_cast<float4*>(dst + (int64_t)p * C1)[q] = o;
  }
}

// ------------------------------------------------------------------------------------------ prep
__global__ void __launch_bounds__(256) prep_conv2_kernel(const float* __restrict__ W, float* __restrict__ wt_fwd,
                                                         float* __restrict__ wt_dgrad) {
  pdl_launch_dependents();
  pdl_wait();
  const int g = blockIdx.x;
  const float* w2 = W + (int64_t)g * P + OFF_C2W;            // [co][ci][tap]
  float* f = wt_fwd + (int64_t)g * 9 * C2 * C1;              // [tap][co][ci]
  float* d = wt_dgrad + (int64_t)g * 9 * C2 * C1;            // [tap][half][ci][co_l]
  for (int i = threadIdx.x; i < 9 * C2 * C1; i += 256) {
    const int tap = i / (C2 * C1), rem = i % (C2 * C1);
    const int co = rem / C1, ci = rem % C1;
    const float v = ddl::to_tf32(w2[(co * C1 + ci) * 9 + tap]);
    f[i] = v;
    d[((tap * 2 + (co >> 5)) * C1 + ci) * 32 + (co & 31)] = v;
  }
}

// ------------------------------------------------------------------------------------------ conv2
constexpr int kStages = 3;
// Two epilogue warpgroups work on alternating tiles / TMEM buffers: one 4-warp epilogue (tcgen05.ld -> staging ->
// pool -> store, ~3000 cycles of latency per tile) could not keep up with the ~1200-cycle MMA phase (ncu: tensor
// pipe 38 % active), see profiles/r1_ncu_conv2_fwd_summary.txt.
constexpr int kEpiGroups = 2;
// A chain of tcgen05.mma into ONE accumulator is latency-bound at small N (measured ~90 cycles per dependent
// MMA at N = 64 on B200 vs a 32-48 cycle issue floor), so the 36 MMAs of a tile alternate between FWD_SUB
// independent TMEM accumulators that the epilogue adds up.
constexpr int FWD_SUB = 1;
constexpr int W_TILE_BYTES = C2 * 128;                 // 64 couts x 32 cins x 4 B = 8 KB per tap
constexpr int W_BYTES = 9 * W_TILE_BYTES;              // 72 KB
constexpr int WIN_BYTES = WIN_ROWS * 128;              // 23 552 B (multiple of 1024)
constexpr int STG_PITCH = 68;                          // floats; conflict-free for row writes and column reads
constexpr int STG_BYTES = TILE_M * STG_PITCH * 4;
constexpr int POOL_PITCH = 25;                          // floats per channel of the pooled staging tile (24 + 1 pad)
constexpr int POOL_BYTES = C2 * POOL_PITCH * 4 + C2 * 24 + C2 * 4;   // pooled values | mask bytes | bias
constexpr int CONV2_SMEM = W_BYTES + kStages * WIN_BYTES + kEpiGroups * (STG_BYTES + POOL_BYTES) + 256 + 1024;
constexpr int kConv2Threads = 64 + kEpiGroups * 128;

struct Conv2Params {
  const float* W;        // weight arena base (bias read from OFF_C2B of each client's row)
  float* pooled;
  uint8_t* pmask;
  int ctas_per_client;
  StepInfo si;
};

__device__ __forceinline__ void epi_bar_sync(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

template <bool TRAIN>
__global__ void __launch_bounds__(kConv2Threads, 1)
conv2_pool_fwd_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_w,
                      const Conv2Params prm) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_w = smem;
  uint8_t* s_win = smem + W_BYTES;
  const int wg = (threadIdx.x >= 64) ? ((threadIdx.x - 64) >> 7) : 0;          // epilogue warpgroup of this thread
  uint8_t* s_epi = s_win + kStages * WIN_BYTES + wg * (STG_BYTES + POOL_BYTES);
  float* s_stage = reinterpret_cast<float*>(s_epi);
  float* s_pool = reinterpret_cast<float*>(s_epi + STG_BYTES);                                 // [64][25]
  uint8_t* s_pmask = reinterpret_cast<uint8_t*>(s_pool + C2 * POOL_PITCH);                     // [64][24]
  float* s_bias = reinterpret_cast<float*>(s_pmask + C2 * 24);                                 // [64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_win + kStages * WIN_BYTES + kEpiGroups * (STG_BYTES + POOL_BYTES));
  uint64_t* full_bar = bars;                 // [kStages] window landed
  uint64_t* empty_bar = bars + kStages;      // [kStages] MMAs reading the window retired
  uint64_t* acc_full = bars + 2 * kStages;   // [2]
  uint64_t* acc_empty = bars + 2 * kStages + 2;  // [2]
  uint64_t* w_bar = bars + 2 * kStages + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const StepInfo& si = prm.si;
  const int g = blockIdx.x / prm.ctas_per_client;
  const int j = blockIdx.x % prm.ctas_per_client;
  const int ntiles = si.bsz[g] * TILES_PER_SAMPLE;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm_a1);
    tma_prefetch_desc(&tm_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
    mbar_init(w_bar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<2 * FWD_SUB * C2>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one() && ntiles > 0) {
      mbar_arrive_expect_tx(w_bar, W_BYTES);
      for (int i = 0; i < 3; ++i) tma_load_2d(s_w + i * 3 * W_TILE_BYTES, &tm_w, w_bar, 0, g * 9 * C2 + i * 3 * C2);
      uint32_t phase = 0;
      int s = 0;
      for (int t = j; t < ntiles; t += prm.ctas_per_client) {
        const int row = g * si.B + t / TILES_PER_SAMPLE;
        const int hb = t % TILES_PER_SAMPLE;
        mbar_wait(&empty_bar[s], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], WIN_BYTES);
        tma_load_2d(s_win + s * WIN_BYTES, &tm_a1, &full_bar[s], 0, row * PIX1 + hb * 4 * H1);
        if (++s == kStages) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (ntiles > 0) {
      constexpr uint32_t idesc = make_idesc(FMT_TF32, TILE_M, C2, false, false);
      mbar_wait(w_bar, 0);
      uint32_t phase = 0, acc_phase = 0;
      int s = 0, buf = 0;
      for (int t = j; t < ntiles; t += prm.ctas_per_client) {
        mbar_wait(&acc_empty[buf], acc_phase ^ 1);
        mbar_wait(&full_bar[s], phase);
        fence_after_thread_sync();
        if (elect_one()) {
          const uint64_t a_base = smem_desc_sw128(smem_u32(s_win + s * WIN_BYTES), 16, 1024);
          const uint64_t b_base = smem_desc_sw128(smem_u32(s_w), 16, 1024);
          const uint32_t acc = tmem_base + buf * (FWD_SUB * C2);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)      // tap (r, s) = the window shifted by r*26 + s rows
              mma_tf32(acc + ((tap * 4 + kk) % FWD_SUB) * C2, desc_advance(a_base, ((tap / 3) * H1 + (tap % 3)) * 128 + kk * 32),
                       desc_advance(b_base, tap * W_TILE_BYTES + kk * 32), idesc, (tap * 4 + kk) >= FWD_SUB);
          }
          mma_commit(&empty_bar[s]);
          mma_commit(&acc_full[buf]);
        }
        __syncwarp();
        if (++s == kStages) { s = 0; phase ^= 1; }
        if (++buf == 2) { buf = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue: bias + ReLU + 2x2 max-pool + dropout =====================
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;            // accumulator row = flat position inside the tile
    const int et = ((warp - 2) & 3) * 32 + lane;  // 0..127 inside this epilogue warpgroup
    const int bar_id = 1 + wg;
    const int pc = et & 63, half = et >> 6;       // pooled channel / pooled row within the tile
    if (et < C2) s_bias[et] = prm.W[(int64_t)g * P + OFF_C2B + et];
    epi_bar_sync(bar_id);
    const float keep_scale = TRAIN ? 1.f / (1.f - si.p1) : 1.f;
    uint32_t acc_phase = 0;
    const int buf = wg;                            // warpgroup w drains TMEM buffer w = every other tile of this CTA
    for (int t = j + wg * prm.ctas_per_client; t < ntiles; t += kEpiGroups * prm.ctas_per_client) {
      const int row = g * si.B + t / TILES_PER_SAMPLE;
      const int hb = t % TILES_PER_SAMPLE;
      mbar_wait(&acc_full[buf], acc_phase);
      fence_after_thread_sync();
#pragma unroll
      for (int c0 = 0; c0 < C2; c0 += 32) {
        float v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * (FWD_SUB * C2) + c0, v);
#pragma unroll
        for (int sub = 1; sub < FWD_SUB; ++sub) {
          float u[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * (FWD_SUB * C2) + sub * C2 + c0, u);
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] += u[i];
        }
        float4* dst = reinterpret_cast<float4*>(s_stage + m * STG_PITCH + c0);
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float4 o;
          const float4 bb = *reinterpret_cast<const float4*>(s_bias + c0 + i);
          o.x = fmaxf(v[i] + bb.x, 0.f);
          o.y = fmaxf(v[i + 1] + bb.y, 0.f);
          o.z = fmaxf(v[i + 2] + bb.z, 0.f);
          o.w = fmaxf(v[i + 3] + bb.w, 0.f);
          dst[i >> 2] = o;
        }
      }
      fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);   // TMEM buffer may be overwritten by tile t+2
      epi_bar_sync(bar_id);
      // pooled row ph = 2*hb + half ; window rows (2*half, 2*half+1) of the 4-row tile.  Results go to
      // a small staging tile first so that the global stores below are contiguous 96 B runs per channel.
      const int64_t tile_base = (int64_t)row * FLAT + hb * 2 * HP;      // + c*144 + [0, 24)
#pragma unroll 4
      for (int pw = 0; pw < HP; ++pw) {
        const int m00 = (2 * half) * H1 + 2 * pw;
        const float v00 = s_stage[m00 * STG_PITCH + pc], v01 = s_stage[(m00 + 1) * STG_PITCH + pc];
        const float v10 = s_stage[(m00 + H1) * STG_PITCH + pc], v11 = s_stage[(m00 + H1 + 1) * STG_PITCH + pc];
        float best = v00;
        int arg = 0;
        if (v01 > best) { best = v01; arg = 1; }
        if (v10 > best) { best = v10; arg = 2; }
        if (v11 > best) { best = v11; arg = 3; }
        bool keep = true;
        const int o = half * HP + pw;
        if (TRAIN) keep = ddl::rng_uniform(si.seed, (uint32_t)(tile_base + pc * (HP * HP) + o)) >= si.p1;
        s_pool[pc * POOL_PITCH + o] = keep ? best * keep_scale : 0.f;
        if (TRAIN) s_pmask[pc * 24 + o] = (uint8_t)(arg | ((keep && best > 0.f) ? 4 : 0));
      }
      epi_bar_sync(bar_id);
      for (int idx = et; idx < C2 * 6; idx += 128) {                    // 64 channels x 6 float4
        const int c = idx / 6, k4 = idx % 6;
        const float* src = s_pool + c * POOL_PITCH + 4 * k4;
        *reinterpret_cast<float4*>(prm.pooled + tile_base + c * (HP * HP) + 4 * k4) = make_float4(src[0], src[1], src[2], src[3]);
      }
      if (TRAIN) {
        for (int idx = et; idx < C2 * 3; idx += 128) {                  // 64 channels x 3 x 8 mask bytes
          const int c = idx / 3, k8 = idx % 3;
          *reinterpret_cast<uint2*>(prm.pmask + tile_base + c * (HP * HP) + 8 * k8) =
              *reinterpret_cast<const uint2*>(s_pmask + c * 24 + 8 * k8);
        }
      }
      acc_phase ^= 1;
    }
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<2 * FWD_SUB * C2>(tmem_base);
  }
}

int check_step(const StepInfo& si) {
  DDL_REQUIRE(si.G >= 1 && si.G <= MAX_G, "G=%d outside [1, %d]", si.G, MAX_G);
  DDL_REQUIRE(si.B >= 1, "B must be positive");
  for (int g = 0; g < si.G; ++g) DDL_REQUIRE(si.bsz[g] >= 0 && si.bsz[g] <= si.B, "bsz[%d]=%d outside [0, B=%d]", g, si.bsz[g], si.B);
  return DDL_OK;
}

}  // namespace

namespace mnist {
StepInfo make_step(int G, int B, const int* bsz, uint32_t seed, float p1, float p2, float lr) {
  StepInfo si{};
  si.G = G; si.B = B; si.seed = seed; si.p1 = p1; si.p2 = p2; si.lr = lr;
  for (int g = 0; g < MAX_G; ++g) si.bsz[g] = (g < G && bsz) ? bsz[g] : 0;
  return si;
}
}  // namespace mnist

extern "C" {

int ddl_mnist_conv1_fwd(const float* W, const float* X, int64_t x_client_stride, const int* perm, int64_t perm_stride,
                        int sample_off, float* a1, int G, int B, const int* bsz, void* stream) {
  StepInfo si = make_step(G, B, bsz, 0, 0.f, 0.f, 0.f);
  if (int e = check_step(si)) return e;
  DDL_CUDA_CHECK(ddl_launch_pdl(conv1_relu_kernel, dim3(G * B), dim3(256), 0, (cudaStream_t)stream, W, X, x_client_stride, perm, perm_stride, sample_off, a1, si));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_mnist_prep_conv2(const float* W, float* wt_fwd, float* wt_dgrad, int G, void* stream) {
  DDL_REQUIRE(G >= 1 && G <= MAX_G, "G=%d outside [1, %d]", G, MAX_G);
  DDL_CUDA_CHECK(ddl_launch_pdl(prep_conv2_kernel, dim3(G), dim3(256), 0, (cudaStream_t)stream, W, wt_fwd, wt_dgrad));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

int ddl_mnist_conv2_pool_fwd(const float* W, const float* wt_fwd, const float* a1, float* pooled, uint8_t* pmask,
                             int G, int B, const int* bsz, int train, float p_drop, uint32_t seed, void* stream) {
  StepInfo si = make_step(G, B, bsz, seed, train ? p_drop : 0.f, 0.f, 0.f);
  if (int e = check_step(si)) return e;
  DDL_REQUIRE(!train || pmask != nullptr, "training mode needs the pool/dropout mask buffer");
  CUtensorMap tm_a, tm_w;
  {
    uint64_t dims[2] = {(uint64_t)C1, (uint64_t)G * B * PIX1};
    uint64_t strides[1] = {128};
    uint32_t box[2] = {(uint32_t)C1, (uint32_t)WIN_ROWS};
    if (int e = ddl_encode_tmap(&tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, a1, dims, strides, box, 1)) return e;
  }
  {
    uint64_t dims[2] = {(uint64_t)C1, (uint64_t)G * 9 * C2};
    uint64_t strides[1] = {128};
    uint32_t box[2] = {(uint32_t)C1, (uint32_t)(3 * C2)};
    if (int e = ddl_encode_tmap(&tm_w, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, wt_fwd, dims, strides, box, 1)) return e;
  }
  Conv2Params prm;
  prm.W = W; prm.pooled = pooled; prm.pmask = pmask; prm.si = si;
  prm.ctas_per_client = kNumSMs / G;
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(conv2_pool_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CONV2_SMEM));
    DDL_CUDA_CHECK(cudaFuncSetAttribute(conv2_pool_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CONV2_SMEM));
    configured.done();
  }
  const int grid = prm.ctas_per_client * G;
  if (train) DDL_CUDA_CHECK(ddl_launch_pdl(conv2_pool_fwd_kernel<true>, dim3(grid), dim3(kConv2Threads), CONV2_SMEM, (cudaStream_t)stream, tm_a, tm_w, prm));
  else DDL_CUDA_CHECK(ddl_launch_pdl(conv2_pool_fwd_kernel<false>, dim3(grid), dim3(kConv2Threads), CONV2_SMEM, (cudaStream_t)stream, tm_a, tm_w, prm));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
