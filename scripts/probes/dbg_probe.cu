// Hardware probe (not on any hot path): does a K-major SWIZZLE_128B operand descriptor whose start
// address is offset by r rows (r*128 B, not a multiple of the 1024 B swizzle atom) read rows
// [r, r+128) of the TMA-written tile?  `mode` selects how descriptor.base_offset is set:
//   0: base_offset = 0      1: base_offset = r & 7      2: base_offset = (8 - r) & 7
// The answer decides whether the conv kernels can serve all 9 filter taps from ONE staged input
// tile (implicit-GEMM "shifted window") instead of re-loading the tile per tap.
#include "common.cuh"
#include "ddl25_kernels.h"
#include "tc05.cuh"
#include "tmap.h"

namespace {
using namespace tc05;

__global__ void __launch_bounds__(128, 1)
probe_shift_kernel(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tb, float* out,
                   int shift, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                 // 256 rows x 128 B
  uint8_t* sB = smem + 256 * 128;     // 64 rows x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + 64 * 128);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc<64>(tslot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tacc = *tslot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bars[0], 256 * 128 + 64 * 128);
    tma_load_2d(sA, &ta, &bars[0], 0, 0);
    tma_load_2d(sB, &tb, &bars[0], 0, 0);
    mbar_wait(&bars[0], 0);
    fence_after_thread_sync();
    constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, 64, false, false);
    const uint32_t a0 = smem_u32(sA) + (uint32_t)shift * 128u;
    const uint32_t bo = mode == 0 ? 0u : (mode == 1 ? (uint32_t)(shift & 7) : (uint32_t)((8 - shift) & 7));
    for (int kk = 0; kk < 4; ++kk) {
      const uint64_t da = smem_desc_sw128(a0 + kk * 32, 16, 1024, LAYOUT_SW128, bo);
      const uint64_t db = smem_desc_sw128(smem_u32(sB) + kk * 32, 16, 1024);
      mma_tf32(tacc, da, db, idesc, kk > 0);
    }
    mma_commit(&bars[1]);
  }
  __syncwarp();
  mbar_wait(&bars[1], 0);
  fence_after_thread_sync();
  const int row = warp * 32 + lane;
  for (int c0 = 0; c0 < 64; c0 += 32) {
    float v[32];
    tmem_ld_32x32(tacc + ((uint32_t)(warp * 32) << 16) + c0, v);
    for (int i = 0; i < 32; ++i) out[row * 64 + c0 + i] = v[i];
  }
  fence_before_thread_sync();
  __syncthreads();
  if (warp == 0) {
    fence_after_thread_sync();
    tmem_dealloc<64>(tacc);
  }
}

// Probe 2: MN-major fp32 operands (SWIZZLE_128B with 32B atoms).  A = a[R][32] staged once; the MMA
// reads it M-major with 4 "atoms" that OVERLAP: atom j starts one row (128 B) after atom j-1
// (LBO = 128), and the whole operand starts `shift` rows into the tile.  B = b[32][64] K rows x 64,
// staged as two 32-column halves.  Expected: D[j*32+c][n] = sum_k a[shift + k + j][c] * b[k][n].
__global__ void __launch_bounds__(128, 1)
probe_mn_kernel(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tb, float* out, int shift) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                  // 128 rows x 128 B
  uint8_t* sB = smem + 128 * 128;      // 2 halves x (32 rows x 128 B)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + 2 * 32 * 128);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bars + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc<64>(tslot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tacc = *tslot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bars[0], 128 * 128 + 2 * 32 * 128);
    tma_load_2d(sA, &ta, &bars[0], 0, 0);
    tma_load_2d(sB, &tb, &bars[0], 0, 0);
    tma_load_2d(sB + 32 * 128, &tb, &bars[0], 32, 0);
    mbar_wait(&bars[0], 0);
    fence_after_thread_sync();
    constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, 64, true, true);
    for (int kk = 0; kk < 4; ++kk) {
      const uint64_t da = smem_desc_sw128(smem_u32(sA) + (shift + kk * 8) * 128, 128, 512, LAYOUT_SW128_BASE32B);
      const uint64_t db = smem_desc_sw128(smem_u32(sB) + kk * 8 * 128, 32 * 128, 512, LAYOUT_SW128_BASE32B);
      mma_tf32(tacc, da, db, idesc, kk > 0);
    }
    mma_commit(&bars[1]);
  }
  __syncwarp();
  mbar_wait(&bars[1], 0);
  fence_after_thread_sync();
  const int row = warp * 32 + lane;
  for (int c0 = 0; c0 < 64; c0 += 32) {
    float v[32];
    tmem_ld_32x32(tacc + ((uint32_t)(warp * 32) << 16) + c0, v);
    for (int i = 0; i < 32; ++i) out[row * 64 + c0 + i] = v[i];
  }
  fence_before_thread_sync();
  __syncthreads();
  if (warp == 0) {
    fence_after_thread_sync();
    tmem_dealloc<64>(tacc);
  }
}
}  // namespace

extern "C" int ddl_dbg_probe_mn(const float* A /*128x32*/, const float* B /*32x64*/, float* out /*128x64*/, int shift,
                                void* stream) {
  CUtensorMap ta, tb;
  uint64_t da[2] = {32, 128}, db[2] = {64, 32};
  uint64_t sa[1] = {128}, sb[1] = {256};
  uint32_t ba[2] = {32, 128}, bb[2] = {32, 32};
  if (int e = ddl_encode_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, A, da, sa, ba, 2)) return e;
  if (int e = ddl_encode_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, B, db, sb, bb, 2)) return e;
  const int smem = 128 * 128 + 2 * 32 * 128 + 64 + 1024;
  probe_mn_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(ta, tb, out, shift);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

extern "C" int ddl_dbg_probe_shift(const float* A /*256x32*/, const float* B /*64x32*/, float* out /*128x64*/,
                                   int shift, int mode, void* stream) {
  CUtensorMap ta, tb;
  uint64_t da[2] = {32, 256}, db[2] = {32, 64};
  uint64_t st[1] = {128};
  uint32_t ba[2] = {32, 256}, bb[2] = {32, 64};
  if (int e = ddl_encode_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, A, da, st, ba, 1)) return e;
  if (int e = ddl_encode_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, B, db, st, bb, 1)) return e;
  const int smem = 256 * 128 + 64 * 128 + 64 + 1024;
  DDL_CUDA_CHECK(cudaFuncSetAttribute(probe_shift_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  probe_shift_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(ta, tb, out, shift, mode);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}
