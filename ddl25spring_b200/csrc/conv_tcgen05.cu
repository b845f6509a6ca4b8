// K9 (general): 2-D convolution forward / input-gradient / weight-gradient as "tap GEMMs" on tcgen05 tensor cores.
//
// Activations are NHWC fp32 ([N, H, W, C], C contiguous), weights are OHWI fp32 ([Cout, T, Cin], T = kh*kw taps).
// A convolution is a sum over taps t of GEMMs whose A operand is the activation tensor read at a spatial offset:
//
//   fprop : Y[n,i,j,:]  = sum_t  X_{map(t)}[n, i+dh(t), j+dw(t), :] * W[:, wt(t), :]^T        (K-major A, K-major B)
//   dgrad : dX[n,h,w,:] = sum_t dY        [n, h+dh(t), w+dw(t), :] * W[:, wt(t), :]           (K-major A, MN-major B)
//   wgrad : dW[:,t,:]   = sum_{n,i,j} dY[n,i,j,:]^T  X_{map(t)}[n, i+dh(t), j+dw(t), :]       (MN-major A and B, K = pixels)
//
// * no im2col buffer and no padded copy: an operand tile is ONE TMA box over the 4-D (5-D for MN-major) NHWC tensor
//   map at coordinates (c0, j0+dw, i0+dh, n0); rows/columns that fall outside the image are zero-filled by the TMA
//   unit, which is exactly the convolution's zero padding;
// * a 128-row M tile is a (bw x bh x bn) box of output pixels, picked on the host so that it tiles the image well
//   (8x8x2 for an 8x8 feature map, 1x1x128 for 1x1, 32x4x1 for 32x32, ...);
// * stride-2 convolutions read the four even/odd "phase" sub-images of the input (conv_phase_split), each with its
//   own tensor map: tap (r, s) -> phase ((r+1)&1, (s+1)&1) at offset (r ? 0 : -1, s ? 0 : -1); their input gradient
//   is four tap-GEMMs (one per output phase) whose epilogue scatters with stride 2; a ConvTranspose2d (DCGAN) is the
//   same four-phase dgrad used as a forward op;
// * taps that lie outside the image for every pixel (a 3x3 kernel on a 1x1 feature map keeps only its centre) are
//   dropped on the host;
// * split-K (over taps x channel blocks, or over pixels for wgrad) writes partial tiles that are summed in a fixed
//   order (deterministic).
//
// Replaces (reference): nn.Conv2d at lab/tutorial_1a/hfl_complete.py:51,53 via ATen/cuDNN, and the ResNet-18 / DCGAN
// convolutions of the north-star configs.
#include "common.cuh"
#include "ddl25_kernels.h"
#include "tc05.cuh"
#include "tmap.h"

namespace {

using namespace tc05;

constexpr int BM = 128;
constexpr int ROW_BYTES = 128;
constexpr int BK = 32;                 // fp32 elements per 128-byte swizzle row
constexpr int UMMA_K = 8;
constexpr int KSTEPS = BK / UMMA_K;
constexpr int kEpiWarps = 4;
constexpr int kThreads = 32 * (2 + kEpiWarps);
constexpr int kMaxTaps = 16;
constexpr int kMaxMaps = 4;

__device__ __forceinline__ void tma_load_5d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3,
                                            int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

struct ConvParams {
  CUtensorMap amap[kMaxMaps];   // activation operand(s): fprop/dgrad 4-D {C, W, H, N}; wgrad 5-D {32, W, H, N, C/32}
  CUtensorMap bmap;             // fprop: weights 3-D {Cin, T, Cout}; dgrad: weights 4-D {32, Cout, Cin/32, T}; wgrad: dY 5-D
  int n_taps;
  int cblocks;                  // fprop/dgrad: channel blocks (of 32) of the reduced dimension per tap
  signed char tap_map[kMaxTaps], tap_dh[kMaxTaps], tap_dw[kMaxTaps], tap_w[kMaxTaps];
  int bw_log2, bh_log2, bn_log2;     // pixel box of one tile / one k-block
  int tiles_w, tiles_h, tiles_n;     // number of pixel boxes along w / h / n
  int Nn, Hh, Ww;                    // extent of the pixel space the boxes tile (output pixels; dY pixels for wgrad)
  // output: element offset of pixel (n, i, j) = o_off + n*o_sn + i*o_sh + j*o_sw; channels contiguous
  int64_t o_off, o_sn, o_sh, o_sw;
  int Nout;                          // output channels (fprop: Cout, dgrad: Cin, wgrad: Cin)
  int Mout;                          // wgrad only: Cout
  float* out;
  const float* bias;
  float alpha, beta;
  int act;
  float slope;
  // fprop/dgrad with several output phases in one launch (stride-2 dgrad): phase f owns taps [ph_begin[f], ph_begin[f+1]) and
  // writes at o_off + ph_off[f]; blockIdx.z = phase (n_split == 1 then).  n_phases == 1: all taps, offset 0.
  int n_phases;
  int ph_begin[5];
  int64_t ph_off[4];
  int n_split;
  int kb_per_split;
  float* scratch;                    // n_split > 1: [split][rows][Nout] partial tiles (rows = tiles*128, or taps*Mout for wgrad)
  int64_t scratch_rows;
};

template <int BN, bool DEEP>
struct ConvPlan {
  static constexpr int A_BYTES = BM * ROW_BYTES;
  static constexpr int B_BYTES = BN * ROW_BYTES;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // DEEP (grids of less than two waves): a CTA's k-loop is bound by the L2 -> shared-memory latency of its TMA loads, not by
  // bandwidth or the tensor core (measured ~350 ns per k-block with 3-4 stages), so the ring is as deep as one CTA per SM
  // allows.  Otherwise <= 98 KB so that two CTAs share an SM and one's epilogue hides under the other's main loop.
  static constexpr int STAGES = DEEP ? (BN >= 128 ? 6 : 8) : (BN >= 128 ? 3 : 4);
  static constexpr int TOTAL = STAGES * STAGE_BYTES + 1024 /*barriers*/ + 1024 /*alignment slack*/;
};

__device__ __forceinline__ float conv_act(float v, int act, float slope) {
  if (act == 1) return fmaxf(v, 0.f);
  if (act == 2) return v > 0.f ? v : v * slope;
  if (act == 3) return tanhf(v);
  return v;
}

// Constant-index dispatch: a dynamically indexed descriptor address would make the compiler copy the parameter block to
// local memory, which the TMA unit cannot read.
__device__ __forceinline__ const CUtensorMap* pick_map(const ConvParams& p, int i) {
  switch (i) {
    case 0: return &p.amap[0];
    case 1: return &p.amap[1];
    case 2: return &p.amap[2];
    default: return &p.amap[3];
  }
}

// Shared by the three kernels: TMEM accumulator (128 lanes x BN columns) -> per-warp transposed staging tile -> global rows.
// `row_off(m_local)` returns the element offset of tile row m (or -1 if the row is not stored).
template <int BN, typename RowOff>
__device__ __forceinline__ void conv_epilogue(uint32_t tmem_acc, float* wsm, int warp, int lane, bool have_acc, float* out,
                                              int n0, int Nout, const float* bias, float alpha, float beta, int act, float slope,
                                              bool raw, RowOff row_off) {
  constexpr int WP = 36;
  const int quarter = warp & 3;
  const int lr = lane >> 3, lc = (lane & 7) * 4;
  int64_t offs[8];
#pragma unroll
  for (int it = 0; it < 8; ++it) offs[it] = row_off(quarter * 32 + it * 4 + lr);
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 32) {
    if (n0 + c0 >= Nout) break;
    float v[32];
    if (have_acc) {
      tmem_ld_32x32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < 32; i += 4)
      *reinterpret_cast<float4*>(wsm + lane * WP + i) = make_float4(v[i] * alpha, v[i + 1] * alpha, v[i + 2] * alpha, v[i + 3] * alpha);
    __syncwarp();
    const int n = n0 + c0 + lc;
    const bool nok = n < Nout;
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!raw && bias && nok) b4 = *reinterpret_cast<const float4*>(bias + n);
    float4 old[8];
    if (!raw && beta != 0.f) {
#pragma unroll
      for (int it = 0; it < 8; ++it)
        old[it] = (nok && offs[it] >= 0) ? *reinterpret_cast<const float4*>(out + offs[it] + n) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = it * 4 + lr;
      float4 x = *reinterpret_cast<const float4*>(wsm + r * WP + lc);
      if (!raw) {
        if (beta != 0.f) {
          x.x = fmaf(beta, old[it].x, x.x); x.y = fmaf(beta, old[it].y, x.y);
          x.z = fmaf(beta, old[it].z, x.z); x.w = fmaf(beta, old[it].w, x.w);
        }
        x.x = conv_act(x.x + b4.x, act, slope); x.y = conv_act(x.y + b4.y, act, slope);
        x.z = conv_act(x.z + b4.z, act, slope); x.w = conv_act(x.w + b4.w, act, slope);
      }
      if (nok && offs[it] >= 0) *reinterpret_cast<float4*>(out + offs[it] + n) = x;
    }
    __syncwarp();
  }
}

struct Bars {
  uint64_t* full;
  uint64_t* empty;
  uint64_t* accum;
  uint32_t* tmem_slot;
};

template <int STAGES>
__device__ __forceinline__ Bars conv_setup(uint8_t* bar_base, int warp, int lane) {
  Bars b;
  uint64_t* bars = reinterpret_cast<uint64_t*>(bar_base);
  b.full = bars;
  b.empty = bars + STAGES;
  b.accum = bars + 2 * STAGES;
  b.tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&b.full[s], 1);
      mbar_init(&b.empty[s], 1);
    }
    mbar_init(b.accum, 1);
    mbar_fence_init();
  }
  return b;
}

// The MMA-issuing warp's loop, shared by the three kernels (the operand majors are compile-time).
template <int BN, int STAGES, int STAGE_BYTES, int A_BYTES, bool A_MN, bool B_MN>
__device__ __forceinline__ void conv_mma_loop(const Bars& b, uint8_t* stage_base, uint32_t tmem_acc, int num_kb) {
  constexpr uint32_t idesc = make_idesc(FMT_TF32, BM, BN, A_MN, B_MN);
  constexpr uint32_t a_lay = A_MN ? LAYOUT_SW128_BASE32B : LAYOUT_SW128, b_lay = B_MN ? LAYOUT_SW128_BASE32B : LAYOUT_SW128;
  constexpr uint32_t a_sbo = A_MN ? 512 : 1024, b_sbo = B_MN ? 512 : 1024;
  constexpr uint32_t a_lbo = A_MN ? BK * ROW_BYTES : 16, b_lbo = B_MN ? BK * ROW_BYTES : 16;
  constexpr uint32_t a_step = A_MN ? UMMA_K * ROW_BYTES : 32, b_step = B_MN ? UMMA_K * ROW_BYTES : 32;
  uint32_t phase = 0;
  int s = 0;
  for (int kb = 0; kb < num_kb; ++kb) {
    mbar_wait(&b.full[s], phase);
    fence_after_thread_sync();
    if (elect_one()) {
      const uint32_t a_addr = smem_u32(stage_base + s * STAGE_BYTES);
      const uint64_t da0 = smem_desc_sw128(a_addr, a_lbo, a_sbo, a_lay);
      const uint64_t db0 = smem_desc_sw128(a_addr + A_BYTES, b_lbo, b_sbo, b_lay);
#pragma unroll
      for (int kk = 0; kk < KSTEPS; ++kk)
        mma_tf32(tmem_acc, desc_advance(da0, kk * a_step), desc_advance(db0, kk * b_step), idesc, (kb > 0) || (kk > 0));
      mma_commit(&b.empty[s]);
      if (kb == num_kb - 1) mma_commit(b.accum);
    }
    __syncwarp();
    if (++s == STAGES) { s = 0; phase ^= 1; }
  }
}

// ---------------------------------------------------------------------------------------- fprop / dgrad
template <int BN, bool B_MN, bool DEEP>
__global__ void __launch_bounds__(kThreads, DEEP ? 1 : 2) conv_tap_gemm_kernel(const __grid_constant__ ConvParams p) {
  using Plan = ConvPlan<BN, DEEP>;
  constexpr int STAGES = Plan::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* stage_base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const Bars b = conv_setup<STAGES>(stage_base + STAGES * Plan::STAGE_BYTES, warp, lane);

  const int tw = blockIdx.x % p.tiles_w, th = (blockIdx.x / p.tiles_w) % p.tiles_h, tn = blockIdx.x / (p.tiles_w * p.tiles_h);
  const int j0 = tw << p.bw_log2, i0 = th << p.bh_log2, n0 = tn << p.bn_log2;
  const int col0 = blockIdx.y * BN;
  const int phase = p.n_phases > 1 ? blockIdx.z : 0;
  const int split = p.n_phases > 1 ? 0 : blockIdx.z;
  const int tap0 = p.ph_begin[phase];
  const int total_kb = (p.ph_begin[phase + 1] - tap0) * p.cblocks;
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(total_kb, kb_begin + p.kb_per_split);
  const int num_kb = max(0, kb_end - kb_begin);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.amap[0]);
    tma_prefetch_desc(&p.bmap);
  }
  if (warp == 2) tmem_alloc<BN>(b.tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_acc = *b.tmem_slot;
  pdl_launch_dependents();     // the next kernel may set up while this one computes ...
  pdl_wait();                  // ... and this one touches global memory only after its predecessor has finished

  if (warp == 0) {
    if (elect_one()) {
      int s = 0;
      uint32_t par = 0;
      int tap = tap0 + kb_begin / p.cblocks, cb = kb_begin % p.cblocks;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&b.empty[s], par ^ 1);
        uint8_t* a_dst = stage_base + s * Plan::STAGE_BYTES;
        uint8_t* b_dst = a_dst + Plan::A_BYTES;
        mbar_arrive_expect_tx(&b.full[s], Plan::STAGE_BYTES);
        tma_load_4d(a_dst, pick_map(p, p.tap_map[tap]), &b.full[s], cb * BK, j0 + p.tap_dw[tap], i0 + p.tap_dh[tap], n0);
        if constexpr (B_MN) tma_load_4d(b_dst, &p.bmap, &b.full[s], 0, cb * BK, col0 / BK, p.tap_w[tap]);
        else tma_load_3d(b_dst, &p.bmap, &b.full[s], cb * BK, p.tap_w[tap], col0);
        if (++cb == p.cblocks) { cb = 0; ++tap; }
        if (++s == STAGES) { s = 0; par ^= 1; }
      }
    }
  } else if (warp == 1) {
    conv_mma_loop<BN, STAGES, Plan::STAGE_BYTES, Plan::A_BYTES, false, B_MN>(b, stage_base, tmem_acc, num_kb);
  } else {
    if (num_kb > 0) {
      mbar_wait(b.accum, 0);
      fence_after_thread_sync();
    }
    float* wsm = reinterpret_cast<float*>(stage_base) + (warp - 2) * (32 * 36);
    const int bwl = p.bw_log2, bhl = p.bh_log2;
    if (p.n_split > 1) {
      float* dst = p.scratch + (int64_t)split * p.scratch_rows * p.Nout;
      const int64_t row0 = (int64_t)blockIdx.x * BM;
      const int Nout = p.Nout;
      conv_epilogue<BN>(tmem_acc, wsm, warp, lane, num_kb > 0, dst, col0, Nout, nullptr, 1.f, 0.f, 0, 0.f, true,
                        [=](int m) -> int64_t { return (row0 + m) * Nout; });
    } else {
      const int Nn = p.Nn, Hh = p.Hh, Ww = p.Ww;
      const int64_t o_off = p.o_off + p.ph_off[phase], o_sn = p.o_sn, o_sh = p.o_sh, o_sw = p.o_sw;
      conv_epilogue<BN>(tmem_acc, wsm, warp, lane, num_kb > 0, p.out, col0, p.Nout, p.bias, p.alpha, p.beta, p.act, p.slope, false,
                        [=](int m) -> int64_t {
                          const int n = n0 + (m >> (bwl + bhl)), i = i0 + ((m >> bwl) & ((1 << bhl) - 1)), j = j0 + (m & ((1 << bwl) - 1));
                          return (n < Nn && i < Hh && j < Ww) ? o_off + n * o_sn + i * o_sh + j * o_sw : -1;
                        });
    }
    fence_before_thread_sync();
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<BN>(tmem_acc);
  }
}

// Sum the split-K partial tiles of conv_tap_gemm_kernel in split order and apply the epilogue (pixel mapping as above).
__global__ void __launch_bounds__(256) conv_splitk_reduce_kernel(const __grid_constant__ ConvParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int nv = p.Nout >> 2;
  const int64_t total = p.scratch_rows * nv;
  const int bwl = p.bw_log2, bhl = p.bh_log2, bnl = p.bn_log2;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = e / nv;
    const int c = (int)(e % nv) * 4;
    const int tile = (int)(row >> 7), m = (int)(row & 127);
    const int tw = tile % p.tiles_w, th = (tile / p.tiles_w) % p.tiles_h, tn = tile / (p.tiles_w * p.tiles_h);
    const int n = (tn << bnl) + (m >> (bwl + bhl)), i = (th << bhl) + ((m >> bwl) & ((1 << bhl) - 1)), j = (tw << bwl) + (m & ((1 << bwl) - 1));
    if (n >= p.Nn || i >= p.Hh || j >= p.Ww) continue;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < p.n_split; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(p.scratch + ((int64_t)s * p.scratch_rows + row) * p.Nout + c);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float* dst = p.out + p.o_off + n * p.o_sn + i * p.o_sh + j * p.o_sw + c;
    float4 o = make_float4(acc.x * p.alpha, acc.y * p.alpha, acc.z * p.alpha, acc.w * p.alpha);
    if (p.beta != 0.f) {
      const float4 old = *reinterpret_cast<const float4*>(dst);
      o.x = fmaf(p.beta, old.x, o.x); o.y = fmaf(p.beta, old.y, o.y); o.z = fmaf(p.beta, old.z, o.z); o.w = fmaf(p.beta, old.w, o.w);
    }
    if (p.bias) {
      const float4 b4 = *reinterpret_cast<const float4*>(p.bias + c);
      o.x += b4.x; o.y += b4.y; o.z += b4.z; o.w += b4.w;
    }
    o.x = conv_act(o.x, p.act, p.slope); o.y = conv_act(o.y, p.act, p.slope);
    o.z = conv_act(o.z, p.act, p.slope); o.w = conv_act(o.w, p.act, p.slope);
    *reinterpret_cast<float4*>(dst) = o;
  }
}

// ---------------------------------------------------------------------------------------- wgrad
// M = Cout (A = dY, M-major), N = Cin (B = X at the tap's offset, N-major), K = pixels in boxes of 32.
// grid = (ceil(Cout/128), ceil(Cin/BN), n_taps * n_split)
template <int BN, bool DEEP>
__global__ void __launch_bounds__(kThreads, DEEP ? 1 : 2) conv_wgrad_kernel(const __grid_constant__ ConvParams p) {
  using Plan = ConvPlan<BN, DEEP>;
  constexpr int STAGES = Plan::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* stage_base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const Bars b = conv_setup<STAGES>(stage_base + STAGES * Plan::STAGE_BYTES, warp, lane);

  const int m0 = blockIdx.x * BM, col0 = blockIdx.y * BN;
  const int tap = blockIdx.z / p.n_split, split = blockIdx.z % p.n_split;
  const int total_kb = p.tiles_w * p.tiles_h * p.tiles_n;
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(total_kb, kb_begin + p.kb_per_split);
  const int num_kb = max(0, kb_end - kb_begin);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.amap[0]);
    tma_prefetch_desc(&p.bmap);
  }
  if (warp == 2) tmem_alloc<BN>(b.tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_acc = *b.tmem_slot;
  pdl_launch_dependents();     // the next kernel may set up while this one computes ...
  pdl_wait();                  // ... and this one touches global memory only after its predecessor has finished

  if (warp == 0) {
    if (elect_one()) {
      uint32_t phase = 0;
      int s = 0;
      const CUtensorMap* xmap = pick_map(p, p.tap_map[tap]);
      const int dh = p.tap_dh[tap], dw = p.tap_dw[tap];
      int tw = kb_begin % p.tiles_w, th = (kb_begin / p.tiles_w) % p.tiles_h, tn = kb_begin / (p.tiles_w * p.tiles_h);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&b.empty[s], phase ^ 1);
        uint8_t* a_dst = stage_base + s * Plan::STAGE_BYTES;
        uint8_t* b_dst = a_dst + Plan::A_BYTES;
        const int j0 = tw << p.bw_log2, i0 = th << p.bh_log2, n0 = tn << p.bn_log2;
        mbar_arrive_expect_tx(&b.full[s], Plan::STAGE_BYTES);
        tma_load_5d(a_dst, &p.bmap, &b.full[s], 0, j0, i0, n0, m0 / BK);
        tma_load_5d(b_dst, xmap, &b.full[s], 0, j0 + dw, i0 + dh, n0, col0 / BK);
        if (++tw == p.tiles_w) { tw = 0; if (++th == p.tiles_h) { th = 0; ++tn; } }
        if (++s == STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    conv_mma_loop<BN, STAGES, Plan::STAGE_BYTES, Plan::A_BYTES, true, true>(b, stage_base, tmem_acc, num_kb);
  } else {
    if (num_kb > 0) {
      mbar_wait(b.accum, 0);
      fence_after_thread_sync();
    }
    float* wsm = reinterpret_cast<float*>(stage_base) + (warp - 2) * (32 * 36);
    const int Mout = p.Mout, Nout = p.Nout;
    if (p.n_split > 1) {
      float* dst = p.scratch + ((int64_t)split * p.scratch_rows + (int64_t)tap * Mout) * Nout;
      conv_epilogue<BN>(tmem_acc, wsm, warp, lane, num_kb > 0, dst, col0, Nout, nullptr, 1.f, 0.f, 0, 0.f, true,
                        [=](int m) -> int64_t { return m0 + m < Mout ? (int64_t)(m0 + m) * Nout : -1; });
    } else {
      const int64_t o_sn = p.o_sn, base = p.o_off + (int64_t)p.tap_w[tap] * p.o_sw;
      conv_epilogue<BN>(tmem_acc, wsm, warp, lane, num_kb > 0, p.out, col0, Nout, nullptr, p.alpha, p.beta, 0, 0.f, false,
                        [=](int m) -> int64_t { return m0 + m < Mout ? base + (int64_t)(m0 + m) * o_sn : -1; });
    }
    fence_before_thread_sync();
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<BN>(tmem_acc);
  }
}

__global__ void __launch_bounds__(256) conv_wgrad_reduce_kernel(const __grid_constant__ ConvParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int nv = p.Nout >> 2;
  const int64_t total = p.scratch_rows * nv;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = e / nv;
    const int c = (int)(e % nv) * 4;
    const int tap = (int)(row / p.Mout), co = (int)(row % p.Mout);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < p.n_split; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(p.scratch + ((int64_t)s * p.scratch_rows + row) * p.Nout + c);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float* dst = p.out + p.o_off + (int64_t)co * p.o_sn + (int64_t)p.tap_w[tap] * p.o_sw + c;
    float4 o = make_float4(acc.x * p.alpha, acc.y * p.alpha, acc.z * p.alpha, acc.w * p.alpha);
    if (p.beta != 0.f) {
      const float4 old = *reinterpret_cast<const float4*>(dst);
      o.x = fmaf(p.beta, old.x, o.x); o.y = fmaf(p.beta, old.y, o.y); o.z = fmaf(p.beta, old.z, o.z); o.w = fmaf(p.beta, old.w, o.w);
    }
    *reinterpret_cast<float4*>(dst) = o;
  }
}

// ------------------------------------------------------------------------------------------ host
int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

// Pick a power-of-two pixel box (bw, bh, bn) with bw*bh*bn == total that wastes the fewest rows on this image extent.
void choose_box(int total, int Nn, int Hh, int Ww, int* bwl, int* bhl, int* bnl) {
  const int L = ilog2(total);
  double best = -1.0;
  for (int a = 0; a <= L; ++a)
    for (int c = 0; a + c <= L; ++c) {
      const int d = L - a - c;
      const int bw = 1 << a, bh = 1 << c, bn = 1 << d;
      const double eff = (double)Ww / (((Ww + bw - 1) / bw) * bw) * (double)Hh / (((Hh + bh - 1) / bh) * bh) * (double)Nn /
                         (((Nn + bn - 1) / bn) * bn);
      // ties: prefer wide boxes (longer contiguous runs in HBM)
      const double score = eff + 1e-6 * a + 1e-7 * c;
      if (score > best) { best = score; *bwl = a; *bhl = c; *bnl = d; }
    }
}

int make_act_map4(CUtensorMap* m, const float* base, int N, int H, int W, int C, int bw, int bh, int bn) {
  uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  uint64_t strides[3] = {(uint64_t)C * 4, (uint64_t)W * C * 4, (uint64_t)H * W * C * 4};
  uint32_t box[4] = {BK, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn};
  return ddl_encode_tmap(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, dims, strides, box, 1);
}

// MN-major view for wgrad: {32 channels of an atom, W, H, N, C/32 atoms}, 128B swizzle with 32B atoms
int make_act_map5(CUtensorMap* m, const float* base, int N, int H, int W, int C, int bw, int bh, int bn, int atoms_box) {
  uint64_t dims[5] = {(uint64_t)(C < BK ? C : BK), (uint64_t)W, (uint64_t)H, (uint64_t)N, (uint64_t)((C + BK - 1) / BK)};
  uint64_t strides[4] = {(uint64_t)C * 4, (uint64_t)W * C * 4, (uint64_t)H * W * C * 4, (uint64_t)ROW_BYTES};
  uint32_t box[5] = {BK, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn, (uint32_t)atoms_box};
  return ddl_encode_tmap(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 5, base, dims, strides, box, 2);
}

bool chan_ok(int C) { return C >= 4 && C % 4 == 0 && (C < BK || C % BK == 0); }

// Copy the taps that can touch the image (phase by phase); returns their number.
int fill_taps(ConvParams& p, const int* tap_map, const int* tap_dh, const int* tap_dw, const int* tap_w, int n_taps, int Hsrc, int Wsrc,
              int Hh, int Ww, int n_phases = 1, const int* phase_begin = nullptr) {
  int k = 0;
  for (int f = 0; f < n_phases; ++f) {
    const int t0 = phase_begin ? phase_begin[f] : 0, t1 = phase_begin ? phase_begin[f + 1] : n_taps;
    p.ph_begin[f] = k;
    for (int t = t0; t < t1; ++t) {
      if (tap_dh[t] >= Hsrc || tap_dh[t] <= -Hh || tap_dw[t] >= Wsrc || tap_dw[t] <= -Ww) continue;
      p.tap_map[k] = (signed char)tap_map[t]; p.tap_dh[k] = (signed char)tap_dh[t];
      p.tap_dw[k] = (signed char)tap_dw[t]; p.tap_w[k] = (signed char)tap_w[t];
      ++k;
    }
  }
  for (int f = n_phases; f <= 4; ++f) p.ph_begin[f] = k;
  p.n_phases = n_phases;
  p.n_taps = k;
  return k;
}

template <typename K>
int set_smem(K kern, int bytes) {
  DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  return DDL_OK;
}

}  // namespace

extern "C" {

// fprop (w_mn_major = 0: reduce over the weight's Cin, produce Cout channels) or dgrad (w_mn_major = 1: reduce over Cout,
// produce Cin channels) of a convolution given as a tap list.  `acts` are n_maps NHWC tensors [N, Ha, Wa, Ca] (the input,
// its four stride-2 phases, or dY).  Output pixel (n, i, j) for n < N, i < Hh, j < Ww is written at
// out + o_off + n*o_sn + i*o_sh + j*o_sw (+ channel).
int ddl_conv_tap_gemm(const float* const* acts, int n_maps, int N, int Ha, int Wa, int Ca, const float* Wt, int Cout, int T, int Cin,
                      int w_mn_major, const int* tap_map, const int* tap_dh, const int* tap_dw, const int* tap_w, int n_taps,
                      int n_phases, const int* phase_begin, const int64_t* phase_off, float* out, int Hh, int Ww, int64_t o_off, int64_t o_sn, int64_t o_sh, int64_t o_sw, const float* bias,
                      float alpha, float beta, int act, float slope, float* scratch, int64_t scratch_floats, void* stream) {
  DDL_REQUIRE(n_maps >= 1 && n_maps <= kMaxMaps && n_taps >= 1 && n_taps <= kMaxTaps, "n_maps=%d / n_taps=%d out of range", n_maps, n_taps);
  DDL_REQUIRE(chan_ok(Cin) && chan_ok(Cout), "channel counts (%d, %d) must be multiples of 4, and of 32 when above 32", Cin, Cout);
  DDL_REQUIRE(Ca == (w_mn_major ? Cout : Cin), "activation channels %d do not match the weight's reduced dimension", Ca);
  if (n_phases < 1 || !phase_begin) n_phases = 1;
  DDL_REQUIRE(n_phases <= 4 && (n_phases == 1 || phase_off), "n_phases=%d outside [1,4] or missing phase offsets", n_phases);
  const int Nout = w_mn_major ? Cin : Cout;
  DDL_REQUIRE(o_off % 4 == 0 && o_sn % 4 == 0 && o_sh % 4 == 0 && o_sw % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "output rows must be 16-byte aligned");
  ConvParams p;
  memset(&p, 0, sizeof(p));
  choose_box(BM, N, Hh, Ww, &p.bw_log2, &p.bh_log2, &p.bn_log2);
  const int bw = 1 << p.bw_log2, bh = 1 << p.bh_log2, bn = 1 << p.bn_log2;
  p.tiles_w = (Ww + bw - 1) / bw; p.tiles_h = (Hh + bh - 1) / bh; p.tiles_n = (N + bn - 1) / bn;
  p.Nn = N; p.Hh = Hh; p.Ww = Ww;
  for (int i = 0; i < n_maps; ++i) {
    DDL_REQUIRE((reinterpret_cast<uintptr_t>(acts[i]) & 15) == 0, "activation map %d is not 16-byte aligned", i);
    if (int e = make_act_map4(&p.amap[i], acts[i], N, Ha, Wa, Ca, bw, bh, bn)) return e;
  }
  for (int t = 0; t < n_taps; ++t)
    DDL_REQUIRE(tap_map[t] >= 0 && tap_map[t] < n_maps && tap_w[t] >= 0 && tap_w[t] < T, "tap %d refers to map %d / weight tap %d", t,
                tap_map[t], tap_w[t]);
  fill_taps(p, tap_map, tap_dh, tap_dw, tap_w, n_taps, Ha, Wa, Hh, Ww, n_phases, n_phases > 1 ? phase_begin : nullptr);
  DDL_REQUIRE(p.n_taps > 0, "no tap touches the image");
  for (int f = 0; f < 4; ++f) {
    p.ph_off[f] = (n_phases > 1 && f < n_phases) ? phase_off[f] : 0;
    DDL_REQUIRE(p.ph_off[f] % 4 == 0, "phase offsets must keep rows 16-byte aligned");
  }
  p.cblocks = (Ca + BK - 1) / BK;
  const int tiles = p.tiles_w * p.tiles_h * p.tiles_n;
  int BN = Nout <= 32 ? 32 : (Nout <= 64 ? 64 : 128);
  if (BN == 128 && tiles * n_phases * ((Nout + 127) / 128) < kNumSMs / 2) BN = 64;       // small problems: more, narrower tiles
  if (w_mn_major) {
    uint64_t dims[4] = {(uint64_t)(Cin < BK ? Cin : BK), (uint64_t)Cout, (uint64_t)((Cin + BK - 1) / BK), (uint64_t)T};
    uint64_t strides[3] = {(uint64_t)T * Cin * 4, (uint64_t)ROW_BYTES, (uint64_t)Cin * 4};
    uint32_t box[4] = {BK, BK, (uint32_t)(BN / BK), 1};
    if (int e = ddl_encode_tmap(&p.bmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, Wt, dims, strides, box, 2)) return e;
  } else {
    uint64_t dims[3] = {(uint64_t)Cin, (uint64_t)T, (uint64_t)Cout};
    uint64_t strides[2] = {(uint64_t)Cin * 4, (uint64_t)T * Cin * 4};
    uint32_t box[3] = {BK, 1, (uint32_t)BN};
    if (int e = ddl_encode_tmap(&p.bmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, Wt, dims, strides, box, 1)) return e;
  }
  p.o_off = o_off; p.o_sn = o_sn; p.o_sh = o_sh; p.o_sw = o_sw;
  p.Nout = Nout; p.Mout = 0; p.out = out; p.bias = bias; p.alpha = alpha; p.beta = beta; p.act = act; p.slope = slope;
  // split-K when the grid cannot fill the machine
  const int total_kb = p.n_taps * p.cblocks;
  const int ctas = tiles * ((Nout + BN - 1) / BN);
  int n_split = 1;
  if (n_phases == 1 && 2 * ctas <= kNumSMs && total_kb >= 8 && scratch) {      // less than half a wave: split the reduction across more SMs
    n_split = kNumSMs / ctas;
    if (n_split > total_kb / 4) n_split = total_kb / 4;
    const int64_t per_split = (int64_t)tiles * BM * Nout;
    if ((int64_t)n_split * per_split > scratch_floats) n_split = (int)(scratch_floats / per_split);
    if (n_split < 1) n_split = 1;
  }
  p.kb_per_split = (total_kb + n_split - 1) / n_split;
  n_split = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.n_split = n_split; p.scratch = scratch; p.scratch_rows = (int64_t)tiles * BM;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(tiles, (Nout + BN - 1) / BN, n_phases > 1 ? n_phases : n_split);
  const bool deep = (int64_t)grid.x * grid.y * grid.z < 2 * kNumSMs;
#define DDL_CONV_CASE3(BN_, MN_, DEEP_)                                                           \
  {                                                                                               \
    auto kern = conv_tap_gemm_kernel<BN_, MN_, DEEP_>;                                            \
    static DdlOncePerDevice configured;                                                               \
    if (configured.need()) { if (int e = set_smem(kern, ConvPlan<BN_, DEEP_>::TOTAL)) return e; configured.done(); } \
    DDL_CUDA_CHECK(ddl_launch_pdl(kern, grid, dim3(kThreads), ConvPlan<BN_, DEEP_>::TOTAL, st, p)); \
  }
#define DDL_CONV_CASE(BN_, MN_) { if (deep) DDL_CONV_CASE3(BN_, MN_, true) else DDL_CONV_CASE3(BN_, MN_, false) }
  if (BN == 32) { if (w_mn_major) DDL_CONV_CASE(32, true) else DDL_CONV_CASE(32, false) }
  else if (BN == 64) { if (w_mn_major) DDL_CONV_CASE(64, true) else DDL_CONV_CASE(64, false) }
  else { if (w_mn_major) DDL_CONV_CASE(128, true) else DDL_CONV_CASE(128, false) }
#undef DDL_CONV_CASE
#undef DDL_CONV_CASE3
  DDL_LAUNCH_CHECK();
  if (n_split > 1) {
    int64_t blocks = (p.scratch_rows * (Nout / 4) + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    DDL_CUDA_CHECK(ddl_launch_pdl(conv_splitk_reduce_kernel, dim3((unsigned)blocks), dim3(256), 0, st, p));
    DDL_LAUNCH_CHECK();
  }
  return DDL_OK;
}

// dW[co, tap_w, ci] = alpha * sum_{n,i,j} dY[n,i,j,co] * X_{map}[n, i+dh, j+dw, ci] (+ beta * dW).  dY: [N, Ho, Wo, Cout];
// every X map: [N, Hx, Wx, Cin]; dW: [Cout, T, Cin].  Taps that touch no pixel get a zero gradient (beta == 0).
int ddl_conv_wgrad(const float* dY, int N, int Ho, int Wo, int Cout, const float* const* xs, int n_maps, int Hx, int Wx, int Cin,
                   const int* tap_map, const int* tap_dh, const int* tap_dw, const int* tap_w, int n_taps, float* dW, int T,
                   float alpha, float beta, float* scratch, int64_t scratch_floats, void* stream) {
  DDL_REQUIRE(n_maps >= 1 && n_maps <= kMaxMaps && n_taps >= 1 && n_taps <= kMaxTaps, "n_maps=%d / n_taps=%d out of range", n_maps, n_taps);
  DDL_REQUIRE(chan_ok(Cin) && chan_ok(Cout), "channel counts (%d, %d) must be multiples of 4, and of 32 when above 32", Cin, Cout);
  ConvParams p;
  memset(&p, 0, sizeof(p));
  choose_box(BK, N, Ho, Wo, &p.bw_log2, &p.bh_log2, &p.bn_log2);
  const int bw = 1 << p.bw_log2, bh = 1 << p.bh_log2, bn = 1 << p.bn_log2;
  p.tiles_w = (Wo + bw - 1) / bw; p.tiles_h = (Ho + bh - 1) / bh; p.tiles_n = (N + bn - 1) / bn;
  p.Nn = N; p.Hh = Ho; p.Ww = Wo;
  int BN = Cin <= 32 ? 32 : (Cin <= 64 ? 64 : 128);
  for (int i = 0; i < n_maps; ++i) {
    DDL_REQUIRE((reinterpret_cast<uintptr_t>(xs[i]) & 15) == 0, "activation map %d is not 16-byte aligned", i);
    if (int e = make_act_map5(&p.amap[i], xs[i], N, Hx, Wx, Cin, bw, bh, bn, BN / BK)) return e;
  }
  if (int e = make_act_map5(&p.bmap, dY, N, Ho, Wo, Cout, bw, bh, bn, BM / BK)) return e;
  for (int t = 0; t < n_taps; ++t)
    DDL_REQUIRE(tap_map[t] >= 0 && tap_map[t] < n_maps && tap_w[t] >= 0 && tap_w[t] < T, "tap %d refers to map %d / weight tap %d", t,
                tap_map[t], tap_w[t]);
  fill_taps(p, tap_map, tap_dh, tap_dw, tap_w, n_taps, Hx, Wx, Ho, Wo);
  cudaStream_t st = (cudaStream_t)stream;
  if (p.n_taps < T && beta == 0.f) DDL_CUDA_CHECK(cudaMemsetAsync(dW, 0, sizeof(float) * (size_t)Cout * T * Cin, st));
  if (p.n_taps == 0) return DDL_OK;
  p.o_off = 0; p.o_sn = (int64_t)T * Cin; p.o_sw = Cin; p.o_sh = 0;
  p.Nout = Cin; p.Mout = Cout; p.out = dW; p.bias = nullptr; p.alpha = alpha; p.beta = beta; p.act = 0; p.slope = 0.f;
  const int total_kb = p.tiles_w * p.tiles_h * p.tiles_n;
  const int ctas = ((Cout + BM - 1) / BM) * ((Cin + BN - 1) / BN) * p.n_taps;
  int n_split = 1;
  if (2 * ctas <= kNumSMs && total_kb >= 8 && scratch) {
    n_split = (kNumSMs + ctas - 1) / ctas;
    if (n_split > total_kb / 4) n_split = total_kb / 4;
    const int64_t per_split = (int64_t)p.n_taps * Cout * Cin;
    if ((int64_t)n_split * per_split > scratch_floats) n_split = (int)(scratch_floats / per_split);
    if (n_split < 1) n_split = 1;
  }
  p.kb_per_split = (total_kb + n_split - 1) / n_split;
  n_split = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.n_split = n_split; p.scratch = scratch; p.scratch_rows = (int64_t)p.n_taps * Cout;
  dim3 grid((Cout + BM - 1) / BM, (Cin + BN - 1) / BN, p.n_taps * n_split);
  const bool deep = (int64_t)grid.x * grid.y * grid.z < 2 * kNumSMs;
#define DDL_WGRAD_CASE3(BN_, DEEP_)                                                               \
  {                                                                                               \
    auto kern = conv_wgrad_kernel<BN_, DEEP_>;                                                    \
    static DdlOncePerDevice configured;                                                               \
    if (configured.need()) { if (int e = set_smem(kern, ConvPlan<BN_, DEEP_>::TOTAL)) return e; configured.done(); } \
    DDL_CUDA_CHECK(ddl_launch_pdl(kern, grid, dim3(kThreads), ConvPlan<BN_, DEEP_>::TOTAL, st, p)); \
  }
#define DDL_WGRAD_CASE(BN_) { if (deep) DDL_WGRAD_CASE3(BN_, true) else DDL_WGRAD_CASE3(BN_, false) }
  if (BN == 32) DDL_WGRAD_CASE(32) else if (BN == 64) DDL_WGRAD_CASE(64) else DDL_WGRAD_CASE(128)
#undef DDL_WGRAD_CASE
#undef DDL_WGRAD_CASE3
  DDL_LAUNCH_CHECK();
  if (n_split > 1) {
    int64_t blocks = (p.scratch_rows * (Cin / 4) + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    DDL_CUDA_CHECK(ddl_launch_pdl(conv_wgrad_reduce_kernel, dim3((unsigned)blocks), dim3(256), 0, st, p));
    DDL_LAUNCH_CHECK();
  }
  return DDL_OK;
}

}  // extern "C"
