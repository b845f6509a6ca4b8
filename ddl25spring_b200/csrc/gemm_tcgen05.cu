// K10 / K3: grouped GEMM on the 5th-generation tensor cores.
//
//   C[g] (M x N, fp32) = epilogue( A[g] (M x K) * B[g]^T (N x K) )          g = 0 .. G-1
//
// * operands are fp32 in HBM and are consumed as TF32 by `tcgen05.mma.kind::tf32` (bf16 operands
//   use kind::f16); each operand can be K-major ([rows, K], K contiguous) or MN-major
//   ([K, rows], rows contiguous), so forward (X W^T), input-gradient (dY W) and weight-gradient
//   (dY^T X) GEMMs of a Linear layer all map onto this kernel without a transpose pass;
// * operand tiles are staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) into a multi-stage
//   shared-memory ring, accumulators live in TMEM, one elected thread issues the MMAs, four
//   epilogue warps read the accumulator back with tcgen05.ld;
// * PRECISE mode ("3xTF32"): the epilogue warps, idle during the main loop, split every landed
//   fp32 tile in shared memory into hi = tf32(x) and lo = x - hi; the MMA thread then issues
//   hi*hi + hi*lo + lo*hi, which recovers fp32-grade products (the reference's nn.Linear runs in
//   true fp32, lab/tutorial_1a/hfl_complete.py:58-61) while staying on the tensor cores;
// * split-K writes per-split partial tiles to a scratch buffer; `splitk_reduce` sums them in a
//   fixed order (deterministic: Krum's selection must not depend on atomics ordering) and applies
//   bias + activation.  The Gram matrix G = U U^T of K3 is this kernel with A = B = U.
//
// Replaces (reference): every nn.Linear and, via ops/gemm.py: gram(), the
// pairwise-distance computation of Krum [north star].
#include <cstdlib>
#include "common.cuh"
#include "ddl25_kernels.h"
#include "tc05.cuh"
#include "tmap.h"

namespace {

using namespace tc05;

constexpr int BM = 128;
constexpr int ROW_BYTES = 128;          // one swizzle row: 32 tf32 or 64 bf16 along the contiguous dim
constexpr int kEpiWarps = 4;
constexpr int kThreads = 32 * (2 + kEpiWarps);           // gram_packed_kernel
// Splitter warps of the general GEMM body (3xTF32 hi / lo split): warps 2 .. 2 + kSplitWarps; the first kEpiWarps of them
// then drain the accumulator.  Measured on B200 (768 x 864 x 288, scripts/dev/gemm_kloop_cost.py): 8 splitter warps are no
// faster than 4 - the 3xTF32 k-loop is bound by shared-memory bandwidth (12 MMAs x 8 KB of operands + 32 KB of TMA writes
// + the splitter's traffic per k-block, at 128 B/clk), which is why not rewriting the hi tile (Epilogue::keep_hi) pays.
constexpr int kSplitWarps = 4;
constexpr int kGemmThreads = 32 * (2 + kSplitWarps);

// 3xTF32: 1 = use the staged fp32 tile as the hi operand without rewriting it (DDL_GEMM_KEEP_HI overrides)
inline int keep_hi_default() {
  static const int v = getenv("DDL_GEMM_KEEP_HI") ? atoi(getenv("DDL_GEMM_KEEP_HI")) : 1;
  return v;
}

struct Epilogue {
  float* C;             // output (or partials scratch when n_split > 1)
  int64_t ldc;
  int64_t c_group_stride;
  // split-K with in-kernel fix-up: every split CTA stores its raw partial tile, bumps the tile's counter, and the CTA that
  // arrives LAST sums the partials in split order (deterministic), applies bias + activation and writes the final tile -
  // no separate reduce launch.  counters == nullptr: the caller runs splitk_reduce_kernel afterwards.
  int* counters;        // one int per output tile (zero before the launch; the last CTA resets it)
  float* C_final;       // final output of the fix-up (C then points at the partials)
  int64_t ldc_final;
  int64_t c_group_stride_final;
  const float* bias;    // per-N bias or null
  int64_t bias_group_stride;
  int act;              // 0 none, 1 relu, 2 leaky relu
  float slope;
  float alpha;
  float beta;           // C = act(alpha*acc + beta*C_old + bias): in the epilogue when n_split == 1, else in splitk_reduce_kernel
  int n_split;          // > 1: write raw partial tiles to C = scratch[split][g][M][N] (ldc = N)
  int keep_hi;          // SPLIT3: the staged fp32 tile serves as the hi operand as it is (the tensor core truncates to TF32)
  int same_ab;          // A and B are the same matrix (Gram / SYRK): stage each tile once, use it as both operands
  const float* center;  // SPLIT3 only, may be null: K-vector subtracted from every row of the staged tiles before the hi/lo
                        // split (Krum scores update DELTAS u_k - global; re-centring in shared memory means the K x D delta
                        // matrix is never materialised in HBM)
  int64_t center_len;   // valid entries of `center` (guards the ragged last k-block)
};

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
  if (act == 1) return fmaxf(v, 0.f);
  if (act == 2) return v > 0.f ? v : v * slope;
  return v;
}

// K7: the A operand is the CONCATENATION along K of up to 8 matrices that live in different buffers —
// the party embeddings of vertical FL, each in its owner GPU's (peer-mapped) memory.  The K loop walks
// the parts; every k-block is TMA-loaded straight from the owning GPU, so the all-gather of embeddings
// (torch.cat at lab/tutorial_2b/vfl.py:36) happens inside the top model's first GEMM.
constexpr int kMaxParts = 8;
struct GatherMaps {
  CUtensorMap maps[kMaxParts];   // A part p: [M, dims[p]] fp32, K-major
  int kb_start[kMaxParts + 1];   // first k-block of part p (prefix sums of ceil(dims[p] / 32))
  int col_off[kMaxParts];        // column of W where part p's slab starts
  int n_parts;
  // row mode (K3 fused all-gather + Gram): the parts are ROW blocks of one tall operand U = [part 0 rows; part 1 rows; ...]
  // living on different GPUs; every k-block stages `rows_per_part` rows from each GPU into consecutive shared-memory rows
  // and the tile is used as BOTH operands (SYRK), so U U^T is accumulated without U ever being gathered in HBM.
  int row_mode;
  int rows_per_part;
};

// SHORT: GEMMs whose K loop is only a few k-blocks long but whose grid is several waves (the per-client fc1 input- and
// weight-gradient GEMMs of the packed engine: K = 100..128, 576 tiles) spend their time in prologue and epilogue, so the
// ring is cut to two stages (<= 98 KB) and two CTAs share an SM: one's epilogue runs under the other's loads.
template <int BN, int EB /*element bytes*/, bool SPLIT3, bool SHORT = false>
struct SmemPlan {
  static constexpr int A_BYTES = BM * ROW_BYTES;
  static constexpr int B_BYTES = BN * ROW_BYTES;
  static constexpr int STAGE_BYTES = (A_BYTES + B_BYTES) * (SPLIT3 ? 2 : 1);
  static constexpr int BUDGET = SHORT ? 98 * 1024 : 200 * 1024;
  static constexpr int RAW_STAGES = BUDGET / STAGE_BYTES;
  static constexpr int STAGES = RAW_STAGES > 8 ? 8 : RAW_STAGES;
  static_assert(STAGES >= 1, "tile does not fit the shared-memory budget");
  static constexpr int BAR_BYTES = 1024;
  static constexpr int TOTAL = STAGES * STAGE_BYTES + BAR_BYTES + 1024 /*alignment slack*/;
};

template <int BN, int EB, bool A_MN, bool B_MN, bool SPLIT3, bool GATHER, bool SHORT = false>
__device__ __forceinline__ void gemm_tcgen05_body(const CUtensorMap& tma_a, const CUtensorMap& tma_b, const GatherMaps* gm,
                                                  const Epilogue& epi, const int M, const int N, const int K,
                                                  const int kblocks_per_split) {
  using Plan = SmemPlan<BN, EB, SPLIT3, SHORT>;
  constexpr int STAGES = Plan::STAGES;
  constexpr int BK = ROW_BYTES / EB;         // elements per k-block (32 tf32 / 64 bf16)
  constexpr int UMMA_K = 32 / EB;            // elements per MMA instruction along K (8 tf32 / 16 bf16)
  constexpr int KSTEPS = BK / UMMA_K;        // 4
  static_assert(!SPLIT3 || EB == 4, "3xTF32 split applies to fp32 operands only");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Plan::STAGE_BYTES);
  uint64_t* full_bar = bars;                  // TMA bytes landed
  uint64_t* empty_bar = bars + STAGES;        // MMAs that read the stage have completed
  uint64_t* split_bar = bars + 2 * STAGES;    // hi/lo split written (SPLIT3 only)
  uint64_t* accum_bar = bars + 3 * STAGES;    // accumulator complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int n_split = epi.n_split > 1 ? epi.n_split : 1;
  const int g = blockIdx.z / n_split;
  const int split = blockIdx.z % n_split;
  const int total_kblocks = GATHER ? (gm->row_mode ? (K + BK - 1) / BK : gm->kb_start[gm->n_parts]) : (K + BK - 1) / BK;
  const int kb_begin = split * kblocks_per_split;
  int kb_end = kb_begin + kblocks_per_split;
  if (kb_end > total_kblocks) kb_end = total_kblocks;
  const int num_kb = kb_end > kb_begin ? kb_end - kb_begin : 0;

  if (warp == 0 && lane == 0) {
    if constexpr (!GATHER) tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&split_bar[s], kSplitWarps);
    }
    mbar_init(accum_bar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<BN>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_acc = *tmem_slot;
  pdl_launch_dependents();     // programmatic dependent launch (common.cuh): the successor may set up now ...
  pdl_wait();                  // ... and this kernel reads / writes global memory only after its predecessor completed

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t phase = 0;
      int s = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[s], phase ^ 1);
        uint8_t* a_dst = stage_base + s * Plan::STAGE_BYTES;
        uint8_t* b_dst = a_dst + Plan::A_BYTES;
        if (GATHER && gm->row_mode) {
          const int gkb = kb_begin + kb;
          const uint32_t part_bytes = (uint32_t)gm->rows_per_part * ROW_BYTES;
          mbar_arrive_expect_tx(&full_bar[s], part_bytes * gm->n_parts);
#pragma unroll
          for (int p = 0; p < kMaxParts; ++p)
            if (p < gm->n_parts) tma_load_3d(a_dst + p * part_bytes, &gm->maps[p], &full_bar[s], gkb * BK, 0, 0);
          if (++s == STAGES) { s = 0; phase ^= 1; }
          continue;
        }
        mbar_arrive_expect_tx(&full_bar[s], epi.same_ab ? Plan::A_BYTES : Plan::A_BYTES + Plan::B_BYTES);
        if constexpr (GATHER) {
          const int gkb = kb_begin + kb;
          int p = 0;
          while (p + 1 < gm->n_parts && gkb >= gm->kb_start[p + 1]) ++p;
          const int kl = (gkb - gm->kb_start[p]) * BK;           // columns past dims[p] are zero-filled by TMA
          // constant indices only: a dynamically indexed descriptor address would make the compiler stage the
          // parameter block through local memory, where the TMA unit cannot read it
          switch (p) {
            case 0: tma_load_3d(a_dst, &gm->maps[0], &full_bar[s], kl, m0, 0); break;
            case 1: tma_load_3d(a_dst, &gm->maps[1], &full_bar[s], kl, m0, 0); break;
            case 2: tma_load_3d(a_dst, &gm->maps[2], &full_bar[s], kl, m0, 0); break;
            case 3: tma_load_3d(a_dst, &gm->maps[3], &full_bar[s], kl, m0, 0); break;
            case 4: tma_load_3d(a_dst, &gm->maps[4], &full_bar[s], kl, m0, 0); break;
            case 5: tma_load_3d(a_dst, &gm->maps[5], &full_bar[s], kl, m0, 0); break;
            case 6: tma_load_3d(a_dst, &gm->maps[6], &full_bar[s], kl, m0, 0); break;
            default: tma_load_3d(a_dst, &gm->maps[7], &full_bar[s], kl, m0, 0); break;
          }
          tma_load_3d(b_dst, &tma_b, &full_bar[s], gm->col_off[p] + kl, n0, g);
          if (++s == STAGES) { s = 0; phase ^= 1; }
          continue;
        }
        const int k0 = (kb_begin + kb) * BK;
        if constexpr (A_MN) tma_load_4d(a_dst, &tma_a, &full_bar[s], 0, k0, m0 / BK, g);
        else tma_load_3d(a_dst, &tma_a, &full_bar[s], k0, m0, g);
        if (!epi.same_ab) {
          if constexpr (B_MN) tma_load_4d(b_dst, &tma_b, &full_bar[s], 0, k0, n0 / BK, g);
          else tma_load_3d(b_dst, &tma_b, &full_bar[s], k0, n0, g);
        }
        if (++s == STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc(EB == 4 ? FMT_TF32 : FMT_BF16, BM, BN, A_MN, B_MN);
    uint32_t phase = 0;
    int s = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      mbar_wait(SPLIT3 ? &split_bar[s] : &full_bar[s], phase);
      fence_after_thread_sync();
      if (elect_one()) {
        // K-major : a k-step advances 32 B inside the 128 B swizzle row; 8-row groups are 1024 B apart.
        // MN-major: a k-step is UMMA_K k-rows of 128 B; MN atoms are BK*128 B apart; 32-bit MN-major
        //           operands must use the 128B swizzle with 32B atoms (pattern period 4 k-rows, SBO 512).
        constexpr uint32_t a_lay = (A_MN && EB == 4) ? LAYOUT_SW128_BASE32B : LAYOUT_SW128;
        constexpr uint32_t b_lay = (B_MN && EB == 4) ? LAYOUT_SW128_BASE32B : LAYOUT_SW128;
        constexpr uint32_t a_sbo = (A_MN && EB == 4) ? 512 : 1024, b_sbo = (B_MN && EB == 4) ? 512 : 1024;
        constexpr uint32_t a_lbo = A_MN ? BK * ROW_BYTES : 16, b_lbo = B_MN ? BK * ROW_BYTES : 16;
        constexpr uint32_t a_step = A_MN ? UMMA_K * ROW_BYTES : 32, b_step = B_MN ? UMMA_K * ROW_BYTES : 32;
        const uint32_t a_hi = smem_u32(stage_base + s * Plan::STAGE_BYTES);
        const uint64_t da_hi0 = smem_desc_sw128(a_hi, a_lbo, a_sbo, a_lay);
        const uint64_t db_hi0 = smem_desc_sw128(epi.same_ab ? a_hi : a_hi + Plan::A_BYTES, b_lbo, b_sbo, b_lay);
        constexpr uint32_t LO_OFF = Plan::A_BYTES + Plan::B_BYTES;       // lo tiles follow the hi tiles
#pragma unroll
        for (int kk = 0; kk < KSTEPS; ++kk) {
          const bool acc = (kb > 0) || (kk > 0);
          const uint64_t da_hi = desc_advance(da_hi0, kk * a_step), db_hi = desc_advance(db_hi0, kk * b_step);
          if constexpr (SPLIT3) {
            const uint64_t da_lo = desc_advance(da_hi, LO_OFF), db_lo = desc_advance(db_hi, LO_OFF);
            mma_tf32(tmem_acc, da_lo, db_hi, idesc, acc);
            mma_tf32(tmem_acc, da_hi, db_lo, idesc, true);
            mma_tf32(tmem_acc, da_hi, db_hi, idesc, true);
          } else if constexpr (EB == 4) {
            mma_tf32(tmem_acc, da_hi, db_hi, idesc, acc);
          } else {
            mma_f16(tmem_acc, da_hi, db_hi, idesc, acc);
          }
        }
        mma_commit(&empty_bar[s]);                       // stage reusable once these MMAs retire
        if (kb == num_kb - 1) mma_commit(accum_bar);     // accumulator complete
      }
      __syncwarp();
      if (++s == STAGES) { s = 0; phase ^= 1; }
    }
  } else {
    // ===================== splitter (PRECISE) + epilogue warps =====================
    const int et = (warp - 2) * 32 + lane;   // index among the splitter threads; 0..127 are the epilogue threads
    if constexpr (SPLIT3) {
      uint32_t phase = 0;
      int s = 0;
      constexpr int NV = (Plan::A_BYTES + Plan::B_BYTES) / 16;   // float4 count of A|B (contiguous)
      const int nv = epi.same_ab ? Plan::A_BYTES / 16 : NV;      // SYRK: only the A tile was staged
      // Re-centring: float4 i of a K-major 128B-swizzled tile is row i/8, physical 16-byte chunk i%8, i.e. logical
      // columns 4*((i%8) ^ (row%8)) .. +3.  With 32 * kSplitWarps splitter threads striding by as many, a thread's chunk and row%8 never
      // change, so it needs ONE float4 of the centre vector per k-block, fetched before the wait on the TMA bytes.
      const int cchunk = ((et & 7) ^ ((et >> 3) & 7)) * 4;
      for (int kb = 0; kb < num_kb; ++kb) {
        float4 cv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (epi.center != nullptr) {
          const int64_t c0 = (int64_t)(kb_begin + kb) * BK + cchunk;
          if (c0 + 3 < epi.center_len) cv = __ldg(reinterpret_cast<const float4*>(epi.center + c0));
          else if (c0 < epi.center_len) {
            cv.x = epi.center[c0];
            if (c0 + 1 < epi.center_len) cv.y = epi.center[c0 + 1];
            if (c0 + 2 < epi.center_len) cv.z = epi.center[c0 + 2];
          }
        }
        mbar_wait(&full_bar[s], phase);
        float4* hi = reinterpret_cast<float4*>(stage_base + s * Plan::STAGE_BYTES);
        float4* lo = hi + NV;
        // The tensor core reads the top 19 bits of an fp32 operand, i.e. the staged tile already IS the hi part; it is
        // rewritten only when the centre was subtracted.  (epi.keep_hi: leave the staged values in place.)
        const bool rewrite_hi = (epi.center != nullptr) || !epi.keep_hi;
#pragma unroll 4
        for (int i = et; i < nv; i += 32 * kSplitWarps) {
          float4 x = hi[i], h, l;
          x.x -= cv.x; x.y -= cv.y; x.z -= cv.z; x.w -= cv.w;
          h.x = __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u); l.x = x.x - h.x;
          h.y = __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u); l.y = x.y - h.y;
          h.z = __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u); l.z = x.z - h.z;
          h.w = __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u); l.w = x.w - h.w;
          if (rewrite_hi) hi[i] = h;
          lo[i] = l;
        }
        fence_proxy_async_smem();            // make the generic-proxy stores visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(&split_bar[s]);
        if (++s == STAGES) { s = 0; phase ^= 1; }
      }
    }
    // ---- epilogue: TMEM -> registers -> global (warps 2..5; the split-only warps 6..9 are done)
    if (warp < 2 + kEpiWarps) {
    const int quarter = warp & 3;            // TMEM lane quarter this warp may read
    if (num_kb > 0) {
      mbar_wait(accum_bar, 0);
      fence_after_thread_sync();
    }
    float* Cg;
    int64_t ldc;
    if (n_split > 1) {
      Cg = epi.C + ((int64_t)split * gridDim.z / n_split + g) * (int64_t)M * N;
      ldc = N;
    } else {
      Cg = epi.C + (int64_t)g * epi.c_group_stride;
      ldc = epi.ldc;
    }
    const float* bias = (epi.bias && n_split == 1) ? epi.bias + (int64_t)g * epi.bias_group_stride : nullptr;
    const bool direct = (n_split == 1);
    // Transpose each 32x32 accumulator chunk through a per-warp shared-memory tile (the operand ring is
    // idle by now) so that every global access of the warp is one contiguous 128-byte row segment:
    // lane = column.  The C reads of the beta path are issued as a batch before they are consumed.
    constexpr int WP = 36;                                   // tile pitch in floats: 16-byte rows, conflict-free for float4
    float* wsm = reinterpret_cast<float*>(stage_base) + (warp - 2) * (32 * WP);
    const int mrow0 = m0 + quarter * 32;
    const bool vec_ok = (ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(Cg) & 15) == 0) && (N % 4 == 0);
    const int lr = lane >> 3, lc = (lane & 7) * 4;           // this lane's row-in-group / first column in the read-back phase
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      if (n0 + c0 >= N) break;
      float v[32];
      if (num_kb > 0) {
        tmem_ld_32x32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<float4*>(wsm + lane * WP + i) =
            make_float4(v[i] * epi.alpha, v[i + 1] * epi.alpha, v[i + 2] * epi.alpha, v[i + 3] * epi.alpha);
      __syncwarp();
      const int n = n0 + c0 + lc;
      if (vec_ok) {
        // 8 steps; each warp instruction touches 4 rows x 128 contiguous bytes
        const bool nok = n < N;
        float4 bn4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bias && nok) bn4 = *reinterpret_cast<const float4*>(bias + n);
        float4 old[8];
        if (direct && epi.beta != 0.f) {
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int m = mrow0 + it * 4 + lr;
            old[it] = (nok && m < M) ? *reinterpret_cast<const float4*>(Cg + (int64_t)m * ldc + n) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + lr, m = mrow0 + r;
          float4 x = *reinterpret_cast<const float4*>(wsm + r * WP + lc);
          if (direct) {
            if (epi.beta != 0.f) {
              x.x = fmaf(epi.beta, old[it].x, x.x); x.y = fmaf(epi.beta, old[it].y, x.y);
              x.z = fmaf(epi.beta, old[it].z, x.z); x.w = fmaf(epi.beta, old[it].w, x.w);
            }
            x.x = apply_act(x.x + bn4.x, epi.act, epi.slope); x.y = apply_act(x.y + bn4.y, epi.act, epi.slope);
            x.z = apply_act(x.z + bn4.z, epi.act, epi.slope); x.w = apply_act(x.w + bn4.w, epi.act, epi.slope);
          }
          if (nok && m < M) *reinterpret_cast<float4*>(Cg + (int64_t)m * ldc + n) = x;
        }
      } else {
        // unaligned output: lane = column, one row per step
        const int nn = n0 + c0 + lane;
        const bool ncol_ok = nn < N;
        const float bn = (bias && ncol_ok) ? bias[nn] : 0.f;
        float* cbase = Cg + (int64_t)mrow0 * ldc + nn;
#pragma unroll 4
        for (int rr = 0; rr < 32; ++rr) {
          float x = wsm[rr * WP + lane];
          if (direct) {
            if (epi.beta != 0.f && ncol_ok && mrow0 + rr < M) x = fmaf(epi.beta, cbase[(int64_t)rr * ldc], x);
            x = apply_act(x + bn, epi.act, epi.slope);
          }
          if (ncol_ok && mrow0 + rr < M) cbase[(int64_t)rr * ldc] = x;
        }
      }
      __syncwarp();
    }
    if (n_split > 1 && epi.counters != nullptr) {
      // ---- cooperative split-K fix-up.  The host only enables it when the whole grid is one wave (<= 148 CTAs, one per SM),
      // so the n_split CTAs of a tile are co-resident: each stores its raw partial tile, they meet at the tile's counter, and
      // then EVERY one of them reduces its 1/n_split share of the tile's rows over all partials (in split order:
      // deterministic), applies bias + activation and writes the final rows.  (A last-arriving-CTA reduction was measured
      // 3x slower than the separate reduce kernel: one CTA cannot keep enough loads in flight for n_split x 64 KB.)
      __threadfence();                                           // this CTA's partial tile is visible device-wide ...
      asm volatile("bar.sync 1, 128;" ::: "memory");             // ... for all four epilogue warps, before the arrival is counted
      const int G = gridDim.z / n_split;
      int* cnt = epi.counters + (g * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
      if (et == 0) {
        atomicAdd(cnt, 1);
        const long long t0 = clock64();
        while (*reinterpret_cast<volatile int*>(cnt) < n_split) {
          if (clock64() - t0 > 4000000000ll) __trap();           // ~2 s: a sibling CTA never arrived
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      __threadfence();
      const int rows = (M - m0) < BM ? (M - m0) : BM;
      const int cols = (N - n0) < BN ? (N - n0) : BN;
      const int per = (rows + n_split - 1) / n_split;
      const int r0 = split * per;
      const int nr = (rows - r0) < per ? (rows - r0) : per;      // may be <= 0 for the last splits of a short tile
      const int64_t sstride = (int64_t)G * M * N;
      const float* pbase = epi.C + (int64_t)g * M * N + (int64_t)(m0 + r0) * N + n0;
      float* fin = epi.C_final + (int64_t)g * epi.c_group_stride_final + (int64_t)(m0 + r0) * epi.ldc_final + n0;
      const float* fbias = epi.bias ? epi.bias + (int64_t)g * epi.bias_group_stride + n0 : nullptr;
      const bool v4 = (N % 4 == 0) && (epi.ldc_final % 4 == 0) && (epi.c_group_stride_final % 4 == 0) &&
                      ((reinterpret_cast<uintptr_t>(epi.C) & 15) == 0) && ((reinterpret_cast<uintptr_t>(epi.C_final) & 15) == 0);
      if (v4) {
        const int c4 = cols / 4;                                 // N % 4 == 0 and BN % 4 == 0 -> cols % 4 == 0
        for (int i = et; i < nr * c4; i += 32 * kEpiWarps) {
          const int r = i / c4, c = (i - r * c4) * 4;
          const float* p0 = pbase + (int64_t)r * N + c;
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          int sp = 0;
          for (; sp + 4 <= n_split; sp += 4) {                   // four independent loads in flight, added in split order
            const float4 x0 = __ldcg(reinterpret_cast<const float4*>(p0 + (sp + 0) * sstride));
            const float4 x1 = __ldcg(reinterpret_cast<const float4*>(p0 + (sp + 1) * sstride));
            const float4 x2 = __ldcg(reinterpret_cast<const float4*>(p0 + (sp + 2) * sstride));
            const float4 x3 = __ldcg(reinterpret_cast<const float4*>(p0 + (sp + 3) * sstride));
            acc.x += x0.x; acc.y += x0.y; acc.z += x0.z; acc.w += x0.w;
            acc.x += x1.x; acc.y += x1.y; acc.z += x1.z; acc.w += x1.w;
            acc.x += x2.x; acc.y += x2.y; acc.z += x2.z; acc.w += x2.w;
            acc.x += x3.x; acc.y += x3.y; acc.z += x3.z; acc.w += x3.w;
          }
          for (; sp < n_split; ++sp) {
            const float4 x = __ldcg(reinterpret_cast<const float4*>(p0 + sp * sstride));
            acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
          }
          if (fbias) { acc.x += fbias[c]; acc.y += fbias[c + 1]; acc.z += fbias[c + 2]; acc.w += fbias[c + 3]; }
          acc.x = apply_act(acc.x, epi.act, epi.slope); acc.y = apply_act(acc.y, epi.act, epi.slope);
          acc.z = apply_act(acc.z, epi.act, epi.slope); acc.w = apply_act(acc.w, epi.act, epi.slope);
          *reinterpret_cast<float4*>(fin + (int64_t)r * epi.ldc_final + c) = acc;
        }
      } else {
        for (int i = et; i < nr * cols; i += 32 * kEpiWarps) {
          const int r = i / cols, c = i - r * cols;
          float acc = 0.f;
          for (int sp = 0; sp < n_split; ++sp) acc += __ldcg(pbase + sp * sstride + (int64_t)r * N + c);
          if (fbias) acc += fbias[c];
          fin[(int64_t)r * epi.ldc_final + c] = apply_act(acc, epi.act, epi.slope);
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (et == 0 && atomicAdd(cnt, 1) == 2 * n_split - 1) *cnt = 0;   // last one out re-arms the counter for the next launch
    }
    fence_before_thread_sync();
    }
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<BN>(tmem_acc);
  }
}

template <int BN, int EB, bool A_MN, bool B_MN, bool SPLIT3>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const Epilogue epi, const int M, const int N, const int K, const int kblocks_per_split) {
  gemm_tcgen05_body<BN, EB, A_MN, B_MN, SPLIT3, false>(tma_a, tma_b, nullptr, epi, M, N, K, kblocks_per_split);
}

template <int BN, bool A_MN, bool B_MN, bool SPLIT3>
__global__ void __launch_bounds__(kGemmThreads, 2)
gemm_tcgen05_short_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                          const Epilogue epi, const int M, const int N, const int K, const int kblocks_per_split) {
  gemm_tcgen05_body<BN, 4, A_MN, B_MN, SPLIT3, false, true>(tma_a, tma_b, nullptr, epi, M, N, K, kblocks_per_split);
}

template <int BN, bool SPLIT3>
__global__ void __launch_bounds__(kGemmThreads, 1)
gather_gemm_tcgen05_kernel(const __grid_constant__ GatherMaps gm, const __grid_constant__ CUtensorMap tma_b,
                           const Epilogue epi, const int M, const int N, const int K, const int kblocks_per_split) {
  gemm_tcgen05_body<BN, 4, false, false, SPLIT3, true>(tma_b, tma_b, &gm, epi, M, N, K, kblocks_per_split);
}

// Sum split-K partial tiles in split order, add bias, apply activation.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ part, int n_split,
                                                            int64_t group_elems /*M*N*/, int G, int N, float* __restrict__ C,
                                                            int64_t ldc, int64_t c_group_stride, const float* __restrict__ bias,
                                                            int64_t bias_group_stride, int act, float slope, float beta) {
  pdl_launch_dependents();
  pdl_wait();
  const int64_t total = group_elems * G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < n_split; ++s) acc += part[(int64_t)s * total + i];
    const int g = (int)(i / group_elems);
    const int64_t r = i % group_elems;
    const int64_t m = r / N;
    const int n = (int)(r % N);
    float* dst = C + (int64_t)g * c_group_stride + m * ldc + n;
    if (beta != 0.f) acc = fmaf(beta, *dst, acc);          // C = act(alpha * A B + beta * C + bias), as on the direct path
    if (bias) acc += bias[(int64_t)g * bias_group_stride + n];
    *dst = apply_act(acc, act, slope);
  }
}


// ------------------------------------------------------------------------------------------ K3: row-packed Gram (SYRK)
// G = U U^T for K <= 64 clients.  A 128-row MMA tile that holds only K rows wastes (128 - K) / 128 of every TMA box, MMA
// and splitter pass, and - worse - the kernel is bound by the per-k-block latency chain (TMA -> split -> MMA), not by
// bytes: with K = 8 a k-block moves 1 KB.  So the tile's 128 rows are filled with S = 128 / Kp DIFFERENT 32-column slices
// of the same Kp = pow2(K) rows (one 3-D TMA box {32, Kp, S} over the view [D/32 slices][K rows][32 cols]); one k-block
// then covers S * 32 columns of every client.  T T^T is a 128 x 128 accumulator whose S diagonal Kp x Kp blocks are the
// slice Grams: the epilogue sums them (fixed order) into the split's partial.  S x fewer k-blocks: 16 x at K = 8.
// SPLIT3 = 3xTF32 with optional re-centring (see the splitter of gemm_tcgen05_body).
struct GramPackedArgs {
  float* partial;        // [n_split][K*K]
  const float* center;   // may be null
  int64_t D;
  int K, Kp, S, kp_shift;
  int total_kb, kb_per_split;
};

template <bool SPLIT3>
struct GramPackedPlan {
  static constexpr int TILE = 128 * ROW_BYTES;                    // 16 KB
  static constexpr int STAGE_BYTES = TILE * (SPLIT3 ? 2 : 1);
  static constexpr int STAGES = SPLIT3 ? 6 : 10;
  static constexpr int TOTAL = STAGES * STAGE_BYTES + 1024 + 1024;
};

template <bool SPLIT3>
__global__ void __launch_bounds__(kThreads, 1) gram_packed_kernel(const __grid_constant__ CUtensorMap tm, const GramPackedArgs a) {
  using Plan = GramPackedPlan<SPLIT3>;
  constexpr int STAGES = Plan::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Plan::STAGE_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* split_bar = bars + 2 * STAGES;
  uint64_t* accum_bar = bars + 3 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x;
  const int kb_begin = split * a.kb_per_split;
  int kb_end = kb_begin + a.kb_per_split;
  if (kb_end > a.total_kb) kb_end = a.total_kb;
  const int num_kb = kb_end > kb_begin ? kb_end - kb_begin : 0;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tm);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&split_bar[s], kEpiWarps);
    }
    mbar_init(accum_bar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<128>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_acc = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      uint32_t phase = 0;
      int s = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[s], phase ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], Plan::TILE);
        tma_load_3d(smem + s * Plan::STAGE_BYTES, &tm, &full_bar[s], 0, 0, (kb_begin + kb) * a.S);
        if (++s == STAGES) { s = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, 128, false, false);
    uint32_t phase = 0;
    int s = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      mbar_wait(SPLIT3 ? &split_bar[s] : &full_bar[s], phase);
      fence_after_thread_sync();
      if (elect_one()) {
        const uint64_t d_hi0 = smem_desc_sw128(smem_u32(smem + s * Plan::STAGE_BYTES), 16, 1024, LAYOUT_SW128);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const bool acc = (kb > 0) || (kk > 0);
          const uint64_t d_hi = desc_advance(d_hi0, kk * 32);
          if constexpr (SPLIT3) {
            const uint64_t d_lo = desc_advance(d_hi, Plan::TILE);
            mma_tf32(tmem_acc, d_lo, d_hi, idesc, acc);
            mma_tf32(tmem_acc, d_hi, d_lo, idesc, true);
            mma_tf32(tmem_acc, d_hi, d_hi, idesc, true);
          } else {
            mma_tf32(tmem_acc, d_hi, d_hi, idesc, acc);
          }
        }
        mma_commit(&empty_bar[s]);
        if (kb == num_kb - 1) mma_commit(accum_bar);
      }
      __syncwarp();
      if (++s == STAGES) { s = 0; phase ^= 1; }
    }
  } else {
    const int et = (warp - 2) * 32 + lane;      // 0..127
    if constexpr (SPLIT3) {
      uint32_t phase = 0;
      int s = 0;
      const int cchunk = ((et & 7) ^ ((et >> 3) & 7)) * 4;          // logical columns of this thread's float4s (see gemm body)
      for (int kb = 0; kb < num_kb; ++kb) {
        // float4 i = et + 128 it lives in row et/8 + 16 it, i.e. slice (row >> kp_shift) of this k-block
        float4 cv[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          cv[it] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a.center != nullptr) {
            const int row = (et >> 3) + 16 * it;
            const int64_t c0 = ((int64_t)(kb_begin + kb) * a.S + (row >> a.kp_shift)) * 32 + cchunk;
            if (c0 + 3 < a.D) cv[it] = __ldg(reinterpret_cast<const float4*>(a.center + c0));
          }
        }
        mbar_wait(&full_bar[s], phase);
        float4* hi = reinterpret_cast<float4*>(smem + s * Plan::STAGE_BYTES);
        float4* lo = hi + Plan::TILE / 16;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int i = et + 128 * it;
          const int row = (et >> 3) + 16 * it;
          float4 x = hi[i], h, l;
          if ((row & (a.Kp - 1)) < a.K) {                            // padding rows (k >= K) stay exactly zero
            x.x -= cv[it].x; x.y -= cv[it].y; x.z -= cv[it].z; x.w -= cv[it].w;
          }
          h.x = __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u); l.x = x.x - h.x;
          h.y = __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u); l.y = x.y - h.y;
          h.z = __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u); l.z = x.z - h.z;
          h.w = __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u); l.w = x.w - h.w;
          hi[i] = h;
          lo[i] = l;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&split_bar[s]);
        if (++s == STAGES) { s = 0; phase ^= 1; }
      }
    }
    // ---- epilogue: diagonal Kp x Kp blocks of the 128 x 128 accumulator -> smem -> ordered sum over the S slices
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;           // accumulator row (TMEM lane) of this thread
    const int sblk = r >> a.kp_shift;            // its slice
    float* diag = reinterpret_cast<float*>(smem);   // [128][Kp]; the operand ring is idle by now
    if (num_kb > 0) {
      mbar_wait(accum_bar, 0);
      fence_after_thread_sync();
#pragma unroll 1
      for (int c0 = 0; c0 < 128; c0 += 32) {
        float v[32];
        tmem_ld_32x32(tmem_acc + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) {
          const int col = c0 + jj;
          if ((col >> a.kp_shift) == sblk) diag[r * a.Kp + (col & (a.Kp - 1))] = v[jj];
        }
      }
    } else {
      for (int j = 0; j < a.Kp; ++j) diag[r * a.Kp + j] = 0.f;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    float* out = a.partial + (int64_t)split * a.K * a.K;
    for (int idx = et; idx < a.K * a.K; idx += 32 * kEpiWarps) {
      const int i = idx / a.K, j = idx - i * a.K;
      float acc = 0.f;
      for (int sl = 0; sl < a.S; ++sl) acc += diag[((sl << a.kp_shift) + i) * a.Kp + j];
      out[idx] = acc;
    }
    fence_before_thread_sync();
  }
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<128>(tmem_acc);
  }
}

// Host side of the packed Gram.  Returns -1 when the shape is not eligible (the caller then uses the generic GEMM).
int launch_gram_packed(const float* U, int64_t ld, int K, int64_t D, const float* center, float* gram, int64_t ldg, int precise,
                       int n_split, float* scratch, cudaStream_t st) {
  if (K > 64 || D % 32 != 0 || D < 32 || (ld * 4) % 16 != 0 || (reinterpret_cast<uintptr_t>(U) & 15) != 0) return -1;
  int Kp = 8, shift = 3;
  while (Kp < K) { Kp <<= 1; ++shift; }
  const int S = 128 / Kp;
  const int64_t slices = D / 32;
  const int total_kb = (int)((slices + S - 1) / S);
  if (n_split < 1) n_split = 1;
  if (n_split > total_kb) n_split = total_kb;          // (3xTF32 callers ask for many splits to keep accumulation chains short)
  const int kbps = (total_kb + n_split - 1) / n_split;
  n_split = (total_kb + kbps - 1) / kbps;
  if (n_split > 1 && scratch == nullptr) return -1;
  CUtensorMap tm;
  uint64_t dims[3] = {32, (uint64_t)K, (uint64_t)slices};
  uint64_t strides[2] = {(uint64_t)ld * 4, (uint64_t)ROW_BYTES};
  uint32_t box[3] = {32, (uint32_t)Kp, (uint32_t)S};
  if (int e = ddl_encode_tmap(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, U, dims, strides, box, 1)) return e;
  GramPackedArgs a;
  // a single split writes its K x K block straight to the output only when the output is dense; otherwise via scratch
  const bool direct = (n_split == 1 && ldg == K);
  if (!direct && scratch == nullptr) return -1;
  a.partial = direct ? gram : scratch;
  a.center = center; a.D = D; a.K = K; a.Kp = Kp; a.S = S; a.kp_shift = shift; a.total_kb = total_kb; a.kb_per_split = kbps;
  const bool split3 = precise != 0 || center != nullptr;
  if (split3) {
    auto kern = gram_packed_kernel<true>;
    static DdlOncePerDevice configured;
    if (configured.need()) { DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GramPackedPlan<true>::TOTAL)); configured.done(); }
    DDL_CUDA_CHECK(ddl_launch_pdl(kern, dim3(n_split), dim3(kThreads), GramPackedPlan<true>::TOTAL, st, tm, a));
  } else {
    auto kern = gram_packed_kernel<false>;
    static DdlOncePerDevice configured;
    if (configured.need()) { DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GramPackedPlan<false>::TOTAL)); configured.done(); }
    DDL_CUDA_CHECK(ddl_launch_pdl(kern, dim3(n_split), dim3(kThreads), GramPackedPlan<false>::TOTAL, st, tm, a));
  }
  DDL_LAUNCH_CHECK();
  if (!direct) {
    const int64_t ge = (int64_t)K * K;
    DDL_CUDA_CHECK(ddl_launch_pdl(splitk_reduce_kernel, dim3((unsigned)((ge + 255) / 256)), dim3(256), 0, st, (const float*)scratch, n_split, ge,
                                  1, K, gram, ldg, (int64_t)0, (const float*)nullptr, (int64_t)0, 0, 0.f, 0.f));
    DDL_LAUNCH_CHECK();
  }
  return DDL_OK;
}

// ------------------------------------------------------------------------------------------ host
struct OperandDesc {
  const void* ptr;
  int64_t rows;          // M (for A) or N (for B)
  int64_t ld;            // leading dimension in elements (K-major: row stride; MN-major: k-row stride)
  int64_t group_stride;  // elements
  int mn_major;
};

int make_operand_map(CUtensorMap* map, const OperandDesc& op, int64_t K, int G, int box_rows, int eb) {
  const int bk = ROW_BYTES / eb;
  const CUtensorMapDataType dt = eb == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  if (!op.mn_major) {
    uint64_t dims[3] = {(uint64_t)K, (uint64_t)op.rows, (uint64_t)G};
    uint64_t strides[2] = {(uint64_t)op.ld * eb, (uint64_t)(G > 1 ? op.group_stride : op.ld * op.rows) * eb};
    uint32_t box[3] = {(uint32_t)bk, (uint32_t)box_rows, 1};
    return ddl_encode_tmap(map, dt, 3, op.ptr, dims, strides, box, /*swizzle128=*/1);
  }
  DDL_REQUIRE(op.rows % bk == 0, "MN-major operand needs rows %% %d == 0 (got %lld)", bk, (long long)op.rows);
  uint64_t dims[4] = {(uint64_t)bk, (uint64_t)K, (uint64_t)(op.rows / bk), (uint64_t)G};
  uint64_t strides[3] = {(uint64_t)op.ld * eb, (uint64_t)ROW_BYTES,
                         (uint64_t)(G > 1 ? op.group_stride : op.ld * K) * eb};
  uint32_t box[4] = {(uint32_t)bk, (uint32_t)bk, (uint32_t)(box_rows / bk), 1};
  return ddl_encode_tmap(map, dt, 4, op.ptr, dims, strides, box, eb == 4 ? 2 : 1);
}

template <int BN, int EB, bool A_MN, bool B_MN, bool SPLIT3>
int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const Epilogue& epi, int M, int N, int K, int G,
                int n_split, int kb_per_split, cudaStream_t st) {
  using Plan = SmemPlan<BN, EB, SPLIT3>;
  auto kern = gemm_tcgen05_kernel<BN, EB, A_MN, B_MN, SPLIT3>;
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::TOTAL));
    configured.done();
  }
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, G * n_split);
  static const int fix_mode = getenv("DDL_FIXUP_LAUNCH") ? atoi(getenv("DDL_FIXUP_LAUNCH")) : 0;
  if (epi.counters != nullptr && fix_mode < 2)   // the split CTAs of a tile wait for each other: cooperative launch (all CTAs co-resident, or an error)
    DDL_CUDA_CHECK(ddl_launch_coop(kern, grid, dim3(kGemmThreads), Plan::TOTAL, st, fix_mode == 1, ta, tb, epi, M, N, K, kb_per_split));
  else
    DDL_CUDA_CHECK(ddl_launch_pdl(kern, grid, dim3(kGemmThreads), Plan::TOTAL, st, ta, tb, epi, M, N, K, kb_per_split));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

template <int BN, bool A_MN, bool B_MN, bool SPLIT3>
int launch_gemm_short(const CUtensorMap& ta, const CUtensorMap& tb, const Epilogue& epi, int M, int N, int K, int G, cudaStream_t st) {
  using Plan = SmemPlan<BN, 4, SPLIT3, true>;
  auto kern = gemm_tcgen05_short_kernel<BN, A_MN, B_MN, SPLIT3>;
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::TOTAL));
    configured.done();
  }
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, G);
  DDL_CUDA_CHECK(ddl_launch_pdl(kern, grid, dim3(kGemmThreads), Plan::TOTAL, st, ta, tb, epi, M, N, K, (K + 31) / 32));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

// short-K fp32 GEMMs with many tiles: BN = 64 (3xTF32) or 128 (TF32), two CTAs per SM
template <bool SPLIT3>
int dispatch_short(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const Epilogue& epi, int M, int N, int K, int G,
                   cudaStream_t st) {
  constexpr int BN = SPLIT3 ? 64 : 128;
  if (!a_mn && !b_mn) return launch_gemm_short<BN, false, false, SPLIT3>(ta, tb, epi, M, N, K, G, st);
  if (!a_mn && b_mn) return launch_gemm_short<BN, false, true, SPLIT3>(ta, tb, epi, M, N, K, G, st);
  if (a_mn && !b_mn) return launch_gemm_short<BN, true, false, SPLIT3>(ta, tb, epi, M, N, K, G, st);
  return launch_gemm_short<BN, true, true, SPLIT3>(ta, tb, epi, M, N, K, G, st);
}


// ------------------------------------------------------------------------------------------ persistent GEMM (many tiles)
// A grid of several waves of one-tile CTAs pays the fixed cost of a CTA (~6 us: first TMA round trip, accumulator drain,
// epilogue) once per tile and runs nothing else on the SM meanwhile: the 768 x 32 000 x 288 vocabulary projection of the LLaMA
// lab took 145 us for ten waves of 9-k-block tiles.  Here ONE CTA per SM walks the tile list: the TMA producer, the splitter
// warps and the MMA issuer run ahead across tile boundaries through the operand ring, and the accumulator is DOUBLE-BUFFERED in
// TMEM (2 x 128 columns) so that the four epilogue warps drain tile i while the tensor core fills tile i + 1.  Tiles are
// numbered so that the CTAs running at the same time share operand slabs through L2 (see n_fast).
// fp32 operands (TF32 or 3xTF32), BN = 128, one group, no split-K.
constexpr int kPersistThreads = 32 * (2 + kEpiWarps + 4);   // TMA, MMA, 4 epilogue warps, 4 splitter warps

template <bool SPLIT3>
struct PersistPlan {
  static constexpr int BN = 128;
  static constexpr int A_BYTES = BM * ROW_BYTES, B_BYTES = BN * ROW_BYTES;
  static constexpr int STAGE_BYTES = (A_BYTES + B_BYTES) * (SPLIT3 ? 2 : 1);
  static constexpr int STAGES = SPLIT3 ? 3 : 6;              // 192 KB of operand ring
  static constexpr int EPI_PITCH = 36;                       // floats; see the epilogue of gemm_tcgen05_body
  static constexpr int EPI_BYTES = kEpiWarps * 32 * EPI_PITCH * 4;
  static constexpr int BAR_BYTES = 1024;
  static constexpr int TOTAL = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024 /*alignment slack*/;
  static_assert(TOTAL <= 227 * 1024, "persistent GEMM plan exceeds the shared memory of an SM");
};

template <bool A_MN, bool B_MN, bool SPLIT3>
__global__ void __launch_bounds__(kPersistThreads, 1)
gemm_persist_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const Epilogue epi,
                    const int M, const int N, const int K) {
  using Plan = PersistPlan<SPLIT3>;
  constexpr int BN = Plan::BN, STAGES = Plan::STAGES, BK = 32, UMMA_K = 8, KSTEPS = 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  float* epi_smem = reinterpret_cast<float*>(smem + STAGES * Plan::STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Plan::STAGE_BYTES + Plan::EPI_BYTES);
  uint64_t* full_bar = bars;                    // TMA bytes landed
  uint64_t* empty_bar = bars + STAGES;          // MMAs that read the stage have completed
  uint64_t* split_bar = bars + 2 * STAGES;      // hi/lo split written (SPLIT3 only)
  uint64_t* acc_full = bars + 3 * STAGES;       // [2] accumulator buffer complete
  uint64_t* acc_empty = bars + 3 * STAGES + 2;  // [2] accumulator buffer drained by the epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = (M + BM - 1) / BM, n_tiles = (N + BN - 1) / BN;
  const int total_tiles = m_tiles * n_tiles;
  const int num_kb = (K + BK - 1) / BK;
  // consecutive tiles (= the tiles in flight on neighbouring SMs) share the slab of the operand with MORE slabs' counterpart:
  // few row tiles -> m fastest (they share a B slab); few column tiles -> n fastest (they share an A slab)
  const bool n_fast = m_tiles > n_tiles;
  auto tile_m0 = [&](int t) { return (n_fast ? t / n_tiles : t % m_tiles) * BM; };
  auto tile_n0 = [&](int t) { return (n_fast ? t % n_tiles : t / m_tiles) * BN; };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&split_bar[s], 4);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], kEpiWarps);
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<2 * BN>(tmem_slot);
  fence_before_thread_sync();
  __syncthreads();
  fence_after_thread_sync();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t phase = 0;
      int s = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int m0 = tile_m0(t), n0 = tile_n0(t);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[s], phase ^ 1);
          uint8_t* a_dst = stage_base + s * Plan::STAGE_BYTES;
          uint8_t* b_dst = a_dst + Plan::A_BYTES;
          mbar_arrive_expect_tx(&full_bar[s], Plan::A_BYTES + Plan::B_BYTES);
          const int k0 = kb * BK;
          if constexpr (A_MN) tma_load_4d(a_dst, &tma_a, &full_bar[s], 0, k0, m0 / BK, 0);
          else tma_load_3d(a_dst, &tma_a, &full_bar[s], k0, m0, 0);
          if constexpr (B_MN) tma_load_4d(b_dst, &tma_b, &full_bar[s], 0, k0, n0 / BK, 0);
          else tma_load_3d(b_dst, &tma_b, &full_bar[s], k0, n0, 0);
          if (++s == STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc(FMT_TF32, BM, BN, A_MN, B_MN);
    constexpr uint32_t a_lay = A_MN ? LAYOUT_SW128_BASE32B : LAYOUT_SW128, b_lay = B_MN ? LAYOUT_SW128_BASE32B : LAYOUT_SW128;
    constexpr uint32_t a_sbo = A_MN ? 512 : 1024, b_sbo = B_MN ? 512 : 1024;
    constexpr uint32_t a_lbo = A_MN ? BK * ROW_BYTES : 16, b_lbo = B_MN ? BK * ROW_BYTES : 16;
    constexpr uint32_t a_step = A_MN ? UMMA_K * ROW_BYTES : 32, b_step = B_MN ? UMMA_K * ROW_BYTES : 32;
    constexpr uint32_t LO_OFF = Plan::A_BYTES + Plan::B_BYTES;
    uint32_t phase = 0, acc_phase = 0;
    int s = 0, buf = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      mbar_wait(&acc_empty[buf], acc_phase ^ 1);          // the epilogue has drained this buffer (tile t - 2)
      fence_after_thread_sync();
      const uint32_t tmem_acc = tmem_base + (uint32_t)(buf * BN);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(SPLIT3 ? &split_bar[s] : &full_bar[s], phase);
        fence_after_thread_sync();
        if (elect_one()) {
          const uint32_t a_hi = smem_u32(stage_base + s * Plan::STAGE_BYTES);
          const uint64_t da_hi0 = smem_desc_sw128(a_hi, a_lbo, a_sbo, a_lay);
          const uint64_t db_hi0 = smem_desc_sw128(a_hi + Plan::A_BYTES, b_lbo, b_sbo, b_lay);
#pragma unroll
          for (int kk = 0; kk < KSTEPS; ++kk) {
            const bool acc = (kb > 0) || (kk > 0);
            const uint64_t da_hi = desc_advance(da_hi0, kk * a_step), db_hi = desc_advance(db_hi0, kk * b_step);
            if constexpr (SPLIT3) {
              const uint64_t da_lo = desc_advance(da_hi, LO_OFF), db_lo = desc_advance(db_hi, LO_OFF);
              mma_tf32(tmem_acc, da_lo, db_hi, idesc, acc);
              mma_tf32(tmem_acc, da_hi, db_lo, idesc, true);
              mma_tf32(tmem_acc, da_hi, db_hi, idesc, true);
            } else {
              mma_tf32(tmem_acc, da_hi, db_hi, idesc, acc);
            }
          }
          mma_commit(&empty_bar[s]);
          if (kb == num_kb - 1) mma_commit(&acc_full[buf]);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; phase ^= 1; }
      }
      if (++buf == 2) { buf = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 2 + kEpiWarps) {
    // ===================== splitter warps (3xTF32): lo = x - trunc_tf32(x); the staged tile is the hi operand =====================
    if constexpr (SPLIT3) {
      const int st = (warp - 2 - kEpiWarps) * 32 + lane;     // 0..127
      constexpr int NV = (Plan::A_BYTES + Plan::B_BYTES) / 16;
      uint32_t phase = 0;
      int s = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[s], phase);
          const float4* hi = reinterpret_cast<const float4*>(stage_base + s * Plan::STAGE_BYTES);
          float4* lo = reinterpret_cast<float4*>(stage_base + s * Plan::STAGE_BYTES) + NV;
#pragma unroll 4
          for (int i = st; i < NV; i += 128) {
            const float4 x = hi[i];
            float4 l;
            l.x = x.x - __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u);
            l.y = x.y - __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u);
            l.z = x.z - __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u);
            l.w = x.w - __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u);
            lo[i] = l;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&split_bar[s]);
          if (++s == STAGES) { s = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue warps: TMEM -> registers -> shared transpose -> global =====================
    const int quarter = warp & 3;
    constexpr int WP = Plan::EPI_PITCH;
    float* wsm = epi_smem + (warp - 2) * (32 * WP);
    float* Cg = epi.C;
    const int64_t ldc = epi.ldc;
    const float* bias = epi.bias;
    const bool vec_ok = (ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(Cg) & 15) == 0) && (N % 4 == 0);
    const int lr = lane >> 3, lc = (lane & 7) * 4;
    uint32_t acc_phase = 0;
    int buf = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int m0 = tile_m0(t), n0 = tile_n0(t);
      const int mrow0 = m0 + quarter * 32;
      mbar_wait(&acc_full[buf], acc_phase);
      fence_after_thread_sync();
      const uint32_t tmem_acc = tmem_base + (uint32_t)(buf * BN) + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        if (n0 + c0 >= N) break;
        float v[32];
        tmem_ld_32x32(tmem_acc + (uint32_t)c0, v);
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          *reinterpret_cast<float4*>(wsm + lane * WP + i) =
              make_float4(v[i] * epi.alpha, v[i + 1] * epi.alpha, v[i + 2] * epi.alpha, v[i + 3] * epi.alpha);
        __syncwarp();
        if (vec_ok) {
          const int n = n0 + c0 + lc;
          const bool nok = n < N;
          float4 bn4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (bias && nok) bn4 = *reinterpret_cast<const float4*>(bias + n);
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + lr, m = mrow0 + r;
            float4 x = *reinterpret_cast<const float4*>(wsm + r * WP + lc);
            if (nok && m < M) {
              float4* dst = reinterpret_cast<float4*>(Cg + (int64_t)m * ldc + n);
              if (epi.beta != 0.f) {
                const float4 old = *dst;
                x.x = fmaf(epi.beta, old.x, x.x); x.y = fmaf(epi.beta, old.y, x.y);
                x.z = fmaf(epi.beta, old.z, x.z); x.w = fmaf(epi.beta, old.w, x.w);
              }
              x.x = apply_act(x.x + bn4.x, epi.act, epi.slope); x.y = apply_act(x.y + bn4.y, epi.act, epi.slope);
              x.z = apply_act(x.z + bn4.z, epi.act, epi.slope); x.w = apply_act(x.w + bn4.w, epi.act, epi.slope);
              *dst = x;
            }
          }
        } else {
          const int nn = n0 + c0 + lane;
          const bool ncol_ok = nn < N;
          const float bn = (bias && ncol_ok) ? bias[nn] : 0.f;
          float* cbase = Cg + (int64_t)mrow0 * ldc + nn;
#pragma unroll 4
          for (int rr = 0; rr < 32; ++rr) {
            float x = wsm[rr * WP + lane];
            if (ncol_ok && mrow0 + rr < M) {
              if (epi.beta != 0.f) x = fmaf(epi.beta, cbase[(int64_t)rr * ldc], x);
              cbase[(int64_t)rr * ldc] = apply_act(x + bn, epi.act, epi.slope);
            }
          }
        }
        __syncwarp();
      }
      fence_before_thread_sync();                   // this warp's tcgen05.ld of the buffer are complete ...
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);  // ... the tensor core may overwrite it (tile t + 2)
      if (++buf == 2) { buf = 0; acc_phase ^= 1; }
    }
  }
  fence_before_thread_sync();
  __syncthreads();
  if (warp == 2) {
    fence_after_thread_sync();
    tmem_dealloc<2 * PersistPlan<SPLIT3>::BN>(tmem_base);
  }
}

template <bool A_MN, bool B_MN, bool SPLIT3>
int launch_gemm_persist(const CUtensorMap& ta, const CUtensorMap& tb, const Epilogue& epi, int M, int N, int K, cudaStream_t st) {
  using Plan = PersistPlan<SPLIT3>;
  auto kern = gemm_persist_kernel<A_MN, B_MN, SPLIT3>;
  static DdlOncePerDevice configured;
  if (configured.need()) {
    DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::TOTAL));
    configured.done();
  }
  const int tiles = ((M + BM - 1) / BM) * ((N + Plan::BN - 1) / Plan::BN);
  const int grid = tiles < kNumSMs ? tiles : kNumSMs;
  DDL_CUDA_CHECK(ddl_launch_pdl(kern, dim3(grid), dim3(kPersistThreads), Plan::TOTAL, st, ta, tb, epi, M, N, K));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

template <bool SPLIT3>
int dispatch_persist(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const Epilogue& epi, int M, int N, int K,
                     cudaStream_t st) {
  if (!a_mn && !b_mn) return launch_gemm_persist<false, false, SPLIT3>(ta, tb, epi, M, N, K, st);
  if (!a_mn && b_mn) return launch_gemm_persist<false, true, SPLIT3>(ta, tb, epi, M, N, K, st);
  if (a_mn && !b_mn) return launch_gemm_persist<true, false, SPLIT3>(ta, tb, epi, M, N, K, st);
  return launch_gemm_persist<true, true, SPLIT3>(ta, tb, epi, M, N, K, st);
}

template <int BN, int EB, bool SPLIT3>
int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const Epilogue& epi, int M,
                   int N, int K, int G, int n_split, int kbps, cudaStream_t st) {
  if (!a_mn && !b_mn) return launch_gemm<BN, EB, false, false, SPLIT3>(ta, tb, epi, M, N, K, G, n_split, kbps, st);
  if (!a_mn && b_mn) return launch_gemm<BN, EB, false, true, SPLIT3>(ta, tb, epi, M, N, K, G, n_split, kbps, st);
  if (a_mn && !b_mn) return launch_gemm<BN, EB, true, false, SPLIT3>(ta, tb, epi, M, N, K, G, n_split, kbps, st);
  return launch_gemm<BN, EB, true, true, SPLIT3>(ta, tb, epi, M, N, K, G, n_split, kbps, st);
}

}  // namespace

extern "C" {

int64_t ddl_gemm_scratch_floats(int M, int N, int G, int n_split) {
  return n_split > 1 ? (int64_t)n_split * G * M * N : 0;
}

static int gemm_impl(const void* A, int64_t lda, int64_t a_group_stride, int a_mn_major, const void* B, int64_t ldb,
                     int64_t b_group_stride, int b_mn_major, float* C, int64_t ldc, int64_t c_group_stride,
                     const float* bias, int64_t bias_group_stride, int act, float slope, float alpha, float beta,
                     int M, int N, int K, int G, int dtype_bytes, int precise, int n_split, float* scratch,
                     const float* center, int* tile_counters, void* stream) {
  DDL_REQUIRE(M > 0 && N > 0 && K > 0 && G > 0, "empty GEMM");
  DDL_REQUIRE(dtype_bytes == 4 || dtype_bytes == 2, "operands must be fp32 (tf32 math) or bf16");
  // Gram / SYRK of at most 64 rows: the row-packed kernel (S slices of the rows share one 128-row tile)
  if (A == B && lda == ldb && G == 1 && M == N && M <= 64 && dtype_bytes == 4 && !a_mn_major && !b_mn_major && bias == nullptr &&
      act == 0 && alpha == 1.f && beta == 0.f && K % 32 == 0) {
    const int rc = launch_gram_packed(static_cast<const float*>(A), lda, M, K, center, C, ldc, precise, n_split, scratch,
                                      (cudaStream_t)stream);
    if (rc >= 0) return rc;
  }
  DDL_REQUIRE(!(precise && dtype_bytes != 4), "precise (3xTF32) mode needs fp32 operands");
  const int eb = dtype_bytes, bk = ROW_BYTES / eb;
  DDL_REQUIRE((lda * eb) % 16 == 0 && (ldb * eb) % 16 == 0, "operand leading dimensions must be multiples of 16 bytes");
  DDL_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0,
              "operands must be 16-byte aligned");
  // a K-major operand's K is the innermost extent of its tensor map: TMA wants whole 16-byte units (a ragged inner extent hangs
  // the copy on B200, see the gather path below) - pad the operand or use the MN-major form
  DDL_REQUIRE((a_mn_major && b_mn_major) || ((int64_t)K * eb) % 16 == 0, "K-major operands need K*%d %% 16 == 0 (K=%d)", eb, K);
  DDL_REQUIRE(K <= 0x7fffffff - 64, "K=%d too large", K);
  if (n_split < 1) n_split = 1;
  const int total_kb = (K + bk - 1) / bk;
  if (n_split > total_kb) n_split = total_kb;
  if (beta != 0.f) tile_counters = nullptr;      // accumulate into C with split-K: the separate reduce kernel applies beta
  DDL_REQUIRE(n_split == 1 || scratch != nullptr, "split-K needs a scratch buffer of ddl_gemm_scratch_floats() floats");
  const int kbps = (total_kb + n_split - 1) / n_split;
  n_split = (total_kb + kbps - 1) / kbps;

  int BN = N <= 32 ? 32 : (N <= 64 ? 64 : 128);
  // short K, several waves of tiles, no split: prologue/epilogue-bound -> two co-resident CTAs per SM (see SmemPlan)
  const int short_bn = precise ? 64 : 128;
  const bool use_short = eb == 4 && n_split == 1 && total_kb <= 6 && N >= short_bn &&
                         (int64_t)((M + BM - 1) / BM) * ((N + short_bn - 1) / short_bn) * G >= 2 * kNumSMs;
  if (use_short) BN = short_bn;
  if (a_mn_major) DDL_REQUIRE(M % bk == 0, "MN-major A needs M %% %d == 0", bk);
  if (b_mn_major) DDL_REQUIRE(N % bk == 0, "MN-major B needs N %% %d == 0", bk);

  CUtensorMap ta, tb;
  OperandDesc oa{A, M, lda, a_group_stride, a_mn_major}, ob{B, N, ldb, b_group_stride, b_mn_major};
  if (int e = make_operand_map(&ta, oa, K, G, BM, eb)) return e;
  if (int e = make_operand_map(&tb, ob, K, G, BN, eb)) return e;

  Epilogue epi;
  epi.keep_hi = keep_hi_default();
  epi.C = n_split > 1 ? scratch : C;
  epi.ldc = ldc; epi.c_group_stride = c_group_stride;
  // the cooperative fix-up needs all CTAs of the grid co-resident: one wave of one CTA per SM
  const int64_t grid_ctas = (int64_t)((M + BM - 1) / BM) * ((N + (N <= 32 ? 32 : (N <= 64 ? 64 : 128)) - 1) / (N <= 32 ? 32 : (N <= 64 ? 64 : 128))) * G * n_split;
  epi.counters = (n_split > 1 && grid_ctas <= kNumSMs) ? tile_counters : nullptr;
  epi.C_final = C; epi.ldc_final = ldc; epi.c_group_stride_final = c_group_stride;
  epi.bias = bias; epi.bias_group_stride = bias_group_stride;
  epi.act = act; epi.slope = slope; epi.alpha = alpha; epi.beta = beta; epi.n_split = n_split;
  // Gram / SYRK: identical operands that fit one tile in both dimensions are staged once
  epi.same_ab = (A == B && lda == ldb && a_group_stride == b_group_stride && a_mn_major == 0 && b_mn_major == 0 && M == N &&
                 M <= BM && N <= BN) ? 1 : 0;
  DDL_REQUIRE(center == nullptr || (precise && eb == 4 && !a_mn_major && !b_mn_major && (reinterpret_cast<uintptr_t>(center) & 15) == 0),
              "re-centring needs the 3xTF32 mode, K-major fp32 operands and a 16-byte aligned centre vector");
  epi.center = center; epi.center_len = K;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  const bool amn = a_mn_major != 0, bmn = b_mn_major != 0;
  // many tiles, one split, fp32: the persistent kernel (accumulator double-buffered in TMEM, epilogue under the next main loop)
  static const int persist_min_tiles = getenv("DDL_GEMM_PERSIST_MIN_TILES") ? atoi(getenv("DDL_GEMM_PERSIST_MIN_TILES")) : 2 * kNumSMs;
  if (!use_short && eb == 4 && G == 1 && n_split == 1 && BN == 128 && center == nullptr && !epi.same_ab &&
      (int64_t)((M + BM - 1) / BM) * ((N + 127) / 128) >= persist_min_tiles) {
    return precise ? dispatch_persist<true>(amn, bmn, ta, tb, epi, M, N, K, st) : dispatch_persist<false>(amn, bmn, ta, tb, epi, M, N, K, st);
  }
  if (use_short) {
    DDL_REQUIRE(center == nullptr, "re-centring is not available on the short-K path");
    epi.same_ab = 0;
    return precise ? dispatch_short<true>(amn, bmn, ta, tb, epi, M, N, K, G, st) : dispatch_short<false>(amn, bmn, ta, tb, epi, M, N, K, G, st);
  }
#define DDL_GEMM_CASE(BN_)                                                                                     \
  if (eb == 4 && precise) rc = dispatch_major<BN_, 4, true>(amn, bmn, ta, tb, epi, M, N, K, G, n_split, kbps, st);   \
  else if (eb == 4) rc = dispatch_major<BN_, 4, false>(amn, bmn, ta, tb, epi, M, N, K, G, n_split, kbps, st);        \
  else rc = dispatch_major<BN_, 2, false>(amn, bmn, ta, tb, epi, M, N, K, G, n_split, kbps, st);
  if (BN == 32) { DDL_GEMM_CASE(32) }
  else if (BN == 64) { DDL_GEMM_CASE(64) }
  else { DDL_GEMM_CASE(128) }
#undef DDL_GEMM_CASE
  if (rc) return rc;
  if (n_split > 1 && epi.counters == nullptr) {
    const int64_t ge = (int64_t)M * N;
    int64_t blocks = (ge * G + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    DDL_CUDA_CHECK(ddl_launch_pdl(splitk_reduce_kernel, dim3((unsigned)blocks), dim3(256), 0, st, scratch, n_split, ge, G, N, C, ldc,
                                  c_group_stride, bias, bias_group_stride, act, slope, beta));
    DDL_LAUNCH_CHECK();
  }
  return DDL_OK;
}

int ddl_gemm_tcgen05(const void* A, int64_t lda, int64_t a_group_stride, int a_mn_major, const void* B, int64_t ldb,
                     int64_t b_group_stride, int b_mn_major, float* C, int64_t ldc, int64_t c_group_stride,
                     const float* bias, int64_t bias_group_stride, int act, float slope, float alpha, float beta,
                     int M, int N, int K, int G, int dtype_bytes, int precise, int n_split, float* scratch,
                     void* stream) {
  return gemm_impl(A, lda, a_group_stride, a_mn_major, B, ldb, b_group_stride, b_mn_major, C, ldc, c_group_stride, bias,
                   bias_group_stride, act, slope, alpha, beta, M, N, K, G, dtype_bytes, precise, n_split, scratch, nullptr, nullptr, stream);
}

// Same GEMM with the split-K reduction done INSIDE the kernel (no second launch) whenever the whole grid is a single wave
// (tiles x G x n_split <= 148): the CTAs of a tile meet at its counter and each reduces its share of the rows.  Larger grids
// fall back to the separate reduce kernel.  tile_counters: ddl_gemm_tile_count(M, N, G) ints, zero before the call and zero
// again after it; launches that may overlap on the GPU must not share them.
int ddl_gemm_tile_count(int M, int N, int G) {
  const int BN = N <= 32 ? 32 : (N <= 64 ? 64 : 128);
  return ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * G;
}
int ddl_gemm_tcgen05_fixup(const void* A, int64_t lda, int64_t a_group_stride, int a_mn_major, const void* B, int64_t ldb,
                           int64_t b_group_stride, int b_mn_major, float* C, int64_t ldc, int64_t c_group_stride,
                           const float* bias, int64_t bias_group_stride, int act, float slope, float alpha, float beta,
                           int M, int N, int K, int G, int dtype_bytes, int precise, int n_split, float* scratch,
                           int* tile_counters, void* stream) {
  return gemm_impl(A, lda, a_group_stride, a_mn_major, B, ldb, b_group_stride, b_mn_major, C, ldc, c_group_stride, bias,
                   bias_group_stride, act, slope, alpha, beta, M, N, K, G, dtype_bytes, precise, n_split, scratch, nullptr, tile_counters,
                   stream);
}

// K3: Gram matrix of update DELTAS, G = (U - 1 c^T)(U - 1 c^T)^T, without materialising the deltas: the 3xTF32 splitter
// subtracts the centre (the global model) from every staged tile in shared memory.  center == NULL: plain U U^T.
int ddl_gram_tcgen05(const float* updates, int64_t ld, int K, int64_t D, const float* center, float* gram, int64_t ldg,
                     int precise, int n_split, float* scratch, void* stream) {
  DDL_REQUIRE(K >= 1 && K <= BM, "K=%d outside [1, %d] (one tile of clients)", K, BM);
  DDL_REQUIRE(D >= 4 && D <= 0x7fffffff - 64, "D=%lld out of range", (long long)D);
  return gemm_impl(updates, ld, 0, 0, updates, ld, 0, 0, gram, ldg, 0, nullptr, 0, 0, 0.f, 1.f, 0.f, K, K, (int)D, 1, 4,
                   center != nullptr ? 1 : precise, n_split, scratch, center, nullptr, stream);
}

// K7 host side: A = cat_p(parts[p]) along K, parts may be peer-mapped device pointers.
int ddl_gather_gemm_tcgen05(const float* const* parts, const int* part_dims, const int64_t* part_lds, int n_parts,
                            const float* W, int64_t ldw, float* C, int64_t ldc, const float* bias, int act, float slope,
                            int M, int N, int precise, int n_split, float* scratch, void* stream) {
  DDL_REQUIRE(n_parts >= 1 && n_parts <= kMaxParts, "n_parts=%d outside [1, %d]", n_parts, kMaxParts);
  DDL_REQUIRE(M > 0 && N > 0, "empty GEMM");
  GatherMaps gm;
  gm.n_parts = n_parts;
  gm.row_mode = 0;
  gm.rows_per_part = 0;
  int kb = 0, col = 0;
  for (int p = 0; p < n_parts; ++p) {
    DDL_REQUIRE(part_dims[p] > 0 && (part_lds[p] * 4) % 16 == 0 && (reinterpret_cast<uintptr_t>(parts[p]) & 15) == 0,
                "part %d: need dim > 0, 16-byte aligned pointer and row stride", p);
    OperandDesc od{parts[p], M, part_lds[p], 0, 0};
    // TMA needs the innermost extent to be a multiple of 16 bytes (a 14-float extent hangs the copy on B200):
    // describe round_up(dim, 4) columns; the caller keeps the up-to-3 padding columns of each part at zero.
    const int kdesc = (part_dims[p] + 3) / 4 * 4;
    DDL_REQUIRE(part_lds[p] >= kdesc, "part %d: row stride %lld < padded width %d", p, (long long)part_lds[p], kdesc);
    if (int e = make_operand_map(&gm.maps[p], od, kdesc, 1, BM, 4)) return e;
    gm.kb_start[p] = kb;
    gm.col_off[p] = col;
    kb += (part_dims[p] + 31) / 32;
    col += kdesc;                 // slabs of W start on 16-byte boundaries (TMA box origins must be 16-byte aligned)
  }
  for (int p = n_parts; p <= kMaxParts; ++p) gm.kb_start[p] = kb;
  const int Ktot = col;
  DDL_REQUIRE((ldw * 4) % 16 == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0, "W must be 16-byte aligned with a 16-byte row stride");
  if (n_split < 1) n_split = 1;
  if (n_split > kb) n_split = kb;
  DDL_REQUIRE(n_split == 1 || scratch != nullptr, "split-K needs scratch");
  const int kbps = (kb + n_split - 1) / n_split;
  n_split = (kb + kbps - 1) / kbps;
  const int BN = N <= 32 ? 32 : (N <= 64 ? 64 : 128);
  CUtensorMap tb;
  OperandDesc ob{W, N, ldw, 0, 0};
  if (int e = make_operand_map(&tb, ob, Ktot, 1, BN, 4)) return e;
  Epilogue epi;
  epi.keep_hi = keep_hi_default();
  epi.C = n_split > 1 ? scratch : C;
  epi.ldc = ldc; epi.c_group_stride = 0; epi.bias = bias; epi.bias_group_stride = 0;
  epi.act = act; epi.slope = slope; epi.alpha = 1.f; epi.beta = 0.f; epi.n_split = n_split; epi.same_ab = 0;
  epi.center = nullptr; epi.center_len = 0;
  epi.counters = nullptr; epi.C_final = C; epi.ldc_final = ldc; epi.c_group_stride_final = 0;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, n_split);
#define DDL_GATHER_CASE(BN_, P_)                                                                              \
  {                                                                                                            \
    auto kern = gather_gemm_tcgen05_kernel<BN_, P_>;                                                           \
    using Plan = SmemPlan<BN_, 4, P_>;                                                                         \
    DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::TOTAL));      \
    kern<<<grid, kGemmThreads, Plan::TOTAL, st>>>(gm, tb, epi, M, N, 0, kbps);                                     \
  }
  if (BN == 32) { if (precise) DDL_GATHER_CASE(32, true) else DDL_GATHER_CASE(32, false) }
  else if (BN == 64) { if (precise) DDL_GATHER_CASE(64, true) else DDL_GATHER_CASE(64, false) }
  else { if (precise) DDL_GATHER_CASE(128, true) else DDL_GATHER_CASE(128, false) }
#undef DDL_GATHER_CASE
  DDL_LAUNCH_CHECK();
  if (n_split > 1) {
    const int64_t ge = (int64_t)M * N;
    int64_t blocks = (ge + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    splitk_reduce_kernel<<<(int)blocks, 256, 0, st>>>(scratch, n_split, ge, 1, N, C, ldc, 0, bias, 0, act, slope, 0.f);
    DDL_LAUNCH_CHECK();
  }
  return DDL_OK;
}

// K3 fused: G = U U^T where the rows of U live on different GPUs (rows_per_part consecutive rows each, row stride ld) and only
// the coordinate slice [d0, d1) is reduced here (every GPU owns one slice; the K x K partials are summed afterwards).
int ddl_gram_rowgather_tcgen05(const float* const* part_ptrs, int n_parts, int rows_per_part, int64_t ld, int64_t d0, int64_t d1,
                               const float* center, float* gram, int64_t ldg, int precise, int n_split, float* scratch, void* stream) {
  DDL_REQUIRE(n_parts >= 1 && n_parts <= kMaxParts, "n_parts=%d outside [1, %d]", n_parts, kMaxParts);
  DDL_REQUIRE(rows_per_part >= 8 && rows_per_part % 8 == 0, "rows_per_part=%d must be a multiple of 8 (1024-byte swizzle atoms)", rows_per_part);
  const int Kc = n_parts * rows_per_part;                      // clients = rows of U
  DDL_REQUIRE(Kc <= BM, "at most %d rows in total", BM);
  DDL_REQUIRE(d0 % 4 == 0 && d1 > d0 && (ld * 4) % 16 == 0, "coordinate slice must start on a float4 boundary");
  const int Kdim = (int)(d1 - d0);
  GatherMaps gm;
  gm.n_parts = n_parts; gm.row_mode = 1; gm.rows_per_part = rows_per_part;
  for (int p = 0; p <= kMaxParts; ++p) gm.kb_start[p] = 0;
  for (int p = 0; p < kMaxParts; ++p) gm.col_off[p] = 0;
  const int kd4 = (Kdim + 3) / 4 * 4;
  for (int p = 0; p < n_parts; ++p) {
    OperandDesc od{part_ptrs[p] + d0, rows_per_part, ld, 0, 0};
    if (int e = make_operand_map(&gm.maps[p], od, kd4, 1, rows_per_part, 4)) return e;
  }
  const int total_kb = (Kdim + 31) / 32;
  if (n_split < 1) n_split = 1;
  if (n_split > total_kb) n_split = total_kb;
  DDL_REQUIRE(n_split == 1 || scratch != nullptr, "split-K needs scratch of n_split * K * K floats");
  const int kbps = (total_kb + n_split - 1) / n_split;
  n_split = (total_kb + kbps - 1) / kbps;
  const int BN = Kc <= 32 ? 32 : (Kc <= 64 ? 64 : 128);
  Epilogue epi;
  epi.keep_hi = keep_hi_default();
  epi.C = n_split > 1 ? scratch : gram;
  epi.ldc = ldg; epi.c_group_stride = 0; epi.bias = nullptr; epi.bias_group_stride = 0;
  epi.act = 0; epi.slope = 0.f; epi.alpha = 1.f; epi.beta = 0.f; epi.n_split = n_split; epi.same_ab = 1;
  DDL_REQUIRE(center == nullptr || (precise && (reinterpret_cast<uintptr_t>(center + d0) & 15) == 0), "re-centring needs precise != 0 and an aligned centre vector");
  epi.center = center ? center + d0 : nullptr; epi.center_len = Kdim;
  epi.counters = nullptr; epi.C_final = gram; epi.ldc_final = ldg; epi.c_group_stride_final = 0;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(1, 1, n_split);
#define DDL_ROWG_CASE2(BN_, P_)                                                                                \
  {                                                                                                            \
    auto kern = gather_gemm_tcgen05_kernel<BN_, P_>;                                                           \
    using Plan = SmemPlan<BN_, 4, P_>;                                                                         \
    DDL_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Plan::TOTAL));      \
    kern<<<grid, kGemmThreads, Plan::TOTAL, st>>>(gm, gm.maps[0], epi, Kc, Kc, Kdim, kbps);                        \
  }
#define DDL_ROWG_CASE(BN_) { if (precise) DDL_ROWG_CASE2(BN_, true) else DDL_ROWG_CASE2(BN_, false) }
  if (BN == 32) DDL_ROWG_CASE(32) else if (BN == 64) DDL_ROWG_CASE(64) else DDL_ROWG_CASE(128)
#undef DDL_ROWG_CASE
#undef DDL_ROWG_CASE2
  DDL_LAUNCH_CHECK();
  if (n_split > 1) {
    const int64_t ge = (int64_t)Kc * Kc;
    int64_t blocks = (ge + 255) / 256;
    splitk_reduce_kernel<<<(int)blocks, 256, 0, st>>>(scratch, n_split, ge, 1, Kc, gram, ldg, 0, nullptr, 0, 0, 0.f, 0.f);
    DDL_LAUNCH_CHECK();
  }
  return DDL_OK;
}

}  // extern "C"
