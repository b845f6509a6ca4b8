// Symmetric peer memory + fused reduce kernels over NVLink (K1 / K2 / K4 / K5 multi-GPU).
//
// One process per GPU.  Every rank cudaMalloc's buffers of the same size, exports them with CUDA IPC and
// maps its peers' buffers (parallel/symm_impl.py does the handle exchange through torch.distributed), so a
// kernel holds a table of peer pointers and reads / writes other GPUs' HBM directly through NVSwitch.
//
// symm_reduce_kernel is ONE launch per GPU that does a whole collective fused with its math:
//   1. entry barrier  : block 0 publishes "my inputs are ready" into every peer's signal pad
//                       (st.release.sys) and waits for all peers (ld.acquire.sys); the other blocks wait
//                       on a device-local flag.
//   2. reduce         : this GPU owns coordinates [v0, v1) and pulls them from EVERY row on EVERY GPU
//                       (coalesced 128-bit peer loads), applying  sum_k w_k x_k  (FedAvg / DP all-reduce,
//                       alpha*base + beta*sum for fused SGD) or the sorted-rank window mean (median /
//                       trimmed mean).
//   3. broadcast      : the result slice is stored to all replicas (peer stores) = the all-gather half.
//   4. exit barrier   : the last block to finish publishes "my stores are done" and waits for all peers, so
//                       kernel completion on a GPU implies every peer's stores into it have landed.
// Traffic per GPU ~ (W-1)/W of its rows out + the same in: the coordinate-sharded exchange of SURVEY 5.8.
// All waits are bounded (5 s on %globaltimer): a missing peer sets an error flag instead of hanging the GPU, and once
// the flag is set every later collective on the arena skips its stores (a timed-out exchange never publishes partial
// results; the host raises at its next check()).
//
// Graph capture: the collective's sequence number ("epoch") and the finished-block counter live in DEVICE memory
// (`state`: int[4] = go flag, finished blocks, epoch, spare) and are advanced by the kernel itself, so a launch carries no
// per-call host value and can be recorded into a CUDA graph and replayed (DataParallel captures its bucket reductions as
// parallel branches of the step graph).  One `state` + signal-pad set is a CHANNEL: collectives of one channel must be
// stream-ordered on every rank and issued in the same order on all ranks.
//
// Optimizer epilogue (K2): with `opt.kind != 0` the weighted sum is treated as the summed gradient of this GPU's
// coordinate shard; the kernel applies SGD(+momentum, weight decay) / Adam / AdamW to the local parameter shard (the
// optimizer state exists only for the owned shard: ZeRO-1 style) and stores the NEW PARAMETERS to every replica, so
// reduce-scatter + optimizer step + all-gather of a bucket are one launch.  With `mc_in` / `mc_out` (NVLS multicast
// addresses) the W peer loads become one multimem.ld_reduce and the W peer stores one multimem.st: the switch adds.
//
// Replaces (reference): the CPU gloo all_reduce of flattened gradients / weights
// (lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:53-67, weight_aggr/intro_DP_WA.py:54-67) and the Python
// list reduction of client messages across "clients" (lab/tutorial_1a/hfl_complete.py:370-383).
#include "common.cuh"
#include "ddl25_kernels.h"
#include "select.cuh"
#include <cstring>

namespace {

constexpr int kMaxWorld = DDL_MAX_WORLD;
constexpr int kMaxRows = DDL_MAX_ROWS;
constexpr unsigned long long kTimeoutNs = 5ull * 1000 * 1000 * 1000;

struct PeerPads {
  int* pad[kMaxWorld];     // pad[p] = signal pad living on GPU p: [0, W) entry flags, [W, 2W) exit flags
};
struct RowTable {
  const float* rows[kMaxRows];
  float w[kMaxRows];
};
struct OutTable {
  float* outs[kMaxWorld];
};

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// threads [0, world) of one block: signal slot `base + rank` on every peer, then wait for every peer's slot here
__device__ void cross_gpu_barrier(const PeerPads& pp, int rank, int world, int base, int epoch, int* err) {
  const int t = threadIdx.x;
  if (t < world) {
    st_release_sys(pp.pad[t] + base + rank, epoch);
    const unsigned long long t0 = gtime();
    while (ld_acquire_sys(pp.pad[rank] + base + t) < epoch) {
      if (gtime() - t0 > kTimeoutNs) { atomicExch(err, 1 + t); break; }
    }
  }
}

__device__ __forceinline__ float4 ld_peer(const float4* p) {
  float4 r;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}

struct ReduceArgs {
  int K, n_out;
  int64_t v0, v1;            // float4 range owned by this GPU
  const float* base;         // optional local base vector (alpha * base + beta * sum); the parameters when opt.kind != 0
  const float* dev_w;        // optional DEVICE weights (override the by-value table: Krum's selection weights)
  float alpha, beta;
  int rank_lo, rank_hi;      // sorted-rank window for MODE 1 / 3
  int rank, world;
  int* state;                // device-local channel state: [0] go flag, [1] finished blocks, [2] epoch
  int* err;
  int one;                   // the integer 1 as a run-time value (selection networks: DDL_CB in select_nets.cuh)
  const float* mc_in;        // NVLS: multicast address of the summed buffer (nullptr = peer loads through the row table)
  float* mc_out;             // NVLS: multicast address of the destination (nullptr = peer stores through the out table)
  ddl_opt_args opt;
};

// One optimizer update of 4 coordinates: g = summed gradient (already scaled by beta = 1/W), p = parameters, mb / vb = the
// moments of these coordinates (loaded by the caller ahead of the peer loads, stored by the caller afterwards).
__device__ __forceinline__ float4 opt_apply(const ddl_opt_args& o, float4 p, float4 g, float4& mb, float4& vb, float bc1, float bc2_rsqrt) {
  float* pp = &p.x;
  float* gg = &g.x;
  float* m4 = &mb.x;
  float* v4 = &vb.x;
  if (o.kind == 1) {                                   // SGD (+ weight decay, + momentum)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float d = fmaf(o.weight_decay, pp[i], gg[i]);
      if (o.m != nullptr) { m4[i] = fmaf(o.momentum, m4[i], d); d = m4[i]; }
      pp[i] = fmaf(-o.lr, d, pp[i]);
    }
  } else {                                             // 2 = Adam (L2 in the gradient), 3 = AdamW (decoupled decay)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float d = gg[i];
      if (o.kind == 3) pp[i] *= (1.f - o.lr * o.weight_decay);
      else d = fmaf(o.weight_decay, pp[i], d);
      m4[i] = fmaf(o.beta1, m4[i], (1.f - o.beta1) * d);
      v4[i] = fmaf(o.beta2, v4[i], (1.f - o.beta2) * d * d);
      pp[i] -= (o.lr / bc1) * (m4[i] / (sqrtf(v4[i]) * bc2_rsqrt + o.eps));
    }
  }
  return p;
}

// MODE 0: weighted sum (float4 per thread), optional optimizer epilogue / NVLS.  MODE 1: sorted-rank window mean
// (one coordinate per thread, KP rows), MODE 3: the same for a window that is exactly the median (pruned network).
// MODE 2: gather of this GPU's coordinate shard of every row.
template <int MODE, int KP>
__global__ void __launch_bounds__(256, MODE == 0 ? 1 : 2) symm_reduce_kernel(RowTable tab, OutTable ot, PeerPads pp, ReduceArgs a) {
  // the channel's epoch is advanced by the LAST block of the previous collective, i.e. it is stable for the whole launch
  const int epoch = ld_acquire_gpu(a.state + 2) + 1;
  // ---- 1. entry barrier
  if (blockIdx.x == 0) {
    cross_gpu_barrier(pp, a.rank, a.world, 0, epoch, a.err);
    __syncthreads();
    if (threadIdx.x == 0) st_release_gpu(a.state, epoch);
  } else {
    if (threadIdx.x == 0) {
      const unsigned long long t0 = gtime();
      while (ld_acquire_gpu(a.state) < epoch) {
        if (gtime() - t0 > kTimeoutNs) { atomicExch(a.err, 100); break; }
      }
    }
    __syncthreads();
  }
  // a collective that timed out (now or earlier) must not publish anything: peers may be absent or out of step
  const bool healthy = (*reinterpret_cast<volatile int*>(a.err) == 0);
  // ---- 2 + 3. reduce my coordinate shard over all rows, store to all replicas
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if constexpr (MODE == 0) {
    float bc1 = 1.f, bc2_rsqrt = 1.f;
    if (a.opt.kind >= 2) {
      const float step = (float)(*a.opt.step_dev + 1);          // completed steps + 1; bumped by the step's last bucket
      bc1 = 1.f - powf(a.opt.beta1, step);
      bc2_rsqrt = 1.f / sqrtf(1.f - powf(a.opt.beta2, step));
    }
    // U coordinates x KC rows in flight per thread: an NVLink round trip is ~2 us, and a bucket kernel runs on few CTAs so
    // that it can overlap the backward pass - with one coordinate per thread in flight it moved ~10 GB/s per CTA.
    constexpr int U = 4, KC = 4;
    const bool has_m = a.opt.kind >= 2 || (a.opt.kind == 1 && a.opt.m != nullptr);
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t v = a.v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; healthy && v < a.v1; v += stride * U) {
      float4 acc[U], pb[U], mb[U], vb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {                    // local operands first: their loads overlap the peer round trip
        const int64_t vv = v + u * stride;
        acc[u] = pb[u] = mb[u] = vb[u] = zero4;
        if (vv < a.v1) {
          if (a.base != nullptr) pb[u] = reinterpret_cast<const float4*>(a.base)[vv];
          if (has_m) mb[u] = reinterpret_cast<const float4*>(a.opt.m)[vv - a.opt.state_off4];
          if (a.opt.kind >= 2) vb[u] = reinterpret_cast<const float4*>(a.opt.v)[vv - a.opt.state_off4];
        }
      }
      if (a.mc_in != nullptr) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int64_t vv = v + u * stride;
          if (vv < a.v1)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(acc[u].x), "=f"(acc[u].y), "=f"(acc[u].z), "=f"(acc[u].w) : "l"(a.mc_in + 4 * vv) : "memory");
        }
      } else {
        // zero-weight rows are neither loaded nor accumulated (Krum: only the selected rows cross NVLink; a rejected
        // row may hold NaN / Inf, which 0 * x would turn into NaN).  Rows are accumulated in ascending order.
        for (int k = 0; k < a.K; k += KC) {
          float4 x[KC][U];
          float w[KC];
#pragma unroll
          for (int c = 0; c < KC; ++c) {
            w[c] = 0.f;
            if (k + c < a.K) w[c] = a.dev_w ? __ldg(a.dev_w + k + c) : tab.w[k + c];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int64_t vv = v + u * stride;
              x[c][u] = zero4;
              if (w[c] != 0.f && vv < a.v1) x[c][u] = ld_peer(reinterpret_cast<const float4*>(tab.rows[k + c]) + vv);
            }
          }
#pragma unroll
          for (int c = 0; c < KC; ++c)
#pragma unroll
            for (int u = 0; u < U; ++u) {
              acc[u].x = fmaf(w[c], x[c][u].x, acc[u].x); acc[u].y = fmaf(w[c], x[c][u].y, acc[u].y);
              acc[u].z = fmaf(w[c], x[c][u].z, acc[u].z); acc[u].w = fmaf(w[c], x[c][u].w, acc[u].w);
            }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t vv = v + u * stride;
        if (vv >= a.v1) continue;
        float4 r = make_float4(a.beta * acc[u].x, a.beta * acc[u].y, a.beta * acc[u].z, a.beta * acc[u].w);
        if (a.opt.kind != 0) {
          r = opt_apply(a.opt, pb[u], r, mb[u], vb[u], bc1, bc2_rsqrt);
          if (has_m) reinterpret_cast<float4*>(a.opt.m)[vv - a.opt.state_off4] = mb[u];
          if (a.opt.kind >= 2) reinterpret_cast<float4*>(a.opt.v)[vv - a.opt.state_off4] = vb[u];
        } else if (a.base != nullptr) {
          r.x = fmaf(a.alpha, pb[u].x, r.x); r.y = fmaf(a.alpha, pb[u].y, r.y); r.z = fmaf(a.alpha, pb[u].z, r.z); r.w = fmaf(a.alpha, pb[u].w, r.w);
        }
        if (a.mc_out != nullptr) {
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a.mc_out + 4 * vv), "f"(r.x), "f"(r.y), "f"(r.z), "f"(r.w) : "memory");
        } else {
          for (int o = 0; o < a.n_out; ++o) reinterpret_cast<float4*>(ot.outs[o])[vv] = r;
        }
      }
    }
  }
  if constexpr (MODE == 2) {
    // all-to-all half of the coordinate-sharded exchange: local[k][j] = rows[k][4*v0 + j], j < 4*(v1 - v0).
    // The caller then runs the Gram GEMM (K3) on the local [K, shard] matrix: every GPU pulls only ITS slice
    // of every client's update over NVLink instead of whole updates.
    const int64_t ld4 = a.v1 - a.v0;            // float4 per local row (a.rank_lo carries extra row padding)
    float4* dst = reinterpret_cast<float4*>(ot.outs[0]);
    for (int64_t v = a.v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; healthy && v < a.v1; v += stride) {
      float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
      if (a.base != nullptr) b = reinterpret_cast<const float4*>(a.base)[v];      // optional: subtract the global model -> deltas
      // eight 128-bit peer loads in flight per thread: an NVLink round trip is several times an HBM one (with four, the
      // exchange ran at 476 GB/s of remote reads on 8 GPUs)
      for (int k = 0; k < a.K; k += 8) {
        float4 x[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (k + u < a.K) x[u] = ld_peer(reinterpret_cast<const float4*>(tab.rows[k + u]) + v);
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (k + u < a.K) {
            float4 y = x[u];
            y.x -= b.x; y.y -= b.y; y.z -= b.z; y.w -= b.w;
            dst[(int64_t)(k + u) * (ld4 + a.rank_lo) + (v - a.v0)] = y;
          }
      }
    }
  }
  if constexpr (MODE == 1 || MODE == 3) {
    // one coordinate per thread (KP values in registers), 128 B coalesced peer loads per warp per row.
    // MODE 3 = median: symmetric -inf / +inf padding keeps the median at the centre ranks, pruned network (select.cuh).
    for (int64_t d = 4 * a.v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; healthy && d < 4 * a.v1; d += stride) {
      float c[KP];
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        c[k] = MODE == 3 ? ddl::median_pad(k, a.K, KP) : INFINITY;
        if (k < a.K) asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(c[k]) : "l"(tab.rows[k] + d) : "memory");
      }
      const float s = MODE == 3 ? ddl::median_of<KP>(c, a.rank_hi - a.rank_lo == 2, a.one) : ddl::window_mean<KP>(c, a.rank_lo, a.rank_hi, a.one);
      for (int o = 0; o < a.n_out; ++o) ot.outs[o][d] = s;
    }
  }
  // ---- 4. exit barrier: the last block of this launch waits until every peer's stores have landed here, then
  //         advances the channel (epoch, finished counter) for the next collective
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(reinterpret_cast<unsigned int*>(a.state + 1), 1u) + 1u == gridDim.x);
  __syncthreads();
  if (s_last) {
    cross_gpu_barrier(pp, a.rank, a.world, a.world, epoch, a.err);
    __syncthreads();
    if (threadIdx.x == 0) {
      if (a.opt.kind != 0 && a.opt.bump_step && a.opt.step_dev != nullptr) atomicAdd(a.opt.step_dev, 1);
      a.state[1] = 0;
      st_release_gpu(a.state + 2, epoch);
    }
  }
}

__global__ void symm_barrier_kernel(PeerPads pp, int rank, int world, int* state, int* err) {
  const int epoch = ld_acquire_gpu(state + 2) + 1;
  cross_gpu_barrier(pp, rank, world, 0, epoch, err);
  __syncthreads();
  cross_gpu_barrier(pp, rank, world, world, epoch, err);
  __syncthreads();
  if (threadIdx.x == 0) st_release_gpu(state + 2, epoch);
}

int fill(RowTable& tab, OutTable& ot, PeerPads& pp, const float* const* rows, const float* w, int K, float* const* outs, int n_out,
         int* const* pads, int world) {
  DDL_REQUIRE(K >= 1 && K <= kMaxRows, "K=%d outside [1, %d]", K, kMaxRows);
  DDL_REQUIRE(world >= 1 && world <= kMaxWorld, "world=%d outside [1, %d]", world, kMaxWorld);
  DDL_REQUIRE(n_out >= 0 && n_out <= kMaxWorld, "n_out=%d outside [0, %d]", n_out, kMaxWorld);
  for (int k = 0; k < K; ++k) { tab.rows[k] = rows ? rows[k] : nullptr; tab.w[k] = w ? w[k] : 1.f; }
  for (int o = 0; o < n_out; ++o) ot.outs[o] = outs[o];
  for (int p = 0; p < world; ++p) pp.pad[p] = pads[p];
  return DDL_OK;
}

}  // namespace

extern "C" {

int ddl_symm_alloc(void** ptr, int64_t bytes) {
  DDL_CUDA_CHECK(cudaMalloc(ptr, (size_t)bytes));
  DDL_CUDA_CHECK(cudaMemset(*ptr, 0, (size_t)bytes));
  return DDL_OK;
}
int ddl_symm_free(void* ptr) {
  DDL_CUDA_CHECK(cudaFree(ptr));
  return DDL_OK;
}
int ddl_ipc_export(void* ptr, unsigned char* handle64) {
  cudaIpcMemHandle_t h;
  DDL_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return DDL_OK;
}
int ddl_ipc_import(const unsigned char* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DDL_CUDA_CHECK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return DDL_OK;
}
int ddl_ipc_close(void* ptr) {
  DDL_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return DDL_OK;
}

int ddl_symm_barrier(int* const* pads, int rank, int world, int* state, int* err, void* stream) {
  PeerPads pp;
  DDL_REQUIRE(world >= 1 && world <= kMaxWorld, "world=%d outside [1, %d]", world, kMaxWorld);
  DDL_REQUIRE(state != nullptr && err != nullptr, "channel state / error flag missing");
  for (int p = 0; p < world; ++p) pp.pad[p] = pads[p];
  symm_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world, state, err);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

// mode 0: out = alpha*base + beta*sum_k w_k rows_k (or the optimizer epilogue); mode 1: mean of sorted ranks
// [rank_lo, rank_hi) (K <= 64); mode 2: outs[0][k][j] = rows[k][d0 + j] - (base ? base[d0 + j] : 0), local row pitch
// (d1 - d0) + 4*rank_lo floats
int ddl_symm_reduce(int mode, const float* const* rows, const float* weights, const float* dev_weights, int K, const float* base,
                    float alpha, float beta, int rank_lo, int rank_hi, float* const* outs, int n_out, int64_t d0, int64_t d1,
                    int* const* pads, int rank, int world, int* state, int* err, int max_ctas, const ddl_opt_args* opt,
                    const float* mc_in, float* mc_out, void* stream) {
  DDL_REQUIRE(d0 % 4 == 0 && d1 % 4 == 0 && d1 >= d0, "coordinate range must be float4 aligned");
  DDL_REQUIRE(mode != 1 || (K <= 64 && rank_lo >= 0 && rank_lo < rank_hi && rank_hi <= K), "bad rank window / K for rank-select mode");
  DDL_REQUIRE(mode >= 0 && mode <= 2, "mode must be 0 (weighted sum), 1 (rank-select mean) or 2 (gather shard)");
  DDL_REQUIRE(state != nullptr && err != nullptr, "channel state / error flag missing");
  DDL_REQUIRE((mc_in == nullptr && mc_out == nullptr) || mode == 0, "multicast operands are for mode 0 only");
  DDL_REQUIRE(mc_in != nullptr || rows != nullptr, "no input rows");
  DDL_REQUIRE(mc_out != nullptr || n_out >= 1, "no outputs");
  RowTable tab;
  OutTable ot;
  PeerPads pp;
  if (int e = fill(tab, ot, pp, rows, weights, K, outs, n_out, pads, world)) return e;
  ReduceArgs a;
  a.K = K; a.n_out = n_out; a.v0 = d0 / 4; a.v1 = d1 / 4; a.base = base; a.dev_w = dev_weights; a.alpha = alpha; a.beta = beta;
  a.rank_lo = rank_lo; a.rank_hi = rank_hi; a.rank = rank; a.world = world; a.state = state; a.err = err;
  a.mc_in = mc_in; a.mc_out = mc_out; a.one = 1;
  memset(&a.opt, 0, sizeof(a.opt));
  if (opt != nullptr && opt->kind != 0) {
    DDL_REQUIRE(mode == 0 && base != nullptr, "the optimizer epilogue needs mode 0 and base = the local parameters");
    DDL_REQUIRE(opt->kind >= 1 && opt->kind <= 3, "opt.kind must be 1 (sgd), 2 (adam) or 3 (adamw)");
    DDL_REQUIRE(opt->kind == 1 || (opt->m != nullptr && opt->v != nullptr && opt->step_dev != nullptr), "Adam needs m, v and a device step counter");
    DDL_REQUIRE(opt->state_off4 >= 0 && opt->state_off4 <= a.v0, "optimizer state offset beyond the shard start");
    a.opt = *opt;
    if (a.opt.kind == 1 && a.opt.momentum == 0.f) a.opt.m = nullptr;
  }
  int64_t g = ((a.v1 - a.v0) + 255) / 256;
  const int cap = max_ctas > 0 ? max_ctas : kNumSMs * 2;     // all blocks must be co-resident (they wait on block 0)
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0) symm_reduce_kernel<0, 1><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else if (mode == 2) symm_reduce_kernel<2, 1><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else {
    // a window that is exactly the median of K values (single centre rank, or the centre pair of an even K) takes the
    // pruned network; any other window the full sort
    const bool med = (rank_lo == (K - 1) / 2 && rank_hi == rank_lo + 1) || (K % 2 == 0 && rank_lo == K / 2 - 1 && rank_hi == K / 2 + 1);
#define DDL_SYMM_SEL(KP)                                                                    \
    do {                                                                                    \
      if (med) symm_reduce_kernel<3, KP><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);           \
      else symm_reduce_kernel<1, KP><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);               \
    } while (0)
    if (K <= 8) DDL_SYMM_SEL(8);
    else if (K <= 16) DDL_SYMM_SEL(16);
    else if (K <= 32) DDL_SYMM_SEL(32);
    else DDL_SYMM_SEL(64);
#undef DDL_SYMM_SEL
  }
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
