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
// All waits are bounded (5 s on %globaltimer): a missing peer sets an error flag instead of hanging the GPU.
//
// Replaces (reference): the CPU gloo all_reduce of flattened gradients / weights
// (lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:53-67, weight_aggr/intro_DP_WA.py:54-67) and the Python
// list reduction of client messages across "clients" (lab/tutorial_1a/hfl_complete.py:370-383).
#include "common.cuh"
#include "ddl25_kernels.h"

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

template <int KP>
__device__ __forceinline__ void sort_regs(float (&a)[KP]) {
#pragma unroll
  for (int k = 2; k <= KP; k <<= 1)
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1)
#pragma unroll
      for (int i = 0; i < KP; ++i) {
        const int l = i ^ j;
        if (l > i) {
          const bool up = ((i & k) == 0);
          const float lo = fminf(a[i], a[l]), hi = fmaxf(a[i], a[l]);
          a[i] = up ? lo : hi;
          a[l] = up ? hi : lo;
        }
      }
}

struct ReduceArgs {
  int K, n_out;
  int64_t v0, v1;            // float4 range owned by this GPU
  const float* base;         // optional local base vector (alpha * base + beta * sum)
  float alpha, beta;
  int rank_lo, rank_hi;      // sorted-rank window for MODE 1
  int rank, world, epoch;
  int* local_flags;          // [0] go flag, [1] finished-block counter (device-local)
  unsigned int done_target;  // counter value that identifies the last block of THIS launch
  int* err;
};

// MODE 0: weighted sum (float4 per thread).  MODE 1: sorted-rank window mean (one coordinate per thread, KP rows).
template <int MODE, int KP>
__global__ void __launch_bounds__(256) symm_reduce_kernel(RowTable tab, OutTable ot, PeerPads pp, ReduceArgs a) {
  // ---- 1. entry barrier
  if (blockIdx.x == 0) {
    cross_gpu_barrier(pp, a.rank, a.world, 0, a.epoch, a.err);
    __syncthreads();
    if (threadIdx.x == 0) st_release_gpu(a.local_flags, a.epoch);
  } else {
    if (threadIdx.x == 0) {
      const unsigned long long t0 = gtime();
      while (ld_acquire_gpu(a.local_flags) < a.epoch) {
        if (gtime() - t0 > kTimeoutNs) { atomicExch(a.err, 100); break; }
      }
    }
    __syncthreads();
  }
  // ---- 2 + 3. reduce my coordinate shard over all rows, store to all replicas
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = a.v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; MODE == 0 && v < a.v1; v += stride) {
    float4 r;
    {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      int k = 0;
      for (; k + 8 <= a.K; k += 8) {
        float4 x[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) x[u] = ld_peer(reinterpret_cast<const float4*>(tab.rows[k + u]) + v);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float w = tab.w[k + u];
          acc.x = fmaf(w, x[u].x, acc.x); acc.y = fmaf(w, x[u].y, acc.y); acc.z = fmaf(w, x[u].z, acc.z); acc.w = fmaf(w, x[u].w, acc.w);
        }
      }
      for (; k < a.K; ++k) {
        const float4 x = ld_peer(reinterpret_cast<const float4*>(tab.rows[k]) + v);
        const float w = tab.w[k];
        acc.x = fmaf(w, x.x, acc.x); acc.y = fmaf(w, x.y, acc.y); acc.z = fmaf(w, x.z, acc.z); acc.w = fmaf(w, x.w, acc.w);
      }
      r = make_float4(a.beta * acc.x, a.beta * acc.y, a.beta * acc.z, a.beta * acc.w);
      if (a.base != nullptr) {
        const float4 b = reinterpret_cast<const float4*>(a.base)[v];
        r.x = fmaf(a.alpha, b.x, r.x); r.y = fmaf(a.alpha, b.y, r.y); r.z = fmaf(a.alpha, b.z, r.z); r.w = fmaf(a.alpha, b.w, r.w);
      }
    }
    for (int o = 0; o < a.n_out; ++o) reinterpret_cast<float4*>(ot.outs[o])[v] = r;
  }
  if constexpr (MODE == 2) {
    // all-to-all half of the coordinate-sharded exchange: local[k][j] = rows[k][4*v0 + j], j < 4*(v1 - v0).
    // The caller then runs the Gram GEMM (K3) on the local [K, shard] matrix: every GPU pulls only ITS slice
    // of every client's update over NVLink instead of whole updates.
    const int64_t ld4 = a.v1 - a.v0;            // float4 per local row (a.rank_lo carries extra row padding)
    float4* dst = reinterpret_cast<float4*>(ot.outs[0]);
    for (int64_t v = a.v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < a.v1; v += stride) {
      for (int k = 0; k < a.K; k += 4) {
        float4 x[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k + u < a.K) x[u] = ld_peer(reinterpret_cast<const float4*>(tab.rows[k + u]) + v);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (k + u < a.K) {
            float4 y = x[u];
            if (a.base != nullptr) {              // optional: subtract the global model -> update deltas
              const float4 b = reinterpret_cast<const float4*>(a.base)[v];
              y.x -= b.x; y.y -= b.y; y.z -= b.z; y.w -= b.w;
            }
            dst[(int64_t)(k + u) * (ld4 + a.rank_lo) + (v - a.v0)] = y;
          }
      }
    }
  }
  if constexpr (MODE == 1) {
    // one coordinate per thread (KP values in registers), 128 B coalesced peer loads per warp per row
    for (int64_t d = 4 * a.v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; d < 4 * a.v1; d += stride) {
      float c[KP];
#pragma unroll
      for (int k = 0; k < KP; ++k) {
        c[k] = INFINITY;
        if (k < a.K) asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(c[k]) : "l"(tab.rows[k] + d) : "memory");
      }
      sort_regs<KP>(c);
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < KP; ++k)
        if (k >= a.rank_lo && k < a.rank_hi) s += c[k];
      s *= 1.f / (float)(a.rank_hi - a.rank_lo);
      for (int o = 0; o < a.n_out; ++o) ot.outs[o][d] = s;
    }
  }
  // ---- 4. exit barrier: the last block of this launch waits until every peer's stores have landed here
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(reinterpret_cast<unsigned int*>(a.local_flags + 1), 1u) + 1u == a.done_target);
  __syncthreads();
  if (s_last) cross_gpu_barrier(pp, a.rank, a.world, a.world, a.epoch, a.err);
}

__global__ void symm_barrier_kernel(PeerPads pp, int rank, int world, int epoch, int* err) {
  cross_gpu_barrier(pp, rank, world, 0, epoch, err);
  __syncthreads();
  cross_gpu_barrier(pp, rank, world, world, epoch, err);
}

int fill(RowTable& tab, OutTable& ot, PeerPads& pp, const float* const* rows, const float* w, int K, float* const* outs, int n_out,
         int* const* pads, int world) {
  DDL_REQUIRE(K >= 1 && K <= kMaxRows, "K=%d outside [1, %d]", K, kMaxRows);
  DDL_REQUIRE(world >= 1 && world <= kMaxWorld, "world=%d outside [1, %d]", world, kMaxWorld);
  DDL_REQUIRE(n_out >= 1 && n_out <= kMaxWorld, "n_out=%d outside [1, %d]", n_out, kMaxWorld);
  for (int k = 0; k < K; ++k) { tab.rows[k] = rows[k]; tab.w[k] = w ? w[k] : 1.f; }
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

int ddl_symm_barrier(int* const* pads, int rank, int world, int epoch, int* err, void* stream) {
  PeerPads pp;
  DDL_REQUIRE(world >= 1 && world <= kMaxWorld, "world=%d outside [1, %d]", world, kMaxWorld);
  for (int p = 0; p < world; ++p) pp.pad[p] = pads[p];
  symm_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world, epoch, err);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

// mode 0: out = alpha*base + beta*sum_k w_k rows_k ; mode 1: mean of sorted ranks [rank_lo, rank_hi) (K <= 64);
// mode 2: outs[0][k][j] = rows[k][d0 + j] - (base ? base[d0 + j] : 0), local row pitch (d1 - d0) + 4*rank_lo floats
int ddl_symm_reduce(int mode, const float* const* rows, const float* weights, int K, const float* base, float alpha, float beta,
                    int rank_lo, int rank_hi, float* const* outs, int n_out, int64_t d0, int64_t d1, int* const* pads, int rank,
                    int world, int epoch, int* local_flags, int64_t* done_counter_host, int* err, int max_ctas, void* stream) {
  DDL_REQUIRE(d0 % 4 == 0 && d1 % 4 == 0 && d1 >= d0, "coordinate range must be float4 aligned");
  DDL_REQUIRE(mode != 1 || (K <= 64 && rank_lo >= 0 && rank_lo < rank_hi && rank_hi <= K), "bad rank window / K for rank-select mode");
  DDL_REQUIRE(mode >= 0 && mode <= 2, "mode must be 0 (weighted sum), 1 (rank-select mean) or 2 (gather shard)");
  RowTable tab;
  OutTable ot;
  PeerPads pp;
  if (int e = fill(tab, ot, pp, rows, weights, K, outs, n_out, pads, world)) return e;
  ReduceArgs a;
  a.K = K; a.n_out = n_out; a.v0 = d0 / 4; a.v1 = d1 / 4; a.base = base; a.alpha = alpha; a.beta = beta;
  a.rank_lo = rank_lo; a.rank_hi = rank_hi; a.rank = rank; a.world = world; a.epoch = epoch; a.local_flags = local_flags; a.err = err;
  int64_t g = ((a.v1 - a.v0) + 255) / 256;
  const int cap = max_ctas > 0 ? max_ctas : kNumSMs * 2;     // all blocks must be co-resident (they wait on block 0)
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  *done_counter_host += g;
  a.done_target = (unsigned int)(*done_counter_host);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0) symm_reduce_kernel<0, 1><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else if (mode == 2) symm_reduce_kernel<2, 1><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else if (K <= 8) symm_reduce_kernel<1, 8><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else if (K <= 16) symm_reduce_kernel<1, 16><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else if (K <= 32) symm_reduce_kernel<1, 32><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  else symm_reduce_kernel<1, 64><<<(int)g, 256, 0, st>>>(tab, ot, pp, a);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}

}  // extern "C"
