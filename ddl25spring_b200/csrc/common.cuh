// Shared device/host helpers for the ddl25spring_b200 sm_100a kernel library.
// Everything here is written for Blackwell B200 only (compile: -gencode arch=compute_100a,code=sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <cstdio>

#define DDL_OK 0
#define DDL_ERR_ARG 1
#define DDL_ERR_CUDA 2
#define DDL_ERR_LIMIT 3

#define DDL_CUDA_CHECK(expr)                                                                    \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      ddl_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));       \
      return DDL_ERR_CUDA;                                                                      \
    }                                                                                           \
  } while (0)

#define DDL_REQUIRE(cond, ...)                                                                  \
  do {                                                                                          \
    if (!(cond)) {                                                                              \
      ddl_set_error(__VA_ARGS__);                                                               \
      return DDL_ERR_ARG;                                                                       \
    }                                                                                           \
  } while (0)

// last-error string (thread local), readable through ddl_last_error()
void ddl_set_error(const char* fmt, ...);
// counts kernel launches issued by this library (bench.py reports it as gpu_launches)
void ddl_count_launch(int n = 1);

#define DDL_LAUNCH_CHECK()                                                                      \
  do {                                                                                          \
    ddl_count_launch();                                                                         \
    cudaError_t _e = cudaGetLastError();                                                        \
    if (_e != cudaSuccess) {                                                                    \
      ddl_set_error("%s:%d kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e));   \
      return DDL_ERR_CUDA;                                                                      \
    }                                                                                           \
  } while (0)

static constexpr int kNumSMs = 148;  // B200

// One-time set-up that CUDA keeps PER DEVICE (cudaFuncSetAttribute for > 48 KB of dynamic shared memory): a plain
// `static bool` would configure only the first device a multi-GPU host process touches and the launch on the second one
// would fail.  Racing threads may both run the set-up; it is idempotent.
struct DdlOncePerDevice {
  std::atomic<unsigned long long> mask{0};
  static unsigned long long bit() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    return 1ull << (dev & 63);
  }
  bool need() const { return (mask.load(std::memory_order_acquire) & bit()) == 0; }
  void done() { mask.fetch_or(bit(), std::memory_order_release); }
};

#ifdef __CUDACC__
// Programmatic dependent launch: a step of the NHWC networks is ~200 dependent kernels of a few microseconds each, so the
// launch / drain gap between two kernels costs as much as the kernels.  Kernels launched through ddl_launch_pdl may start
// while their predecessor in the stream is still running; they execute pdl_wait() before touching global memory (it returns
// once the predecessor has completed and its writes are visible) and call pdl_launch_dependents() as soon as their own
// successor may begin its prologue.  Both are no-ops for a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t ddl_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Cooperative launch: for grids whose CTAs wait for one another (the split-K fix-up of gemm_tcgen05.cu).  The runtime either
// makes every CTA of the grid co-resident or fails the launch - a spin on a sibling CTA can then never wait for a block
// that has not been scheduled.
template <typename... KArgs, typename... Args>
inline cudaError_t ddl_launch_coop(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

#ifdef __CUDACC__
namespace ddl {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; `red` must hold >= 32 floats. Result valid in every thread.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (lane < nw) ? red[lane] : 0.f;
  v = warp_sum(v);
  return v;
}

// Streaming 128-bit accesses: client updates / gradients are touched once per kernel, keep them
// out of L1 (guide: G13/G14).
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
// Plain (coherent) 128-bit load: used for peer-mapped rows that another GPU has just written.
__device__ __forceinline__ float4 ld_coherent(const float4* p) {
  float4 r;
  asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// Counter-based RNG for dropout masks: one 32-bit hash per (seed, index).  Cheap, stateless,
// reproducible across launches (same (seed, idx) -> same mask in forward and backward).
__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t rng_u32(uint32_t seed, uint32_t idx) {
  return hash_u32(idx * 0x9E3779B9U + hash_u32(seed ^ 0x85ebca6bU));
}
__device__ __forceinline__ float rng_uniform(uint32_t seed, uint32_t idx) {
  return (rng_u32(seed, idx) >> 8) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

}  // namespace ddl
#endif  // __CUDACC__
