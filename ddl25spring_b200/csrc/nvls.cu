// NVLS (NVLink SHARP) all-reduce: the NVSwitch does the reduction.
//
// `mc` is a MULTICAST address bound to the same symmetric buffer on every GPU of the group (set up by
// torch.distributed._symmetric_memory in parallel/nvls.py).  Each GPU owns the coordinate shard [v0, v1):
//   multimem.ld_reduce  pulls that shard from ALL GPUs and returns the element-wise sum computed inside the switch,
//   multimem.st         broadcasts the (scaled) result back to ALL GPUs through the switch.
// One load + one store per 16 bytes per GPU: the traffic of a reduce-scatter + all-gather without any SM doing
// the adds.  Ordering across GPUs (inputs ready / results visible) is provided by the caller's barrier kernels.
// Replaces (reference): the CPU gloo all_reduce of lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:63.
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

__global__ void __launch_bounds__(256) nvls_allreduce_kernel(float* mc, int64_t v0, int64_t v1, float scale) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v = v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < v1; v += stride) {
    float* p = mc + 4 * v;
    float x, y, z, w;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(x), "=f"(y), "=f"(z), "=f"(w)
                 : "l"(p)
                 : "memory");
    x *= scale; y *= scale; z *= scale; w *= scale;
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
  }
}

}  // namespace

extern "C" int ddl_nvls_allreduce(float* multicast_ptr, int64_t d0, int64_t d1, float scale, int max_ctas, void* stream) {
  DDL_REQUIRE(multicast_ptr != nullptr, "null multicast pointer (NVLS not available for this group)");
  DDL_REQUIRE(d0 % 4 == 0 && d1 % 4 == 0 && d1 >= d0, "coordinate range must be float4 aligned");
  if (d1 == d0) return DDL_OK;
  int64_t g = ((d1 - d0) / 4 + 255) / 256;
  const int cap = max_ctas > 0 ? max_ctas : kNumSMs * 4;
  if (g > cap) g = cap;
  nvls_allreduce_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(multicast_ptr, d0 / 4, d1 / 4, scale);
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}
