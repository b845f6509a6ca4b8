// Input pipeline: raw uint8 pixels (MNIST's storage format, 1 byte per pixel over PCIe instead of 4) -> normalised fp32 on
// the device, bit-identical to the host transform x.float().div(255).sub(mean).div(std) (round-to-nearest division and
// subtraction, no FMA contraction), so a model trained from uint8 shards equals one trained from pre-normalised floats.
// Replaces (reference): transforms.ToTensor() + Normalize((0.1307,), (0.3081,)) run per sample on the CPU by the DataLoader
// (lab/tutorial_1a/hfl_complete.py:19-24).
#include "common.cuh"
#include "ddl25_kernels.h"

namespace {

__global__ void __launch_bounds__(256) u8_normalize_kernel(const uchar4* __restrict__ in, float4* __restrict__ out, int64_t n4, float mean,
                                                           float std) {
  pdl_launch_dependents();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const uchar4 u = in[i];
    float4 o;
    o.x = __fdiv_rn(__fsub_rn(__fdiv_rn((float)u.x, 255.f), mean), std);
    o.y = __fdiv_rn(__fsub_rn(__fdiv_rn((float)u.y, 255.f), mean), std);
    o.z = __fdiv_rn(__fsub_rn(__fdiv_rn((float)u.z, 255.f), mean), std);
    o.w = __fdiv_rn(__fsub_rn(__fdiv_rn((float)u.w, 255.f), mean), std);
    out[i] = o;
  }
}

}  // namespace

extern "C" int ddl_u8_normalize(const unsigned char* in, float* out, int64_t n, float mean, float std, void* stream) {
  DDL_REQUIRE(n > 0 && n % 4 == 0, "u8_normalize: n must be a positive multiple of 4");
  int64_t blocks = (n / 4 + 255) / 256;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  DDL_CUDA_CHECK(ddl_launch_pdl(u8_normalize_kernel, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream,
                                reinterpret_cast<const uchar4*>(in), reinterpret_cast<float4*>(out), n / 4, mean, std));
  DDL_LAUNCH_CHECK();
  return DDL_OK;
}
