// Gram matrix G = U U^T (the pairwise-distance step of Krum) for callers that must not allocate.
//
// The regular entry point, ddl_gemm_tcgen05(..., n_split, scratch), needs the caller to hand in a
// scratch buffer for the split-K partial tiles.  Some users of this library cannot produce one at
// call time: a plain C/C++ host that links libddl25_kernels.so without PyTorch has no caching
// allocator, and calling cudaMalloc from inside its round loop would stall the whole device
// (cudaMalloc/cudaFree synchronise) and can fail late when HBM is already planned out; an
// aggregation service with a fixed start-up memory budget is in the same position.  For them this
// file owns the scratch itself: a fixed-size array in device global memory that is part of the
// library image, so nothing is allocated or freed here, ever - not by cudaMalloc, not by torch.
//
// Cost of that convenience: every GPU that loads the library carries the array
// (DDL_GRAM_STATIC_SCRATCH_FLOATS * 4 bytes = 4 MiB) until the process exits, and the problem size
// is capped (n_split * K * K floats must fit).
//
// Sharing rules.  There is exactly ONE array per device and every call on that device uses it:
//  * two calls whose kernels can overlap on the GPU (different streams, or different host threads
//    on different streams) race on it and corrupt each other's result - serialise them on one
//    stream or with events;
//  * the array keeps whatever the previous call left behind.  Nothing in it has to start at zero:
//    each split's partial tile is written completely (all K x K entries) by the GEMM kernel before
//    the ordered reduction reads it, and entries beyond n_split * K * K are never read.
//
// Set-up.  Host code cannot use the array's name as a pointer (in host code the symbol is only a
// stand-in; a kernel given it would fault or touch the wrong memory), so the first call on each
// device resolves the real device address with cudaGetSymbolAddress while that device is current
// and caches it (per device ordinal, for the life of the process: the cached address does not survive a
// cudaDeviceReset).  That first call - which also triggers the usual one-time CUDA module load and
// function-attribute set-up - should be made before any stream capture is begun on that device.
//
// This entry point has no counterpart in the reference (the reference contains no native code and
// no Krum); it is a variant of ddl_gemm_tcgen05 and changes none of the existing code paths.
#include "common.cuh"
#include "ddl25_kernels.h"
#include <mutex>

namespace {

static __device__ __align__(256) float g_gram_scratch[DDL_GRAM_STATIC_SCRATCH_FLOATS];

constexpr int kMaxDevices = 64;
float* g_scratch_ptr[kMaxDevices] = {nullptr};
std::mutex g_setup_mutex;

// Device address of the static scratch on the CURRENT device (resolved once per device).
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  DDL_CUDA_CHECK(cudaGetDevice(&dev));
  DDL_REQUIRE(dev >= 0 && dev < kMaxDevices, "device index %d outside the supported range [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (g_scratch_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    DDL_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_gram_scratch));
    DDL_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch");
    g_scratch_ptr[dev] = ptr;
  }
  *out = g_scratch_ptr[dev];
  return DDL_OK;
}

}  // namespace

extern "C" int ddl_gram_static_scratch(const float* updates, int64_t ld, int K, int64_t D, float* gram, int64_t ldg,
                                       int n_split, int precise, void* stream) {
  DDL_REQUIRE(updates != nullptr && gram != nullptr, "updates and gram must be device pointers supplied by the caller");
  DDL_REQUIRE(K >= 1 && K <= 128, "K=%d outside [1, 128] (one 128-row tile of clients)", K);
  DDL_REQUIRE(D >= 4 && D <= 0x7fffffff - 64 && D % 4 == 0, "D=%lld must be a multiple of 4 in [4, 2^31 - 64)", (long long)D);
  DDL_REQUIRE((ld * 4) % 16 == 0 && (reinterpret_cast<uintptr_t>(updates) & 15) == 0, "updates must be 16-byte aligned with a 16-byte row stride");
  DDL_REQUIRE(n_split >= 2, "n_split must be >= 2 (with one split no scratch is needed: call ddl_gemm_tcgen05)");
  const int64_t need = (int64_t)n_split * K * K;
  DDL_REQUIRE(need <= (int64_t)DDL_GRAM_STATIC_SCRATCH_FLOATS,
              "n_split * K * K = %lld floats exceeds the built-in scratch of %d floats (4 MiB): lower n_split or use "
              "ddl_gemm_tcgen05 with a caller-provided scratch buffer",
              (long long)need, DDL_GRAM_STATIC_SCRATCH_FLOATS);
  float* scratch = nullptr;
  if (int e = static_scratch_for_current_device(&scratch)) return e;
  // A = B = updates (K-major), C = gram; the partial tiles land in the static array.
  return ddl_gemm_tcgen05(updates, ld, 0, 0, updates, ld, 0, 0, gram, ldg, 0, nullptr, 0, /*act=*/0, 0.f, /*alpha=*/1.f,
                          /*beta=*/0.f, K, K, (int)D, /*G=*/1, /*dtype_bytes=*/4, precise, n_split, scratch, stream);
}
