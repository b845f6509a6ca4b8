// Host helper: build a CUtensorMap (TMA descriptor) without linking libcuda directly – the
// driver entry point is resolved through the runtime (cudaGetDriverEntryPoint).
#pragma once
#include <cuda.h>
#include <cstdint>

// dims/box are innermost-first; strides_bytes has rank-1 entries (stride of dim 1..rank-1).
// swizzle: 0 none, 1 = CU_TENSOR_MAP_SWIZZLE_128B, 2 = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
// (box[0]*elem must be <= 128 bytes when swizzling).
int ddl_encode_tmap(CUtensorMap* out, CUtensorMapDataType dtype, int rank, const void* base, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, int swizzle);
