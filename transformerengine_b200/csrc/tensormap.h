// Host-side TMA descriptor helper shared by the GEMM and the NVFP4 quantizer.
#pragma once

#include <cuda.h>

#include <cstdint>

namespace teb {
namespace gemm {
// Byte tensor [outer, inner_bytes] (inner contiguous), box [box_outer, box_inner_bytes],
// 128-byte swizzle, OOB zero fill. Descriptors are cached by (pointer, shape, box).
CUtensorMap tensor_map_bytes_2d(const void *ptr, uint64_t inner_bytes, uint64_t outer,
                                uint32_t box_inner_bytes, uint32_t box_outer);
CUtensorMap tensor_map_bytes_2d_sw64(const void *ptr, uint64_t inner_bytes, uint64_t outer,
                                     uint32_t box_inner_bytes, uint32_t box_outer);
}  // namespace gemm
}  // namespace teb
