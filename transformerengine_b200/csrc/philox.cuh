// Counter-based random bits for NVFP4 stochastic rounding (reference cast/nvfp4/core_nvfp4.cuh:97-110
// draws them from a Philox4x32 state seeded by the quantization config's rng_state = {seed, offset}).
// One Philox4x32 block = 4 x 32 bits = the random bits of one 1x16 NVFP4 block
// (cvt.rs.satfinite.e2m1x4.f32 takes one 32-bit word per 4 values). The counter is the block's
// position in the tensor, so the stream does not depend on how the kernel tiles the tensor; like
// the reference's, the result is only specified statistically (its tests check the mean error).
#pragma once

#include <cstdint>

namespace teb {
namespace philox {

constexpr int kRounds = 7;   // Random123: Philox4x32-7 passes BigCrush; curand's default is 10

__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// random bits of block `block` (a 64-bit linear index) in direction `dir` (0 rowwise, 1 columnwise)
__device__ __forceinline__ uint4 block_bits(uint64_t seed, uint64_t offset, uint64_t block, uint32_t dir) {
  const uint4 c = make_uint4(static_cast<uint32_t>(block), static_cast<uint32_t>(block >> 32),
                             static_cast<uint32_t>(offset) ^ (dir * 0x85EBCA6Bu), static_cast<uint32_t>(offset >> 32));
  return philox4x32(c, make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32)));
}

// four floats -> four E2M1 nibbles (value 0 lowest), stochastic rounding with 32 random bits
__device__ __forceinline__ uint32_t cvt_rs_e2m1x4(float f0, float f1, float f2, float f3, uint32_t rbits) {
  uint16_t h;
  asm("cvt.rs.satfinite.e2m1x4.f32 %0, {%4, %3, %2, %1}, %5;" : "=h"(h) : "f"(f0), "f"(f1), "f"(f2), "f"(f3), "r"(rbits));
  return h;
}

}  // namespace philox
}  // namespace teb
