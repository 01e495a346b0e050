// sm_100a PTX helpers shared by the quantize kernels.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

namespace teb {
namespace ptx {

// ---- E8M0 ----------------------------------------------------------------
// float -> E8M0 biased exponent, rounding towards +inf, saturating to 0xFE
// (NaN -> 0xFF). Same semantics as the reference's float_to_e8m0
// (reference util/ptx.cuh:338-400), done by the sm_100 conversion instruction.
__device__ __forceinline__ uint32_t float_to_e8m0(float v) {
  uint16_t out;
  asm("cvt.rp.satfinite.ue8m0x2.f32 %0, 0.0, %1;" : "=h"(out) : "f"(v));
  return out & 0xffu;
}
// 2^-(e-127) as fp32 bits (reference util/ptx.cuh exp2f_rcp).
__device__ __forceinline__ float e8m0_rcp(uint32_t e) {
  const uint32_t bits = (e == 255u) ? 0x7fffffffu : (e == 254u) ? 0x00400000u : ((254u - e) << 23);
  return __uint_as_float(bits);
}
__device__ __forceinline__ float e8m0_to_float(uint32_t e) {
  const uint32_t bits = (e == 255u) ? 0x7fffffffu : (e == 0u) ? 0x00400000u : (e << 23);
  return __uint_as_float(bits);
}

// ---- packed scale + FP8 conversion ---------------------------------------
// (x0*s, x1*s) -> two FP8 bytes, x0 in the low byte. RN, saturate-to-finite.
template <bool E5M2>
__device__ __forceinline__ uint32_t mul_cvt_fp8x2(float x0, float x1, float s) {
  uint16_t out;
  if constexpr (E5M2) {
    asm("{\n\t.reg .b64 p, q;\n\t.reg .f32 a, b;\n\t"
        "mov.b64 p, {%1, %2};\n\tmov.b64 q, {%3, %3};\n\t"
        "mul.f32x2 p, p, q;\n\tmov.b64 {a, b}, p;\n\t"
        "cvt.rn.satfinite.e5m2x2.f32 %0, b, a;\n\t}"
        : "=h"(out) : "f"(x0), "f"(x1), "f"(s));
  } else {
    asm("{\n\t.reg .b64 p, q;\n\t.reg .f32 a, b;\n\t"
        "mov.b64 p, {%1, %2};\n\tmov.b64 q, {%3, %3};\n\t"
        "mul.f32x2 p, p, q;\n\tmov.b64 {a, b}, p;\n\t"
        "cvt.rn.satfinite.e4m3x2.f32 %0, b, a;\n\t}"
        : "=h"(out) : "f"(x0), "f"(x1), "f"(s));
  }
  return out;
}
// (x0*s0, x1*s1) -> two FP8 bytes, x0 in the low byte.
template <bool E5M2>
__device__ __forceinline__ uint32_t mul2_cvt_fp8x2(float x0, float x1, float s0, float s1) {
  uint16_t out;
  if constexpr (E5M2) {
    asm("{\n\t.reg .b64 p, q;\n\t.reg .f32 a, b;\n\t"
        "mov.b64 p, {%1, %2};\n\tmov.b64 q, {%3, %4};\n\t"
        "mul.f32x2 p, p, q;\n\tmov.b64 {a, b}, p;\n\t"
        "cvt.rn.satfinite.e5m2x2.f32 %0, b, a;\n\t}"
        : "=h"(out) : "f"(x0), "f"(x1), "f"(s0), "f"(s1));
  } else {
    asm("{\n\t.reg .b64 p, q;\n\t.reg .f32 a, b;\n\t"
        "mov.b64 p, {%1, %2};\n\tmov.b64 q, {%3, %4};\n\t"
        "mul.f32x2 p, p, q;\n\tmov.b64 {a, b}, p;\n\t"
        "cvt.rn.satfinite.e4m3x2.f32 %0, b, a;\n\t}"
        : "=h"(out) : "f"(x0), "f"(x1), "f"(s0), "f"(s1));
  }
  return out;
}
template <bool E5M2>
__device__ __forceinline__ uint32_t cvt_fp8x2(float x0, float x1) {
  uint16_t out;
  if constexpr (E5M2) {
    asm("cvt.rn.satfinite.e5m2x2.f32 %0, %2, %1;" : "=h"(out) : "f"(x0), "f"(x1));
  } else {
    asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(out) : "f"(x0), "f"(x1));
  }
  return out;
}
// two FP8 bytes (low byte first) -> two floats
template <bool E5M2>
__device__ __forceinline__ float2 fp8x2_to_float2(uint32_t v) {
  uint32_t h2;
  const uint16_t in = static_cast<uint16_t>(v);
  if constexpr (E5M2) {
    asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(h2) : "h"(in));
  } else {
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(in));
  }
  return __half22float2(*reinterpret_cast<__half2 *>(&h2));
}

// (x0, x1) -> one byte holding two E2M1 nibbles, x0 in the low nibble.
__device__ __forceinline__ uint32_t cvt_e2m1x2(float x0, float x1) {
  uint32_t out;
  asm("{\n\t.reg .b8 b;\n\tcvt.rn.satfinite.e2m1x2.f32 b, %2, %1;\n\tcvt.u32.u8 %0, b;\n\t}"
      : "=r"(out) : "f"(x0), "f"(x1));
  return out;
}

// ---- packed abs/max on raw register words ---------------------------------
// One 32-bit word holds either one fp32 or two 16-bit floats. abs is a mask;
// max runs on the non-negative values.
template <typename T>
struct Word;
template <>
struct Word<float> {
  static constexpr int kElems = 1;
  static __device__ __forceinline__ uint32_t abs(uint32_t v) { return v & 0x7fffffffu; }
  static __device__ __forceinline__ uint32_t max(uint32_t a, uint32_t b) {
    return __float_as_uint(fmaxf(__uint_as_float(a), __uint_as_float(b)));
  }
  static __device__ __forceinline__ float hmax(uint32_t a) { return __uint_as_float(a); }
  static __device__ __forceinline__ float lo(uint32_t v) { return __uint_as_float(v); }
  static __device__ __forceinline__ float hi(uint32_t v) { return __uint_as_float(v); }
};
template <>
struct Word<__nv_bfloat16> {
  static constexpr int kElems = 2;
  static __device__ __forceinline__ uint32_t abs(uint32_t v) { return v & 0x7fff7fffu; }
  static __device__ __forceinline__ uint32_t max(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  static __device__ __forceinline__ float lo(uint32_t v) { return __uint_as_float(v << 16); }
  static __device__ __forceinline__ float hi(uint32_t v) {
    return __uint_as_float(v & 0xffff0000u);
  }
  static __device__ __forceinline__ float hmax(uint32_t a) { return fmaxf(lo(a), hi(a)); }
};
template <>
struct Word<__half> {
  static constexpr int kElems = 2;
  static __device__ __forceinline__ uint32_t abs(uint32_t v) { return v & 0x7fff7fffu; }
  static __device__ __forceinline__ uint32_t max(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
  }
  static __device__ __forceinline__ float lo(uint32_t v) {
    return __half2float(__ushort_as_half(static_cast<unsigned short>(v & 0xffffu)));
  }
  static __device__ __forceinline__ float hi(uint32_t v) {
    return __half2float(__ushort_as_half(static_cast<unsigned short>(v >> 16)));
  }
  static __device__ __forceinline__ float hmax(uint32_t a) { return fmaxf(lo(a), hi(a)); }
};

// ---- global memory ---------------------------------------------------------
// Streaming 8/16-byte loads (read once: do not allocate in L1).
__device__ __forceinline__ uint2 ld_stream_v2(const void *p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];"
               : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_stream_v4(const void *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// Non-negative floats order like their bit patterns: atomic max on the bits.
__device__ __forceinline__ void atomic_max_nonneg(float *addr, float v) {
  atomicMax(reinterpret_cast<int *>(addr), __float_as_int(v));
}

// ---- flags in peer-mapped (NVLink / symmetric) memory ------------------------
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t ld_acquire_sys_u64(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t *p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_release_sys_u64(uint64_t *p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add_u32(uint32_t *p, uint32_t v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// NVSwitch multicast (NVLS): one store / reduction to a multicast address reaches every GPU of the group
__device__ __forceinline__ void multimem_st_v4(void *mc, const uint4 &v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ void multimem_red_release_add_u32(uint32_t *mc, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Spin until *flag == want. A peer that never arrives must not hang the GPU: after 5 s the error word is
// set and the wait gives up (the result is then wrong, and the host side reports it).
__device__ __forceinline__ bool wait_flag_u32(const uint32_t *flag, uint32_t want, uint32_t *error) {
  if (ld_acquire_sys_u32(flag) == want) return true;
  const uint64_t t0 = globaltimer_ns();
  while (ld_acquire_sys_u32(flag) != want) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > 5000000000ull) {
      if (error != nullptr) *error = 1u;
      return false;
    }
  }
  return true;
}

// The same for a counter that only grows (modulo 2^32): until it has reached `want`.
__device__ __forceinline__ bool wait_count_u32(const uint32_t *flag, uint32_t want, uint32_t *error) {
  if (static_cast<int32_t>(ld_acquire_sys_u32(flag) - want) >= 0) return true;
  const uint64_t t0 = globaltimer_ns();
  while (static_cast<int32_t>(ld_acquire_sys_u32(flag) - want) < 0) {
    __nanosleep(64);
    if (globaltimer_ns() - t0 > 5000000000ull) {
      if (error != nullptr) *error = 1u;
      return false;
    }
  }
  return true;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace ptx
}  // namespace teb
