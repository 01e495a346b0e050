/* TEST INFRASTRUCTURE — CPU oracle, never shipped or timed as product.
 *
 * Bit-level number-format conversions used by the oracle. Everything here is a
 * plain-C restatement of published formats (OCP FP8 E4M3/E5M2, E2M1, E8M0,
 * bfloat16, IEEE half) with the rounding/saturation behaviour the reference
 * relies on:
 *   - fp8/fp4 casts: round-to-nearest-even, saturate-to-finite
 *     (reference device path: util/ptx.cuh cvt.rn.satfinite.*; host twin:
 *     tests/cpp/test_common.h static_cast<fp8e4m3>(float)).
 *   - E8M0 encode: round towards +inf with the reference's special cases
 *     (tests/cpp/test_common.h:480-500, util/ptx.cuh:338-400).
 *   - E8M0 reciprocal: tests/cpp/test_common.h:502-513.
 */
#ifndef TEO_CONVERT_H_
#define TEO_CONVERT_H_

#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

static inline uint32_t teo_f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float teo_u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* ---- bfloat16 ---------------------------------------------------------- */
static inline float teo_bf16_to_f32(uint16_t h) { return teo_u2f((uint32_t)h << 16); }
static inline uint16_t teo_f32_to_bf16(float f) {
  uint32_t u = teo_f2u(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fff; /* NaN -> canonical */
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

/* ---- IEEE half --------------------------------------------------------- */
static inline float teo_f16_to_f32(uint16_t h) {
  uint32_t s = (uint32_t)(h & 0x8000u) << 16;
  uint32_t e = (h >> 10) & 0x1fu, m = h & 0x3ffu;
  if (e == 0) {
    if (m == 0) return teo_u2f(s);
    float v = ldexpf((float)m, -24);
    return (s ? -v : v);
  }
  if (e == 31) return teo_u2f(s | 0x7f800000u | (m << 13));
  return teo_u2f(s | ((e + 112u) << 23) | (m << 13));
}

/* Generic float -> small float (1 sign, E exponent bits, M mantissa bits),
 * round-to-nearest-even, saturating to +-max_code. Returns magnitude code
 * (without sign) in the low bits; caller adds the sign. `min_norm_exp` is the
 * unbiased exponent of the smallest normal. */
static inline unsigned teo_round_small(float ax, int mbits, int bias, unsigned max_code) {
  /* ax is finite, >= 0 */
  if (ax == 0.0f) return 0;
  int e;
  float fr = frexpf(ax, &e); /* ax = fr * 2^e, fr in [0.5,1) */
  int ue = e - 1;            /* unbiased exponent, ax = 1.x * 2^ue */
  int min_norm = 1 - bias;
  int qexp = ue < min_norm ? min_norm : ue; /* exponent of the quantum grid */
  /* quantum = 2^(qexp - mbits) ; n = ax / quantum, exact in double */
  double n = ldexp((double)ax, mbits - qexp);
  double fl = floor(n);
  double rem = n - fl;
  uint64_t k = (uint64_t)fl;
  if (rem > 0.5 || (rem == 0.5 && (k & 1u))) k += 1;
  (void)fr;
  /* k counts quanta; in the normal range k in [2^mbits, 2^(mbits+1)] */
  unsigned code;
  if (ue < min_norm) {
    code = (unsigned)k; /* subnormal (k may reach 2^mbits = smallest normal) */
  } else {
    unsigned be = (unsigned)(qexp + bias);
    code = (be << mbits) + (unsigned)(k - ((uint64_t)1 << mbits));
  }
  if (code > max_code) code = max_code;
  return code;
}

/* float -> FP8 E4M3 (OCP "fn": no inf, max 448 = 0x7E, NaN = 0x7F) */
static inline uint8_t teo_f32_to_e4m3(float f) {
  uint32_t u = teo_f2u(f);
  uint8_t sign = (uint8_t)((u >> 24) & 0x80u);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7f;              /* NaN */
  float ax = fabsf(f);
  if (isinf(ax) || ax > 464.0f) return (uint8_t)(sign | 0x7e);   /* satfinite */
  return (uint8_t)(sign | teo_round_small(ax, 3, 7, 0x7e));
}
static inline float teo_e4m3_to_f32(uint8_t b) {
  unsigned e = (b >> 3) & 0xf, m = b & 7;
  float v;
  if (e == 0xf && m == 7) v = NAN;
  else if (e == 0) v = ldexpf((float)m, -9);
  else v = ldexpf((float)(8 + m), (int)e - 10);
  return (b & 0x80) ? -v : v;
}

/* float -> FP8 E5M2 (IEEE-like; satfinite max 57344 = 0x7B) */
static inline uint8_t teo_f32_to_e5m2(float f) {
  uint32_t u = teo_f2u(f);
  uint8_t sign = (uint8_t)((u >> 24) & 0x80u);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint8_t)(sign | 0x7f);
  float ax = fabsf(f);
  if (isinf(ax) || ax >= 61440.0f) return (uint8_t)(sign | 0x7b);
  return (uint8_t)(sign | teo_round_small(ax, 2, 15, 0x7b));
}
static inline float teo_e5m2_to_f32(uint8_t b) {
  unsigned e = (b >> 2) & 0x1f, m = b & 3;
  float v;
  if (e == 0x1f) v = m ? NAN : INFINITY;
  else if (e == 0) v = ldexpf((float)m, -16);
  else v = ldexpf((float)(4 + m), (int)e - 17);
  return (b & 0x80) ? -v : v;
}

/* float -> FP4 E2M1 nibble; grid {0,.5,1,1.5,2,3,4,6}; NaN -> +6 (0x7) */
static inline uint8_t teo_f32_to_e2m1(float f) {
  uint32_t u = teo_f2u(f);
  uint8_t sign = (uint8_t)((u >> 28) & 0x8u);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7;
  float ax = fabsf(f);
  if (isinf(ax) || ax > 6.0f) return (uint8_t)(sign | 0x7);
  return (uint8_t)(sign | teo_round_small(ax, 1, 1, 0x7));
}
static inline float teo_e2m1_to_f32(uint8_t nib) {
  static const float tbl[8] = {0.f, .5f, 1.f, 1.5f, 2.f, 3.f, 4.f, 6.f};
  float v = tbl[nib & 7];
  return (nib & 8) ? -v : v;
}

/* ---- E8M0 (tests/cpp/test_common.h:480-513) ---------------------------- */
static inline uint8_t teo_f32_to_e8m0(float v) {
  if (isnan(v)) return 0xff;
  if (isinf(v)) return 0xfe;
  if (v == 0.0f) return 0x00;
  uint32_t u = teo_f2u(v);
  uint8_t e = (uint8_t)(u >> 23);
  uint32_t m = u & 0x7fffffu;
  if ((m > 0 && e != 0xfe) && !(e == 0 && m <= 0x400000u)) ++e;
  return e;
}
static inline float teo_e8m0_rcp(uint8_t e) {
  if (e == 255) return teo_u2f(0x7fffffffu);
  if (e == 254) return teo_u2f(0x00400000u);
  return teo_u2f((uint32_t)(254 - e) << 23);
}
static inline float teo_e8m0_to_f32(uint8_t e) {
  if (e == 255) return NAN;
  if (e == 0) return teo_u2f(0x00400000u); /* 2^-127 */
  return teo_u2f((uint32_t)e << 23);
}

/* ---- typed loads: dtype 0 = fp32, 1 = bf16, 2 = fp16 ------------------- */
enum { TEO_F32 = 0, TEO_BF16 = 1, TEO_F16 = 2 };
static inline float teo_load(const void *p, int dtype, size_t i) {
  switch (dtype) {
    case TEO_F32: return ((const float *)p)[i];
    case TEO_BF16: return teo_bf16_to_f32(((const uint16_t *)p)[i]);
    default: return teo_f16_to_f32(((const uint16_t *)p)[i]);
  }
}
static inline uint8_t teo_to_fp8(float v, int kind) {
  return kind == 0 ? teo_f32_to_e4m3(v) : teo_f32_to_e5m2(v);
}
static inline float teo_from_fp8(uint8_t b, int kind) {
  return kind == 0 ? teo_e4m3_to_f32(b) : teo_e5m2_to_f32(b);
}
static inline float teo_fp8_max(int kind) { return kind == 0 ? 448.0f : 57344.0f; }

#endif /* TEO_CONVERT_H_ */
