/* TEST INFRASTRUCTURE — CPU oracle, never shipped or timed as product
 * (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg use it).
 *
 * Plain-C restatement of the reference's quantizers for the low-precision
 * Linear hot path. Each function cites the reference code it follows
 * (paths relative to the reference checkout).
 *
 * Parity pin: see oracle/README.md — checked against the reference's own host
 * compute_ref functions compiled in place (oracle/_ref) and against the
 * reference's pure-torch NVFP4 quantizer (tests/golden/).
 */
#include "teo_convert.h"

#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TEO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------ *
 * amax  (recipe/current_scaling.cu:32-70 — max |x| over the tensor)
 * ------------------------------------------------------------------------ */
TEO_API float teo_amax(const void *in, int dtype, size_t n) {
  float m = 0.0f;
#pragma omp parallel for reduction(max : m) if (n > (1u << 16))
  for (size_t i = 0; i < n; ++i) {
    float a = fabsf(teo_load(in, dtype, i));
    if (a > m) m = a; /* NaN is ignored like fmaxf */
  }
  return m;
}

/* ------------------------------------------------------------------------ *
 * scale from amax  (recipe/recipe_common.cuh:14-50;
 * tests/pytorch/references/quantize_scale_calc.py:9-60)
 * ------------------------------------------------------------------------ */
TEO_API float teo_scale_from_amax(float amax, float max_fp8, int force_pow2, float eps) {
  if (amax < eps) amax = eps;
  if (isinf(amax) || amax == 0.0f || isnan(amax)) return 1.0f;
  float scale = max_fp8 / amax; /* IEEE RN divide (== __fdiv_rn) */
  if (isinf(scale)) scale = 3.402823466e+38f;
  if (force_pow2) scale = teo_u2f(teo_f2u(scale) & 0xFF800000u);
  return scale;
}

/* ------------------------------------------------------------------------ *
 * per-tensor FP8 cast  (cast/fp8/quantize_fp8.cuh:305-348;
 * tests/cpp/operator/test_cast_transpose.cu:24-40)
 *   out = satfinite_rn(x * scale); amax = max(amax_in, |x|); scale_inv = 1/scale
 * out_t (optional) receives the transposed copy [cols, rows].
 * ------------------------------------------------------------------------ */
TEO_API void teo_fp8_quantize(const void *in, int dtype, size_t rows, size_t cols, float scale,
                              int kind, uint8_t *out, uint8_t *out_t, float *amax,
                              float *scale_inv) {
  float m = amax ? *amax : 0.0f;
#pragma omp parallel for reduction(max : m) if (rows * cols > (1u << 16))
  for (size_t i = 0; i < rows; ++i) {
    for (size_t j = 0; j < cols; ++j) {
      float x = teo_load(in, dtype, i * cols + j);
      float a = fabsf(x);
      if (a > m) m = a;
      uint8_t q = teo_to_fp8(x * scale, kind);
      if (out) out[i * cols + j] = q;
      if (out_t) out_t[j * rows + i] = q;
    }
  }
  if (amax) *amax = m;
  if (scale_inv) *scale_inv = 1.0f / scale; /* __frcp_rn */
}

/* ------------------------------------------------------------------------ *
 * delayed-scaling update  (recipe/delayed_scaling.cu:152-253;
 * closed form in tests/pytorch/test_recipe.py:110-180)
 * history: [hist_len, num_scales] row-major, updated in place.
 * reduced_amax (optional): replaces history[0][*] before the roll when
 * non-NULL (the "after_reduction" variant).
 * algo: 0 = max, 1 = most_recent.
 * ------------------------------------------------------------------------ */
TEO_API void teo_delayed_scaling_update(float *history, float *scale, const float *reduced_amax,
                                        size_t hist_len, size_t num_scales, int algo,
                                        float fp8_max, float margin, int bulk) {
  /* bulk != 0 follows kernel_bulk (delayed_scaling.cu:152-253): the reduced
   * amax replaces history[0] when non-zero, and a zero most-recent amax skips
   * the roll entirely. bulk == 0 follows the single-tensor kernel (:77-150). */
  const float scaled_max = fp8_max * powf(2.0f, -margin);
  for (size_t s = 0; s < num_scales; ++s) {
    float last = history[s];
    if (bulk && reduced_amax && reduced_amax[s] != 0.0f) last = reduced_amax[s];
    float amax = 0.0f;
    if (!bulk || last != 0.0f) {
      for (size_t i = 0; i < hist_len; ++i) {
        float a = (i + 1 < hist_len) ? history[(i + 1) * num_scales + s] : last;
        amax = fmaxf(amax, a);
        history[i * num_scales + s] = (i > 0) ? a : 0.0f;
      }
    }
    if (algo == 1) amax = last;
    float sc = scale[s];
    if (isfinite(amax) && amax > 0.0f) sc = scaled_max / amax;
    if (isinf(sc)) sc = 3.402823466e+38f;
    scale[s] = sc;
  }
}

/* ------------------------------------------------------------------------ *
 * MXFP8  (cast/mxfp8/quantize_mxfp8.cuh:211-461;
 * host twin tests/cpp/operator/test_cast_mxfp8.cu:39-166)
 * Rowwise: 1x32 blocks along cols, scales [rows, stride_row] (E8M0 bytes).
 * Colwise: 32x1 blocks along rows, scales [ceil(rows/32), stride_col];
 *          colwise data is NOT transposed.
 * Either output set may be NULL.
 * ------------------------------------------------------------------------ */
TEO_API void teo_mxfp8_quantize(const void *in, int dtype, size_t rows, size_t cols, int kind,
                                uint8_t *out_row, uint8_t *sc_row, size_t stride_row,
                                uint8_t *out_col, uint8_t *sc_col, size_t stride_col,
                                float *amax) {
  const float max_rcp = (float)(1.0 / (double)teo_fp8_max(kind));
  const size_t ty = (rows + 31) / 32, tx = (cols + 31) / 32;
  float gm = 0.0f;
#pragma omp parallel for schedule(static) reduction(max : gm)
  for (size_t t = 0; t < ty * tx; ++t) {
    const size_t i0 = (t / tx) * 32, j0 = (t % tx) * 32;
    const size_t i1 = i0 + 32 < rows ? i0 + 32 : rows;
    const size_t j1 = j0 + 32 < cols ? j0 + 32 : cols;
    float tile[32][32];
    for (size_t i = i0; i < i1; ++i)
      for (size_t j = j0; j < j1; ++j) tile[i - i0][j - j0] = teo_load(in, dtype, i * cols + j);
    if (out_row) {
      for (size_t i = i0; i < i1; ++i) {
        float bm = 0.0f;
        for (size_t j = j0; j < j1; ++j) bm = fmaxf(bm, fabsf(tile[i - i0][j - j0]));
        gm = fmaxf(gm, bm);
        uint8_t e = teo_f32_to_e8m0(bm * max_rcp);
        sc_row[i * stride_row + j0 / 32] = e;
        float r = teo_e8m0_rcp(e);
        for (size_t j = j0; j < j1; ++j)
          out_row[i * cols + j] = teo_to_fp8(tile[i - i0][j - j0] * r, kind);
      }
    }
    if (out_col) {
      for (size_t j = j0; j < j1; ++j) {
        float bm = 0.0f;
        for (size_t i = i0; i < i1; ++i) bm = fmaxf(bm, fabsf(tile[i - i0][j - j0]));
        gm = fmaxf(gm, bm);
        uint8_t e = teo_f32_to_e8m0(bm * max_rcp);
        sc_col[(i0 / 32) * stride_col + j] = e;
        float r = teo_e8m0_rcp(e);
        for (size_t i = i0; i < i1; ++i)
          out_col[i * cols + j] = teo_to_fp8(tile[i - i0][j - j0] * r, kind);
      }
    }
  }
  if (amax) *amax = fmaxf(*amax, gm);
}

/* MXFP8 dequantize (cast/mxfp8/dequantize_mxfp8.cuh): x = fp8 * 2^(e-127) */
TEO_API void teo_mxfp8_dequantize(const uint8_t *q, const uint8_t *sc, size_t stride, int colwise,
                                  size_t rows, size_t cols, int kind, float *out) {
#pragma omp parallel for
  for (size_t i = 0; i < rows; ++i)
    for (size_t j = 0; j < cols; ++j) {
      uint8_t e = colwise ? sc[(i / 32) * stride + j] : sc[i * stride + j / 32];
      out[i * cols + j] = teo_from_fp8(q[i * cols + j], kind) * teo_e8m0_to_f32(e);
    }
}

/* ------------------------------------------------------------------------ *
 * NVFP4 1-D  (cast/nvfp4/core_nvfp4.cuh:40-95,
 * cast/nvfp4/quantize_transpose_nvfp4.cuh:353-363,466-497,639-713;
 * host twin tests/cpp/operator/test_cast_nvfp4_transpose.cu:65-82,183-284;
 * torch twin pytorch/custom_recipes/quantization_ref_nvfp4.py:655-781)
 * in: [rows, cols]; out: [rows, cols/2] bytes (element 2j in the low nibble);
 * scales: [rows, stride] E4M3 bytes, one per 1x16 block.
 * ------------------------------------------------------------------------ */
TEO_API float teo_nvfp4_global_encode_scale(float global_amax) {
  float s = 448.0f * 6.0f / global_amax;
  s = fminf(s, 3.402823466e+38f);
  if (global_amax == 0.0f || s == 0.0f) return 1.0f;
  return s;
}

/* host_twin != 0 reproduces the C++ host reference's special case (encode scale 0
 * when the E4M3 block scale underflows to 0, test_cast_nvfp4_transpose.cu:262-263);
 * host_twin == 0 is the kernel / pure-torch behaviour (1/0 -> inf -> FLT_MAX,
 * core_nvfp4.cuh + quantization_ref_nvfp4.py:760-770). They differ only for a
 * non-zero block whose scale underflows (global amax ~1e38). */
TEO_API void teo_nvfp4_quantize_1d(const void *in, int dtype, size_t rows, size_t cols,
                                   float global_amax, uint8_t *out, uint8_t *scales,
                                   size_t stride, int host_twin) {
  const float S_enc = teo_nvfp4_global_encode_scale(global_amax);
  const float S_dec = 1.0f / S_enc;
  const float mult = S_enc * (1.0f / 6.0f);
  const size_t nb = cols / 16;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < rows; ++i) {
    for (size_t b = 0; b < nb; ++b) {
      float v[16], bm = 0.0f;
      for (int k = 0; k < 16; ++k) {
        v[k] = teo_load(in, dtype, i * cols + b * 16 + k);
        bm = fmaxf(bm, fabsf(v[k]));
      }
      float sdec = fminf(bm * mult, 3.402823466e+38f);
      uint8_t s8 = teo_f32_to_e4m3(sdec);
      scales[i * stride + b] = s8;
      float sf = teo_e4m3_to_f32(s8);
      float enc = (host_twin && sf == 0.0f) ? 0.0f : fminf(1.0f / (sf * S_dec), 3.402823466e+38f);
      for (int k = 0; k < 16; k += 2) {
        uint8_t lo = teo_f32_to_e2m1(v[k] * enc), hi = teo_f32_to_e2m1(v[k + 1] * enc);
        out[(i * cols + b * 16 + k) / 2] = (uint8_t)(lo | (hi << 4));
      }
    }
  }
}

/* NVFP4 2-D (16x16 tiles, weights): cast/nvfp4/quantize_transpose_nvfp4.cuh:780-1312;
 * host twin tests/cpp/operator/test_cast_nvfp4_transpose.cu:287-440.
 * One E4M3 scale per 16x16 tile, replicated to the tile's 16 scale rows:
 *   S_dec_b = e4m3(block_amax / 6 * S_enc);  enc_b = S_enc / S_dec_b  (0 if S_dec_b == 0)
 * Quantizing the transpose with the same call gives the columnwise output. */
TEO_API void teo_nvfp4_quantize_2d(const void *in, int dtype, size_t rows, size_t cols,
                                   float global_amax, uint8_t *out, uint8_t *scales,
                                   size_t stride, int host_twin) {
  /* host_twin != 0: the C++ host reference's arithmetic (above). host_twin == 0: the
   * device kernel's, which reuses the 1-D scale math on the 16x16 tile amax
   * (quantize_transpose_nvfp4.cuh:1077-1089,1202-1221 + core_nvfp4.cuh:40-58), as does the
   * pure-torch reference (quantization_ref_nvfp4.py:741-770). */
  const float S_enc = teo_nvfp4_global_encode_scale(global_amax);
  const float S_dec = 1.0f / S_enc;
  const size_t by = (rows + 15) / 16, bx = cols / 16;
#pragma omp parallel for schedule(static) collapse(2)
  for (size_t Y = 0; Y < by; ++Y)
    for (size_t X = 0; X < bx; ++X) {
      const size_t i1 = (Y * 16 + 16 < rows) ? Y * 16 + 16 : rows;
      float bm = 0.0f;
      for (size_t i = Y * 16; i < i1; ++i)
        for (size_t j = X * 16; j < X * 16 + 16; ++j)
          bm = fmaxf(bm, fabsf(teo_load(in, dtype, i * cols + j)));
      const uint8_t s8 = host_twin ? teo_f32_to_e4m3(bm / 6.0f * S_enc)
                                   : teo_f32_to_e4m3(fminf(bm * (S_enc * (1.0f / 6.0f)), 3.402823466e+38f));
      const float sf = teo_e4m3_to_f32(s8);
      const float enc = host_twin ? (sf == 0.0f ? 0.0f : S_enc / sf)
                                  : fminf(1.0f / (sf * S_dec), 3.402823466e+38f);
      for (size_t i = Y * 16; i < i1; ++i) {
        if (scales) scales[i * stride + X] = s8;
        for (size_t j = X * 16; j < X * 16 + 16; j += 2) {
          uint8_t lo = teo_f32_to_e2m1(teo_load(in, dtype, i * cols + j) * enc);
          uint8_t hi = teo_f32_to_e2m1(teo_load(in, dtype, i * cols + j + 1) * enc);
          out[(i * cols + j) / 2] = (uint8_t)(lo | (hi << 4));
        }
      }
    }
}

/* NVFP4 dequantize (cast/nvfp4/dequantize_nvfp4.cuh):
 * x = e2m1 * e4m3(block scale) * (global_amax / (6*448)) */
TEO_API void teo_nvfp4_dequantize(const uint8_t *q, const uint8_t *scales, size_t stride,
                                  size_t rows, size_t cols, float global_amax, float *out) {
  const float factor_inv = 1.0f / (6.0f * 448.0f);
#pragma omp parallel for
  for (size_t i = 0; i < rows; ++i)
    for (size_t j = 0; j < cols; ++j) {
      uint8_t byte = q[(i * cols + j) / 2];
      uint8_t nib = (j & 1) ? (byte >> 4) : (byte & 0xf);
      out[i * cols + j] =
          teo_e2m1_to_f32(nib) * (teo_e4m3_to_f32(scales[i * stride + j / 16]) * global_amax * factor_inv);
    }
}

/* ------------------------------------------------------------------------ *
 * transpose helper and random Hadamard transform of the transpose
 * (hadamard_transform/hadamard_transform_utils.cuh:22,98-176;
 *  pytorch/tensor/nvfp4_tensor.py:48-110;
 *  torch twin quantization_ref_nvfp4.py:399-451)
 * out[c, g*16 + j] = bf16( 0.25 * sum_i x[g*16+i, c] * s_i * H16[i][j] )
 * with s_i = -1 where bit i of sign_mask is set, H16 Sylvester
 * (H[i][j] = (-1)^popcount(i & j)). The sum is formed exactly (double) and
 * rounded once to bf16.
 * ------------------------------------------------------------------------ */
TEO_API void teo_transpose_bf16(const uint16_t *in, size_t rows, size_t cols, uint16_t *out) {
#pragma omp parallel for
  for (size_t j = 0; j < cols; ++j)
    for (size_t i = 0; i < rows; ++i) out[j * rows + i] = in[i * cols + j];
}

TEO_API void teo_rht_transposed_bf16(const uint16_t *in, size_t rows, size_t cols,
                                     unsigned sign_mask, uint16_t *out) {
#pragma omp parallel for schedule(static)
  for (size_t c = 0; c < cols; ++c) {
    for (size_t g = 0; g < rows / 16; ++g) {
      double x[16];
      for (int i = 0; i < 16; ++i) {
        double v = (double)teo_bf16_to_f32(in[(g * 16 + i) * cols + c]);
        x[i] = ((sign_mask >> i) & 1u) ? -v : v;
      }
      for (int j = 0; j < 16; ++j) {
        double acc = 0.0;
        for (int i = 0; i < 16; ++i)
          acc += (__builtin_popcount((unsigned)(i & j)) & 1) ? -x[i] : x[i];
        out[c * rows + g * 16 + j] = teo_f32_to_bf16((float)(acc * 0.25));
      }
    }
  }
}

/* ------------------------------------------------------------------------ *
 * GEMM-swizzled scale layout  (cast/mxfp8/swizzle.cuh:30-42;
 * host twin tests/cpp/operator/test_swizzle.cu:30-57;
 * torch twin tests/pytorch/mxfp8/mxfp8_utils.py:25-62)
 * Logical scale matrix [M, K] (M, K padded to 128, 4). Source element (m, k)
 * lives at in[m*ld + k] (rowwise) or in[k*ld + m] (transposed source, the
 * MXFP8 columnwise case).
 * ------------------------------------------------------------------------ */
TEO_API size_t teo_swizzle_index(size_t m, size_t k, size_t K) {
  return ((m / 128) * (K / 4) + k / 4) * 512 + (m % 32) * 16 + ((m % 128) / 32) * 4 + (k % 4);
}
TEO_API void teo_swizzle_scales(const uint8_t *in, size_t ld, int transposed_src, size_t M,
                                size_t K, uint8_t *out) {
  for (size_t m = 0; m < M; ++m)
    for (size_t k = 0; k < K; ++k)
      out[teo_swizzle_index(m, k, K)] = transposed_src ? in[k * ld + m] : in[m * ld + k];
}
TEO_API void teo_unswizzle_scales(const uint8_t *in, size_t ld, int transposed_dst, size_t M,
                                  size_t K, uint8_t *out) {
  for (size_t m = 0; m < M; ++m)
    for (size_t k = 0; k < K; ++k) {
      uint8_t v = in[teo_swizzle_index(m, k, K)];
      if (transposed_dst) out[k * ld + m] = v; else out[m * ld + k] = v;
    }
}

/* ------------------------------------------------------------------------ *
 * NVFP4 GEMM alpha  (recipe/nvfp4.cu:72-78; gemm/cublaslt_gemm.cu:364-404)
 * ------------------------------------------------------------------------ */
TEO_API float teo_nvfp4_alpha(float alpha, float amax_a, float amax_b) {
  const float factor_inv = 1.0f / (6.0f * 6.0f * 448.0f * 448.0f);
  return alpha * amax_a * amax_b * factor_inv;
}

/* ------------------------------------------------------------------------ *
 * Reference GEMM on dequantised operands, fp64 accumulate:
 * D[n, m] (row-major [N, M]) = alpha * sum_k A[m, k] * B[n, k] + beta * C[n, m]
 * (TN convention of gemm/cublaslt_gemm.cu:324-335 — both operands K-major.)
 * ------------------------------------------------------------------------ */
TEO_API void teo_gemm_tn_ref(const float *A, const float *B, size_t M, size_t N, size_t K,
                             float alpha, float beta, const float *C, const float *bias,
                             float *D) {
#pragma omp parallel for schedule(static)
  for (size_t n = 0; n < N; ++n)
    for (size_t m = 0; m < M; ++m) {
      double acc = 0.0;
      const float *a = A + m * K, *b = B + n * K;
      for (size_t k = 0; k < K; ++k) acc += (double)a[k] * (double)b[k];
      double r = (double)alpha * acc;
      if (bias) r += bias[m];
      if (C && beta != 0.0f) r += (double)beta * C[n * M + m];
      D[n * M + m] = (float)r;
    }
}

/* OpenMP team size for every later parallel region of this process (the reference's compute_ref functions in
 * oracle/_ref share the same libgomp): torchrun exports OMP_NUM_THREADS=1, which made the N >= 2 CPU arm 20x slower. */
TEO_API void teo_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

TEO_API int teo_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* scalar conversion exports for unit tests */
TEO_API uint8_t teo_cvt_e4m3(float f) { return teo_f32_to_e4m3(f); }
TEO_API uint8_t teo_cvt_e5m2(float f) { return teo_f32_to_e5m2(f); }
TEO_API uint8_t teo_cvt_e2m1(float f) { return teo_f32_to_e2m1(f); }
TEO_API uint8_t teo_cvt_e8m0(float f) { return teo_f32_to_e8m0(f); }
TEO_API float teo_cvt_e8m0_rcp(uint8_t e) { return teo_e8m0_rcp(e); }
TEO_API uint16_t teo_cvt_bf16(float f) { return teo_f32_to_bf16(f); }
