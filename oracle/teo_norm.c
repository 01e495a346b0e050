/* CPU oracle for LayerNorm / RMSNorm forward and backward.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): imported by tests/, smoke() and bench.py's
 * cpu_baseline leg, never by the product.
 *
 * Restates the reference's own CPU check, tests/cpp/operator/test_normalization.h:
 *   teo_norm_stats    <- compute_ref_stats     (:39-63)   sequential fp32 sums, two-pass variance
 *   teo_norm_output   <- compute_ref_output    (:100-135) elementwise, given mu / rsigma
 *   teo_norm_backward <- compute_ref_backward  (:140-195)
 * One deliberate difference in teo_norm_output: LayerNorm's `y * g + beta` is ONE fused
 * multiply-add, which is what the reference's GPU kernel computes (ln_fwd_kernels.cuh:98-103,
 * `g_ij * y_ij + b_ij` compiled with nvcc's default -fmad=true) while its CPU check rounds twice
 * and absorbs the difference in its tolerance. With the fused form the quantized bytes of the
 * fused norm -> cast kernels can be checked exactly for given statistics.
 * Parity: the statistics are order-of-summation dependent => tolerance (the reference's tests use
 * rtol/atol 1e-5..5e-3 by dtype, test_normalization.cu:205-240); everything downstream of given
 * statistics is exact. Pinned against the reference's torch LayerNorm/RMSNorm modules through
 * tests/golden/norm_*.npz (oracle/gen_golden.py).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>

#include "teo_convert.h"

#define TEO_API __attribute__((visibility("default")))

static inline float teo_gamma(const void *gamma, int wtype, size_t j, int zero_centered) {
  float g = teo_load(gamma, wtype, j);
  if (zero_centered) g += 1.0f;
  return g;
}

TEO_API void teo_norm_stats(const void *x, int dtype, size_t N, size_t H, double eps, int layernorm, float *mu,
                            float *rsigma) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < N; ++i) {
    float sum = 0.0f;
    for (size_t j = 0; j < H; ++j) sum += teo_load(x, dtype, i * H + j);
    float m = 0.0f;
    if (layernorm) {
      mu[i] = sum / (float)H;
      m = mu[i];
    }
    float sum_sq = 0.0f;
    for (size_t j = 0; j < H; ++j) {
      const float c = teo_load(x, dtype, i * H + j);
      sum_sq += (c - m) * (c - m);
    }
    /* the reference adds its `double epsilon` in double and rounds once (test_normalization.h:61) */
    /* and the host rsqrtf of the CUDA headers is (float)(1.0 / sqrt((double)a)) (crt/math_functions.hpp) */
    rsigma[i] = (float)(1.0 / sqrt((double)(float)((double)(sum_sq / (float)H) + eps)));
  }
}

/* y[i, j] in fp32 (the value the reference then scales / casts / quantizes) */
TEO_API void teo_norm_output(const void *x, int dtype, const void *gamma, const void *beta, int wtype, size_t N,
                             size_t H, const float *mu, const float *rsigma, int layernorm, int zero_centered,
                             float *y) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < N; ++i) {
    for (size_t j = 0; j < H; ++j) {
      const float c = teo_load(x, dtype, i * H + j);
      const float g = teo_gamma(gamma, wtype, j, zero_centered);
      float t;
      if (layernorm) {
        volatile float d = c - mu[i];
        volatile float n = d * rsigma[i];
        t = fmaf(n, g, teo_load(beta, wtype, j));
      } else {
        volatile float n = c * rsigma[i];
        t = n * g;
      }
      y[i * H + j] = t;
    }
  }
}

TEO_API void teo_norm_backward(const void *dz, const void *x, const void *add, int dtype, const float *mu,
                               const float *rsigma, const void *gamma, int wtype, size_t N, size_t H, int layernorm,
                               int zero_centered, float *dx, float *dgamma, float *dbeta) {
  for (size_t j = 0; j < H; ++j) {
    dgamma[j] = 0.0f;
    if (dbeta) dbeta[j] = 0.0f;
  }
  for (size_t i = 0; i < N; ++i) {
    const float lmu = layernorm ? mu[i] : 0.0f;
    float mdy = 0.0f, mdyy = 0.0f;
    for (size_t j = 0; j < H; ++j) {
      const float xv = teo_load(x, dtype, i * H + j);
      /* `local_mu` is a double in the reference (`cond ? mu[i] : 0.`): y is rounded once */
      const float y = (float)(((double)xv - (double)lmu) * (double)rsigma[i]);
      const float g = teo_gamma(gamma, wtype, j, zero_centered);
      const float z = teo_load(dz, dtype, i * H + j);
      const float dy = g * z;
      dgamma[j] += y * z;
      if (layernorm) {
        dbeta[j] += z;
        mdy += dy;
      }
      mdyy += dy * y;
    }
    mdy /= (float)H;
    mdyy /= (float)H;
    for (size_t j = 0; j < H; ++j) {
      const float xv = teo_load(x, dtype, i * H + j);
      /* `local_mu` is a double in the reference (`cond ? mu[i] : 0.`): y is rounded once */
      const float y = (float)(((double)xv - (double)lmu) * (double)rsigma[i]);
      const float g = teo_gamma(gamma, wtype, j, zero_centered);
      const float z = teo_load(dz, dtype, i * H + j);
      const float dy = g * z;
      const float a = add ? teo_load(add, dtype, i * H + j) : 0.0f;
      dx[i * H + j] = a + rsigma[i] * (dy - mdyy * y - mdy);
    }
  }
}
