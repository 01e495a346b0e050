// TEST INFRASTRUCTURE. The reference's host LayerNorm / RMSNorm references included in place
// (tests/cpp/operator/test_normalization.h:39-63 compute_ref_stats, :100-135 compute_ref_output,
// :140-195 compute_ref_backward), bf16 tensors, fp32 normalized output (scale 1).
#include REF_TEST_FILE
using namespace test;

extern "C" void ref_norm_stats(const void *x, int layernorm, size_t N, size_t H, double eps, float *mu, float *rsigma) {
  compute_ref_stats<bf16>(layernorm ? LayerNorm : RMSNorm, reinterpret_cast<const bf16 *>(x), mu, rsigma, N, H, eps);
}
extern "C" void ref_norm_output(const void *x, const void *gamma, const void *beta, int layernorm, size_t N, size_t H,
                                const float *mu, const float *rsigma, int zero_centered, float *y, float *amax) {
  compute_ref_output<bf16, float>(layernorm ? LayerNorm : RMSNorm, reinterpret_cast<const bf16 *>(x),
                                  reinterpret_cast<const bf16 *>(gamma), reinterpret_cast<const bf16 *>(beta), y, mu,
                                  rsigma, N, H, amax, 1.0f, zero_centered != 0, false, false);
}
extern "C" void ref_norm_backward(const void *dz, const void *add, const void *x, int layernorm, size_t N, size_t H,
                                  const float *mu, const float *rsigma, const void *gamma, int zero_centered,
                                  void *dx, void *dgamma, void *dbeta) {
  compute_ref_backward<bf16, bf16>(layernorm ? LayerNorm : RMSNorm, reinterpret_cast<const bf16 *>(dz),
                                   reinterpret_cast<const bf16 *>(add), reinterpret_cast<const bf16 *>(x), mu, rsigma,
                                   reinterpret_cast<const bf16 *>(gamma), reinterpret_cast<bf16 *>(dx),
                                   reinterpret_cast<bf16 *>(dgamma), reinterpret_cast<bf16 *>(dbeta), N, H,
                                   zero_centered != 0, false, false);
}
