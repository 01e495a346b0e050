// TEST INFRASTRUCTURE. Reference host current-scaling amax/scale reference
// (tests/cpp/operator/test_cast_current_scaling.cu:39-85) included in place.
#include REF_TEST_FILE
using namespace test;

extern "C" void ref_amax_scale(const void *data, int in_bf16, size_t size, float *amax,
                               float *scale, float *scale_inv, float max_fp8, float epsilon) {
  if (in_bf16)
    compute_amax_scale_ref<bf16, fp8e4m3>(reinterpret_cast<const bf16 *>(data), size, amax, scale,
                                          scale_inv, max_fp8, epsilon);
  else
    compute_amax_scale_ref<fp32, fp8e4m3>(reinterpret_cast<const fp32 *>(data), size, amax, scale,
                                          scale_inv, max_fp8, epsilon);
}
