// TEST INFRASTRUCTURE. Reference host swizzle reference
// (tests/cpp/operator/test_swizzle.cu:30-86) included in place.
#include REF_TEST_FILE
using namespace test;

extern "C" void ref_swizzle(const unsigned char *in, unsigned char *out, size_t M, size_t K,
                            int row_scaling) {
  if (row_scaling) compute_ref_swizzle<128, 4, true>(in, out, M, K);
  else compute_ref_swizzle<128, 4, false>(in, out, M, K);
}
extern "C" void ref_unswizzle(const unsigned char *in, unsigned char *out, size_t M, size_t K,
                              int row_scaling) {
  if (row_scaling) compute_ref_unswizzle<128, 4, true>(in, out, M, K);
  else compute_ref_unswizzle<128, 4, false>(in, out, M, K);
}
