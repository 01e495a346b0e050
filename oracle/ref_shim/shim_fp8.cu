// TEST INFRASTRUCTURE. Reference host per-tensor FP8 references included in
// place: cast + transpose (tests/cpp/operator/test_cast_transpose.cu:24-40).
#include REF_TEST_FILE
using namespace test;

extern "C" void ref_fp8_cast_transpose(const void *data, int in_bf16, int e5m2,
                                       unsigned char *out_c, unsigned char *out_t, size_t N,
                                       size_t H, float *amax, float scale) {
#define CALL(IT, OT)                                                                  \
  compute_ref<IT, OT>(reinterpret_cast<const IT *>(data), reinterpret_cast<OT *>(out_c), \
                      reinterpret_cast<OT *>(out_t), N, H, amax, scale)
  if (in_bf16) { if (e5m2) CALL(bf16, fp8e5m2); else CALL(bf16, fp8e4m3); }
  else { if (e5m2) CALL(fp32, fp8e5m2); else CALL(fp32, fp8e4m3); }
#undef CALL
}
