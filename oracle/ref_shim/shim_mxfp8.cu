// TEST INFRASTRUCTURE. Compiles the reference's own host reference for MXFP8
// (tests/cpp/operator/test_cast_mxfp8.cu:39-166, `compute_ref`) from the
// source where it lies under /root/reference and exposes it through a C
// symbol, so the oracle restatement can be checked against it on the CPU.
// Nothing is copied: the reference file is included in place.
#include REF_TEST_FILE

extern "C" void ref_mxfp8_compute_ref_bf16(const void *input, int e5m2, int rowwise, int colwise,
                                           unsigned char *out_row, unsigned char *out_col,
                                           unsigned char *sc_row, unsigned char *sc_col,
                                           size_t rows, size_t cols, size_t stride_row,
                                           size_t stride_col) {
  std::vector<bf16> dbias(cols);
  const bf16 *in = reinterpret_cast<const bf16 *>(input);
  if (e5m2) {
    compute_ref<bf16, fp8e5m2>(CAST_ONLY, &identity, rowwise, colwise, in, in,
                               reinterpret_cast<fp8e5m2 *>(out_row),
                               reinterpret_cast<fp8e5m2 *>(out_col), sc_row, sc_col, dbias.data(),
                               rows, cols, stride_row, stride_col);
  } else {
    compute_ref<bf16, fp8e4m3>(CAST_ONLY, &identity, rowwise, colwise, in, in,
                               reinterpret_cast<fp8e4m3 *>(out_row),
                               reinterpret_cast<fp8e4m3 *>(out_col), sc_row, sc_col, dbias.data(),
                               rows, cols, stride_row, stride_col);
  }
}
extern "C" void ref_mxfp8_compute_ref_fp32(const void *input, int rowwise, int colwise,
                                           unsigned char *out_row, unsigned char *out_col,
                                           unsigned char *sc_row, unsigned char *sc_col,
                                           size_t rows, size_t cols, size_t stride_row,
                                           size_t stride_col) {
  std::vector<float> dbias(cols);
  const float *in = reinterpret_cast<const float *>(input);
  compute_ref<float, fp8e4m3>(CAST_ONLY, &identity, rowwise, colwise, in, in,
                              reinterpret_cast<fp8e4m3 *>(out_row),
                              reinterpret_cast<fp8e4m3 *>(out_col), sc_row, sc_col, dbias.data(),
                              rows, cols, stride_row, stride_col);
}
