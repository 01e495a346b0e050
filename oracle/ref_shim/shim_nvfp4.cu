// TEST INFRASTRUCTURE. Reference host NVFP4 quantizer
// (tests/cpp/operator/test_cast_nvfp4_transpose.cu:183-284 1-D, :287-440 2-D,
// :461-555 compute_ref) included in place and exported through C symbols.
#include REF_TEST_FILE

extern "C" void ref_nvfp4_quantize_bf16(const void *input, unsigned char *output,
                                        unsigned char *scales, size_t rows, size_t cols,
                                        size_t scales_stride, float global_amax, int use_2d) {
  quantize_nvfp4<bf16>(&identity, reinterpret_cast<const bf16 *>(input),
                       reinterpret_cast<fp4e2m1x2 *>(output),
                       reinterpret_cast<fp8e4m3 *>(scales), rows, cols, scales_stride,
                       global_amax, /*use_fast_math=*/false, use_2d != 0);
}
extern "C" void ref_nvfp4_compute_ref_bf16(const void *input, unsigned char *output,
                                           unsigned char *output_t, unsigned char *scales,
                                           unsigned char *scales_t, float amax, size_t rows,
                                           size_t cols, size_t stride, size_t stride_t,
                                           int use_2d) {
  compute_ref<bf16>(&identity, reinterpret_cast<const bf16 *>(input),
                    reinterpret_cast<fp4e2m1x2 *>(output), reinterpret_cast<fp4e2m1x2 *>(output_t),
                    reinterpret_cast<fp8e4m3 *>(scales), reinterpret_cast<fp8e4m3 *>(scales_t),
                    &amax, rows, cols, stride, stride_t, /*use_fast_math=*/false, use_2d != 0);
}
