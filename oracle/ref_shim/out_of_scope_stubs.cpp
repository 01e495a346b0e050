// TEST INFRASTRUCTURE. The reference's test translation units reference C API
// entry points that are outside the Linear hot path (activations, dbias
// fusions, grouped tensors). They are never called by the shims; these
// definitions only satisfy the linker and abort loudly if reached.
#include <cstdio>
#include <cstdlib>
#define TE_REF_STUB(name)                                                         \
  extern "C" void name() {                                                        \
    std::fprintf(stderr, #name " is outside the low-precision Linear path\n");    \
    std::abort();                                                                 \
  }
TE_REF_STUB(nvte_create_grouped_tensor)
TE_REF_STUB(nvte_destroy_grouped_tensor)
TE_REF_STUB(nvte_set_grouped_tensor_param)
TE_REF_STUB(nvte_get_grouped_tensor_param)
TE_REF_STUB(nvte_grouped_tensor_num_tensors)
TE_REF_STUB(nvte_grouped_tensor_type)
TE_REF_STUB(nvte_grouped_tensor_scaling_mode)
TE_REF_STUB(nvte_get_grouped_tensor_logical_shape)
TE_REF_STUB(nvte_gelu)
TE_REF_STUB(nvte_dgelu)
TE_REF_STUB(nvte_qgelu)
TE_REF_STUB(nvte_dqgelu)
TE_REF_STUB(nvte_relu)
TE_REF_STUB(nvte_drelu)
TE_REF_STUB(nvte_silu)
TE_REF_STUB(nvte_dsilu)
TE_REF_STUB(nvte_srelu)
TE_REF_STUB(nvte_dsrelu)
TE_REF_STUB(nvte_quantize_dbias)
TE_REF_STUB(nvte_quantize_dbias_dgelu)
TE_REF_STUB(nvte_quantize_dbias_dqgelu)
TE_REF_STUB(nvte_quantize_dbias_drelu)
TE_REF_STUB(nvte_quantize_dbias_dsilu)
TE_REF_STUB(nvte_quantize_dbias_dsrelu)
TE_REF_STUB(nvte_swizzle_grouped_scaling_factors)
TE_REF_STUB(nvte_unswizzle_grouped_scaling_factors)
