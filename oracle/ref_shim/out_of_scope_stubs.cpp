// TEST INFRASTRUCTURE. The reference's test translation units reference C API entry points that are outside
// the low-precision Linear path (activations and their fused dbias variants, grouped swizzles). They are never
// called by the oracle shims; in the reference's own gtest binary (make tests) a test that reaches one fails
// with this exception and the run continues. Everything the product library exports itself (tensor and
// grouped-tensor handles, nvte_quantize_dbias, ...) is NOT stubbed here.
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#ifdef TE_REF_STUB_THROWS
#define TE_REF_STUB(name)                                                                      \
  extern "C" void name() { throw std::runtime_error(#name " is outside the low-precision Linear path"); }
#else
#define TE_REF_STUB(name)                                                         \
  extern "C" void name() {                                                        \
    std::fprintf(stderr, #name " is outside the low-precision Linear path\n");    \
    std::abort();                                                                 \
  }
#endif
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
TE_REF_STUB(nvte_quantize_dbias_dgelu)
TE_REF_STUB(nvte_quantize_dbias_dqgelu)
TE_REF_STUB(nvte_quantize_dbias_drelu)
TE_REF_STUB(nvte_quantize_dbias_dsilu)
TE_REF_STUB(nvte_quantize_dbias_dsrelu)
TE_REF_STUB(nvte_swizzle_grouped_scaling_factors)
TE_REF_STUB(nvte_unswizzle_grouped_scaling_factors)
