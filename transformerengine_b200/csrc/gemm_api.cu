// extern "C" GEMM entry points (reference gemm/cublaslt_gemm.cu:803-920) and the
// NVFP4 per-tensor alpha helper (reference recipe/nvfp4.cu:72-78,912-943).
#include "core.h"

using namespace teb;

namespace {
template <typename F>
int guarded(F &&f) {
  try {
    f();
    return 0;
  } catch (const std::exception &e) {
    set_last_error(e.what());
    return 1;
  } catch (...) {
    set_last_error("unknown C++ exception");
    return 1;
  }
}

__global__ void nvfp4_alpha_kernel(const float *amax_a, const float *amax_b, float alpha_in,
                                   float *alpha_out) {
  constexpr float factor_inv = 1.0f / (6.0f * 6.0f * 448.0f * 448.0f);
  *alpha_out = alpha_in * (*amax_a) * (*amax_b) * factor_inv;
}

void gemm_legacy(const NVTETensor A, const NVTETensor B, NVTETensor D, const NVTETensor bias,
                 NVTETensor pre_gelu_out, bool transa, bool transb, bool grad,
                 NVTETensor workspace, float alpha, float beta, bool use_split_accumulator,
                 int math_sm_count, cudaStream_t stream) {
  const Tensor *pg = lookup(pre_gelu_out);
  TEB_CHECK(pg == nullptr || pg->data.dptr == nullptr,
            "GELU epilogues are outside the Linear path and not implemented");
  MatmulConfig cfg;
  const Tensor *b = lookup(bias);
  if (b != nullptr && b->data.dptr != nullptr) {
    TEB_CHECK(!grad, "The bias-gradient epilogue is not implemented");
    cfg.bias_tensor = bias;
  }
  cfg.use_split_accumulator = use_split_accumulator;
  cfg.sm_count = math_sm_count;
  Tensor &d = lookup_checked(D);
  gemm::run(transa, transb, alpha, lookup_checked(A), lookup_checked(B), beta, &d, d,
            lookup(workspace), cfg, stream);
}
}  // namespace

extern "C" {

void nvte_cublas_gemm(const NVTETensor A, const NVTETensor B, NVTETensor D, const NVTETensor bias,
                      NVTETensor pre_gelu_out, bool transa, bool transb, bool grad,
                      NVTETensor workspace, bool accumulate, bool use_split_accumulator,
                      int math_sm_count, cudaStream_t stream) {
  gemm_legacy(A, B, D, bias, pre_gelu_out, transa, transb, grad, workspace, 1.0f,
              accumulate ? 1.0f : 0.0f, use_split_accumulator, math_sm_count, stream);
}
void nvte_cublas_gemm_scaled(const NVTETensor A, const NVTETensor B, NVTETensor D,
                             const NVTETensor bias, NVTETensor pre_gelu_out, bool transa,
                             bool transb, bool grad, NVTETensor workspace, float alpha, float beta,
                             bool use_split_accumulator, int math_sm_count, cudaStream_t stream) {
  gemm_legacy(A, B, D, bias, pre_gelu_out, transa, transb, grad, workspace, alpha, beta,
              use_split_accumulator, math_sm_count, stream);
}
void nvte_cublas_gemm_v2(int transa, int transb, const float *alpha, const NVTETensor A,
                         const NVTETensor B, const float *beta, const NVTETensor C, NVTETensor D,
                         NVTETensor workspace, NVTEMatmulConfig config, cudaStream_t stream) {
  TEB_CHECK(alpha != nullptr && beta != nullptr, "alpha and beta must be host pointers");
  const MatmulConfig cfg = config ? *reinterpret_cast<const MatmulConfig *>(config) : MatmulConfig{};
  Tensor &d = lookup_checked(D);
  const Tensor *c = lookup(C);
  // reference cublaslt_gemm.cu:845-846: in-place accumulation only
  TEB_CHECK(c == nullptr || c->data.dptr == nullptr || c->data.dptr == d.data.dptr,
            "nvte_cublas_gemm_v2 requires C == D");
  gemm::run(transa, transb, *alpha, lookup_checked(A), lookup_checked(B), *beta, c, d,
            lookup(workspace), cfg, stream);
}
// reference gemm.h:295 / cublaslt_gemm.cu:970-1061 (GroupedLinear: one GEMM per expert). The
// reference loops cuBLASLt calls over its compute streams; here the whole list is one persistent
// launch whenever the members share a scaling mode (gemm::run_grouped), else a loop on `stream`.
void nvte_multi_tensor_gemm(const NVTETensor *A, const NVTETensor *B, NVTETensor *D,
                            const NVTETensor *bias, NVTETensor *pre_gelu_out, const int num_gemms,
                            bool transa, bool transb, bool grad, NVTETensor *workspace,
                            bool accumulate, bool use_split_accumulator, int math_sm_count,
                            cudaStream_t stream) {
  if (num_gemms <= 0) return;
  TEB_CHECK(A != nullptr && B != nullptr && D != nullptr, "Tensor lists must not be null");
  std::vector<gemm::GroupedMember> members(num_gemms);
  std::vector<MatmulConfig> cfgs(num_gemms);
  for (int i = 0; i < num_gemms; ++i) {
    const Tensor *pg = pre_gelu_out ? lookup(pre_gelu_out[i]) : nullptr;
    TEB_CHECK(pg == nullptr || pg->data.dptr == nullptr,
              "GELU epilogues are outside the Linear path and not implemented");
    const Tensor *b = bias ? lookup(bias[i]) : nullptr;
    if (b != nullptr && b->data.dptr != nullptr) {
      TEB_CHECK(!grad, "The bias-gradient epilogue is not implemented");
      cfgs[i].bias_tensor = bias[i];
    }
    cfgs[i].use_split_accumulator = use_split_accumulator;
    cfgs[i].sm_count = math_sm_count;
    members[i] = {&lookup_checked(A[i]), &lookup_checked(B[i]), &lookup_checked(D[i]), 1.0f,
                  accumulate ? 1.0f : 0.0f};
  }
  Tensor *ws = workspace ? lookup(workspace[0]) : nullptr;
  if (gemm::run_grouped(transa, transb, members, ws, cfgs, stream)) return;
  for (int i = 0; i < num_gemms; ++i)
    gemm::run(transa, transb, 1.0f, *members[i].A, *members[i].B, members[i].beta, members[i].D,
              *members[i].D, ws, cfgs[i], stream);
}
void nvte_nvfp4_compute_per_tensor_scale(const NVTETensor inpA, const bool use_rowwise_amax_A,
                                         const NVTETensor inpB, const bool use_rowwise_amax_B,
                                         float alpha_in, NVTETensor alpha_out,
                                         cudaStream_t stream) {
  const Tensor &a = lookup_checked(inpA);
  const Tensor &b = lookup_checked(inpB);
  Tensor &o = lookup_checked(alpha_out);
  const float *pa = (use_rowwise_amax_A ? a.amax : a.columnwise_amax).as<float>();
  const float *pb = (use_rowwise_amax_B ? b.amax : b.columnwise_amax).as<float>();
  TEB_CHECK(pa != nullptr && pb != nullptr && o.data.dptr != nullptr,
            "nvfp4 per-tensor scale needs both amaxes and an output buffer");
  nvfp4_alpha_kernel<<<1, 1, 0, stream>>>(pa, pb, alpha_in, o.data.as<float>());
  TEB_CHECK_CUDA(cudaGetLastError());
}

int nvte_cublas_gemm_v2_status(int transa, int transb, const float *alpha, const NVTETensor A,
                               const NVTETensor B, const float *beta, const NVTETensor C,
                               NVTETensor D, NVTETensor workspace, NVTEMatmulConfig config,
                               cudaStream_t stream) {
  return guarded([&] { nvte_cublas_gemm_v2(transa, transb, alpha, A, B, beta, C, D, workspace, config, stream); });
}
int nvte_multi_tensor_gemm_status(const NVTETensor *A, const NVTETensor *B, NVTETensor *D,
                                  const NVTETensor *bias, NVTETensor *pre_gelu_out, int num_gemms,
                                  int transa, int transb, int grad, NVTETensor *workspace,
                                  int accumulate, int use_split_accumulator, int math_sm_count,
                                  cudaStream_t stream) {
  return guarded([&] {
    nvte_multi_tensor_gemm(A, B, D, bias, pre_gelu_out, num_gemms, transa != 0, transb != 0, grad != 0,
                           workspace, accumulate != 0, use_split_accumulator != 0, math_sm_count, stream);
  });
}
int nvte_nvfp4_compute_per_tensor_scale_status(const NVTETensor inpA, const bool ra,
                                               const NVTETensor inpB, const bool rb, float alpha_in,
                                               NVTETensor alpha_out, cudaStream_t stream) {
  return guarded([&] { nvte_nvfp4_compute_per_tensor_scale(inpA, ra, inpB, rb, alpha_in, alpha_out, stream); });
}

}  // extern "C"
