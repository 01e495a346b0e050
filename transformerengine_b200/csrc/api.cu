// extern "C" compute entry points: argument checking, dispatch on the output
// tensor's scaling mode (reference cast/dispatch/quantize.cuh:30-195), and the
// FFI-safe `_status` twins.
#include <nvtx3/nvToolsExt.h>

#include "core.h"

using namespace teb;

namespace {

// One NVTX range per API call, like the reference's NVTE_API_CALL
// (reference common/common.h:1174-1175).
struct ApiRange {
  explicit ApiRange(const char *name) { nvtxRangePushA(name); }
  ~ApiRange() { nvtxRangePop(); }
};
#define TEB_API_CALL(name) ApiRange teb_api_range_(#name)

const Tensor *noop_of(const QuantConfig &cfg) {
  return cfg.noop_tensor ? lookup(cfg.noop_tensor) : nullptr;
}

void quantize_dispatch(const NVTETensor input, NVTETensor output, const QuantConfig &cfg,
                       cudaStream_t stream) {
  const Tensor &in = lookup_checked(input);
  Tensor &out = lookup_checked(output);
  TEB_CHECK(out.has_data() || out.has_columnwise_data(),
            "Either rowwise or columnwise output data need to be allocated.");
  TEB_CHECK(!cfg.use_fast_math,
            "use_fast_math changes quantization numerics and is not implemented.");
  const Tensor *noop = noop_of(cfg);
  TEB_CHECK((cfg.peer_rowwise[0] == 0 && cfg.symm.world <= 1) || out.scaling_mode == NVTE_DELAYED_TENSOR_SCALING,
            "Peer destinations are implemented for per-tensor FP8 outputs only");
  switch (out.scaling_mode) {
    case NVTE_DELAYED_TENSOR_SCALING:
      if (cfg.symm.world > 1) fp8::quantize_exchange(in, out, cfg.symm, stream);
      else if (cfg.peer_rowwise[0] != 0) fp8::quantize_push(in, out, cfg.peer_rowwise, noop, stream);
      else fp8::quantize(in, out, noop, stream);
      break;
    case NVTE_MXFP8_1D_SCALING: mxfp8::quantize(in, out, noop, stream); break;
    case NVTE_NVFP4_1D_SCALING:
      nvfp4::quantize(in, out, cfg, noop, /*with_rht=*/false, 0, nullptr, stream);
      break;
    default:
      TEB_ERROR("Not implemented scaling mode: ", static_cast<int>(out.scaling_mode), ".");
  }
}

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

}  // namespace

extern "C" {

void nvte_quantize(const NVTETensor input, NVTETensor output, cudaStream_t stream) {
  TEB_API_CALL(nvte_quantize);
  quantize_dispatch(input, output, QuantConfig{}, stream);
}
void nvte_quantize_noop(const NVTETensor input, NVTETensor output, NVTETensor noop,
                        cudaStream_t stream) {
  TEB_API_CALL(nvte_quantize_noop);
  QuantConfig cfg;
  cfg.noop_tensor = noop;
  quantize_dispatch(input, output, cfg, stream);
}
void nvte_quantize_v2(const NVTETensor input, NVTETensor output,
                      const NVTEQuantizationConfig quant_config, cudaStream_t stream) {
  TEB_API_CALL(nvte_quantize_v2);
  const QuantConfig cfg = quant_config ? *reinterpret_cast<const QuantConfig *>(quant_config)
                                       : QuantConfig{};
  quantize_dispatch(input, output, cfg, stream);
}
// reference cast/cast.cu:57-95: the same cast for a list of tensors (MoE experts, split_quantize).
// The reference spreads the launches over its pool of compute streams; the kernels here are
// persistent and chained by programmatic dependent launch, so they stay on the caller's stream.
void nvte_multi_tensor_quantize(const NVTETensor *inputs, NVTETensor *outputs,
                                const NVTEQuantizationConfig quant_config, const size_t num_tensors,
                                cudaStream_t stream) {
  TEB_API_CALL(nvte_multi_tensor_quantize);
  TEB_CHECK(num_tensors == 0 || (inputs != nullptr && outputs != nullptr), "Tensor lists must not be null");
  const QuantConfig cfg = quant_config ? *reinterpret_cast<const QuantConfig *>(quant_config)
                                       : QuantConfig{};
  // MXFP8 lists (x / W / dy of a Linear step, the experts of a GroupedLinear): one persistent launch over
  // the concatenated tile list of all members (mxfp8::quantize_multi); anything else: one cast each
  bool all_mx = num_tensors > 1;
  for (size_t i = 0; i < num_tensors && all_mx; ++i) {
    const Tensor *o = lookup(outputs[i]);
    all_mx = o != nullptr && lookup(inputs[i]) != nullptr && o->scaling_mode == NVTE_MXFP8_1D_SCALING;
  }
  if (all_mx) {
    TEB_CHECK(!cfg.use_fast_math, "use_fast_math changes quantization numerics and is not implemented.");
    std::vector<const Tensor *> ins(num_tensors);
    std::vector<Tensor *> outs(num_tensors);
    for (size_t i = 0; i < num_tensors; ++i) {
      ins[i] = &lookup_checked(inputs[i]);
      outs[i] = &lookup_checked(outputs[i]);
    }
    mxfp8::quantize_multi(ins.data(), outs.data(), num_tensors, noop_of(cfg), stream);
    return;
  }
  for (size_t i = 0; i < num_tensors; ++i) quantize_dispatch(inputs[i], outputs[i], cfg, stream);
}
void nvte_quantize_dbias(const NVTETensor input, NVTETensor output, NVTETensor dbias,
                         NVTETensor workspace, cudaStream_t stream) {
  TEB_API_CALL(nvte_quantize_dbias);
  dbias::quantize_dbias(lookup_checked(input), lookup_checked(output), lookup_checked(dbias),
                        lookup_checked(workspace), stream);
}
// ---- normalization (reference include/transformer_engine/normalization.h:44-165) ----
void nvte_layernorm_fwd(const NVTETensor x, const NVTETensor gamma, const NVTETensor beta, const float epsilon,
                        NVTETensor z, NVTETensor mu, NVTETensor rsigma, NVTETensor workspace,
                        const int multiprocessorCount, const bool zero_centered_gamma, cudaStream_t stream) {
  TEB_API_CALL(nvte_layernorm_fwd);
  (void)multiprocessorCount;   // the slab/cluster decomposition does not depend on it
  norm::forward(true, lookup_checked(x), lookup_checked(gamma), &lookup_checked(beta), epsilon, lookup_checked(z),
                &lookup_checked(mu), lookup_checked(rsigma), lookup_checked(workspace), zero_centered_gamma, stream);
}
void nvte_rmsnorm_fwd(const NVTETensor x, const NVTETensor gamma, const float epsilon, NVTETensor z,
                      NVTETensor rsigma, NVTETensor workspace, const int multiprocessorCount,
                      const bool zero_centered_gamma, cudaStream_t stream) {
  TEB_API_CALL(nvte_rmsnorm_fwd);
  (void)multiprocessorCount;
  norm::forward(false, lookup_checked(x), lookup_checked(gamma), nullptr, epsilon, lookup_checked(z), nullptr,
                lookup_checked(rsigma), lookup_checked(workspace), zero_centered_gamma, stream);
}
void nvte_layernorm_bwd(const NVTETensor dz, const NVTETensor x, const NVTETensor mu, const NVTETensor rsigma,
                        const NVTETensor gamma, NVTETensor dx, NVTETensor dgamma, NVTETensor dbeta,
                        NVTETensor workspace, const int multiprocessorCount, const bool zero_centered_gamma,
                        cudaStream_t stream) {
  TEB_API_CALL(nvte_layernorm_bwd);
  (void)multiprocessorCount;
  norm::backward(true, lookup_checked(dz), lookup_checked(x), nullptr, &lookup_checked(mu), lookup_checked(rsigma),
                 lookup_checked(gamma), lookup_checked(dx), lookup_checked(dgamma), &lookup_checked(dbeta),
                 lookup_checked(workspace), zero_centered_gamma, stream);
}
void nvte_rmsnorm_bwd(const NVTETensor dz, const NVTETensor x, const NVTETensor rsigma, const NVTETensor gamma,
                      NVTETensor dx, NVTETensor dgamma, NVTETensor workspace, const int multiprocessorCount,
                      const bool zero_centered_gamma, cudaStream_t stream) {
  TEB_API_CALL(nvte_rmsnorm_bwd);
  (void)multiprocessorCount;
  norm::backward(false, lookup_checked(dz), lookup_checked(x), nullptr, nullptr, lookup_checked(rsigma),
                 lookup_checked(gamma), lookup_checked(dx), lookup_checked(dgamma), nullptr,
                 lookup_checked(workspace), zero_centered_gamma, stream);
}
void nvte_rmsnorm_bwd_add(const NVTETensor dz, const NVTETensor x, const NVTETensor add, const NVTETensor rsigma,
                          const NVTETensor gamma, NVTETensor dx, NVTETensor dgamma, NVTETensor workspace,
                          const int multiprocessorCount, const bool zero_centered_gamma, cudaStream_t stream) {
  TEB_API_CALL(nvte_rmsnorm_bwd_add);
  (void)multiprocessorCount;
  norm::backward(false, lookup_checked(dz), lookup_checked(x), &lookup_checked(add), nullptr, lookup_checked(rsigma),
                 lookup_checked(gamma), lookup_checked(dx), lookup_checked(dgamma), nullptr,
                 lookup_checked(workspace), zero_centered_gamma, stream);
}
// The reference switches between its own kernels and cuDNN graphs with these; there is one
// implementation here, the calls are accepted and ignored.
void nvte_enable_cudnn_norm_fwd(bool) {}
void nvte_enable_cudnn_norm_bwd(bool) {}
void nvte_enable_zero_centered_gamma_in_weight_dtype(bool) {}

void nvte_dequantize(const NVTETensor input, NVTETensor output, cudaStream_t stream) {
  TEB_API_CALL(nvte_dequantize);
  const Tensor &in = lookup_checked(input);
  Tensor &out = lookup_checked(output);
  switch (in.scaling_mode) {
    case NVTE_DELAYED_TENSOR_SCALING: fp8::dequantize(in, out, stream); break;
    case NVTE_MXFP8_1D_SCALING: mxfp8::dequantize(in, out, stream); break;
    case NVTE_NVFP4_1D_SCALING: nvfp4::dequantize(in, out, stream); break;
    default: TEB_ERROR("Not implemented scaling mode: ", static_cast<int>(in.scaling_mode), ".");
  }
}

void nvte_compute_amax(const NVTETensor input, NVTETensor output, cudaStream_t stream) {
  TEB_API_CALL(nvte_compute_amax);
  fp8::compute_amax(lookup_checked(input), lookup_checked(output), nullptr, stream);
}
void nvte_compute_amax_with_config(const NVTETensor input, NVTETensor output,
                                   const NVTEQuantizationConfig config, cudaStream_t stream) {
  TEB_API_CALL(nvte_compute_amax_with_config);
  const Tensor *noop = config ? noop_of(*reinterpret_cast<const QuantConfig *>(config)) : nullptr;
  fp8::compute_amax(lookup_checked(input), lookup_checked(output), noop, stream);
  // extension: the amax also goes to peer-mapped slots (kNVTEQuantizationConfigB200PeerRowwiseData)
  if (config != nullptr && reinterpret_cast<const QuantConfig *>(config)->peer_rowwise[0] != 0)
    fp8::push_amax(lookup_checked(output), reinterpret_cast<const QuantConfig *>(config)->peer_rowwise, noop,
                   stream);
  // extension: symmetric exchange -- (amax, step) to every rank's slot array (kNVTEQuantizationConfigB200SymmetricExchange)
  if (config != nullptr && reinterpret_cast<const QuantConfig *>(config)->symm.world > 1)
    fp8::push_amax_step(lookup_checked(output), reinterpret_cast<const QuantConfig *>(config)->symm, stream);
}
void nvte_compute_scale_from_amax(NVTETensor output, const NVTEQuantizationConfig config,
                                  cudaStream_t stream) {
  TEB_API_CALL(nvte_compute_scale_from_amax);
  const QuantConfig cfg = config ? *reinterpret_cast<const QuantConfig *>(config) : QuantConfig{};
  fp8::compute_scale_from_amax(lookup_checked(output), cfg, stream);
}
void nvte_delayed_scaling_recipe_amax_and_scale_update(
    const NVTETensor amax_history, const NVTETensor scale, NVTETensor updated_amax_history,
    NVTETensor updated_scale, const char *amax_compute_algo, NVTEDType fp8_dtype, float margin,
    cudaStream_t stream) {
  TEB_API_CALL(nvte_delayed_scaling_recipe_amax_and_scale_update);
  fp8::delayed_update(lookup_checked(amax_history), lookup_checked(scale),
                      lookup_checked(updated_amax_history), lookup_checked(updated_scale),
                      amax_compute_algo, fp8_dtype, margin, stream);
}
void nvte_b200_delayed_scaling_update_after_reduction(
    const NVTETensor amax_reduction_buffer, const NVTETensor *amax_histories,
    const NVTETensor *scales, size_t num_tensors, const char *amax_compute_algo,
    NVTEDType fp8_dtype, float margin, cudaStream_t stream) {
  TEB_API_CALL(nvte_delayed_scaling_recipe_amax_and_scale_update_after_reduction);
  fp8::delayed_update_after_reduction(lookup(amax_reduction_buffer), amax_histories, scales,
                                      num_tensors, amax_compute_algo, fp8_dtype, margin, stream);
}
// The reference's own signature (recipe.h:71-74, declared inside `extern "C"`: the symbol is unmangled
// and takes the two std::vector<NVTETensor> by value, which is what the reference's
// pytorch/csrc/extensions/recipe.cpp binds).
void nvte_delayed_scaling_recipe_amax_and_scale_update_after_reduction(
    const NVTETensor amax_reduction_buffer, std::vector<NVTETensor> amax_histories,
    std::vector<NVTETensor> scales, const char *amax_compute_algo, NVTEDType fp8_dtype, float margin,
    cudaStream_t stream) {
  TEB_CHECK(amax_histories.size() == scales.size(), "Number of amax histories (", amax_histories.size(),
            ") does not match number of scales (", scales.size(), ")");
  nvte_b200_delayed_scaling_update_after_reduction(amax_reduction_buffer, amax_histories.data(),
                                                   scales.data(), amax_histories.size(),
                                                   amax_compute_algo, fp8_dtype, margin, stream);
}

void nvte_hadamard_transform(const NVTETensor input, NVTETensor output, int random_sign_mask,
                             int random_sign_mask_t, cudaStream_t stream) {
  TEB_API_CALL(nvte_hadamard_transform);
  nvfp4::rht(lookup_checked(input), lookup_checked(output), random_sign_mask, random_sign_mask_t,
             stream);
}
void nvte_hadamard_transform_amax(const NVTETensor input, NVTETensor output, int random_sign_mask,
                                  int random_sign_mask_t, cudaStream_t stream) {
  TEB_API_CALL(nvte_hadamard_transform_amax);
  nvfp4::rht_amax(lookup_checked(input), lookup_checked(output), random_sign_mask,
                  random_sign_mask_t, stream);
}

void nvte_quantize_with_hadamard_transform(const NVTETensor input, NVTETensor output,
                                           const NVTETensor hadamard_matrix,
                                           const NVTEQuantizationConfig quant_config,
                                           cudaStream_t stream) {
  TEB_API_CALL(nvte_quantize_with_hadamard_transform);
  const QuantConfig cfg = quant_config ? *reinterpret_cast<const QuantConfig *>(quant_config)
                                       : QuantConfig{};
  const Tensor &m = lookup_checked(hadamard_matrix);
  nvfp4::quantize(lookup_checked(input), lookup_checked(output), cfg, noop_of(cfg),
                  /*with_rht=*/true, 0, &m, stream);
}
void nvte_hadamard_transform_cast_fusion_columnwise(const NVTETensor input, NVTETensor output,
                                                    const NVTETensor hadamard_matrix,
                                                    const NVTEQuantizationConfig quant_config,
                                                    cudaStream_t stream) {
  TEB_API_CALL(nvte_hadamard_transform_cast_fusion_columnwise);
  const QuantConfig cfg = quant_config ? *reinterpret_cast<const QuantConfig *>(quant_config)
                                       : QuantConfig{};
  // columnwise-only view of the output: the rowwise buffers are left untouched
  Tensor out = lookup_checked(output);
  out.data.clear();
  out.scale_inv.clear();
  const Tensor &m = lookup_checked(hadamard_matrix);
  nvfp4::quantize(lookup_checked(input), out, cfg, noop_of(cfg), /*with_rht=*/true, 0, &m, stream);
}

void nvte_swizzle_scaling_factors(const NVTETensor input, NVTETensor output, cudaStream_t stream) {
  TEB_API_CALL(nvte_swizzle_scaling_factors);
  swizzle::swizzle_scales(lookup_checked(input), lookup_checked(output), false, stream);
}
void nvte_unswizzle_scaling_factors(const NVTETensor input, NVTETensor output,
                                    cudaStream_t stream) {
  TEB_API_CALL(nvte_unswizzle_scaling_factors);
  swizzle::swizzle_scales(lookup_checked(input), lookup_checked(output), true, stream);
}
// reference swizzle/swizzle.cu (multi-tensor variants, swizzle.h:45,84): the reference batches the
// list into one kernel to save launches; the quantizers here write swizzled scales directly, so
// this is only reached for tensors quantized with compact scales and stays a plain loop.
void nvte_multi_tensor_swizzle_scaling_factors(const NVTETensor *inputs, NVTETensor *outputs,
                                               const size_t num_tensors, cudaStream_t stream) {
  TEB_API_CALL(nvte_multi_tensor_swizzle_scaling_factors);
  TEB_CHECK(num_tensors == 0 || (inputs != nullptr && outputs != nullptr), "Tensor lists must not be null");
  for (size_t i = 0; i < num_tensors; ++i)
    swizzle::swizzle_scales(lookup_checked(inputs[i]), lookup_checked(outputs[i]), false, stream);
}
void nvte_multi_tensor_unswizzle_scaling_factors(const NVTETensor *inputs, NVTETensor *outputs,
                                                 const size_t num_tensors, cudaStream_t stream) {
  TEB_API_CALL(nvte_multi_tensor_unswizzle_scaling_factors);
  TEB_CHECK(num_tensors == 0 || (inputs != nullptr && outputs != nullptr), "Tensor lists must not be null");
  for (size_t i = 0; i < num_tensors; ++i)
    swizzle::swizzle_scales(lookup_checked(inputs[i]), lookup_checked(outputs[i]), true, stream);
}

// ---- FFI-safe twins ------------------------------------------------------------
int nvte_multi_tensor_swizzle_scaling_factors_status(const NVTETensor *i, NVTETensor *o, size_t n,
                                                     cudaStream_t s) {
  return guarded([&] { nvte_multi_tensor_swizzle_scaling_factors(i, o, n, s); });
}
int nvte_quantize_v2_status(const NVTETensor i, NVTETensor o, const NVTEQuantizationConfig c,
                            cudaStream_t s) {
  return guarded([&] { nvte_quantize_v2(i, o, c, s); });
}
int nvte_multi_tensor_quantize_status(const NVTETensor *i, NVTETensor *o, const NVTEQuantizationConfig c, size_t n,
                                      cudaStream_t s) {
  return guarded([&] { nvte_multi_tensor_quantize(i, o, c, n, s); });
}
int nvte_quantize_dbias_status(const NVTETensor i, NVTETensor o, NVTETensor d, NVTETensor w, cudaStream_t s) {
  return guarded([&] { nvte_quantize_dbias(i, o, d, w, s); });
}
int nvte_b200_reduce_partials(const void *partials, uint32_t world, size_t count, NVTEDType dtype, void *out,
                              cudaStream_t s) {
  return guarded([&] {
    TEB_API_CALL(nvte_b200_reduce_partials);
    rs::reduce_partials(partials, world, count, dtype, out, s);
  });
}
// nvte_quantize_dbias without a workspace argument and without any allocation: the partial sums go to a
// static device array of the library (csrc/dbias.cu). Returns a status like the *_status twins.
int nvte_b200_quantize_dbias_static_scratch(const NVTETensor i, NVTETensor o, NVTETensor d, cudaStream_t s) {
  return guarded([&] {
    TEB_API_CALL(nvte_b200_quantize_dbias_static_scratch);
    dbias::quantize_dbias_static_scratch(lookup_checked(i), lookup_checked(o), lookup_checked(d), s);
  });
}
int nvte_layernorm_fwd_status(const NVTETensor x, const NVTETensor g, const NVTETensor b, float eps, NVTETensor z,
                              NVTETensor mu, NVTETensor rs, NVTETensor ws, int sms, int zc, cudaStream_t s) {
  return guarded([&] { nvte_layernorm_fwd(x, g, b, eps, z, mu, rs, ws, sms, zc != 0, s); });
}
int nvte_rmsnorm_fwd_status(const NVTETensor x, const NVTETensor g, float eps, NVTETensor z, NVTETensor rs,
                            NVTETensor ws, int sms, int zc, cudaStream_t s) {
  return guarded([&] { nvte_rmsnorm_fwd(x, g, eps, z, rs, ws, sms, zc != 0, s); });
}
int nvte_layernorm_bwd_status(const NVTETensor dz, const NVTETensor x, const NVTETensor mu, const NVTETensor rs,
                              const NVTETensor g, NVTETensor dx, NVTETensor dg, NVTETensor db, NVTETensor ws, int sms,
                              int zc, cudaStream_t s) {
  return guarded([&] { nvte_layernorm_bwd(dz, x, mu, rs, g, dx, dg, db, ws, sms, zc != 0, s); });
}
int nvte_rmsnorm_bwd_status(const NVTETensor dz, const NVTETensor x, const NVTETensor rs, const NVTETensor g,
                            NVTETensor dx, NVTETensor dg, NVTETensor ws, int sms, int zc, cudaStream_t s) {
  return guarded([&] { nvte_rmsnorm_bwd(dz, x, rs, g, dx, dg, ws, sms, zc != 0, s); });
}
int nvte_rmsnorm_bwd_add_status(const NVTETensor dz, const NVTETensor x, const NVTETensor add, const NVTETensor rs,
                                const NVTETensor g, NVTETensor dx, NVTETensor dg, NVTETensor ws, int sms, int zc,
                                cudaStream_t s) {
  return guarded([&] { nvte_rmsnorm_bwd_add(dz, x, add, rs, g, dx, dg, ws, sms, zc != 0, s); });
}
int nvte_dequantize_status(const NVTETensor i, NVTETensor o, cudaStream_t s) {
  return guarded([&] { nvte_dequantize(i, o, s); });
}
int nvte_compute_amax_with_config_status(const NVTETensor i, NVTETensor o,
                                         const NVTEQuantizationConfig c, cudaStream_t s) {
  return guarded([&] { nvte_compute_amax_with_config(i, o, c, s); });
}
int nvte_compute_scale_from_amax_status(NVTETensor o, const NVTEQuantizationConfig c,
                                        cudaStream_t s) {
  return guarded([&] { nvte_compute_scale_from_amax(o, c, s); });
}
int nvte_delayed_scaling_recipe_amax_and_scale_update_status(
    const NVTETensor h, const NVTETensor sc, NVTETensor ho, NVTETensor so, const char *algo,
    NVTEDType t, float margin, cudaStream_t s) {
  return guarded(
      [&] { nvte_delayed_scaling_recipe_amax_and_scale_update(h, sc, ho, so, algo, t, margin, s); });
}
int nvte_b200_delayed_scaling_update_after_reduction_status(
    const NVTETensor buf, const NVTETensor *hs, const NVTETensor *ss, size_t n, const char *algo,
    NVTEDType t, float margin, cudaStream_t s) {
  return guarded(
      [&] { nvte_b200_delayed_scaling_update_after_reduction(buf, hs, ss, n, algo, t, margin, s); });
}
int nvte_hadamard_transform_status(const NVTETensor i, NVTETensor o, int m, int mt,
                                   cudaStream_t s) {
  return guarded([&] { nvte_hadamard_transform(i, o, m, mt, s); });
}
int nvte_hadamard_transform_amax_status(const NVTETensor i, NVTETensor o, int m, int mt,
                                        cudaStream_t s) {
  return guarded([&] { nvte_hadamard_transform_amax(i, o, m, mt, s); });
}
int nvte_quantize_with_hadamard_transform_status(const NVTETensor i, NVTETensor o,
                                                 const NVTETensor m,
                                                 const NVTEQuantizationConfig c, cudaStream_t s) {
  return guarded([&] { nvte_quantize_with_hadamard_transform(i, o, m, c, s); });
}
int nvte_swizzle_scaling_factors_status(const NVTETensor i, NVTETensor o, cudaStream_t s) {
  return guarded([&] { nvte_swizzle_scaling_factors(i, o, s); });
}
int nvte_unswizzle_scaling_factors_status(const NVTETensor i, NVTETensor o, cudaStream_t s) {
  return guarded([&] { nvte_unswizzle_scaling_factors(i, o, s); });
}

}  // extern "C"
