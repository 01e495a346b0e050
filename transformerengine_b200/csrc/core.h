// Internal core types of the B200 low-precision Linear library.
// Host-side view structs behind the NVTETensor handles, error macros and
// small helpers shared by every translation unit.
#pragma once

#include <cuda_runtime.h>

#include <array>
#include <cstdint>
#include <cstring>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/transformerengine_b200.h"

namespace teb {

// ---- errors: same convention as the reference (throw std::runtime_error
// through the C boundary, reference util/logging.h:38-56) --------------------
template <typename... Ts>
inline std::string cat(const Ts &...xs) {
  std::ostringstream os;
  (void)std::initializer_list<int>{((os << xs), 0)...};
  return os.str();
}

#define TEB_ERROR(...)                                                                    \
  do {                                                                                    \
    throw ::std::runtime_error(::teb::cat(__FILE__, ":", __LINE__, " in function ",       \
                                          __func__, ": ", ::teb::cat(__VA_ARGS__)));      \
  } while (false)

#define TEB_CHECK(cond, ...)                                                              \
  do {                                                                                    \
    if (!(cond)) TEB_ERROR("Assertion failed: " #cond ". ", ::teb::cat(__VA_ARGS__));     \
  } while (false)

#define TEB_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    const cudaError_t teb_status_ = (expr);                                               \
    if (teb_status_ != cudaSuccess)                                                       \
      TEB_ERROR("CUDA Error: ", cudaGetErrorString(teb_status_));                         \
  } while (false)

// ---- dtype helpers -----------------------------------------------------------
inline size_t dtype_bits(NVTEDType t) {
  switch (t) {
    case kNVTEByte: return 8;
    case kNVTEInt16: return 16;
    case kNVTEInt32: return 32;
    case kNVTEInt64: return 64;
    case kNVTEFloat32: return 32;
    case kNVTEFloat16: return 16;
    case kNVTEBFloat16: return 16;
    case kNVTEFloat8E4M3: return 8;
    case kNVTEFloat8E5M2: return 8;
    case kNVTEFloat8E8M0: return 8;
    case kNVTEFloat4E2M1: return 4;
    default: TEB_ERROR("Invalid dtype ", static_cast<int>(t));
  }
}
inline bool is_fp8(NVTEDType t) { return t == kNVTEFloat8E4M3 || t == kNVTEFloat8E5M2; }
inline bool is_fp4(NVTEDType t) { return t == kNVTEFloat4E2M1; }
inline bool is_high_precision(NVTEDType t) {
  return t == kNVTEFloat32 || t == kNVTEFloat16 || t == kNVTEBFloat16;
}
inline const char *dtype_name(NVTEDType t) {
  static const char *names[] = {"Byte", "Int16", "Int32", "Int64", "Float32", "Float16",
                                "BFloat16", "Float8E4M3", "Float8E5M2", "Float8E8M0",
                                "Float4E2M1"};
  return (static_cast<int>(t) >= 0 && t < kNVTENumTypes) ? names[t] : "Invalid";
}

inline size_t shape_numel(const NVTEShape &s) {
  size_t n = 1;
  for (size_t i = 0; i < s.ndim; ++i) n *= s.data[i];
  return n;
}

// ---- tensor views ------------------------------------------------------------
// A non-owning buffer view. The "unset" state is {nullptr, shape {0}} exactly
// as in the reference (common.h:243-281) because callers probe it through
// nvte_get_tensor_param.
struct Buffer {
  void *dptr = nullptr;
  NVTEShape shape{};
  NVTEDType dtype = kNVTEFloat32;

  Buffer() { clear(); }
  explicit Buffer(const NVTEBasicTensor &b) : dptr(b.data_ptr), shape(b.shape), dtype(b.dtype) {}
  void clear() {
    dptr = nullptr;
    std::memset(&shape, 0, sizeof(shape));
    shape.ndim = 1;
    shape.data[0] = 0;
    dtype = kNVTEFloat32;
  }
  bool has_data() const { return !(dptr == nullptr && shape.ndim == 1 && shape.data[0] == 0); }
  size_t numel() const { return shape_numel(shape); }
  size_t bytes() const { return (numel() * dtype_bits(dtype) + 7) / 8; }
  NVTEBasicTensor basic() const { return NVTEBasicTensor{dptr, dtype, shape}; }
  template <typename T>
  T *as() const { return reinterpret_cast<T *>(dptr); }
};

struct Tensor {
  Buffer data, columnwise_data, amax, columnwise_amax, scale, scale_inv, columnwise_scale_inv;
  NVTEScalingMode scaling_mode = NVTE_DELAYED_TENSOR_SCALING;
  bool with_gemm_swizzled_scales = false;
  bool row_scaled_nvfp4 = false;
  int nvfp4_e4m3_max = 448;

  void clear() { *this = Tensor(); }
  bool has_data() const { return data.has_data(); }
  bool has_columnwise_data() const { return columnwise_data.has_data(); }
  NVTEDType dtype() const {
    return (!has_data() && has_columnwise_data()) ? columnwise_data.dtype : data.dtype;
  }
  // Logical shape (reference common.h:383-420): per-tensor FP8 and NVFP4 store
  // columnwise data transposed, MXFP8 does not.
  NVTEShape shape() const;
  std::array<size_t, 2> flat_2d() const {
    const NVTEShape s = shape();
    if (s.ndim == 0) return {1, 1};
    size_t first = 1;
    for (size_t i = 0; i + 1 < s.ndim; ++i) first *= s.data[i];
    return {first, s.data[s.ndim - 1]};
  }
  size_t rows() const { return flat_2d()[0]; }
  size_t cols() const { return flat_2d()[1]; }
  size_t numel() const { return shape_numel(shape()); }
};

struct QuantConfig {  // reference common.h:599-622
  bool force_pow_2_scales = false;
  float amax_epsilon = 0.0f;
  NVTETensor noop_tensor = nullptr;
  int block_scale_format = 0;  // deprecated attribute, stored only
  NVTETensor rng_state = nullptr;
  bool nvfp4_2d_quantization = false;
  bool stochastic_rounding = false;
  bool use_fast_math = false;
  uint8_t nvfp4_4over6_mode = 0;
  bool nvfp4_4over6_err_use_fast_math = false;
  uint64_t peer_rowwise[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // extension: {count, pointers...}
  NVTEB200SymmetricExchange symm{};                          // extension: world == 0 -> unset
};

struct MatmulConfig {  // reference gemm/config.h:17-37
  NVTETensor bias_tensor = nullptr;
  NVTETensor dbias_tensor = nullptr;
  bool with_gelu_epilogue = false;
  bool with_dgelu_epilogue = false;
  NVTETensor epilogue_aux_tensor = nullptr;
  bool use_split_accumulator = false;
  int32_t sm_count = 0;
  // extension: output rows go to D in blocks {rows per block, block stride in rows, first row}
  int32_t out_row_blocks[3] = {0, 0, 0};
  NVTEB200ArrivalFlags arrival{};   // extension: world == 0 -> unset
  NVTEB200ScatterOutput scatter{};  // extension: world == 0 -> unset
};

// Handle table lookups (tensor_api.cpp).
Tensor *lookup(NVTETensor t);         // nullptr if invalid
Tensor &lookup_checked(NVTETensor t); // throws if invalid

// Device queries, cached per device (runtime.cpp).
int sm_arch(int device = -1);
int sm_count(int device = -1);
void require_blackwell(const char *what);

// FFI-safe error slot.
void set_last_error(const char *msg);

inline size_t ceil_div(size_t a, size_t b) { return (a + b - 1) / b; }
inline size_t round_up(size_t a, size_t b) { return ceil_div(a, b) * b; }

// ---- kernel launch entry points, one group per .cu ---------------------------
namespace mxfp8 {
void quantize(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream);
// every (ins[i] -> outs[i]) in one persistent launch when the list allows it, else one launch each
void quantize_multi(const Tensor *const *ins, Tensor *const *outs, size_t n, const Tensor *noop,
                    cudaStream_t stream);
void dequantize(const Tensor &in, Tensor &out, cudaStream_t stream);
// one tensor as the TMA kernel (quantize_mxfp8_tma.cu) sees it
struct TmaTensor {
  NVTEDType in_dtype;
  const void *in;
  size_t rows, cols;
  uint8_t *out_row, *out_col;   // null = direction not produced
  uint8_t *sc_row, *sc_col;
  size_t ld_row, ld_col;
  float *amax;
  float *dbias_ws;
};
bool quantize_tma(const TmaTensor &t, bool e5m2, bool swz, const float *noop, cudaStream_t stream);
bool quantize_tma_multi(const TmaTensor *ts, size_t n, bool e5m2, bool swz, const float *noop,
                        cudaStream_t stream);
}  // namespace mxfp8
namespace fp8 {
void quantize(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream);
void quantize_push(const Tensor &in, Tensor &out, const uint64_t *peer_rowwise, const Tensor *noop,
                   cudaStream_t stream);
void dequantize(const Tensor &in, Tensor &out, cudaStream_t stream);
void compute_amax(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream);
void push_amax(const Tensor &out, const uint64_t *peer_slots, const Tensor *noop, cudaStream_t stream);
void compute_scale_from_amax(Tensor &out, const QuantConfig &cfg, cudaStream_t stream);
// symmetric exchange (NVTEB200SymmetricExchange): amax push with the step number, cast that raises arrival flags
void push_amax_step(const Tensor &out, const NVTEB200SymmetricExchange &x, cudaStream_t stream);
void quantize_exchange(const Tensor &in, Tensor &out, const NVTEB200SymmetricExchange &x, cudaStream_t stream);
void delayed_update(const Tensor &hist, const Tensor &scale, Tensor &hist_out, Tensor &scale_out,
                    const char *algo, NVTEDType fp8_dtype, float margin, cudaStream_t stream);
void delayed_update_after_reduction(const Tensor *reduction, const NVTETensor *hists,
                                    const NVTETensor *scales, size_t num_tensors, const char *algo,
                                    NVTEDType fp8_dtype, float margin, cudaStream_t stream);
}  // namespace fp8
namespace nvfp4 {
void quantize(const Tensor &in, Tensor &out, const QuantConfig &cfg, const Tensor *noop,
              bool with_rht, int sign_mask_t, const Tensor *rht_matrix, cudaStream_t stream);
void dequantize(const Tensor &in, Tensor &out, cudaStream_t stream);
void rht_amax(const Tensor &in, Tensor &out, int sign_mask, int sign_mask_t, cudaStream_t stream);
void rht(const Tensor &in, Tensor &out, int sign_mask, int sign_mask_t, cudaStream_t stream);
}  // namespace nvfp4
namespace dbias {
void quantize_dbias(const Tensor &in, Tensor &out, Tensor &dbias, Tensor &workspace, cudaStream_t stream);
// the same with the partial sums in the library's static device array (no allocation; bounded size)
void quantize_dbias_static_scratch(const Tensor &in, Tensor &out, Tensor &dbias, cudaStream_t stream);
}
namespace rs {   // reduce_partials.cu: second half of the reduce-scatter whose first half is the GEMM epilogue
void reduce_partials(const void *partials, uint32_t world, size_t count, NVTEDType dtype, void *out,
                     cudaStream_t stream);
}
namespace norm {
void forward(bool layernorm, const Tensor &x, const Tensor &gamma, const Tensor *beta, float epsilon, Tensor &z,
             Tensor *mu, Tensor &rsigma, Tensor &workspace, bool zero_centered_gamma, cudaStream_t stream);
void backward(bool layernorm, const Tensor &dz, const Tensor &x, const Tensor *add, const Tensor *mu,
              const Tensor &rsigma, const Tensor &gamma, Tensor &dx, Tensor &dgamma, Tensor *dbeta,
              Tensor &workspace, bool zero_centered_gamma, cudaStream_t stream);
}
namespace swizzle {
void swizzle_scales(const Tensor &in, Tensor &out, bool inverse, cudaStream_t stream);
}
namespace gemm {
unsigned long long *debug_counters();
void run(int transa, int transb, float alpha, const Tensor &A, const Tensor &B, float beta,
         const Tensor *C, Tensor &D, Tensor *workspace, const MatmulConfig &cfg,
         cudaStream_t stream);
// One member of nvte_multi_tensor_gemm: D = alpha * op(A) op(B) + beta * D
struct GroupedMember {
  const Tensor *A, *B;
  Tensor *D;
  float alpha, beta;
};
bool run_grouped(int transa, int transb, const std::vector<GroupedMember> &members, Tensor *workspace,
                 const std::vector<MatmulConfig> &cfgs, cudaStream_t stream);
}

}  // namespace teb
