// Handle table, tensor parameter get/set, config objects and device queries.
// Behaviour follows the reference's C API contract
// (transformer_engine/common/transformer_engine.cpp:396-1145): NVTETensor is a
// 1-based index, 0 is invalid; handles are non-owning views.
#include <atomic>
#include <mutex>

#include "core.h"

namespace teb {

namespace {

// Chunked table: chunks are never moved or freed, so lookups are lock-free
// (they read `size_`, published with release semantics after the slot and its
// chunk pointer exist). Allocation and free take the mutex.
class HandleTable {
 public:
  static HandleTable &instance() {
    static HandleTable t;
    return t;
  }
  void allocate(NVTEScalingMode mode, NVTETensor *out, size_t n) {
    std::lock_guard<std::mutex> lock(mu_);
    for (size_t i = 0; i < n; ++i) {
      uintptr_t idx;
      if (!free_.empty()) {
        idx = free_.back();
        free_.pop_back();
      } else {
        const size_t cur = size_.load(std::memory_order_relaxed);
        TEB_CHECK(cur < kMaxTensors, "Cannot allocate a new NVTETensor: ", kMaxTensors,
                  " handles are live. There is probably a handle leak in the application.");
        if (cur % kChunk == 0) chunks_[cur / kChunk] = new Tensor[kChunk];
        idx = cur + 1;
        size_.store(idx, std::memory_order_release);
      }
      slot(idx).clear();
      slot(idx).scaling_mode = mode;
      out[i] = reinterpret_cast<NVTETensor>(idx);
    }
  }
  void release(NVTETensor *t, size_t n) {
    std::lock_guard<std::mutex> lock(mu_);
    for (size_t i = 0; i < n; ++i) {
      const uintptr_t idx = reinterpret_cast<uintptr_t>(t[i]);
      if (idx == 0) continue;
      TEB_CHECK(idx <= size_.load(std::memory_order_relaxed), "Invalid tensor.");
      slot(idx).clear();
      free_.push_back(idx);
    }
  }
  Tensor *get(NVTETensor t) {
    const uintptr_t idx = reinterpret_cast<uintptr_t>(t);
    if (idx != 0 && idx <= size_.load(std::memory_order_acquire)) return &slot(idx);
    return nullptr;
  }

 private:
  static constexpr size_t kChunk = 4096;
  static constexpr size_t kMaxTensors = 1u << 20;
  Tensor &slot(uintptr_t idx) { return chunks_[(idx - 1) / kChunk][(idx - 1) % kChunk]; }
  std::mutex mu_;
  std::atomic<size_t> size_{0};
  Tensor *chunks_[kMaxTensors / kChunk] = {};
  std::vector<uintptr_t> free_;
};

thread_local std::string g_last_error;

}  // namespace

Tensor *lookup(NVTETensor t) { return HandleTable::instance().get(t); }
Tensor &lookup_checked(NVTETensor t) {
  Tensor *p = lookup(t);
  TEB_CHECK(p != nullptr, "Invalid tensor.");
  return *p;
}
void set_last_error(const char *msg) { g_last_error = msg ? msg : ""; }

NVTEShape Tensor::shape() const {
  switch (scaling_mode) {
    case NVTE_DELAYED_TENSOR_SCALING:
    case NVTE_NVFP4_1D_SCALING: {
      if (!has_data() && has_columnwise_data()) {
        // columnwise data is stored transposed: logical = dims[1:] + dims[0]
        NVTEShape r{};
        const NVTEShape &c = columnwise_data.shape;
        for (size_t i = 1; i < c.ndim; ++i) r.data[r.ndim++] = c.data[i];
        if (c.ndim > 0) r.data[r.ndim++] = c.data[0];
        return r;
      }
      return data.shape;
    }
    case NVTE_MXFP8_1D_SCALING:
      if (!has_data() && has_columnwise_data()) return columnwise_data.shape;
      return data.shape;
    default:
      TEB_ERROR("Scaling mode ", static_cast<int>(scaling_mode),
                " is not supported by the B200 low-precision Linear library.");
  }
}

// ---- device queries ----------------------------------------------------------
namespace {
struct DeviceInfo {
  std::once_flag once;
  int arch = 0, sms = 0;
};
DeviceInfo g_devs[64];
DeviceInfo *dev_info(int device) {
  if (device < 0) {
    if (cudaGetDevice(&device) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
  }
  if (device >= 64) return nullptr;
  DeviceInfo &d = g_devs[device];
  std::call_once(d.once, [&] {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) == cudaSuccess) {
      d.arch = 10 * p.major + p.minor;
      d.sms = p.multiProcessorCount;
    } else {
      cudaGetLastError();
    }
  });
  return &d;
}
}  // namespace
int sm_arch(int device) {
  DeviceInfo *d = dev_info(device);
  return d ? d->arch : 0;
}
int sm_count(int device) {
  DeviceInfo *d = dev_info(device);
  return d ? d->sms : 0;
}
void require_blackwell(const char *what) {
  const int arch = sm_arch();
  TEB_CHECK(arch >= 100 && arch < 110, what,
            " requires an sm_100-class GPU (B200); there is no fallback path. Found sm_", arch,
            ".");
}

}  // namespace teb

using namespace teb;

extern "C" {

NVTETensor nvte_create_tensor(NVTEScalingMode scaling_mode) {
  NVTETensor t;
  HandleTable::instance().allocate(scaling_mode, &t, 1);
  return t;
}
void nvte_create_tensors(NVTEScalingMode scaling_mode, NVTETensor *tensors, size_t N) {
  HandleTable::instance().allocate(scaling_mode, tensors, N);
}
void nvte_destroy_tensor(NVTETensor tensor) { HandleTable::instance().release(&tensor, 1); }
void nvte_destroy_tensors(NVTETensor *tensors, size_t N) {
  HandleTable::instance().release(tensors, N);
}

NVTEDType nvte_tensor_type(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? t->dtype() : kNVTEFloat32;
}
NVTEShape nvte_make_shape(const size_t *data, size_t ndim) {
  NVTEShape r{};
  TEB_CHECK(ndim <= 15, "Too many dims for NVTEShape (requested: ", ndim, ", max: 15)");
  for (size_t i = 0; i < ndim; ++i) r.data[i] = data ? data[i] : 0;
  r.ndim = ndim;
  return r;
}
NVTEShape nvte_tensor_shape(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  if (!t) TEB_ERROR("Invalid tensor: received null pointer in nvte_tensor_shape");
  return t->shape();
}
NVTEShape nvte_tensor_columnwise_shape(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  if (!t) TEB_ERROR("Invalid tensor: received null pointer in nvte_tensor_columnwise_shape");
  return t->columnwise_data.shape;
}
size_t nvte_tensor_ndims(const NVTETensor tensor) { return nvte_tensor_shape(tensor).ndim; }
size_t nvte_tensor_size(const NVTETensor tensor, const size_t dim) {
  const NVTEShape s = nvte_tensor_shape(tensor);
  TEB_CHECK(dim < s.ndim, "Attempted to access index ", dim, " in a shape array with ", s.ndim,
            " entries");
  return s.data[dim];
}
size_t nvte_tensor_numel(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? t->numel() : 0;
}
size_t nvte_tensor_element_size_bits(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? dtype_bits(t->dtype()) : 8 * sizeof(float);
}
size_t nvte_tensor_element_size(const NVTETensor tensor) {
  return (nvte_tensor_element_size_bits(tensor) + 7) / 8;
}
size_t nvte_tensor_size_bytes(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? (t->numel() * dtype_bits(t->dtype()) + 7) / 8 : 0;
}
void *nvte_tensor_data(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? t->data.dptr : nullptr;
}
void *nvte_tensor_columnwise_data(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? t->columnwise_data.dptr : nullptr;
}
float *nvte_tensor_amax(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  if (!t) return nullptr;
  TEB_CHECK(t->amax.dtype == kNVTEFloat32, "Tensor's amax must have Float32 type!");
  return t->amax.as<float>();
}
float *nvte_tensor_scale(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  if (!t) return nullptr;
  TEB_CHECK(t->scale.dtype == kNVTEFloat32, "Tensor's scale must have Float32 type!");
  return t->scale.as<float>();
}
float *nvte_tensor_scale_inv(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  return t ? t->scale_inv.as<float>() : nullptr;
}
NVTEShape nvte_tensor_scale_inv_shape(const NVTETensor tensor) {
  Tensor *t = lookup(tensor);
  if (!t) return nvte_make_shape(nullptr, 0);
  return t->scale_inv.shape;
}
NVTEScalingMode nvte_tensor_scaling_mode(const NVTETensor tensor) {
  if (tensor == nullptr) return NVTE_DELAYED_TENSOR_SCALING;
  return lookup_checked(tensor).scaling_mode;
}

static Buffer *buffer_param(Tensor &t, NVTETensorParam p) {
  switch (p) {
    case kNVTERowwiseData: return &t.data;
    case kNVTEColumnwiseData: return &t.columnwise_data;
    case kNVTEScale: return &t.scale;
    case kNVTEAmax: return &t.amax;
    case kNVTERowwiseScaleInv: return &t.scale_inv;
    case kNVTEColumnwiseScaleInv: return &t.columnwise_scale_inv;
    case kNVTEColumnwiseAmax: return &t.columnwise_amax;
    default: return nullptr;
  }
}
static size_t tensor_param_size(NVTETensorParam p) {
  if (p <= kNVTEColumnwiseAmax) return sizeof(NVTEBasicTensor);
  if (p == kNVTENVFP4E4M3Max) return sizeof(int);
  return sizeof(uint8_t);
}

void nvte_set_tensor_param(NVTETensor *tensor, NVTETensorParam param_name,
                           const NVTEBasicTensor *param) {
  TEB_CHECK(tensor != nullptr, "Tensor pointer can't be NULL.");
  Tensor &t = lookup_checked(*tensor);
  Buffer *b = buffer_param(t, param_name);
  if (!b)
    TEB_ERROR("Unsupported tensor parameter (", static_cast<int>(param_name),
              "). Consider using nvte_set_tensor_param_v2 instead.");
  *b = Buffer(*param);
}
NVTEBasicTensor nvte_get_tensor_param(const NVTETensor tensor, NVTETensorParam param_name) {
  if (tensor == nullptr) return NVTEBasicTensor{nullptr, kNVTEFloat32, nvte_make_shape(nullptr, 1)};
  Tensor &t = lookup_checked(tensor);
  Buffer *b = buffer_param(t, param_name);
  if (!b)
    TEB_ERROR("Unsupported tensor parameter (", static_cast<int>(param_name),
              "). Consider using nvte_get_tensor_param_v2 instead.");
  return b->basic();
}
void nvte_set_tensor_param_v2(NVTETensor tensor, NVTETensorParam param, const void *buf,
                              size_t size_in_bytes) {
  TEB_CHECK(static_cast<int>(param) >= 0 && param < kNVTENumTensorParams,
            "Invalid NVTETensorParam (got ", static_cast<int>(param), ")");
  TEB_CHECK(tensor != nullptr, "Tensor pointer can't be NULL.");
  Tensor &t = lookup_checked(tensor);
  const size_t need = tensor_param_size(param);
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for tensor parameter (parameter ",
            static_cast<int>(param), " needs ", need, " bytes, but buffer has ", size_in_bytes,
            " bytes)");
  TEB_CHECK(buf != nullptr, "Invalid buffer (got NULL)");
  if (Buffer *b = buffer_param(t, param)) {
    *b = Buffer(*reinterpret_cast<const NVTEBasicTensor *>(buf));
    return;
  }
  switch (param) {
    case kNVTEWithGEMMSwizzledScales:
      t.with_gemm_swizzled_scales = *reinterpret_cast<const uint8_t *>(buf) != 0;
      break;
    case kNVTERowScaledNVFP4:
      t.row_scaled_nvfp4 = *reinterpret_cast<const uint8_t *>(buf) != 0;
      break;
    case kNVTENVFP4E4M3Max:
      std::memcpy(&t.nvfp4_e4m3_max, buf, sizeof(int));
      TEB_CHECK(t.nvfp4_e4m3_max == 448 || t.nvfp4_e4m3_max == 256,
                "Unsupported NVFP4 E4M3 max (got ", t.nvfp4_e4m3_max, ")");
      break;
    default:
      TEB_ERROR("Unsupported tensor parameter (", static_cast<int>(param), ")");
  }
}
void nvte_get_tensor_param_v2(const NVTETensor tensor, NVTETensorParam param, void *buf,
                              size_t size_in_bytes, size_t *size_written) {
  TEB_CHECK(static_cast<int>(param) >= 0 && param < kNVTENumTensorParams,
            "Invalid NVTETensorParam (got ", static_cast<int>(param), ")");
  const size_t need = tensor_param_size(param);
  if (size_written) *size_written = need;
  if (buf == nullptr) return;
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for tensor parameter (parameter ",
            static_cast<int>(param), " needs ", need, " bytes, but buffer has ", size_in_bytes,
            " bytes)");
  Tensor dummy;
  Tensor *t = lookup(tensor);
  if (!t) t = &dummy;
  if (Buffer *b = buffer_param(*t, param)) {
    *reinterpret_cast<NVTEBasicTensor *>(buf) = b->basic();
    return;
  }
  switch (param) {
    case kNVTEWithGEMMSwizzledScales:
      *reinterpret_cast<uint8_t *>(buf) = t->with_gemm_swizzled_scales;
      break;
    case kNVTERowScaledNVFP4:
      *reinterpret_cast<uint8_t *>(buf) = t->row_scaled_nvfp4;
      break;
    case kNVTENVFP4E4M3Max:
      std::memcpy(buf, &t->nvfp4_e4m3_max, sizeof(int));
      break;
    default:
      TEB_ERROR("Unsupported tensor parameter (", static_cast<int>(param), ")");
  }
}

void nvte_zero_tensor(const NVTETensor tensor, cudaStream_t stream) {
  if (tensor == nullptr) return;
  Tensor &t = lookup_checked(tensor);
  if (t.data.dptr) TEB_CHECK_CUDA(cudaMemsetAsync(t.data.dptr, 0, t.data.bytes(), stream));
  if (t.amax.dptr) TEB_CHECK_CUDA(cudaMemsetAsync(t.amax.dptr, 0, t.amax.bytes(), stream));
}
void nvte_memset(void *ptr, int value, size_t size_in_bytes, cudaStream_t stream) {
  if (ptr == nullptr || size_in_bytes == 0) return;
  TEB_CHECK_CUDA(cudaMemsetAsync(ptr, value, size_in_bytes, stream));
}
int nvte_is_non_tn_fp8_gemm_supported(void) {
  // Blackwell tensor cores take MN-major FP8 operands, so per-tensor FP8 needs
  // no transposed copy (reference transformer_engine.cpp:1131-1145).
  const int arch = sm_arch();
  return (arch >= 100 && arch < 120) ? 1 : 0;
}

// ---- quantization config -----------------------------------------------------
static const size_t kQuantAttrSize[kNVTEQuantizationConfigNumAttributes] = {
    1, sizeof(float), sizeof(NVTETensor), sizeof(int), sizeof(NVTETensor), 1, 1, 1, 1, 1};
// extension attributes (0x1000 range) carry raw device addresses: exact size or nothing
static const size_t kPeerAttrSize = 9 * sizeof(uint64_t);

NVTEQuantizationConfig nvte_create_quantization_config(void) { return new QuantConfig; }
void nvte_destroy_quantization_config(NVTEQuantizationConfig config) {
  delete reinterpret_cast<QuantConfig *>(config);
}
void nvte_get_quantization_config_attribute(NVTEQuantizationConfig config,
                                            NVTEQuantizationConfigAttribute attr, void *buf,
                                            size_t size_in_bytes, size_t *size_written) {
  if (attr == kNVTEQuantizationConfigB200PeerRowwiseData) {
    TEB_CHECK(size_written != nullptr, "Invalid size_written (got NULL)");
    *size_written = kPeerAttrSize;
    if (buf == nullptr) return;
    TEB_CHECK(size_in_bytes == kPeerAttrSize && config != nullptr,
              "kNVTEQuantizationConfigB200PeerRowwiseData needs a ", kPeerAttrSize, "-byte buffer");
    std::memcpy(buf, reinterpret_cast<const QuantConfig *>(config)->peer_rowwise, kPeerAttrSize);
    return;
  }
  if (attr == kNVTEQuantizationConfigB200SymmetricExchange) {
    TEB_CHECK(size_written != nullptr, "Invalid size_written (got NULL)");
    *size_written = sizeof(NVTEB200SymmetricExchange);
    if (buf == nullptr) return;
    TEB_CHECK(size_in_bytes == sizeof(NVTEB200SymmetricExchange) && config != nullptr,
              "kNVTEQuantizationConfigB200SymmetricExchange needs a ", sizeof(NVTEB200SymmetricExchange), "-byte buffer");
    std::memcpy(buf, &reinterpret_cast<const QuantConfig *>(config)->symm, sizeof(NVTEB200SymmetricExchange));
    return;
  }
  TEB_CHECK(static_cast<int>(attr) >= 0 && attr < kNVTEQuantizationConfigNumAttributes,
            "Invalid NVTEQuantizationConfigAttribute (got ", static_cast<int>(attr), ")");
  TEB_CHECK(size_written != nullptr, "Invalid size_written (got NULL)");
  const size_t need = kQuantAttrSize[attr];
  *size_written = need;
  if (buf == nullptr) return;
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for quantization config attribute ",
            "(attribute ", static_cast<int>(attr), " needs ", need, " bytes, but buffer has ",
            size_in_bytes, " bytes)");
  TEB_CHECK(config != nullptr, "Invalid NVTEQuantizationConfig (got NULL)");
  const QuantConfig &c = *reinterpret_cast<const QuantConfig *>(config);
  auto put8 = [&](bool v) { *reinterpret_cast<uint8_t *>(buf) = v; };
  switch (attr) {
    case kNVTEQuantizationConfigForcePow2Scales: put8(c.force_pow_2_scales); break;
    case kNVTEQuantizationConfigAmaxEpsilon: std::memcpy(buf, &c.amax_epsilon, need); break;
    case kNVTEQuantizationConfigNoopTensor: std::memcpy(buf, &c.noop_tensor, need); break;
    case kNVTEQuantizationConfigFloat8BlockScaleTensorFormat:
      std::memcpy(buf, &c.block_scale_format, need);
      break;
    case kNVTEQuantizationConfigRNGState: std::memcpy(buf, &c.rng_state, need); break;
    case kNVTEQuantizationConfigNVFP42DQuantization: put8(c.nvfp4_2d_quantization); break;
    case kNVTEQuantizationConfigStochasticRounding: put8(c.stochastic_rounding); break;
    case kNVTEQuantizationConfigUseFastMath: put8(c.use_fast_math); break;
    case kNVTEQuantizationConfigNVFP44Over6Mode:
      *reinterpret_cast<uint8_t *>(buf) = c.nvfp4_4over6_mode;
      break;
    case kNVTEQuantizationConfigNVFP44Over6ErrUseFastMath:
      put8(c.nvfp4_4over6_err_use_fast_math);
      break;
    default: break;
  }
}
void nvte_set_quantization_config_attribute(NVTEQuantizationConfig config,
                                            NVTEQuantizationConfigAttribute attr,
                                            const void *buf, size_t size_in_bytes) {
  if (attr == kNVTEQuantizationConfigB200PeerRowwiseData) {
    TEB_CHECK(buf != nullptr && config != nullptr && size_in_bytes == kPeerAttrSize,
              "kNVTEQuantizationConfigB200PeerRowwiseData takes exactly ", kPeerAttrSize,
              " bytes (uint64_t[9] = count, addresses), got ", size_in_bytes);
    uint64_t v[9];
    std::memcpy(v, buf, kPeerAttrSize);
    TEB_CHECK(v[0] <= 8, "At most 8 peer destinations (got ", v[0], ")");
    std::memcpy(reinterpret_cast<QuantConfig *>(config)->peer_rowwise, v, kPeerAttrSize);
    return;
  }
  if (attr == kNVTEQuantizationConfigB200SymmetricExchange) {
    TEB_CHECK(buf != nullptr && config != nullptr && size_in_bytes == sizeof(NVTEB200SymmetricExchange),
              "kNVTEQuantizationConfigB200SymmetricExchange takes exactly ", sizeof(NVTEB200SymmetricExchange),
              " bytes (NVTEB200SymmetricExchange), got ", size_in_bytes);
    NVTEB200SymmetricExchange x;
    std::memcpy(&x, buf, sizeof(x));
    TEB_CHECK(x.world <= 8 && x.rank < (x.world ? x.world : 1), "Symmetric exchange: world <= 8, rank < world");
    // (flags_per_rank == 0: amax exchange only -- the scale kernel advances the step, the GEMM pushes the rows)
    TEB_CHECK(x.world == 0 || (x.rows_per_flag % 256 == 0 && x.rows_per_flag > 0 &&
                               x.step != 0 && x.done != 0 && x.error != 0),
              "Symmetric exchange: rows_per_flag must be a positive multiple of 256 and step / done / error set");
    for (uint32_t r = 0; r < x.world; ++r)
      TEB_CHECK(x.gather[r] != 0 && x.gather[r] % 16 == 0 && x.flags[r] != 0 && x.amax_slots[r] % 8 == 0 &&
                    x.amax_slots[r] != 0,
                "Symmetric exchange: null or misaligned address for rank ", r);
    reinterpret_cast<QuantConfig *>(config)->symm = x;
    return;
  }
  TEB_CHECK(static_cast<int>(attr) >= 0 && attr < kNVTEQuantizationConfigNumAttributes,
            "Invalid NVTEQuantizationConfigAttribute (got ", static_cast<int>(attr), ")");
  const size_t need = kQuantAttrSize[attr];
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for quantization config attribute ",
            "(attribute ", static_cast<int>(attr), " needs ", need, " bytes, but buffer has ",
            size_in_bytes, " bytes)");
  TEB_CHECK(buf != nullptr, "Invalid buffer (got NULL)");
  TEB_CHECK(config != nullptr, "Invalid NVTEQuantizationConfig (got NULL)");
  QuantConfig &c = *reinterpret_cast<QuantConfig *>(config);
  auto get8 = [&]() { return *reinterpret_cast<const uint8_t *>(buf) != 0; };
  switch (attr) {
    case kNVTEQuantizationConfigForcePow2Scales: c.force_pow_2_scales = get8(); break;
    case kNVTEQuantizationConfigAmaxEpsilon: std::memcpy(&c.amax_epsilon, buf, need); break;
    case kNVTEQuantizationConfigNoopTensor: std::memcpy(&c.noop_tensor, buf, need); break;
    case kNVTEQuantizationConfigFloat8BlockScaleTensorFormat:
      std::memcpy(&c.block_scale_format, buf, need);
      break;
    case kNVTEQuantizationConfigRNGState: std::memcpy(&c.rng_state, buf, need); break;
    case kNVTEQuantizationConfigNVFP42DQuantization: c.nvfp4_2d_quantization = get8(); break;
    case kNVTEQuantizationConfigStochasticRounding: c.stochastic_rounding = get8(); break;
    case kNVTEQuantizationConfigUseFastMath: c.use_fast_math = get8(); break;
    case kNVTEQuantizationConfigNVFP44Over6Mode:
      c.nvfp4_4over6_mode = *reinterpret_cast<const uint8_t *>(buf);
      break;
    case kNVTEQuantizationConfigNVFP44Over6ErrUseFastMath:
      c.nvfp4_4over6_err_use_fast_math = get8();
      break;
    default: break;
  }
}

// ---- matmul config -------------------------------------------------------------
static const size_t kMatmulAttrSize[kNVTEMatmulConfigNumAttributes] = {
    sizeof(NVTETensor), sizeof(NVTETensor), 1, 1, sizeof(NVTETensor), 1, sizeof(int32_t)};
static const size_t kRowBlocksAttrSize = 3 * sizeof(int32_t);

NVTEMatmulConfig nvte_create_matmul_config(void) { return new MatmulConfig; }
void nvte_destroy_matmul_config(NVTEMatmulConfig config) {
  delete reinterpret_cast<MatmulConfig *>(config);
}
void nvte_get_matmul_config_attribute(NVTEMatmulConfig config, NVTEMatmulConfigAttribute attr,
                                      void *buf, size_t size_in_bytes, size_t *size_written) {
  if (attr == kNVTEMatmulConfigB200OutputRowBlocks) {
    if (size_written) *size_written = kRowBlocksAttrSize;
    if (buf == nullptr) return;
    TEB_CHECK(size_in_bytes == kRowBlocksAttrSize && config != nullptr,
              "kNVTEMatmulConfigB200OutputRowBlocks needs a ", kRowBlocksAttrSize, "-byte buffer");
    std::memcpy(buf, reinterpret_cast<const MatmulConfig *>(config)->out_row_blocks, kRowBlocksAttrSize);
    return;
  }
  if (attr == kNVTEMatmulConfigB200ArrivalFlags) {
    if (size_written) *size_written = sizeof(NVTEB200ArrivalFlags);
    if (buf == nullptr) return;
    TEB_CHECK(size_in_bytes == sizeof(NVTEB200ArrivalFlags) && config != nullptr,
              "kNVTEMatmulConfigB200ArrivalFlags needs a ", sizeof(NVTEB200ArrivalFlags), "-byte buffer");
    std::memcpy(buf, &reinterpret_cast<const MatmulConfig *>(config)->arrival, sizeof(NVTEB200ArrivalFlags));
    return;
  }
  if (attr == kNVTEMatmulConfigB200ScatterOutput) {
    if (size_written) *size_written = sizeof(NVTEB200ScatterOutput);
    if (buf == nullptr) return;
    TEB_CHECK(size_in_bytes == sizeof(NVTEB200ScatterOutput) && config != nullptr,
              "kNVTEMatmulConfigB200ScatterOutput needs a ", sizeof(NVTEB200ScatterOutput), "-byte buffer");
    std::memcpy(buf, &reinterpret_cast<const MatmulConfig *>(config)->scatter, sizeof(NVTEB200ScatterOutput));
    return;
  }
  TEB_CHECK(static_cast<int>(attr) >= 0 && attr < kNVTEMatmulConfigNumAttributes,
            "Invalid NVTEMatmulConfigAttribute (got ", static_cast<int>(attr), ")");
  const size_t need = kMatmulAttrSize[attr];
  if (size_written) *size_written = need;
  if (buf == nullptr) return;
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for matmul config attribute (attribute ",
            static_cast<int>(attr), " needs ", need, " bytes, but buffer has ", size_in_bytes,
            " bytes)");
  TEB_CHECK(config != nullptr, "Invalid NVTEMatmulConfig (got NULL)");
  const MatmulConfig &c = *reinterpret_cast<const MatmulConfig *>(config);
  switch (attr) {
    case kNVTEMatmulConfigBiasTensor: std::memcpy(buf, &c.bias_tensor, need); break;
    case kNVTEMatmulConfigDBiasTensor: std::memcpy(buf, &c.dbias_tensor, need); break;
    case kNVTEMatmulConfigWithGELUEpilogue:
      *reinterpret_cast<uint8_t *>(buf) = c.with_gelu_epilogue;
      break;
    case kNVTEMatmulConfigWithDGELUEpilogue:
      *reinterpret_cast<uint8_t *>(buf) = c.with_dgelu_epilogue;
      break;
    case kNVTEMatmulConfigEpilogueAuxTensor: std::memcpy(buf, &c.epilogue_aux_tensor, need); break;
    case kNVTEMatmulConfigUseSplitAccumulator:
      *reinterpret_cast<uint8_t *>(buf) = c.use_split_accumulator;
      break;
    case kNVTEMatmulConfigSMCount: std::memcpy(buf, &c.sm_count, need); break;
    default: break;
  }
}
void nvte_set_matmul_config_attribute(NVTEMatmulConfig config, NVTEMatmulConfigAttribute attr,
                                      const void *buf, size_t size_in_bytes) {
  if (attr == kNVTEMatmulConfigB200OutputRowBlocks) {
    TEB_CHECK(buf != nullptr && config != nullptr && size_in_bytes == kRowBlocksAttrSize,
              "kNVTEMatmulConfigB200OutputRowBlocks takes exactly ", kRowBlocksAttrSize,
              " bytes (int32_t[3]), got ", size_in_bytes);
    std::memcpy(reinterpret_cast<MatmulConfig *>(config)->out_row_blocks, buf, kRowBlocksAttrSize);
    return;
  }
  if (attr == kNVTEMatmulConfigB200ArrivalFlags) {
    TEB_CHECK(buf != nullptr && config != nullptr && size_in_bytes == sizeof(NVTEB200ArrivalFlags),
              "kNVTEMatmulConfigB200ArrivalFlags takes exactly ", sizeof(NVTEB200ArrivalFlags), " bytes, got ",
              size_in_bytes);
    NVTEB200ArrivalFlags a;
    std::memcpy(&a, buf, sizeof(a));
    TEB_CHECK(a.world <= 8 && (a.world == 0 || (a.rank < a.world && a.flags != 0 && a.step != 0 && a.error != 0 &&
                                                a.rows_per_flag % 256 == 0 && a.rows_per_flag > 0 &&
                                                a.flags_per_rank > 0)),
              "Arrival flags: world <= 8, rank < world, rows_per_flag a positive multiple of 256, addresses set");
    reinterpret_cast<MatmulConfig *>(config)->arrival = a;
    return;
  }
  if (attr == kNVTEMatmulConfigB200ScatterOutput) {
    TEB_CHECK(buf != nullptr && config != nullptr && size_in_bytes == sizeof(NVTEB200ScatterOutput),
              "kNVTEMatmulConfigB200ScatterOutput takes exactly ", sizeof(NVTEB200ScatterOutput), " bytes, got ",
              size_in_bytes);
    NVTEB200ScatterOutput s;
    std::memcpy(&s, buf, sizeof(s));
    TEB_CHECK(s.world <= 8 && (s.world == 0 || (s.rank < s.world && s.rows_per_rank > 0)),
              "Scatter output: world <= 8, rank < world, rows_per_rank > 0");
    for (uint32_t q = 0; q < s.world; ++q)
      TEB_CHECK(s.dst[q] != 0 && s.dst[q] % 32 == 0, "Scatter output: null or misaligned destination for rank ", q);
    reinterpret_cast<MatmulConfig *>(config)->scatter = s;
    return;
  }
  TEB_CHECK(static_cast<int>(attr) >= 0 && attr < kNVTEMatmulConfigNumAttributes,
            "Invalid NVTEMatmulConfigAttribute (got ", static_cast<int>(attr), ")");
  const size_t need = kMatmulAttrSize[attr];
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for matmul config attribute (attribute ",
            static_cast<int>(attr), " needs ", need, " bytes, but buffer has ", size_in_bytes,
            " bytes)");
  TEB_CHECK(buf != nullptr, "Invalid buffer (got NULL)");
  TEB_CHECK(config != nullptr, "Invalid NVTEMatmulConfig (got NULL)");
  MatmulConfig &c = *reinterpret_cast<MatmulConfig *>(config);
  switch (attr) {
    case kNVTEMatmulConfigBiasTensor: std::memcpy(&c.bias_tensor, buf, need); break;
    case kNVTEMatmulConfigDBiasTensor: std::memcpy(&c.dbias_tensor, buf, need); break;
    case kNVTEMatmulConfigWithGELUEpilogue:
      c.with_gelu_epilogue = *reinterpret_cast<const uint8_t *>(buf) != 0;
      break;
    case kNVTEMatmulConfigWithDGELUEpilogue:
      c.with_dgelu_epilogue = *reinterpret_cast<const uint8_t *>(buf) != 0;
      break;
    case kNVTEMatmulConfigEpilogueAuxTensor: std::memcpy(&c.epilogue_aux_tensor, buf, need); break;
    case kNVTEMatmulConfigUseSplitAccumulator:
      c.use_split_accumulator = *reinterpret_cast<const uint8_t *>(buf) != 0;
      break;
    case kNVTEMatmulConfigSMCount: std::memcpy(&c.sm_count, buf, need); break;
    default: break;
  }
}


// ---- grouped tensors (TE/transformer_engine.h:536-644) -------------------------------------------
// Host-side views only: the grouped (MoE) quantize / GEMM kernels are a "next" row,
// but the handle API is part of the boundary (the reference's tests/cpp/test_common.cu links it).
namespace {
struct GroupedTensor {
  NVTEScalingMode scaling_mode;
  size_t num_tensors;
  NVTEShape logical_shape;
  NVTEBasicTensor params[kNVTENumGroupedTensorParams];
  bool with_gemm_swizzled_scales = false;
};
GroupedTensor &grouped(NVTEGroupedTensor t) {
  TEB_CHECK(t != nullptr, "Invalid grouped tensor (got NULL)");
  return *reinterpret_cast<GroupedTensor *>(t);
}
}  // namespace

NVTEGroupedTensor nvte_create_grouped_tensor(NVTEScalingMode scaling_mode, size_t num_tensors,
                                             NVTEShape logical_shape) {
  TEB_CHECK(num_tensors > 0, "Number of tensors must be greater than 0");
  TEB_CHECK(logical_shape.ndim == 2, "Logical shape must be 2D");
  auto *g = new GroupedTensor{};
  g->scaling_mode = scaling_mode;
  g->num_tensors = num_tensors;
  g->logical_shape = logical_shape;
  for (auto &b : g->params) b = NVTEBasicTensor{nullptr, kNVTEFloat32, nvte_make_shape(nullptr, 1)};
  return g;
}
void nvte_destroy_grouped_tensor(NVTEGroupedTensor tensor) { delete reinterpret_cast<GroupedTensor *>(tensor); }
void nvte_set_grouped_tensor_param(NVTEGroupedTensor tensor, NVTEGroupedTensorParam param, const void *buf,
                                   size_t size_in_bytes) {
  GroupedTensor &g = grouped(tensor);
  TEB_CHECK(static_cast<int>(param) >= 0 && param < kNVTENumGroupedTensorParams,
            "Invalid NVTEGroupedTensorParam (got ", static_cast<int>(param), ")");
  TEB_CHECK(buf != nullptr, "Invalid buffer (got NULL)");
  if (param == kNVTEGroupedWithGEMMSwizzledScales) {
    TEB_CHECK(size_in_bytes >= 1, "Buffer is too small for grouped tensor parameter");
    g.with_gemm_swizzled_scales = *reinterpret_cast<const uint8_t *>(buf) != 0;
    return;
  }
  TEB_CHECK(size_in_bytes >= sizeof(NVTEBasicTensor), "Buffer is too small for grouped tensor parameter (needs ",
            sizeof(NVTEBasicTensor), " bytes, but buffer has ", size_in_bytes, " bytes)");
  g.params[param] = *reinterpret_cast<const NVTEBasicTensor *>(buf);
}
void nvte_get_grouped_tensor_param(const NVTEGroupedTensor tensor, NVTEGroupedTensorParam param, void *buf,
                                   size_t size_in_bytes, size_t *size_written) {
  const GroupedTensor &g = grouped(tensor);
  TEB_CHECK(static_cast<int>(param) >= 0 && param < kNVTENumGroupedTensorParams,
            "Invalid NVTEGroupedTensorParam (got ", static_cast<int>(param), ")");
  const size_t need = param == kNVTEGroupedWithGEMMSwizzledScales ? 1 : sizeof(NVTEBasicTensor);
  if (size_written != nullptr) *size_written = need;
  if (buf == nullptr) return;
  TEB_CHECK(size_in_bytes >= need, "Buffer is too small for grouped tensor parameter");
  if (param == kNVTEGroupedWithGEMMSwizzledScales)
    *reinterpret_cast<uint8_t *>(buf) = g.with_gemm_swizzled_scales ? 1 : 0;
  else
    *reinterpret_cast<NVTEBasicTensor *>(buf) = g.params[param];
}
size_t nvte_grouped_tensor_num_tensors(const NVTEGroupedTensor tensor) {
  return tensor == nullptr ? 0 : grouped(tensor).num_tensors;
}
NVTEDType nvte_grouped_tensor_type(const NVTEGroupedTensor tensor) {
  if (tensor == nullptr) return kNVTEFloat32;
  const GroupedTensor &g = grouped(tensor);
  const NVTEBasicTensor &r = g.params[kNVTEGroupedRowwiseData];
  return r.data_ptr != nullptr ? r.dtype : g.params[kNVTEGroupedColumnwiseData].dtype;
}
NVTEScalingMode nvte_grouped_tensor_scaling_mode(const NVTEGroupedTensor tensor) {
  return tensor == nullptr ? NVTE_DELAYED_TENSOR_SCALING : grouped(tensor).scaling_mode;
}
NVTEShape nvte_get_grouped_tensor_logical_shape(const NVTEGroupedTensor tensor) {
  if (tensor == nullptr) return nvte_make_shape(nullptr, 1);
  return grouped(tensor).logical_shape;
}

// ---- streams (TE/multi_stream.h:21) ---------------------------------------------------------------
// This path enqueues everything on the caller's stream; the helper streams of the reference's
// multi-tensor entry points do not exist here.
int nvte_get_num_compute_streams(void) { return 1; }

const char *nvte_b200_last_error(void) { return g_last_error.c_str(); }
int nvte_b200_sm_arch(void) { return sm_arch(); }
int nvte_b200_sm_count(void) { return sm_count(); }

// FFI fast path: a buffer parameter from scalars (no NVTEBasicTensor to marshal). Returns 0, or 1
// with the message in nvte_b200_last_error(). Up to 4 dimensions; more go through the struct form.
int nvte_b200_set_tensor_param_flat(NVTETensor tensor, int param, void *ptr, int dtype, int ndim, size_t d0,
                                    size_t d1, size_t d2, size_t d3) {
  try {
    TEB_CHECK(ndim >= 0 && ndim <= 4, "nvte_b200_set_tensor_param_flat takes at most 4 dimensions");
    NVTEBasicTensor b{};
    b.data_ptr = ptr;
    b.dtype = static_cast<NVTEDType>(dtype);
    b.shape.ndim = static_cast<size_t>(ndim);
    const size_t d[4] = {d0, d1, d2, d3};
    for (int i = 0; i < ndim; ++i) b.shape.data[i] = d[i];
    nvte_set_tensor_param_v2(tensor, static_cast<NVTETensorParam>(param), &b, sizeof(b));
    return 0;
  } catch (const std::exception &e) {
    set_last_error(e.what());
    return 1;
  }
}
// nvte_create_tensor(NVTE_DELAYED_TENSOR_SCALING) + rowwise data, in one call; nullptr on error
NVTETensor nvte_b200_create_plain_tensor(void *ptr, int dtype, int ndim, size_t d0, size_t d1, size_t d2, size_t d3) {
  NVTETensor t = nullptr;
  try {
    t = nvte_create_tensor(NVTE_DELAYED_TENSOR_SCALING);
  } catch (const std::exception &e) {
    set_last_error(e.what());
    return nullptr;
  }
  if (nvte_b200_set_tensor_param_flat(t, kNVTERowwiseData, ptr, dtype, ndim, d0, d1, d2, d3) != 0) {
    nvte_destroy_tensor(t);
    return nullptr;
  }
  return t;
}

}  // extern "C"

// ---- C++ helpers the reference's torch extension imports (libtransformer_engine.version:5-19) ------
// Same namespaces, names and parameter types as the reference (util/cuda_runtime.h:22-95,
// common.h:1160), so the mangled symbols match.
namespace transformer_engine {
enum class DType { kByte = 0, kInt16, kInt32, kInt64, kFloat32, kFloat16, kBFloat16, kFloat8E4M3, kFloat8E5M2,
                   kFloat8E8M0, kFloat4E2M1, kNumTypes };   // TE/transformer_engine.h:660-673
namespace cuda {
__attribute__((visibility("default"))) int current_device() {
  int d = 0;
  TEB_CHECK_CUDA(cudaGetDevice(&d));
  return d;
}
__attribute__((visibility("default"))) int sm_arch(int device_id) {
  (void)device_id;
  return teb::sm_arch();
}
__attribute__((visibility("default"))) int sm_count(int device_id) {
  (void)device_id;
  return teb::sm_count();
}
__attribute__((visibility("default"))) bool supports_multicast(int device_id) {
  (void)device_id;
  return false;   // no userbuffers / multicast objects on this path
}
__attribute__((visibility("default"))) void stream_priority_range(int *low_priority, int *high_priority,
                                                                  int device_id) {
  (void)device_id;
  TEB_CHECK_CUDA(cudaDeviceGetStreamPriorityRange(low_priority, high_priority));
}
__attribute__((visibility("default"))) size_t cublas_version() { return 0; }   // no cuBLAS on this path
}  // namespace cuda
__attribute__((visibility("default"))) size_t typeToSize(const DType type) {
  return teb::dtype_bits(static_cast<NVTEDType>(type)) / 8;
}
}  // namespace transformer_engine
