// nvte_quantize_dbias: cast + column reduction of the same tensor (dbias[c] = sum_r x[r, c]),
// the backward of a Linear's bias (reference cast/cast.cu nvte_quantize_dbias,
// cast/core/common.cuh:60-97 reduce_dbias, cast/mxfp8/quantize_mxfp8.cuh:496-541).
// Like the reference: per 32-row slab fp32 partial sums go to a caller-provided workspace
// [ceil(rows / 32), cols] (queried by calling with an empty workspace), a second small kernel
// reduces the slabs and rounds once to the input dtype.
//   * MXFP8, 16-bit input, 128-aligned shape, columnwise output requested: the partial sums come
//     out of the TMA quantize kernel's column pass (one thread already walks 2 columns x 32 rows);
//   * per-tensor FP8: fp8::quantize_dbias (slab kernel: cast + sums in one pass);
//   * anything else: column sums by a separate slab kernel, then the ordinary quantizer.
#include <mutex>

#include "core.h"
#include "ptx.cuh"

// Scratch of nvte_b200_quantize_dbias_static_scratch (below): the per-slab partial sums live in this
// file-scope device array instead of a caller-allocated workspace, so that entry point allocates nothing.
// Every GPU that loads the library holds these bytes until the process exits. In HOST code this name is
// only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress (static_scratch()).
#define TEB_DBIAS_STATIC_SCRATCH_BYTES NVTE_B200_DBIAS_STATIC_SCRATCH_BYTES
static __device__ __align__(256) float teb_dbias_static_scratch[TEB_DBIAS_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace teb {
namespace mxfp8 {
void quantize_with_dbias_ws(const Tensor &in, Tensor &out, const Tensor *noop, float *dbias_ws,
                            bool *fused, cudaStream_t stream);
}
namespace fp8 {
void quantize_dbias_slabs(const Tensor &in, Tensor &out, float *ws, cudaStream_t stream);
}
namespace dbias {

namespace {

constexpr int kSlab = 32;

template <typename IT>
__global__ void __launch_bounds__(256) partial_kernel(const IT *__restrict__ in, float *__restrict__ ws,
                                                      size_t rows, size_t cols) {
  const size_t c = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
  if (c >= cols) return;
  const size_t r0 = static_cast<size_t>(blockIdx.y) * kSlab;
  const size_t r1 = r0 + kSlab < rows ? r0 + kSlab : rows;
  float s = 0.0f;
  for (size_t r = r0; r < r1; ++r) s += static_cast<float>(in[r * cols + c]);
  ws[static_cast<size_t>(blockIdx.y) * cols + c] = s;
}

template <typename OT>
__global__ void __launch_bounds__(256) reduce_kernel(const float *__restrict__ ws, OT *__restrict__ out,
                                                     size_t slabs, size_t cols) {
  const size_t c = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
  if (c >= cols) return;
  float s = 0.0f;
  for (size_t k = 0; k < slabs; ++k) s += ws[k * cols + c];
  out[c] = static_cast<OT>(s);
}

}  // namespace

void quantize_dbias(const Tensor &in, Tensor &out, Tensor &dbias, Tensor &workspace,
                    cudaStream_t stream) {
  TEB_CHECK(in.has_data(), "Cannot quantize tensor without rowwise data.");
  TEB_CHECK(is_high_precision(in.data.dtype), "Input must be in higher precision.");
  const size_t rows = in.rows(), cols = in.cols();
  const size_t slabs = ceil_div(rows, static_cast<size_t>(kSlab));
  if (workspace.data.dptr == nullptr) {   // size query (reference quantize_mxfp8.cuh:627-631)
    workspace.data.shape.ndim = 2;
    workspace.data.shape.data[0] = slabs;
    workspace.data.shape.data[1] = cols;
    workspace.data.dtype = kNVTEFloat32;
    return;
  }
  TEB_CHECK(workspace.data.dtype == kNVTEFloat32 && workspace.data.numel() >= slabs * cols,
            "dbias workspace must be fp32 [", slabs, ", ", cols, "]");
  TEB_CHECK(dbias.data.dptr != nullptr && dbias.data.numel() == cols, "dbias must have ", cols, " elements");
  TEB_CHECK(dbias.data.dtype == in.data.dtype, "dbias must have the input dtype");
  if (rows == 0 || cols == 0) return;
  float *ws = workspace.data.as<float>();
  bool fused = false;
  switch (out.scaling_mode) {
    case NVTE_DELAYED_TENSOR_SCALING:
      if (!out.has_columnwise_data() && cols % 8 == 0 && reinterpret_cast<uintptr_t>(in.data.dptr) % 16 == 0 &&
          reinterpret_cast<uintptr_t>(out.data.dptr) % 8 == 0) {
        fp8::quantize_dbias_slabs(in, out, ws, stream);
        fused = true;
      }
      break;
    case NVTE_MXFP8_1D_SCALING: mxfp8::quantize_with_dbias_ws(in, out, nullptr, ws, &fused, stream); break;
    default: TEB_ERROR("nvte_quantize_dbias supports per-tensor FP8 and MXFP8 outputs");
  }
  if (!fused) {
    const dim3 grid(static_cast<unsigned>(ceil_div(cols, static_cast<size_t>(256))), static_cast<unsigned>(slabs));
    switch (in.data.dtype) {
      case kNVTEBFloat16: partial_kernel<<<grid, 256, 0, stream>>>(in.data.as<__nv_bfloat16>(), ws, rows, cols); break;
      case kNVTEFloat16: partial_kernel<<<grid, 256, 0, stream>>>(in.data.as<__half>(), ws, rows, cols); break;
      default: partial_kernel<<<grid, 256, 0, stream>>>(in.data.as<float>(), ws, rows, cols); break;
    }
    TEB_CHECK_CUDA(cudaGetLastError());
    if (out.scaling_mode == NVTE_DELAYED_TENSOR_SCALING) fp8::quantize(in, out, nullptr, stream);
  }
  const unsigned rgrid = static_cast<unsigned>(ceil_div(cols, static_cast<size_t>(256)));
  switch (in.data.dtype) {
    case kNVTEBFloat16: reduce_kernel<<<rgrid, 256, 0, stream>>>(ws, dbias.data.as<__nv_bfloat16>(), slabs, cols); break;
    case kNVTEFloat16: reduce_kernel<<<rgrid, 256, 0, stream>>>(ws, dbias.data.as<__half>(), slabs, cols); break;
    default: reduce_kernel<<<rgrid, 256, 0, stream>>>(ws, dbias.data.as<float>(), slabs, cols); break;
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}

// ---- the same operation with NO allocation on the call path ---------------------------------------
// (no counterpart in the reference; added for callers that
// cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it)
namespace {

// Device address of teb_dbias_static_scratch on the CURRENT device, looked up once per device.
float *static_scratch() {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float *ptrs[kMaxDevices] = {};
  int dev = -1;
  TEB_CHECK_CUDA(cudaGetDevice(&dev));
  TEB_CHECK(dev >= 0 && dev < kMaxDevices, "device ordinal ", dev, " out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void *p = nullptr;
    TEB_CHECK_CUDA(cudaGetSymbolAddress(&p, teb_dbias_static_scratch));
    TEB_CHECK(p != nullptr && reinterpret_cast<uintptr_t>(p) % 256 == 0, "static scratch address is not usable");
    ptrs[dev] = static_cast<float *>(p);
  }
  return ptrs[dev];
}

}  // namespace

// quantize_dbias() with the partial sums in the library's static device array. Every slab row of the
// scratch that the reduction reads is fully overwritten by the producing kernel of the same call, so
// nothing has to start at zero; the array keeps the last call's contents. All calls on a device share
// it: two calls that overlap on the GPU (different streams, no ordering between them) race.
void quantize_dbias_static_scratch(const Tensor &in, Tensor &out, Tensor &dbias, cudaStream_t stream) {
  TEB_CHECK(in.has_data(), "Cannot quantize tensor without rowwise data.");
  const size_t rows = in.rows(), cols = in.cols();
  const size_t slabs = ceil_div(rows, static_cast<size_t>(kSlab));
  const size_t need = slabs * cols * sizeof(float);
  TEB_CHECK(need <= TEB_DBIAS_STATIC_SCRATCH_BYTES,
            "nvte_b200_quantize_dbias_static_scratch: input [", rows, ", ", cols, "] needs ", need,
            " bytes of scratch (ceil(rows / 32) * cols * 4), the static array holds ",
            static_cast<size_t>(TEB_DBIAS_STATIC_SCRATCH_BYTES),
            "; use nvte_quantize_dbias with a caller-allocated workspace");
  Tensor ws;
  ws.data.dptr = static_scratch();
  ws.data.dtype = kNVTEFloat32;
  ws.data.shape.ndim = 2;
  ws.data.shape.data[0] = slabs > 0 ? slabs : 1;
  ws.data.shape.data[1] = cols > 0 ? cols : 1;
  quantize_dbias(in, out, dbias, ws, stream);
}

}  // namespace dbias
}  // namespace teb
