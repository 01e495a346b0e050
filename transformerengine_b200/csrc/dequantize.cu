// Block-scaled dequantization (test support and `backward_override="dequantized"`).
// Replaces reference cast/mxfp8/dequantize_mxfp8.cuh and
// cast/nvfp4/dequantize_nvfp4.cuh:30-140. Streaming kernels: one thread per
// 16 (MXFP8: 8) consecutive elements, 8-byte loads, vector stores.
#include "core.h"
#include "ptx.cuh"

namespace teb {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ size_t swz(size_t m, size_t k, size_t K) {
  return ((m >> 7) * (K >> 2) + (k >> 2)) * 512 + (m & 31) * 16 + ((m & 127) >> 5) * 4 + (k & 3);
}

template <typename OT>
__device__ __forceinline__ void store8(OT *dst, const float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) dst[i] = static_cast<OT>(f[i]);
}

// MXFP8: thread handles 8 consecutive elements of a row.
template <typename OT, bool E5M2>
__global__ void __launch_bounds__(kThreads)
    mxfp8_dequantize_kernel(const uint8_t *__restrict__ q, const uint8_t *__restrict__ sc,
                            OT *__restrict__ out, size_t rows, size_t cols, size_t ld,
                            bool colwise, bool swizzled) {
  const size_t vec_per_row = cols / 8;
  const size_t idx = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
  if (idx >= rows * vec_per_row) return;
  const size_t r = idx / vec_per_row, c = (idx % vec_per_row) * 8;
  const uint2 v = *reinterpret_cast<const uint2 *>(q + r * cols + c);
  const uint32_t w[4] = {v.x & 0xffffu, v.x >> 16, v.y & 0xffffu, v.y >> 16};
  float f[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 p = ptx::fp8x2_to_float2<E5M2>(w[i]);
    f[2 * i] = p.x;
    f[2 * i + 1] = p.y;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    size_t si;
    if (colwise) si = swizzled ? swz(c + i, r / 32, ld) : (r / 32) * ld + c + i;
    else si = swizzled ? swz(r, c / 32, ld) : r * ld + c / 32;
    f[i] *= ptx::e8m0_to_float(sc[si]);
  }
  store8(out + r * cols + c, f);
}

// NVFP4: thread handles one 1x16 block (8 bytes) (reference dequantize_nvfp4.cuh:35-80).
template <typename OT>
__global__ void __launch_bounds__(kThreads)
    nvfp4_dequantize_kernel(const uint8_t *__restrict__ q, const uint8_t *__restrict__ sc,
                            const float *__restrict__ amax, OT *__restrict__ out, size_t rows,
                            size_t blocks_per_row, size_t ld, bool swizzled) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
  if (idx >= rows * blocks_per_row) return;
  const size_t r = idx / blocks_per_row, b = idx % blocks_per_row;
  const uint2 v = *reinterpret_cast<const uint2 *>(q + idx * 8);
  const uint8_t s8 = sc[swizzled ? swz(r, b, ld) : r * ld + b];
  const float s = ptx::fp8x2_to_float2<false>(s8).x;
  constexpr float factor_inv = 1.0f / (6.0f * 448.0f);
  const float final_scale = s * amax[0] * factor_inv;
  const float lut[8] = {0.f, .5f, 1.f, 1.5f, 2.f, 3.f, 4.f, 6.f};
  OT *dst = out + idx * 16;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const uint32_t word = i < 8 ? v.x : v.y;
    const uint32_t nib = (word >> (4 * (i & 7))) & 0xfu;
    const float mag = lut[nib & 7];
    dst[i] = static_cast<OT>((nib & 8 ? -mag : mag) * final_scale);
  }
}

template <typename F>
void dispatch_out(NVTEDType t, F &&f) {
  switch (t) {
    case kNVTEBFloat16: f(__nv_bfloat16{}); break;
    case kNVTEFloat16: f(__half{}); break;
    case kNVTEFloat32: f(float{}); break;
    default: TEB_ERROR("Output must be in higher precision, got ", dtype_name(t));
  }
}

}  // namespace

namespace mxfp8 {
void dequantize(const Tensor &in, Tensor &out, cudaStream_t stream) {
  const bool colwise = !in.has_data();
  const Buffer &q = colwise ? in.columnwise_data : in.data;
  const Buffer &sc = colwise ? in.columnwise_scale_inv : in.scale_inv;
  if (in.numel() == 0 || out.data.numel() == 0) return;   // empty tensors have no buffers (reference tests: shape {0, 256})
  TEB_CHECK(q.dptr != nullptr, "Cannot dequantize tensor without data.");
  TEB_CHECK(is_fp8(q.dtype), "Input must have FP8 type.");
  TEB_CHECK(sc.dptr != nullptr, "Input needs scaling factors.");
  TEB_CHECK(out.data.numel() == q.numel(), "Input and output shapes need to match.");
  const size_t rows = in.rows(), cols = in.cols();
  TEB_CHECK(cols % 8 == 0, "MXFP8 dequantize needs the last dimension to be a multiple of 8");
  if (rows * cols == 0) return;
  const bool swizzled = in.with_gemm_swizzled_scales;
  size_t ld;
  if (swizzled) ld = colwise ? round_up(ceil_div(rows, 32), 4) : round_up(ceil_div(cols, 32), 4);
  else ld = sc.shape.data[sc.shape.ndim - 1];
  const bool e5m2 = q.dtype == kNVTEFloat8E5M2;
  const int grid = static_cast<int>(ceil_div(rows * cols / 8, kThreads));
  dispatch_out(out.data.dtype, [&](auto tag) {
    using OT = decltype(tag);
    if (e5m2)
      mxfp8_dequantize_kernel<OT, true><<<grid, kThreads, 0, stream>>>(
          q.as<uint8_t>(), sc.as<uint8_t>(), out.data.as<OT>(), rows, cols, ld, colwise, swizzled);
    else
      mxfp8_dequantize_kernel<OT, false><<<grid, kThreads, 0, stream>>>(
          q.as<uint8_t>(), sc.as<uint8_t>(), out.data.as<OT>(), rows, cols, ld, colwise, swizzled);
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}
}  // namespace mxfp8

namespace nvfp4 {
void dequantize(const Tensor &in, Tensor &out, cudaStream_t stream) {
  if (in.numel() == 0 || out.data.numel() == 0) return;   // empty tensors have no buffers
  TEB_CHECK(in.data.dptr != nullptr && in.data.dtype == kNVTEFloat4E2M1,
            "Input must have FP4 type.");
  TEB_CHECK(is_high_precision(out.data.dtype), "Output must be in higher precision.");
  TEB_CHECK(!in.row_scaled_nvfp4, "Row-scaled NVFP4 is not implemented.");
  TEB_CHECK(in.nvfp4_e4m3_max == 448, "Only the standard NVFP4 E4M3 bound (448) is implemented.");
  const size_t rows = in.rows(), cols = in.cols();
  TEB_CHECK(cols % 16 == 0, "Last dimension of FP4 tensors needs to be divisible by 16, but got ",
            cols, ".");
  TEB_CHECK(out.data.numel() == rows * cols, "Input and output shapes need to match.");
  TEB_CHECK(in.scale_inv.dptr != nullptr && in.amax.dptr != nullptr,
            "Input needs block scales and amax.");
  if (rows * cols == 0) return;
  const size_t bpr = cols / 16;
  const bool swizzled = in.with_gemm_swizzled_scales;
  const size_t ld = swizzled ? round_up(bpr, 4)
                             : in.scale_inv.shape.data[in.scale_inv.shape.ndim - 1];
  const int grid = static_cast<int>(ceil_div(rows * bpr, kThreads));
  dispatch_out(out.data.dtype, [&](auto tag) {
    using OT = decltype(tag);
    nvfp4_dequantize_kernel<OT><<<grid, kThreads, 0, stream>>>(
        in.data.as<uint8_t>(), in.scale_inv.as<uint8_t>(), in.amax.as<float>(), out.data.as<OT>(),
        rows, bpr, ld, swizzled);
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}
}  // namespace nvfp4

}  // namespace teb
