// Scale-factor swizzle: compact [M, K] block scales -> the 128x4-tile
// interleaved order tcgen05 block-scaled MMA (and cuBLASLt) expects, and back.
//   dst(m, k) = ((m/128)*(K/4) + k/4)*512 + (m%32)*16 + ((m%128)/32)*4 + k%4
// Replaces reference swizzle/swizzle.cu:104-565,959-1200 for single tensors.
// Pure byte permutation, HBM-bound and tiny (<= a few MB): every thread moves
// one 16-byte swizzled group (4 rows x 4 k-blocks) with a single 128-bit
// access on the swizzled side.
#include "core.h"

namespace teb {
namespace swizzle {

namespace {

constexpr int kThreads = 256;

// Rowwise-stored compact scales in[m*K + k]. CTA = 128 rows x 8 k-tiles.
template <bool INVERSE>
__global__ void __launch_bounds__(kThreads)
    swizzle_rowmajor_kernel(const uint8_t *__restrict__ in, uint8_t *__restrict__ out, int K) {
  const int tk = blockIdx.x * 8 + (threadIdx.x & 7);
  const int m32 = threadIdx.x >> 3;
  const int tm = blockIdx.y;
  if (tk * 4 >= K) return;
  const size_t sw = (static_cast<size_t>(tm) * (K / 4) + tk) * 512 + m32 * 16;
  const size_t cp = (static_cast<size_t>(tm) * 128 + m32) * K + tk * 4;
  if (!INVERSE) {
    uint4 v;
    v.x = *reinterpret_cast<const uint32_t *>(in + cp);
    v.y = *reinterpret_cast<const uint32_t *>(in + cp + 32 * static_cast<size_t>(K));
    v.z = *reinterpret_cast<const uint32_t *>(in + cp + 64 * static_cast<size_t>(K));
    v.w = *reinterpret_cast<const uint32_t *>(in + cp + 96 * static_cast<size_t>(K));
    *reinterpret_cast<uint4 *>(out + sw) = v;
  } else {
    const uint4 v = *reinterpret_cast<const uint4 *>(in + sw);
    *reinterpret_cast<uint32_t *>(out + cp) = v.x;
    *reinterpret_cast<uint32_t *>(out + cp + 32 * static_cast<size_t>(K)) = v.y;
    *reinterpret_cast<uint32_t *>(out + cp + 64 * static_cast<size_t>(K)) = v.z;
    *reinterpret_cast<uint32_t *>(out + cp + 96 * static_cast<size_t>(K)) = v.w;
  }
}

// Transposed compact scales in[k*M + m] (MXFP8 columnwise). CTA = 128 m x 8 k-tiles,
// lanes run along m so the byte accesses on the compact side coalesce.
template <bool INVERSE>
__global__ void __launch_bounds__(kThreads)
    swizzle_colmajor_kernel(const uint8_t *__restrict__ in, uint8_t *__restrict__ out, int M,
                            int K) {
  const int m32 = threadIdx.x & 31;
  const int tk = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int tm = blockIdx.y;
  if (tk * 4 >= K) return;
  const size_t sw = (static_cast<size_t>(tm) * (K / 4) + tk) * 512 + m32 * 16;
  uint32_t w[4];
  if (!INVERSE) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      w[j] = 0;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        w[j] |= static_cast<uint32_t>(in[static_cast<size_t>(tk * 4 + kk) * M + tm * 128 + m32 + 32 * j])
                << (8 * kk);
    }
    *reinterpret_cast<uint4 *>(out + sw) = make_uint4(w[0], w[1], w[2], w[3]);
  } else {
    const uint4 v = *reinterpret_cast<const uint4 *>(in + sw);
    w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        out[static_cast<size_t>(tk * 4 + kk) * M + tm * 128 + m32 + 32 * j] =
            static_cast<uint8_t>(w[j] >> (8 * kk));
  }
}

}  // namespace

void swizzle_scales(const Tensor &in, Tensor &out, bool inverse, cudaStream_t stream) {
  const NVTEScalingMode mode = in.scaling_mode;
  TEB_CHECK(mode == NVTE_MXFP8_1D_SCALING || mode == NVTE_NVFP4_1D_SCALING,
            "Input tensor has invalid scaling mode (", static_cast<int>(mode), ").");
  if (!inverse) {
    TEB_CHECK(!in.with_gemm_swizzled_scales, "Expected input tensor with scales in compact format.");
    TEB_CHECK(out.with_gemm_swizzled_scales,
              "Expected output tensor with scales in GEMM swizzled format.");
  } else {
    TEB_CHECK(in.with_gemm_swizzled_scales,
              "Expected input tensor with scales in GEMM swizzled format.");
    TEB_CHECK(!out.with_gemm_swizzled_scales,
              "Expected output tensor with scales in compact format.");
  }
  const bool has_row = in.scale_inv.has_data(), has_col = in.columnwise_scale_inv.has_data();
  TEB_CHECK(!has_row || !has_col, "Input tensor has both row-wise and column-wise scaling factors");
  if (!has_row && !has_col) return;
  const Buffer &src = has_row ? in.scale_inv : in.columnwise_scale_inv;
  const Buffer &dst = has_row ? out.scale_inv : out.columnwise_scale_inv;
  TEB_CHECK(src.shape.ndim == 2, "Expected 2D scaling factors, got ", src.shape.ndim, "D.");
  const bool transposed = (mode == NVTE_MXFP8_1D_SCALING) && has_col;
  const size_t m = transposed ? src.shape.data[1] : src.shape.data[0];
  const size_t k = transposed ? src.shape.data[0] : src.shape.data[1];
  TEB_CHECK(m % 128 == 0, "Input should be padded in M/N dimension!");
  TEB_CHECK(k % 4 == 0, "Input should be padded in K dimension!");
  TEB_CHECK(dst.has_data() && dst.dptr != nullptr, "Output tensor does not have ",
            has_row ? "row-wise" : "column-wise", " scaling factors.");
  TEB_CHECK(m * k == dst.numel(), "Expected output tensor to have ", m * k,
            " scaling factors, but got ", dst.numel(), ".");
  if (m == 0 || k == 0) return;
  TEB_CHECK(reinterpret_cast<uintptr_t>(src.dptr) % 16 == 0 &&
                reinterpret_cast<uintptr_t>(dst.dptr) % 16 == 0,
            "Scale buffers must be 16-byte aligned");
  const dim3 grid(ceil_div(k / 4, 8), m / 128);
  const uint8_t *ip = src.as<uint8_t>();
  uint8_t *op = dst.as<uint8_t>();
  if (transposed) {
    if (inverse) swizzle_colmajor_kernel<true><<<grid, kThreads, 0, stream>>>(ip, op, m, k);
    else swizzle_colmajor_kernel<false><<<grid, kThreads, 0, stream>>>(ip, op, m, k);
  } else {
    if (inverse) swizzle_rowmajor_kernel<true><<<grid, kThreads, 0, stream>>>(ip, op, k);
    else swizzle_rowmajor_kernel<false><<<grid, kThreads, 0, stream>>>(ip, op, k);
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}

}  // namespace swizzle
}  // namespace teb
