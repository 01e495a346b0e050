// MXFP8 quantization: 1x32 (rowwise) and 32x1 (columnwise) blocks with E8M0
// scales. Replaces the reference's mxfp8::quantize
// (transformer_engine/common/cast/mxfp8/quantize_mxfp8.cuh:558-845 and
// specialized/quantize_mxfp8.cuh); numerics per SURVEY.md section 3A.
//
// Design (HBM-bound; 2 B read + up to 2 B written per element):
//   One warp owns a 32-row x 128-column tile (16-bit inputs; 64 columns for
//   fp32). Lane l owns 4 (2) adjacent columns for all 32 rows, so
//     * every global access is a fully coalesced 256 B (input) / 128 B (fp8
//       output) row segment per warp, 32 independent loads in flight per lane;
//     * the columnwise 32x1 block maximum is a private register reduction,
//     * the rowwise 1x32 block maximum is a 3-step (4-step) xor-shuffle inside
//       8 (16) lanes,
//   and the tile never touches shared memory: no barriers, no bank conflicts.
//   Block scales leave as packed 4-byte words (rowwise: lane r stores the 4
//   scales of row r after a 4-shuffle transpose; columnwise: lane stores its 4
//   columns' scales). With `with_gemm_swizzled_scales` the same bytes go
//   straight into the 128x4 tcgen05 tile order, so no separate swizzle pass.
#include "core.h"
#include "ptx.cuh"

namespace teb {
namespace mxfp8 {

namespace {

constexpr int kWarpsPerCta = 4;
constexpr int kTileRows = 32;

// offset of scale (m, k) in the GEMM-swizzled layout; K = padded scale columns
__device__ __forceinline__ size_t swizzled_offset(size_t m, size_t k, size_t K) {
  return ((m >> 7) * (K >> 2) + (k >> 2)) * 512 + (m & 31) * 16 + ((m & 127) >> 5) * 4 + (k & 3);
}

struct Args {
  const void *in;
  uint8_t *out_row, *out_col;
  uint8_t *sc_row, *sc_col;
  size_t rows, cols;
  size_t ld_row;   // compact rowwise scale row stride; swizzled: padded K (= scale cols)
  size_t ld_col;   // compact colwise scale row stride; swizzled: padded K (= scale rows)
  float *amax;     // optional tensor amax (atomic max)
  const float *noop;
};

template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ, bool WITH_AMAX>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
    quantize_kernel(const Args a) {
  using W = ptx::Word<IT>;
  constexpr int CPL = 2 * W::kElems;     // columns per lane (8 bytes for 16-bit, 2 x fp32)
  constexpr int LPB = 32 / CPL;          // lanes per 1x32 block
  constexpr int TILE_COLS = 32 * CPL;
  constexpr float kMaxRcp = E5M2 ? (1.0f / 57344.0f) : (1.0f / 448.0f);

  if (a.noop != nullptr && a.noop[0] == 1.0f) return;

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const size_t tile_c = static_cast<size_t>(blockIdx.x) * kWarpsPerCta + warp;
  const size_t r0 = static_cast<size_t>(blockIdx.y) * kTileRows;
  const size_t c0 = tile_c * TILE_COLS;
  if (c0 >= a.cols) return;
  const size_t c = c0 + lane * CPL;
  const bool col_ok = c < a.cols;
  const IT *src = reinterpret_cast<const IT *>(a.in) + r0 * a.cols + c;

  // ---- load the whole tile column strip into registers ----------------------
  uint2 v[kTileRows];
#pragma unroll
  for (int r = 0; r < kTileRows; ++r) {
    v[r] = make_uint2(0u, 0u);
    if (col_ok && r0 + r < a.rows) v[r] = ptx::ld_stream_v2(src + static_cast<size_t>(r) * a.cols);
  }

  uint32_t cm0 = 0u, cm1 = 0u;  // running |x| maxima of my columns
  uint32_t row_sc = 0u;         // my share of the rowwise scale bytes
  float tmax = 0.0f;

#pragma unroll
  for (int r = 0; r < kTileRows; ++r) {
    const uint32_t a0 = W::abs(v[r].x), a1 = W::abs(v[r].y);
    if (COL || WITH_AMAX) {
      cm0 = W::max(cm0, a0);
      cm1 = W::max(cm1, a1);
    }
    if (ROW) {
      uint32_t t = W::max(a0, a1);
#pragma unroll
      for (int o = 1; o < LPB; o <<= 1) t = W::max(t, __shfl_xor_sync(0xffffffffu, t, o));
      const float bmax = W::hmax(t);
      const uint32_t e = ptx::float_to_e8m0(bmax * kMaxRcp);
      const float s = ptx::e8m0_rcp(e);
      uint32_t q;
      if (W::kElems == 2) {
        q = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[r].x), W::hi(v[r].x), s) |
            (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[r].y), W::hi(v[r].y), s) << 16);
      } else {
        q = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[r].x), W::lo(v[r].y), s);
      }
      if (col_ok && r0 + r < a.rows) {
        uint8_t *dst = a.out_row + (r0 + r) * a.cols + c;
        if (CPL == 4) *reinterpret_cast<uint32_t *>(dst) = q;
        else *reinterpret_cast<uint16_t *>(dst) = static_cast<uint16_t>(q);
      }
      // Lane (g, j) of block-group g keeps the scale of rows with r % LPB == j.
      if ((r % LPB) == (lane % LPB)) row_sc |= e << (8 * (r / LPB));
    }
  }

  if (ROW) {
    // Transpose: lane r gathers the CPL scale bytes of row r (k-blocks 0..CPL-1).
    uint32_t word = 0u;
#pragma unroll
    for (int g = 0; g < CPL; ++g) {
      const uint32_t x = __shfl_sync(0xffffffffu, row_sc, g * LPB + (lane % LPB));
      word |= ((x >> (8 * (lane / LPB))) & 0xffu) << (8 * g);
    }
    const size_t row = r0 + lane;
    const size_t k0 = tile_c * CPL;
    const size_t nblk = (a.cols + 31) / 32;
    if (row < a.rows) {
      uint8_t *dst = SWZ ? a.sc_row + swizzled_offset(row, k0, a.ld_row)
                         : a.sc_row + row * a.ld_row + k0;
      if (CPL == 4 && k0 + 4 <= nblk) {
        *reinterpret_cast<uint32_t *>(dst) = word;
      } else if (CPL == 2 && k0 + 2 <= nblk) {
        *reinterpret_cast<uint16_t *>(dst) = static_cast<uint16_t>(word);
      } else {
#pragma unroll
        for (int g = 0; g < CPL; ++g)
          if (k0 + g < nblk) dst[g] = static_cast<uint8_t>(word >> (8 * g));
      }
    }
  }

  if (COL || WITH_AMAX) {
    float cmax[CPL];
    if (W::kElems == 2) {
      cmax[0] = W::lo(cm0); cmax[1] = W::hi(cm0);
      cmax[CPL - 2] = W::lo(cm1); cmax[CPL - 1] = W::hi(cm1);
    } else {
      cmax[0] = W::lo(cm0); cmax[CPL - 1] = W::lo(cm1);
    }
#pragma unroll
    for (int k = 0; k < CPL; ++k) tmax = fmaxf(tmax, cmax[k]);

    if (COL) {
      float s[CPL];
      uint32_t word = 0u;
#pragma unroll
      for (int k = 0; k < CPL; ++k) {
        const uint32_t e = ptx::float_to_e8m0(cmax[k] * kMaxRcp);
        s[k] = ptx::e8m0_rcp(e);
        word |= e << (8 * k);
      }
      if (col_ok) {
        const size_t kb = blockIdx.y;  // 32-row block index
        if (SWZ) {
#pragma unroll
          for (int k = 0; k < CPL; ++k)
            a.sc_col[swizzled_offset(c + k, kb, a.ld_col)] = static_cast<uint8_t>(word >> (8 * k));
        } else {
          uint8_t *dst = a.sc_col + kb * a.ld_col + c;
          if (CPL == 4) *reinterpret_cast<uint32_t *>(dst) = word;
          else *reinterpret_cast<uint16_t *>(dst) = static_cast<uint16_t>(word);
        }
      }
#pragma unroll
      for (int r = 0; r < kTileRows; ++r) {
        uint32_t q;
        if (W::kElems == 2) {
          q = ptx::cvt_fp8x2<E5M2>(W::lo(v[r].x) * s[0], W::hi(v[r].x) * s[1]) |
              (ptx::cvt_fp8x2<E5M2>(W::lo(v[r].y) * s[CPL - 2], W::hi(v[r].y) * s[CPL - 1]) << 16);
        } else {
          q = ptx::cvt_fp8x2<E5M2>(W::lo(v[r].x) * s[0], W::lo(v[r].y) * s[CPL - 1]);
        }
        if (col_ok && r0 + r < a.rows) {
          uint8_t *dst = a.out_col + (r0 + r) * a.cols + c;
          if (CPL == 4) *reinterpret_cast<uint32_t *>(dst) = q;
          else *reinterpret_cast<uint16_t *>(dst) = static_cast<uint16_t>(q);
        }
      }
    }
    if (WITH_AMAX) {
      tmax = ptx::warp_max(tmax);
      if (lane == 0 && tmax > 0.0f) ptx::atomic_max_nonneg(a.amax, tmax);
    }
  }
}

// Generic path for shapes/pointers the vector kernel cannot take (columns not a
// multiple of the lane vector, unaligned bases). One 32x32 tile per CTA,
// one element per thread, block maxima through shared memory. Same numerics.
template <typename IT, bool E5M2>
__global__ void __launch_bounds__(1024) quantize_generic_kernel(const Args a, bool row, bool col,
                                                                bool swz) {
  constexpr float kMaxRcp = E5M2 ? (1.0f / 57344.0f) : (1.0f / 448.0f);
  if (a.noop != nullptr && a.noop[0] == 1.0f) return;
  __shared__ float tile[32][33];
  __shared__ uint32_t ecol[32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const size_t r = static_cast<size_t>(blockIdx.y) * 32 + ty;
  const size_t c = static_cast<size_t>(blockIdx.x) * 32 + tx;
  const bool ok = r < a.rows && c < a.cols;
  float x = 0.0f;
  if (ok) x = static_cast<float>(reinterpret_cast<const IT *>(a.in)[r * a.cols + c]);
  tile[ty][tx] = fabsf(x);
  __syncthreads();
  const float rmax = ptx::warp_max(tile[ty][tx]);  // warp ty holds row ty
  if (row) {
    const uint32_t e = ptx::float_to_e8m0(rmax * kMaxRcp);
    if (ok)
      a.out_row[r * a.cols + c] =
          static_cast<uint8_t>(ptx::cvt_fp8x2<E5M2>(x * ptx::e8m0_rcp(e), 0.0f));
    if (tx == 0 && r < a.rows) {
      const size_t k = blockIdx.x;
      a.sc_row[swz ? swizzled_offset(r, k, a.ld_row) : r * a.ld_row + k] = static_cast<uint8_t>(e);
    }
  }
  if (col) {
    const float cmax = ptx::warp_max(tile[tx][ty]);  // warp ty reduces column ty
    if (tx == 0) ecol[ty] = ptx::float_to_e8m0(cmax * kMaxRcp);
    __syncthreads();
    const uint32_t ec = ecol[tx];
    if (ok)
      a.out_col[r * a.cols + c] =
          static_cast<uint8_t>(ptx::cvt_fp8x2<E5M2>(x * ptx::e8m0_rcp(ec), 0.0f));
    if (ty == 0 && c < a.cols) {
      const size_t k = blockIdx.y;
      a.sc_col[swz ? swizzled_offset(c, k, a.ld_col) : k * a.ld_col + c] = static_cast<uint8_t>(ec);
    }
  }
  if (a.amax != nullptr && tx == 0 && rmax > 0.0f) ptx::atomic_max_nonneg(a.amax, rmax);
}

template <typename IT, bool E5M2, bool SWZ, bool WITH_AMAX>
void launch_vec(const Args &a, bool row, bool col, dim3 grid, cudaStream_t stream) {
  const dim3 block(kWarpsPerCta * 32);
  if (row && col) quantize_kernel<IT, true, true, E5M2, SWZ, WITH_AMAX><<<grid, block, 0, stream>>>(a);
  else if (row) quantize_kernel<IT, true, false, E5M2, SWZ, WITH_AMAX><<<grid, block, 0, stream>>>(a);
  else quantize_kernel<IT, false, true, E5M2, SWZ, WITH_AMAX><<<grid, block, 0, stream>>>(a);
}

template <typename IT>
void launch(const Args &a, bool row, bool col, bool e5m2, bool swz, cudaStream_t stream) {
  constexpr int CPL = 2 * ptx::Word<IT>::kElems;
  const bool aligned = (a.cols % CPL == 0) && (reinterpret_cast<uintptr_t>(a.in) % 8 == 0) &&
                       (!row || reinterpret_cast<uintptr_t>(a.out_row) % 4 == 0) &&
                       (!col || reinterpret_cast<uintptr_t>(a.out_col) % 4 == 0) &&
                       (!row || (reinterpret_cast<uintptr_t>(a.sc_row) % 4 == 0 &&
                                 (swz || a.ld_row % 4 == 0))) &&
                       (!col || swz || (reinterpret_cast<uintptr_t>(a.sc_col) % 4 == 0 &&
                                        a.ld_col % 4 == 0));
  if (!aligned) {
    const dim3 grid(ceil_div(a.cols, 32), ceil_div(a.rows, 32));
    if (e5m2) quantize_generic_kernel<IT, true><<<grid, 1024, 0, stream>>>(a, row, col, swz);
    else quantize_generic_kernel<IT, false><<<grid, 1024, 0, stream>>>(a, row, col, swz);
    return;
  }
  const dim3 grid(ceil_div(a.cols, static_cast<size_t>(32 * CPL * kWarpsPerCta)),
                  ceil_div(a.rows, static_cast<size_t>(kTileRows)));
  const bool with_amax = a.amax != nullptr;
#define TEB_MX_DISPATCH(E5, SW, AM) launch_vec<IT, E5, SW, AM>(a, row, col, grid, stream)
  if (e5m2) {
    if (swz) { if (with_amax) TEB_MX_DISPATCH(true, true, true); else TEB_MX_DISPATCH(true, true, false); }
    else { if (with_amax) TEB_MX_DISPATCH(true, false, true); else TEB_MX_DISPATCH(true, false, false); }
  } else {
    if (swz) { if (with_amax) TEB_MX_DISPATCH(false, true, true); else TEB_MX_DISPATCH(false, true, false); }
    else { if (with_amax) TEB_MX_DISPATCH(false, false, true); else TEB_MX_DISPATCH(false, false, false); }
  }
#undef TEB_MX_DISPATCH
}

}  // namespace

void quantize(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream) {
  require_blackwell("MXFP8 quantization");
  const bool row = out.has_data(), col = out.has_columnwise_data();
  TEB_CHECK(row || col, "Either rowwise or columnwise output data need to be allocated.");
  TEB_CHECK(in.has_data(), "Cannot quantize tensor without rowwise data.");
  TEB_CHECK(is_high_precision(in.data.dtype), "Input must be in higher precision.");
  const NVTEDType otype = out.dtype();
  TEB_CHECK(is_fp8(otype), "Output must have FP8 type.");
  const size_t rows = in.rows(), cols = in.cols();
  if (row) {
    TEB_CHECK(out.scale_inv.dptr != nullptr, "Scaling tensor must be allocated");
    TEB_CHECK(out.data.numel() == rows * cols, "Input and output shapes need to match.");
  }
  if (col) {
    TEB_CHECK(out.columnwise_scale_inv.dptr != nullptr,
              "Columnwise scaling tensor must be allocated");
    TEB_CHECK(out.columnwise_data.numel() == rows * cols, "Input and output shapes need to match.");
  }
  if (rows == 0 || cols == 0) return;
  const bool swz = out.with_gemm_swizzled_scales;

  Args a{};
  a.in = in.data.dptr;
  a.out_row = out.data.as<uint8_t>();
  a.out_col = out.columnwise_data.as<uint8_t>();
  a.sc_row = out.scale_inv.as<uint8_t>();
  a.sc_col = out.columnwise_scale_inv.as<uint8_t>();
  a.rows = rows;
  a.cols = cols;
  auto last_dim = [](const Buffer &b) { return b.shape.ndim ? b.shape.data[b.shape.ndim - 1] : 0; };
  if (row) {
    a.ld_row = last_dim(out.scale_inv);
    TEB_CHECK(a.ld_row >= ceil_div(cols, 32), "Rowwise scale tensor is too narrow: ", a.ld_row);
    if (swz) {
      TEB_CHECK(a.ld_row % 4 == 0 && out.scale_inv.numel() >= round_up(rows, 128) * a.ld_row,
                "GEMM-swizzled rowwise scales need a [roundup(rows,128), roundup(cols/32,4)] buffer");
      if (rows % 128 != 0 || ceil_div(cols, 32) % 4 != 0)  // padding must read as zero
        TEB_CHECK_CUDA(cudaMemsetAsync(a.sc_row, 0, out.scale_inv.numel(), stream));
    }
  }
  if (col) {
    // compact: [roundup(rows/32,4), roundup(cols,128)]; swizzled: logical [cols(pad 128), rows/32(pad 4)]
    if (swz) {
      a.ld_col = out.columnwise_scale_inv.shape.ndim >= 2 ? out.columnwise_scale_inv.shape.data[0]
                                                          : round_up(ceil_div(rows, 32), 4);
      TEB_CHECK(a.ld_col % 4 == 0 && a.ld_col >= ceil_div(rows, 32) &&
                    out.columnwise_scale_inv.numel() >= round_up(cols, 128) * a.ld_col,
                "GEMM-swizzled columnwise scales need a [roundup(rows/32,4), roundup(cols,128)] buffer");
      if (cols % 128 != 0 || ceil_div(rows, 32) % 4 != 0)
        TEB_CHECK_CUDA(cudaMemsetAsync(a.sc_col, 0, out.columnwise_scale_inv.numel(), stream));
    } else {
      a.ld_col = last_dim(out.columnwise_scale_inv);
      TEB_CHECK(a.ld_col >= cols, "Columnwise scale tensor is too narrow: ", a.ld_col);
    }
  }
  a.amax = out.amax.as<float>();
  a.noop = noop ? noop->data.as<float>() : nullptr;
  const bool e5m2 = otype == kNVTEFloat8E5M2;
  switch (in.data.dtype) {
    case kNVTEBFloat16: launch<__nv_bfloat16>(a, row, col, e5m2, swz, stream); break;
    case kNVTEFloat16: launch<__half>(a, row, col, e5m2, swz, stream); break;
    case kNVTEFloat32: launch<float>(a, row, col, e5m2, swz, stream); break;
    default: TEB_ERROR("Unsupported input dtype ", dtype_name(in.data.dtype));
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}

}  // namespace mxfp8
}  // namespace teb
