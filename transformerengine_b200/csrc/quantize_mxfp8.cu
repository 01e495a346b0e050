// MXFP8 quantization: 1x32 (rowwise) and 32x1 (columnwise) blocks with E8M0
// scales. Replaces the reference's mxfp8::quantize
// (transformer_engine/common/cast/mxfp8/quantize_mxfp8.cuh:558-845 and
// specialized/quantize_mxfp8.cuh); numerics per SURVEY.md section 3A.
//
// Design (HBM-bound; 2 B read + up to 2 B written per element):
//   One warp owns a 32-row x 128-column tile (16-bit inputs; 64 columns for
//   fp32). Half-warp h owns rows 16h..16h+15; lane j of a half owns 8 (4)
//   adjacent columns, i.e. one 16-byte vector per row, so
//     * every global access is a fully coalesced 256 B (input) / 128 B (fp8
//       output) row segment per half-warp, 16 independent 128-bit loads in
//       flight per lane;
//     * the columnwise 32x1 block maximum is a private register reduction plus
//       one exchange between the two halves,
//     * the rowwise 1x32 block maximum is a 2-step (3-step) xor-shuffle inside
//       4 (8) lanes,
//   and the tile never touches shared memory: no barriers, no bank conflicts.
//   Block scales leave as packed words (rowwise: lane (h, j) stores the 4
//   scales of row 16h+j after a 4-shuffle transpose; columnwise: lane j stores
//   its 8 columns' scales). With `with_gemm_swizzled_scales` the same bytes go
//   straight into the 128x4 tcgen05 tile order, so no separate swizzle pass.
#include "core.h"
#include "ptx.cuh"

namespace teb {
namespace mxfp8 {

// quantize_mxfp8_tma.cu (declared in core.h): TMA / thread-owned-block kernel for 16-bit inputs with
// 128-aligned shapes, one tensor or a list of tensors per launch

void quantize_with_dbias_ws(const Tensor &in, Tensor &out, const Tensor *noop, float *dbias_ws,
                            bool *fused, cudaStream_t stream);

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

// Unpack word w of a lane's 16-byte vector into floats f[0..kElems).
template <typename W>
__device__ __forceinline__ void unpack_word(uint32_t w, float &f0, float &f1) {
  f0 = W::lo(w);
  f1 = W::hi(w);
}

// One warp quantizes a 32-row x TILE_COLS tile. Lane (h = lane/16, j = lane%16)
// holds rows h*16 .. h*16+15 of columns j*CPL .. j*CPL+CPL-1 (one 16-byte
// vector per row).
template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ, bool WITH_AMAX, bool FULL>
__device__ __forceinline__ void quantize_tile(const Args &a, const size_t tile_c, const size_t r0,
                                              const int lane) {
  using W = ptx::Word<IT>;
  constexpr int CPL = 4 * W::kElems;   // columns per lane: 8 (16-bit) or 4 (fp32)
  constexpr int LPB = 32 / CPL;        // lanes per 1x32 block: 4 or 8
  constexpr int G = 16 / LPB;          // k-blocks per tile row: 4 or 2
  constexpr int TILE_COLS = 16 * CPL;  // 128 or 64
  constexpr int HR = 16;               // rows per lane
  constexpr float kMaxRcp = E5M2 ? (1.0f / 57344.0f) : (1.0f / 448.0f);

  const int h = lane >> 4, j = lane & 15;
  const size_t c = tile_c * TILE_COLS + j * CPL;
  const size_t rbase = r0 + h * HR;
  const bool col_ok = FULL || c < a.cols;
  const IT *src = reinterpret_cast<const IT *>(a.in) + rbase * a.cols + c;

  uint4 v[HR];
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    if (FULL || (col_ok && rbase + i < a.rows)) v[i] = ptx::ld_stream_v4(src + static_cast<size_t>(i) * a.cols);
    else v[i] = make_uint4(0u, 0u, 0u, 0u);
  }

  uint32_t cm[4] = {0u, 0u, 0u, 0u};  // running |x| maxima of my columns (packed)
  uint32_t row_sc = 0u;               // my share of the rowwise scale bytes
  uint8_t *orow = a.out_row + rbase * a.cols + c;

#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const uint32_t a0 = W::abs(v[i].x), a1 = W::abs(v[i].y), a2 = W::abs(v[i].z), a3 = W::abs(v[i].w);
    if (COL || WITH_AMAX) {
      cm[0] = W::max(cm[0], a0);
      cm[1] = W::max(cm[1], a1);
      cm[2] = W::max(cm[2], a2);
      cm[3] = W::max(cm[3], a3);
    }
    if (ROW) {
      uint32_t t = W::max(W::max(a0, a1), W::max(a2, a3));
#pragma unroll
      for (int o = 1; o < LPB; o <<= 1) t = W::max(t, __shfl_xor_sync(0xffffffffu, t, o));
      const uint32_t e = ptx::float_to_e8m0(W::hmax(t) * kMaxRcp);
      const float s = ptx::e8m0_rcp(e);
      uint32_t q0, q1;
      if (W::kElems == 2) {
        q0 = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[i].x), W::hi(v[i].x), s) |
             (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[i].y), W::hi(v[i].y), s) << 16);
        q1 = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[i].z), W::hi(v[i].z), s) |
             (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[i].w), W::hi(v[i].w), s) << 16);
      } else {
        q0 = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[i].x), W::lo(v[i].y), s) |
             (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v[i].z), W::lo(v[i].w), s) << 16);
        q1 = 0u;
      }
      if (FULL || (col_ok && rbase + i < a.rows)) {
        uint8_t *dst = orow + static_cast<size_t>(i) * a.cols;
        if (CPL == 8) *reinterpret_cast<uint2 *>(dst) = make_uint2(q0, q1);
        else *reinterpret_cast<uint32_t *>(dst) = q0;
      }
      // Lane q of a block group keeps the scales of rows with i % LPB == q.
      if ((i % LPB) == (j % LPB)) row_sc |= e << (8 * (i / LPB));
    }
  }

  if (ROW) {
    // Lane (h, j) gathers the G scale bytes of row h*16 + j.
    uint32_t word = 0u;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const uint32_t x = __shfl_sync(0xffffffffu, row_sc, h * 16 + g * LPB + (j % LPB));
      word |= ((x >> (8 * (j / LPB))) & 0xffu) << (8 * g);
    }
    const size_t row = rbase + j;
    const size_t k0 = tile_c * G;
    if (FULL || row < a.rows) {
      uint8_t *dst = SWZ ? a.sc_row + swizzled_offset(row, k0, a.ld_row)
                         : a.sc_row + row * a.ld_row + k0;
      const size_t nblk = (a.cols + 31) / 32;
      if (FULL || k0 + G <= nblk) {
        if (G == 4) *reinterpret_cast<uint32_t *>(dst) = word;
        else *reinterpret_cast<uint16_t *>(dst) = static_cast<uint16_t>(word);
      } else {
#pragma unroll
        for (int g = 0; g < G; ++g)
          if (k0 + g < nblk) dst[g] = static_cast<uint8_t>(word >> (8 * g));
      }
    }
  }

  if (COL || WITH_AMAX) {
    // Combine the two 16-row halves: lane j and lane j+16 own the same columns.
#pragma unroll
    for (int k = 0; k < 4; ++k) cm[k] = W::max(cm[k], __shfl_xor_sync(0xffffffffu, cm[k], 16));

    if (COL) {
      float s[CPL];
      uint32_t ew[2] = {0u, 0u};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (W::kElems == 2) {
          const uint32_t e0 = ptx::float_to_e8m0(W::lo(cm[k]) * kMaxRcp);
          const uint32_t e1 = ptx::float_to_e8m0(W::hi(cm[k]) * kMaxRcp);
          s[2 * k] = ptx::e8m0_rcp(e0);
          s[2 * k + 1] = ptx::e8m0_rcp(e1);
          ew[k >> 1] |= (e0 | (e1 << 8)) << (16 * (k & 1));
        } else {
          const uint32_t e0 = ptx::float_to_e8m0(W::lo(cm[k]) * kMaxRcp);
          s[k] = ptx::e8m0_rcp(e0);
          ew[0] |= e0 << (8 * k);
        }
      }
      if (h == 0 && col_ok) {
        const size_t kb = r0 / 32;
        if (SWZ) {
#pragma unroll
          for (int k = 0; k < CPL; ++k)
            a.sc_col[swizzled_offset(c + k, kb, a.ld_col)] =
                static_cast<uint8_t>(ew[k >> 2] >> (8 * (k & 3)));
        } else {
          uint8_t *dst = a.sc_col + kb * a.ld_col + c;
          if (CPL == 8) *reinterpret_cast<uint2 *>(dst) = make_uint2(ew[0], ew[1]);
          else *reinterpret_cast<uint32_t *>(dst) = ew[0];
        }
      }
      uint8_t *ocol = a.out_col + rbase * a.cols + c;
#pragma unroll
      for (int i = 0; i < HR; ++i) {
        uint32_t q0, q1;
        if (W::kElems == 2) {
          q0 = ptx::mul2_cvt_fp8x2<E5M2>(W::lo(v[i].x), W::hi(v[i].x), s[0], s[1]) |
               (ptx::mul2_cvt_fp8x2<E5M2>(W::lo(v[i].y), W::hi(v[i].y), s[2], s[3]) << 16);
          q1 = ptx::mul2_cvt_fp8x2<E5M2>(W::lo(v[i].z), W::hi(v[i].z), s[CPL - 4], s[CPL - 3]) |
               (ptx::mul2_cvt_fp8x2<E5M2>(W::lo(v[i].w), W::hi(v[i].w), s[CPL - 2], s[CPL - 1]) << 16);
        } else {
          q0 = ptx::mul2_cvt_fp8x2<E5M2>(W::lo(v[i].x), W::lo(v[i].y), s[0], s[1]) |
               (ptx::mul2_cvt_fp8x2<E5M2>(W::lo(v[i].z), W::lo(v[i].w), s[2], s[3]) << 16);
          q1 = 0u;
        }
        if (FULL || (col_ok && rbase + i < a.rows)) {
          uint8_t *dst = ocol + static_cast<size_t>(i) * a.cols;
          if (CPL == 8) *reinterpret_cast<uint2 *>(dst) = make_uint2(q0, q1);
          else *reinterpret_cast<uint32_t *>(dst) = q0;
        }
      }
    }
    if (WITH_AMAX) {
      float tmax = fmaxf(fmaxf(W::hmax(cm[0]), W::hmax(cm[1])), fmaxf(W::hmax(cm[2]), W::hmax(cm[3])));
      tmax = ptx::warp_max(tmax);
      if (lane == 0 && tmax > 0.0f) ptx::atomic_max_nonneg(a.amax, tmax);
    }
  }
}

template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ, bool WITH_AMAX>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 4)
    quantize_kernel(const Args a) {
  constexpr int TILE_COLS = 64 * ptx::Word<IT>::kElems;
  if (a.noop != nullptr && a.noop[0] == 1.0f) return;
  const int lane = threadIdx.x & 31;
  const size_t tile_c = static_cast<size_t>(blockIdx.x) * kWarpsPerCta + (threadIdx.x >> 5);
  const size_t r0 = static_cast<size_t>(blockIdx.y) * kTileRows;
  if (tile_c * TILE_COLS >= a.cols) return;
  if (r0 + kTileRows <= a.rows && (tile_c + 1) * TILE_COLS <= a.cols)
    quantize_tile<IT, ROW, COL, E5M2, SWZ, WITH_AMAX, true>(a, tile_c, r0, lane);
  else
    quantize_tile<IT, ROW, COL, E5M2, SWZ, WITH_AMAX, false>(a, tile_c, r0, lane);
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
  constexpr int CPL = 4 * ptx::Word<IT>::kElems;
  const bool aligned = (a.cols % CPL == 0) && (reinterpret_cast<uintptr_t>(a.in) % 16 == 0) &&
                       (!row || reinterpret_cast<uintptr_t>(a.out_row) % 8 == 0) &&
                       (!col || reinterpret_cast<uintptr_t>(a.out_col) % 8 == 0) &&
                       (!row || (reinterpret_cast<uintptr_t>(a.sc_row) % 4 == 0 &&
                                 (swz || a.ld_row % 4 == 0))) &&
                       (!col || swz || (reinterpret_cast<uintptr_t>(a.sc_col) % 8 == 0 &&
                                        a.ld_col % 8 == 0));
  if (!aligned) {
    const dim3 grid(ceil_div(a.cols, 32), ceil_div(a.rows, 32));
    if (e5m2) quantize_generic_kernel<IT, true><<<grid, 1024, 0, stream>>>(a, row, col, swz);
    else quantize_generic_kernel<IT, false><<<grid, 1024, 0, stream>>>(a, row, col, swz);
    return;
  }
  const dim3 grid(ceil_div(a.cols, static_cast<size_t>(16 * CPL * kWarpsPerCta)),
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

// Padding of a GEMM-swizzled scale buffer must read as zero. Honours the noop flag (a skipped quantize must
// leave the cached operand's scales alone: reference cast/mxfp8/quantize_mxfp8.cuh noop handling).
__global__ void zero_scales_kernel(uint8_t *p, size_t n, const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * 16;
  for (size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 16; i < n; i += stride) {
    if (i + 16 <= n && (reinterpret_cast<uintptr_t>(p) & 15u) == 0) {
      *reinterpret_cast<uint4 *>(p + i) = make_uint4(0, 0, 0, 0);
    } else {
      for (size_t j = i; j < n && j < i + 16; ++j) p[j] = 0;
    }
  }
}
void zero_scales(uint8_t *p, size_t n, const float *noop, cudaStream_t stream) {
  if (noop == nullptr) {
    TEB_CHECK_CUDA(cudaMemsetAsync(p, 0, n, stream));
    return;
  }
  const int blocks = static_cast<int>(std::min<size_t>(ceil_div(n, static_cast<size_t>(256 * 16)), 296));
  zero_scales_kernel<<<blocks, 256, 0, stream>>>(p, n, noop);
}

}  // namespace

void quantize(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream) {
  quantize_with_dbias_ws(in, out, noop, nullptr, nullptr, stream);
}

namespace {
// Argument checks + launch arguments of one (input, output) pair; false = empty tensor, nothing to do.
bool prepare(const Tensor &in, Tensor &out, const Tensor *noop, Args &a, bool &row, bool &col, bool &e5m2, bool &swz,
             cudaStream_t stream) {
  row = out.has_data();
  col = out.has_columnwise_data();
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
  if (rows == 0 || cols == 0) return false;
  swz = out.with_gemm_swizzled_scales;
  const float *noop_flag = noop ? noop->data.as<float>() : nullptr;
  a = Args{};
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
        zero_scales(a.sc_row, out.scale_inv.numel(), noop_flag, stream);
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
        zero_scales(a.sc_col, out.columnwise_scale_inv.numel(), noop_flag, stream);
    } else {
      a.ld_col = last_dim(out.columnwise_scale_inv);
      TEB_CHECK(a.ld_col >= cols, "Columnwise scale tensor is too narrow: ", a.ld_col);
    }
  }
  a.amax = out.amax.as<float>();
  a.noop = noop_flag;
  e5m2 = otype == kNVTEFloat8E5M2;
  return true;
}

TmaTensor as_tma(const Tensor &in, const Args &a, bool row, bool col, float *dbias_ws) {
  return TmaTensor{in.data.dtype, a.in, a.rows, a.cols, row ? a.out_row : nullptr, col ? a.out_col : nullptr,
                   a.sc_row, a.sc_col, a.ld_row, a.ld_col, a.amax, dbias_ws};
}

void launch_fallback(const Tensor &in, const Args &a, bool row, bool col, bool e5m2, bool swz, cudaStream_t stream) {
  switch (in.data.dtype) {
    case kNVTEBFloat16: launch<__nv_bfloat16>(a, row, col, e5m2, swz, stream); break;
    case kNVTEFloat16: launch<__half>(a, row, col, e5m2, swz, stream); break;
    case kNVTEFloat32: launch<float>(a, row, col, e5m2, swz, stream); break;
    default: TEB_ERROR("Unsupported input dtype ", dtype_name(in.data.dtype));
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}
}  // namespace

// `dbias_ws` != nullptr: also leave per-32-row-slab column sums in dbias_ws[rows / 32][cols] when
// the TMA kernel's column pass can produce them (*fused = true); otherwise only quantize.
void quantize_with_dbias_ws(const Tensor &in, Tensor &out, const Tensor *noop, float *dbias_ws,
                            bool *fused, cudaStream_t stream) {
  require_blackwell("MXFP8 quantization");
  Args a;
  bool row, col, e5m2, swz;
  if (!prepare(in, out, noop, a, row, col, e5m2, swz, stream)) return;
  const bool want_fused = dbias_ws != nullptr && col;
  if (quantize_tma(as_tma(in, a, row, col, want_fused ? dbias_ws : nullptr), e5m2, swz, a.noop, stream)) {
    if (fused != nullptr) *fused = want_fused;
    return;
  }
  launch_fallback(in, a, row, col, e5m2, swz, stream);
}

// reference cast/cast.cu:57-95 (nvte_multi_tensor_quantize / tex.split_quantize): the same cast for a list.
// When every member is eligible for the TMA kernel and they agree on dtype / directions / FP8 type / scale
// layout, the whole list is ONE persistent launch over the concatenated tile list.
void quantize_multi(const Tensor *const *ins, Tensor *const *outs, size_t n, const Tensor *noop,
                    cudaStream_t stream) {
  require_blackwell("MXFP8 quantization");
  std::vector<Args> args;
  std::vector<TmaTensor> ts;
  std::vector<size_t> idx;
  bool row0 = false, col0 = false, e0 = false, s0 = false, uniform = true;
  for (size_t i = 0; i < n; ++i) {
    Args a;
    bool row, col, e5m2, swz;
    if (!prepare(*ins[i], *outs[i], noop, a, row, col, e5m2, swz, stream)) continue;
    if (idx.empty()) { row0 = row; col0 = col; e0 = e5m2; s0 = swz; }
    uniform = uniform && row == row0 && col == col0 && e5m2 == e0 && swz == s0;
    args.push_back(a);
    ts.push_back(as_tma(*ins[i], a, row, col, nullptr));
    idx.push_back(i);
  }
  if (idx.empty()) return;
  if (uniform && idx.size() > 1 &&
      quantize_tma_multi(ts.data(), ts.size(), e0, s0, noop ? noop->data.as<float>() : nullptr, stream))
    return;
  for (size_t j = 0; j < idx.size(); ++j) {
    const bool row = ts[j].out_row != nullptr, col = ts[j].out_col != nullptr;
    const bool e5m2 = outs[idx[j]]->dtype() == kNVTEFloat8E5M2, swz = outs[idx[j]]->with_gemm_swizzled_scales;
    if (quantize_tma(ts[j], e5m2, swz, args[j].noop, stream)) continue;
    launch_fallback(*ins[idx[j]], args[j], row, col, e5m2, swz, stream);
  }
}

}  // namespace mxfp8
}  // namespace teb
