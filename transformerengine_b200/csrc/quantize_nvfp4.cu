// NVFP4 quantization: rowwise 1x16 blocks of x and columnwise 1x16 blocks of x^T
// (stored transposed), E4M3 block scales + FP32 global amax, optional 16-point
// random Hadamard transform (RHT) on the columnwise operand, optional 16x16 (2-D)
// scale tiles for weights; plus the amax / RHT-amax pass that feeds it.
// Replaces, for the Linear path, the reference's
//   cast/nvfp4/quantize_transpose_nvfp4.cuh:237-1470 (+ specialized tuned 1-D kernel)
//   hadamard_transform/hadamard_transform.cu:171-680 (amax, unfused RHT)
//   hadamard_transform/row_cast_col_hadamard_transform_cast_fusion.cu:1042-1374
// Numerics per SURVEY.md section 3A (scale math of cast/nvfp4/core_nvfp4.cuh:40-95).
//
// Design (HBM-bound, 2 B read + 0.5+0.5 B written per element):
//   CTA = 4 warps stacked vertically = 128 rows x 128 columns (64 for fp32). Inside
//   a warp, half-warp h owns 16 rows and lane j one 16-byte vector per row, exactly
//   like the MXFP8 kernel, so the 16 rows a lane holds for one column ARE one
//   columnwise 1x16 block: its maximum, the 16-point Hadamard butterfly (64 fp32
//   add/sub in registers, no tensor-core round trip, no shared memory) and its
//   8 output bytes are private to the lane. Rowwise blocks span 2 (4) lanes: one
//   (two) xor-shuffles. The transposed bytes are staged in 8 KB of XOR-swizzled
//   shared memory so every column leaves as one contiguous 64-byte row segment.
#include <cfloat>

#include "core.h"
#include "philox.cuh"
#include "ptx.cuh"

namespace teb {
namespace nvfp4 {

// tensor-core fast path for eligible bf16 shapes (quantize_nvfp4_tc.cu)
bool quantize_tc(const Tensor &in, size_t rows, size_t cols, uint8_t *out_row, uint8_t *out_col,
                 uint8_t *sc_row, uint8_t *sc_col, size_t ld_row, size_t ld_col,
                 const float *amax_row, const float *amax_col, const uint16_t *matrix,
                 uint32_t sign_mask, bool swz, const float *noop, const uint64_t *rng, cudaStream_t stream);
bool amax_tc(const Tensor &in, size_t rows, size_t cols, float *amax_row, float *amax_col,
             const uint16_t *matrix, uint32_t sign_mask, cudaStream_t stream);

namespace {

constexpr int kWarps = 4;
constexpr int kTileRows = 32 * kWarps;  // 128

__device__ __forceinline__ size_t swz(size_t m, size_t k, size_t K) {
  return ((m >> 7) * (K >> 2) + (k >> 2)) * 512 + (m & 31) * 16 + ((m & 127) >> 5) * 4 + (k & 3);
}

// reference core_nvfp4.cuh: compute_global_encode_scaling_factor_FP4
__device__ __forceinline__ float global_encode_scale(float amax) {
  float s = fminf(448.0f * 6.0f / amax, FLT_MAX);
  return (amax == 0.0f || s == 0.0f) ? 1.0f : s;
}
// Per-block E4M3 decode scale (byte) and the matching encode multiplier.
__device__ __forceinline__ void block_scales(float block_amax, float S_enc, float S_dec,
                                             uint32_t &sf_byte, float &enc) {
  const float sdec = fminf(block_amax * (S_enc * (1.0f / 6.0f)), FLT_MAX);
  sf_byte = ptx::cvt_fp8x2<false>(sdec, 0.0f) & 0xffu;
  const float sf = ptx::fp8x2_to_float2<false>(sf_byte).x;
  enc = fminf(1.0f / (sf * S_dec), FLT_MAX);  // IEEE division, like the reference
}
// 8 floats -> 4 bytes (element 0 in the low nibble of byte 0)
__device__ __forceinline__ uint32_t quant8(const float (&f)[8], float enc) {
  return ptx::cvt_e2m1x2(f[0] * enc, f[1] * enc) | (ptx::cvt_e2m1x2(f[2] * enc, f[3] * enc) << 8) |
         (ptx::cvt_e2m1x2(f[4] * enc, f[5] * enc) << 16) |
         (ptx::cvt_e2m1x2(f[6] * enc, f[7] * enc) << 24);
}
// the same with stochastic rounding: `half` = which 8 values of the 16-value block these are
__device__ __forceinline__ uint32_t quant8_sr(const float (&f)[8], float enc, const uint64_t *rng, uint64_t block,
                                              uint32_t dir, int half) {
  const uint4 rb = philox::block_bits(rng[0], rng[1], block, dir);
  const uint32_t r0 = half ? rb.z : rb.x, r1 = half ? rb.w : rb.y;
  return philox::cvt_rs_e2m1x4(f[0] * enc, f[1] * enc, f[2] * enc, f[3] * enc, r0) |
         (philox::cvt_rs_e2m1x4(f[4] * enc, f[5] * enc, f[6] * enc, f[7] * enc, r1) << 16);
}
__device__ __forceinline__ float bf16_round(float v) {
  return __bfloat162float(__float2bfloat16_rn(v));
}

// In-place 16-point randomized Hadamard transform of one column's 16 rows:
// out_j = bf16(0.25 * sum_i s_i x_i (-1)^popcount(i & j)), s_i = -1 where bit i of mask.
__device__ __forceinline__ void rht16(float (&x)[16], uint32_t mask) {
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if ((mask >> i) & 1u) x[i] = -x[i];
#pragma unroll
  for (int len = 1; len < 16; len <<= 1) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if ((i & len) == 0) {
        const float a = x[i], b = x[i + len];
        x[i] = a + b;
        x[i + len] = a - b;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = bf16_round(x[i] * 0.25f);
}

struct Args {
  const void *in;
  uint8_t *out_row, *out_col;   // [rows, cols/2], [cols, rows/2]
  uint8_t *sc_row, *sc_col;     // E4M3 scale bytes
  size_t rows, cols;
  size_t ld_row, ld_col;        // scale row strides (compact) / padded K (swizzled)
  const float *amax_row, *amax_col;
  float *amax_out_row, *amax_out_col;  // amax pass only
  uint32_t sign_mask;           // RHT signs for the columnwise operand
  const uint16_t *rht_matrix;   // optional 16x16 bf16 RHT matrix: signs are read from its column 0
  const float *noop;
  const uint64_t *rng;          // {seed, offset}: stochastic rounding when not null
};

template <typename W>
__device__ __forceinline__ float elem(const uint4 &v, int k) {
  // k-th element of the lane's 16-byte vector
  const uint32_t w = (&v.x)[k / W::kElems];
  return (W::kElems == 2 && (k & 1)) ? W::hi(w) : W::lo(w);
}

// MODE 0: quantize; MODE 1: amax pass (amax(x) and amax(bf16(RHT(x^T))), nothing else written).
template <typename IT, int MODE, bool ROW, bool COL, bool RHT, bool TWO_D, bool SWZ>
__global__ void __launch_bounds__(kWarps * 32, 3) nvfp4_kernel(const Args a) {
  using W = ptx::Word<IT>;
  constexpr int CPL = 4 * W::kElems;    // columns per lane: 8 (16-bit) / 4 (fp32)
  constexpr int LPB = 16 / CPL;         // lanes per rowwise 1x16 block: 2 / 4
  constexpr int TILE_COLS = 16 * CPL;   // 128 / 64
  constexpr int NB = TILE_COLS / 16;    // rowwise blocks per tile row: 8 / 4
  __shared__ __align__(16) uint8_t col_s[TILE_COLS][64];   // transposed bytes, 8-byte chunks XOR-swizzled
  __shared__ __align__(8) uint8_t sct_s[TILE_COLS][8];      // columnwise scales [col][row-block]
  __shared__ __align__(8) uint8_t scr_s[kTileRows][8];      // rowwise scales [row][col-block]

  if (a.noop != nullptr && a.noop[0] == 1.0f) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int h = lane >> 4, j = lane & 15;
  const size_t r0 = static_cast<size_t>(blockIdx.y) * kTileRows;
  const size_t c0 = static_cast<size_t>(blockIdx.x) * TILE_COLS;
  const size_t rbase = r0 + warp * 32 + h * 16;
  const size_t c = c0 + j * CPL;
  const bool col_ok = c < a.cols;
  const IT *src = reinterpret_cast<const IT *>(a.in) + rbase * a.cols + c;

  uint4 v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    if (col_ok && rbase + i < a.rows) v[i] = ptx::ld_stream_v4(src + static_cast<size_t>(i) * a.cols);
    else v[i] = make_uint4(0u, 0u, 0u, 0u);
  }

  uint32_t sign_mask = a.sign_mask;
  if (RHT && a.rht_matrix != nullptr) {
    // matrix = diag(s) . H16 . 0.25 with H16[i][0] = +1: row i's sign is the sign of entry [i][0]
    sign_mask = 0u;
#pragma unroll
    for (int i = 0; i < 16; ++i) sign_mask |= static_cast<uint32_t>(a.rht_matrix[i * 16] >> 15) << i;
  }
  float S_enc_r = 1.0f, S_dec_r = 1.0f, S_enc_c = 1.0f, S_dec_c = 1.0f;
  if (MODE == 0) {
    if (ROW) { S_enc_r = global_encode_scale(*a.amax_row); S_dec_r = 1.0f / S_enc_r; }
    if (COL) {
      // reference quantize_transpose_nvfp4.cuh:353-363: no columnwise amax -> reuse the rowwise one
      S_enc_c = global_encode_scale(a.amax_col ? *a.amax_col : *a.amax_row);
      S_dec_c = 1.0f / S_enc_c;
    }
  }

  // per-column |x| maxima over my 16 rows (packed words), needed by COL (no RHT), 2-D and amax
  uint32_t cm[4] = {0u, 0u, 0u, 0u};
  if ((COL && !RHT) || TWO_D || MODE == 1) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      cm[0] = W::max(cm[0], W::abs(v[i].x));
      cm[1] = W::max(cm[1], W::abs(v[i].y));
      cm[2] = W::max(cm[2], W::abs(v[i].z));
      cm[3] = W::max(cm[3], W::abs(v[i].w));
    }
  }
  float tile2d = 0.0f;  // 16x16 tile amax (2-D): my 16 rows x the LPB lanes' 16 columns
  if (TWO_D) {
    uint32_t t = W::max(W::max(cm[0], cm[1]), W::max(cm[2], cm[3]));
#pragma unroll
    for (int o = 1; o < LPB; o <<= 1) t = W::max(t, __shfl_xor_sync(0xffffffffu, t, o));
    tile2d = W::hmax(t);
  }

  if (MODE == 1) {
    float ra = fmaxf(fmaxf(W::hmax(cm[0]), W::hmax(cm[1])), fmaxf(W::hmax(cm[2]), W::hmax(cm[3])));
    float ca = 0.0f;
    if (RHT) {
#pragma unroll
      for (int k = 0; k < CPL; ++k) {
        float x[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = elem<W>(v[i], k);
        rht16(x, sign_mask);
#pragma unroll
        for (int i = 0; i < 16; ++i) ca = fmaxf(ca, fabsf(x[i]));
      }
    }
    ra = ptx::warp_max(ra);
    ca = ptx::warp_max(ca);
    if (lane == 0) {
      if (a.amax_out_row != nullptr && ra > 0.0f) ptx::atomic_max_nonneg(a.amax_out_row, ra);
      if (a.amax_out_col != nullptr) {
        const float cv = RHT ? ca : ra;
        if (cv > 0.0f) ptx::atomic_max_nonneg(a.amax_out_col, cv);
      }
    }
    return;
  }

  // ---------------- rowwise: 1x16 blocks along the columns ----------------
  if (ROW) {
    uint8_t *orow = a.out_row + (rbase * a.cols + c) / 2;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float bmax;
      if (TWO_D) {
        bmax = tile2d;
      } else {
        uint32_t t = W::max(W::max(W::abs(v[i].x), W::abs(v[i].y)), W::max(W::abs(v[i].z), W::abs(v[i].w)));
#pragma unroll
        for (int o = 1; o < LPB; o <<= 1) t = W::max(t, __shfl_xor_sync(0xffffffffu, t, o));
        bmax = W::hmax(t);
      }
      uint32_t sf;
      float enc;
      block_scales(bmax, S_enc_r, S_dec_r, sf, enc);
      float f[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = k < CPL ? elem<W>(v[i], k) : 0.0f;
      uint32_t q;
      if (a.rng != nullptr) {
        // block = (row, 16-column block); my CPL values are its `part`-th piece
        const uint64_t blk = (rbase + i) * (a.cols >> 4) + (c >> 4);
        if (CPL == 8) {
          q = quant8_sr(f, enc, a.rng, blk, 0u, static_cast<int>((c >> 3) & 1));
        } else {   // 4 values per lane: one random word per lane
          const uint4 rb = philox::block_bits(a.rng[0], a.rng[1], blk, 0u);
          const uint32_t part = static_cast<uint32_t>((c >> 2) & 3);
          const uint32_t r = part == 0 ? rb.x : part == 1 ? rb.y : part == 2 ? rb.z : rb.w;
          q = philox::cvt_rs_e2m1x4(f[0] * enc, f[1] * enc, f[2] * enc, f[3] * enc, r);
        }
      } else {
        q = quant8(f, enc);
      }
      if (col_ok && rbase + i < a.rows) {
        uint8_t *dst = orow + static_cast<size_t>(i) * (a.cols / 2);
        if (CPL == 8) *reinterpret_cast<uint32_t *>(dst) = q;
        else *reinterpret_cast<uint16_t *>(dst) = static_cast<uint16_t>(q);
      }
      if ((j % LPB) == 0) scr_s[warp * 32 + h * 16 + i][j / LPB] = static_cast<uint8_t>(sf);
    }
  }

  // ---------------- columnwise: my 16 rows of each column are one block ----------------
  if (COL) {
    const int chunk = warp * 2 + h;  // 8-byte chunk (16 rows) inside the 64-byte column row
#pragma unroll
    for (int k = 0; k < CPL; ++k) {
      float x[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = elem<W>(v[i], k);
      float bmax;
      if (RHT) {
        rht16(x, sign_mask);
        bmax = 0.0f;
#pragma unroll
        for (int i = 0; i < 16; ++i) bmax = fmaxf(bmax, fabsf(x[i]));
      } else if (TWO_D) {
        bmax = tile2d;
      } else {
        const uint32_t w = cm[k / W::kElems];
        bmax = (W::kElems == 2 && (k & 1)) ? W::hi(w) : W::lo(w);
      }
      uint32_t sf;
      float enc;
      block_scales(bmax, S_enc_c, S_dec_c, sf, enc);
      const float (&lo)[8] = *reinterpret_cast<const float (*)[8]>(&x[0]);
      const float (&hi)[8] = *reinterpret_cast<const float (*)[8]>(&x[8]);
      const int cl = j * CPL + k;
      *reinterpret_cast<uint2 *>(&col_s[cl][8 * (chunk ^ ((cl >> 3) & 7))]) =
          a.rng != nullptr
              ? make_uint2(quant8_sr(lo, enc, a.rng, (c0 + cl) * (a.rows >> 4) + (rbase >> 4), 1u, 0),
                           quant8_sr(hi, enc, a.rng, (c0 + cl) * (a.rows >> 4) + (rbase >> 4), 1u, 1))
              : make_uint2(quant8(lo, enc), quant8(hi, enc));
      sct_s[cl][chunk] = static_cast<uint8_t>(sf);
    }
  }
  __syncthreads();

  // ---------------- write-out of the staged scales / transposed bytes ----------------
  const int t = threadIdx.x;
  if (ROW) {
    const size_t row = r0 + t;
    if (row < a.rows) {
      const size_t k0 = static_cast<size_t>(blockIdx.x) * NB;
      const size_t nblk = a.cols / 16;
#pragma unroll
      for (int g = 0; g < NB; g += 4) {
        const uint32_t word = *reinterpret_cast<const uint32_t *>(&scr_s[t][g]);
        uint8_t *dst = SWZ ? a.sc_row + swz(row, k0 + g, a.ld_row) : a.sc_row + row * a.ld_row + k0 + g;
        if (k0 + g + 4 <= nblk) {
          *reinterpret_cast<uint32_t *>(dst) = word;
        } else {
#pragma unroll
          for (int b = 0; b < 4; ++b)
            if (k0 + g + b < nblk) dst[b] = static_cast<uint8_t>(word >> (8 * b));
        }
      }
    }
  }
  if (COL && t < TILE_COLS) {
    const size_t col = c0 + t;
    if (col < a.cols) {
      const size_t rows_here = a.rows - r0 < kTileRows ? a.rows - r0 : kTileRows;  // multiple of 16
      const int nchunks = static_cast<int>(rows_here / 16);
      uint8_t *dst = a.out_col + (col * a.rows + r0) / 2;
      const int sw = (t >> 3) & 7;
      if (nchunks == 8 && (a.rows & 31) == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint2 lo = *reinterpret_cast<const uint2 *>(&col_s[t][8 * ((2 * q) ^ sw)]);
          const uint2 hi = *reinterpret_cast<const uint2 *>(&col_s[t][8 * ((2 * q + 1) ^ sw)]);
          *reinterpret_cast<uint4 *>(dst + 16 * q) = make_uint4(lo.x, lo.y, hi.x, hi.y);
        }
      } else {
        for (int q = 0; q < nchunks; ++q)
          *reinterpret_cast<uint2 *>(dst + 8 * q) = *reinterpret_cast<const uint2 *>(&col_s[t][8 * (q ^ sw)]);
      }
      const size_t kb0 = r0 / 16;
      const size_t nblk = a.rows / 16;
#pragma unroll
      for (int g = 0; g < 8; g += 4) {
        const uint32_t word = *reinterpret_cast<const uint32_t *>(&sct_s[t][g]);
        uint8_t *sd = SWZ ? a.sc_col + swz(col, kb0 + g, a.ld_col) : a.sc_col + col * a.ld_col + kb0 + g;
        if (kb0 + g + 4 <= nblk) {
          *reinterpret_cast<uint32_t *>(sd) = word;
        } else {
#pragma unroll
          for (int b = 0; b < 4; ++b)
            if (kb0 + g + b < nblk) sd[b] = static_cast<uint8_t>(word >> (8 * b));
        }
      }
    }
  }
}

// Unfused RHT (reference hadamard_transform.cu:325-565): out_t[c, r] = bf16(RHT(x^T)), and/or
// out[r, c] = bf16(RHT along the last dimension of x). One thread per 16-element group.
__global__ void __launch_bounds__(256)
    rht_kernel(const __nv_bfloat16 *in, __nv_bfloat16 *out, __nv_bfloat16 *out_t, size_t rows,
               size_t cols, uint32_t mask, uint32_t mask_t) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x;
  if (out_t != nullptr) {
    const size_t groups = rows / 16;
    if (idx < groups * cols) {
      const size_t g = idx / cols, c = idx % cols;  // adjacent threads -> adjacent columns
      float x[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = __bfloat162float(in[(g * 16 + i) * cols + c]);
      rht16(x, mask_t);
#pragma unroll
      for (int i = 0; i < 16; ++i) out_t[c * rows + g * 16 + i] = __float2bfloat16_rn(x[i]);
    }
  }
  if (out != nullptr) {
    const size_t groups = cols / 16;
    if (idx < rows * groups) {
      float x[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = __bfloat162float(in[idx * 16 + i]);
      rht16(x, mask);
#pragma unroll
      for (int i = 0; i < 16; ++i) out[idx * 16 + i] = __float2bfloat16_rn(x[i]);
    }
  }
}

__global__ void zero2_kernel(float *a, float *b, const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  if (a != nullptr) *a = 0.0f;
  if (b != nullptr) *b = 0.0f;
}

template <typename IT, int MODE>
void launch(const Args &a, bool row, bool col, bool rht, bool two_d, bool swz, cudaStream_t stream) {
  constexpr int TILE_COLS = 64 * ptx::Word<IT>::kElems;
  const dim3 grid(ceil_div(a.cols, TILE_COLS), ceil_div(a.rows, kTileRows));
  const dim3 block(kWarps * 32);
#define TEB_NVFP4(ROW, COL, RHT, TWOD, SWZ) \
  nvfp4_kernel<IT, MODE, ROW, COL, RHT, TWOD, SWZ><<<grid, block, 0, stream>>>(a)
  if (MODE == 1) {
    if (rht) TEB_NVFP4(false, false, true, false, false); else TEB_NVFP4(false, false, false, false, false);
    return;
  }
#define TEB_NVFP4_DIR(RHT, TWOD, SWZ)                                  \
  do {                                                                  \
    if (row && col) TEB_NVFP4(true, true, RHT, TWOD, SWZ);              \
    else if (row) TEB_NVFP4(true, false, false, TWOD, SWZ);             \
    else TEB_NVFP4(false, true, RHT, TWOD, SWZ);                        \
  } while (false)
  if (two_d) { if (swz) TEB_NVFP4_DIR(false, true, true); else TEB_NVFP4_DIR(false, true, false); }
  else if (rht) { if (swz) TEB_NVFP4_DIR(true, false, true); else TEB_NVFP4_DIR(true, false, false); }
  else { if (swz) TEB_NVFP4_DIR(false, false, true); else TEB_NVFP4_DIR(false, false, false); }
#undef TEB_NVFP4_DIR
#undef TEB_NVFP4
}

template <int MODE>
void dispatch(const Tensor &in, const Args &a, bool row, bool col, bool rht, bool two_d, bool swz,
              cudaStream_t stream) {
  switch (in.data.dtype) {
    case kNVTEBFloat16: launch<__nv_bfloat16, MODE>(a, row, col, rht, two_d, swz, stream); break;
    case kNVTEFloat16: launch<__half, MODE>(a, row, col, rht, two_d, swz, stream); break;
    case kNVTEFloat32: launch<float, MODE>(a, row, col, rht, two_d, swz, stream); break;
    default: TEB_ERROR("Unsupported input dtype ", dtype_name(in.data.dtype));
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}

void check_input(const Tensor &in, size_t &rows, size_t &cols) {
  TEB_CHECK(in.has_data() && in.data.dptr != nullptr, "Cannot quantize tensor without rowwise data.");
  TEB_CHECK(is_high_precision(in.data.dtype), "Input must be in higher precision.");
  rows = in.rows();
  cols = in.cols();
  const size_t vec = 16 / (dtype_bits(in.data.dtype) / 8);
  TEB_CHECK(cols % 16 == 0, "NVFP4 needs the last dimension to be a multiple of 16 (got ", cols, ")");
  TEB_CHECK(cols % vec == 0 && reinterpret_cast<uintptr_t>(in.data.dptr) % 16 == 0,
            "NVFP4 kernels need 16-byte aligned rows");
}

}  // namespace

void quantize(const Tensor &in, Tensor &out, const QuantConfig &cfg, const Tensor *noop,
              bool with_rht, int sign_mask_t, const Tensor *rht_matrix, cudaStream_t stream) {
  require_blackwell("NVFP4 quantization");
  // stochastic rounding draws its bits from {seed, offset} in cfg.rng_state
  // (reference quantize_transpose_nvfp4.cuh:1395-1404)
  const uint64_t *rng = nullptr;
  if (cfg.stochastic_rounding) {
    TEB_CHECK(cfg.rng_state != nullptr, "NVFP4 stochastic rounding needs an RNG state in the quantization config");
    const Tensor &rs = lookup_checked(cfg.rng_state);
    TEB_CHECK(rs.data.dtype == kNVTEInt64, "RNG state should contain 2 64-bit values.");
    TEB_CHECK(rs.data.numel() == 2 && rs.data.dptr != nullptr, "Shape of the RNG state should be [2]");
    rng = rs.data.as<uint64_t>();
  }
  TEB_CHECK(cfg.nvfp4_4over6_mode == 0, "NVFP4 4over6 quantization is not implemented");
  TEB_CHECK(!out.row_scaled_nvfp4, "Row-scaled NVFP4 is not implemented");
  TEB_CHECK(out.nvfp4_e4m3_max == 448, "Only the standard NVFP4 E4M3 bound (448) is implemented");
  TEB_CHECK(!(with_rht && cfg.nvfp4_2d_quantization), "RHT and 2-D quantization cannot be combined");
  size_t rows, cols;
  check_input(in, rows, cols);
  const bool row = out.data.dptr != nullptr, col = out.columnwise_data.dptr != nullptr;
  TEB_CHECK(row || col, "Either rowwise or columnwise output data need to be allocated.");
  if (rows == 0 || cols == 0) return;
  Args a{};
  a.in = in.data.dptr;
  a.rows = rows;
  a.cols = cols;
  a.noop = noop ? noop->data.as<float>() : nullptr;
  a.rng = rng;
  const bool swz = out.with_gemm_swizzled_scales;
  auto last = [](const Buffer &b) { return b.shape.ndim ? b.shape.data[b.shape.ndim - 1] : 0; };
  if (row) {
    TEB_CHECK(out.data.dtype == kNVTEFloat4E2M1, "Output must have FP4 type.");
    TEB_CHECK(out.data.numel() == rows * cols, "Input and output shapes need to match.");
    TEB_CHECK(out.scale_inv.dptr != nullptr, "Scaling tensor must be allocated");
    TEB_CHECK(out.amax.dptr != nullptr, "NVFP4 rowwise quantization needs the global amax");
    a.out_row = out.data.as<uint8_t>();
    a.sc_row = out.scale_inv.as<uint8_t>();
    a.amax_row = out.amax.as<float>();
    a.ld_row = swz ? round_up(cols / 16, 4) : last(out.scale_inv);
    TEB_CHECK(a.ld_row >= cols / 16 && a.ld_row % 4 == 0, "Rowwise scale tensor has a bad width");
    TEB_CHECK(out.scale_inv.numel() >= (swz ? round_up(rows, 128) : rows) * a.ld_row,
              "Rowwise scale tensor is too small");
    if (swz && (rows % 128 != 0 || (cols / 16) % 4 != 0))
      TEB_CHECK_CUDA(cudaMemsetAsync(a.sc_row, 0, out.scale_inv.numel(), stream));
  }
  if (col) {
    TEB_CHECK(rows % 16 == 0, "NVFP4 columnwise output needs rows % 16 == 0 (got ", rows, ")");
    TEB_CHECK(out.columnwise_data.dtype == kNVTEFloat4E2M1, "Output must have FP4 type.");
    TEB_CHECK(out.columnwise_data.numel() == rows * cols, "Input and output shapes need to match.");
    TEB_CHECK(out.columnwise_scale_inv.dptr != nullptr, "Columnwise scaling tensor must be allocated");
    a.out_col = out.columnwise_data.as<uint8_t>();
    a.sc_col = out.columnwise_scale_inv.as<uint8_t>();
    a.amax_col = out.columnwise_amax.as<float>();
    if (a.amax_col == nullptr) a.amax_col = out.amax.as<float>();
    TEB_CHECK(a.amax_col != nullptr, "NVFP4 columnwise quantization needs a global amax");
    if (a.amax_row == nullptr) a.amax_row = a.amax_col;
    a.ld_col = swz ? round_up(rows / 16, 4) : last(out.columnwise_scale_inv);
    TEB_CHECK(a.ld_col >= rows / 16 && a.ld_col % 4 == 0, "Columnwise scale tensor has a bad width");
    TEB_CHECK(out.columnwise_scale_inv.numel() >= (swz ? round_up(cols, 128) : cols) * a.ld_col,
              "Columnwise scale tensor is too small");
    if (swz && (cols % 128 != 0 || (rows / 16) % 4 != 0))
      TEB_CHECK_CUDA(cudaMemsetAsync(a.sc_col, 0, out.columnwise_scale_inv.numel(), stream));
  }
  a.sign_mask = static_cast<uint32_t>(sign_mask_t) & 0xffffu;
  if (with_rht && rht_matrix != nullptr && rht_matrix->data.dptr != nullptr) {
    TEB_CHECK(rht_matrix->data.dtype == kNVTEBFloat16 && rht_matrix->data.numel() == 256,
              "The Hadamard matrix must be a 16x16 bf16 tensor");
    a.rht_matrix = rht_matrix->data.as<uint16_t>();
  }
  // Fast path: 1-D scales, bf16 input, full 128x128 tiles; the transform is the caller's matrix
  // when given, else built from the sign mask (identity without RHT). 2-D scaling stays generic.
  const bool rht = with_rht && col;
  if (!cfg.nvfp4_2d_quantization &&
      quantize_tc(in, rows, cols, a.out_row, a.out_col, a.sc_row, a.sc_col, a.ld_row, a.ld_col,
                  a.amax_row, a.amax_col, rht ? a.rht_matrix : nullptr,
                  rht ? a.sign_mask : 0x10000u, swz, a.noop, a.rng, stream))
    return;
  dispatch<0>(in, a, row, col, rht, cfg.nvfp4_2d_quantization, swz, stream);
}

void rht_amax(const Tensor &in, Tensor &out, int sign_mask, int sign_mask_t, cudaStream_t stream) {
  (void)sign_mask;
  require_blackwell("The RHT amax kernel");
  size_t rows, cols;
  check_input(in, rows, cols);
  Args a{};
  a.in = in.data.dptr;
  a.rows = rows;
  a.cols = cols;
  a.amax_out_row = out.amax.as<float>();
  a.amax_out_col = out.columnwise_amax.as<float>();
  TEB_CHECK(a.amax_out_row != nullptr || a.amax_out_col != nullptr,
            "Output tensor has no amax buffers");
  a.sign_mask = static_cast<uint32_t>(sign_mask_t) & 0xffffu;
  const bool rht = a.amax_out_col != nullptr;
  if (rht) TEB_CHECK(rows % 16 == 0, "RHT along the rows needs rows % 16 == 0 (got ", rows, ")");
  zero2_kernel<<<1, 1, 0, stream>>>(a.amax_out_row, a.amax_out_col, nullptr);
  if (rows == 0 || cols == 0) return;
  if (amax_tc(in, rows, cols, a.amax_out_row, a.amax_out_col, nullptr, rht ? a.sign_mask : 0x10000u, stream))
    return;
  dispatch<1>(in, a, false, false, rht, false, false, stream);
}

void rht(const Tensor &in, Tensor &out, int sign_mask, int sign_mask_t, cudaStream_t stream) {
  require_blackwell("The Hadamard transform");
  TEB_CHECK(in.data.dtype == kNVTEBFloat16 && in.data.dptr != nullptr,
            "The unfused Hadamard transform takes a bf16 input");
  const size_t rows = in.rows(), cols = in.cols();
  __nv_bfloat16 *o = out.data.as<__nv_bfloat16>();
  __nv_bfloat16 *ot = out.columnwise_data.as<__nv_bfloat16>();
  TEB_CHECK(o != nullptr || ot != nullptr, "Output tensor has no data buffers");
  if (o) TEB_CHECK(cols % 16 == 0 && out.data.dtype == kNVTEBFloat16, "bad rowwise RHT output");
  if (ot) TEB_CHECK(rows % 16 == 0 && out.columnwise_data.dtype == kNVTEBFloat16, "bad columnwise RHT output");
  const size_t n = rows * cols / 16;
  if (n == 0) return;
  rht_kernel<<<static_cast<unsigned>(ceil_div(n, 256)), 256, 0, stream>>>(
      in.data.as<__nv_bfloat16>(), o, ot, rows, cols, static_cast<uint32_t>(sign_mask) & 0xffffu,
      static_cast<uint32_t>(sign_mask_t) & 0xffffu);
  TEB_CHECK_CUDA(cudaGetLastError());
}

}  // namespace nvfp4
}  // namespace teb
