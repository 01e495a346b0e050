// NVFP4 quantization, tensor-core variant for bf16 activations / gradients (the shapes the
// Linear path produces: rows and cols multiples of 128). Same bytes as quantize_nvfp4.cu, about
// 2.4x fewer instructions per element: that kernel was issue-bound (ncu: 115 M warp instructions
// for 16384x8192, 56 % issue-slot utilisation, DRAM at 24 %), so the work was restructured:
//   * TMA brings a 128 x 128 bf16 tile into 128B-swizzled shared memory (no address arithmetic,
//     32 KB in flight per CTA, 5 CTAs per SM);
//   * the 16-point random Hadamard transform of the column-wise operand is one
//     ldmatrix.x4.trans + two mma.sync.m16n8k16 (bf16 in, fp32 accumulate) per 16 x 16 block
//     against the 16 x 16 matrix the caller passes -- like the reference, which also runs it on
//     the tensor cores (hadamard_transform/*.cu); the result is rounded to bf16 before amax and
//     cast, as the reference does (row_cast_col_hadamard_transform_cast_fusion.cu:776-788).
//     The B operand's columns are permuted so that a thread ends up with 4 consecutive
//     transformed rows = one 16-bit store of packed FP4. Without RHT the same path runs with
//     the identity matrix (exact);
//   * block amax in packed bf16x2 (max.xorsign.abs.bf16x2: 1 instruction per 2 elements);
//   * the per-block encode multiplier 1 / (e4m3(S_dec_b) * S_dec) depends only on the E4M3 byte:
//     a 128-entry table per direction is built once per CTA with exactly the reference's
//     arithmetic (core_nvfp4.cuh:40-95), replacing an IEEE division per block.
// Replaces reference cast/nvfp4/quantize_transpose_nvfp4.cuh, hadamard_transform.cu:567-680 (amax)
// and row_cast_col_hadamard_transform_cast_fusion.cu:1042-1374 for eligible shapes.
#include <cfloat>

#include "core.h"
#include "philox.cuh"
#include "ptx.cuh"
#include "sm100.cuh"
#include "tensormap.h"

namespace teb {
namespace nvfp4 {

namespace {

constexpr int kTile = 128;
constexpr int kThreads = 256;

struct TcArgs {
  uint8_t *out_row, *out_col;   // [rows, cols/2], [cols, rows/2]
  uint8_t *sc_row, *sc_col;     // E4M3 scale bytes (compact or GEMM-swizzled)
  size_t rows, cols;
  size_t ld_row, ld_col;
  const float *amax_row, *amax_col;
  float *amax_out_row, *amax_out_col;   // amax mode
  const uint16_t *matrix;               // 16x16 bf16 transform; nullptr: built from sign_mask
  uint32_t sign_mask;                   // bit i set: row i of 0.25*H16 negated; 0x10000 = identity
  const float *noop;
  const uint64_t *rng;                  // {seed, offset} for stochastic rounding
};

__device__ __forceinline__ size_t swz(size_t m, size_t k, size_t K) {
  return ((m >> 7) * (K >> 2) + (k >> 2)) * 512 + (m & 31) * 16 + ((m & 127) >> 5) * 4 + (k & 3);
}
__device__ __forceinline__ float global_encode_scale(float amax) {
  float s = fminf(448.0f * 6.0f / amax, FLT_MAX);
  return (amax == 0.0f || s == 0.0f) ? 1.0f : s;
}
// |max| of packed bf16 pairs; the sign bits of the result are garbage (xor of the inputs')
__device__ __forceinline__ uint32_t amax2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.xorsign.abs.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ float amax_of_pair(uint32_t m) {
  return fmaxf(__uint_as_float((m << 16) & 0x7fffffffu), __uint_as_float(m & 0x7fff0000u));
}
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16x2_rn(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// two packed bf16 words (4 values) * enc -> two bytes of packed E2M1 (value 0 in the low nibble
// of byte 0); the products are fp32 (packed FMUL2), rounded once by the conversion
__device__ __forceinline__ uint32_t quant4(uint32_t w0, uint32_t w1, float enc) {
  uint32_t r;
  asm("{\n\t.reg .b8 b0, b1;\n\t.reg .b16 h;\n\t.reg .b64 p0, p1, e;\n\t.reg .f32 a0, a1, a2, a3;\n\t"
      "mov.b64 p0, {%1, %2};\n\tmov.b64 p1, {%3, %4};\n\tmov.b64 e, {%5, %5};\n\t"
      "mul.f32x2 p0, p0, e;\n\tmul.f32x2 p1, p1, e;\n\t"
      "mov.b64 {a0, a1}, p0;\n\tmov.b64 {a2, a3}, p1;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b0, a1, a0;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b1, a3, a2;\n\t"
      "mov.b16 h, {b0, b1};\n\tcvt.u32.u16 %0, h;\n\t}"
      : "=r"(r)
      : "r"(w0 << 16), "r"(w0 & 0xffff0000u), "r"(w1 << 16), "r"(w1 & 0xffff0000u), "f"(enc));
  return r;
}
// E4M3 byte of a non-negative scale, zero-extended (RN, saturate-to-finite)
__device__ __forceinline__ uint32_t e4m3_byte(float v) {
  uint32_t r;
  asm("{\n\t.reg .b16 h;\n\tcvt.rn.satfinite.e4m3x2.f32 h, 0f00000000, %1;\n\tcvt.u32.u16 %0, h;\n\t}"
      : "=r"(r) : "f"(v));
  return r;
}
// max(|lo|, |hi|) of a packed amax word whose sign bits are garbage
__device__ __forceinline__ float amax_of_pair_abs(uint32_t m) {
  return fmaxf(fabsf(__uint_as_float(m << 16)), fabsf(__uint_as_float(m & 0xffff0000u)));
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
// eight packed bf16 words (16 values) * enc -> two words of packed E2M1 (value 0 in the low nibble
// of byte 0)
__device__ __forceinline__ uint32_t quant8(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, float enc) {
  uint32_t r;
  asm("{\n\t.reg .b8 b0, b1, b2, b3;\n\t.reg .b64 p0, p1, p2, p3, e;\n\t"
      ".reg .f32 a0, a1, a2, a3, a4, a5, a6, a7;\n\t"
      "mov.b64 p0, {%1, %2};\n\tmov.b64 p1, {%3, %4};\n\tmov.b64 p2, {%5, %6};\n\tmov.b64 p3, {%7, %8};\n\t"
      "mov.b64 e, {%9, %9};\n\t"
      "mul.f32x2 p0, p0, e;\n\tmul.f32x2 p1, p1, e;\n\tmul.f32x2 p2, p2, e;\n\tmul.f32x2 p3, p3, e;\n\t"
      "mov.b64 {a0, a1}, p0;\n\tmov.b64 {a2, a3}, p1;\n\tmov.b64 {a4, a5}, p2;\n\tmov.b64 {a6, a7}, p3;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b0, a1, a0;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b1, a3, a2;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b2, a5, a4;\n\t"
      "cvt.rn.satfinite.e2m1x2.f32 b3, a7, a6;\n\t"
      "mov.b32 %0, {b0, b1, b2, b3};\n\t}"
      : "=r"(r)
      : "r"(w0 << 16), "r"(w0 & 0xffff0000u), "r"(w1 << 16), "r"(w1 & 0xffff0000u), "r"(w2 << 16),
        "r"(w2 & 0xffff0000u), "r"(w3 << 16), "r"(w3 & 0xffff0000u), "f"(enc));
  return r;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_v2(uint32_t addr, uint32_t x, uint32_t y) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void sts_u8(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// eight packed bf16 words * enc -> two words of packed E2M1, stochastic rounding (4 random words)
__device__ __forceinline__ uint32_t quant8_sr(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, float enc,
                                              uint32_t r0, uint32_t r1) {
  const uint32_t lo = philox::cvt_rs_e2m1x4(bf_lo(w0) * enc, bf_hi(w0) * enc, bf_lo(w1) * enc, bf_hi(w1) * enc, r0);
  const uint32_t hi = philox::cvt_rs_e2m1x4(bf_lo(w2) * enc, bf_hi(w2) * enc, bf_lo(w3) * enc, bf_hi(w3) * enc, r1);
  return lo | (hi << 16);
}
// One 1x16 block held by one thread as two 16-byte vectors: block amax -> E4M3 scale byte ->
// encode multiplier from the table -> 8 bytes of packed E2M1 (round to nearest even, or stochastic
// rounding with the block's Philox bits).
struct SrState {
  uint64_t seed, offset;
};
template <bool SR>
__device__ __forceinline__ uint2 quant_block16(const uint4 &v0, const uint4 &v1, float k, uint32_t lut_base,
                                               uint32_t &sf, const SrState &sr, uint64_t block, uint32_t dir) {
  const uint32_t m = amax2(amax2(amax2(v0.x, v0.y), amax2(v0.z, v0.w)),
                           amax2(amax2(v1.x, v1.y), amax2(v1.z, v1.w)));
  sf = e4m3_byte(fminf(amax_of_pair_abs(m) * k, FLT_MAX));
  const float enc = lds_f32(lut_base + sf * 4);
  if (SR) {
    const uint4 rb = philox::block_bits(sr.seed, sr.offset, block, dir);
    return make_uint2(quant8_sr(v0.x, v0.y, v0.z, v0.w, enc, rb.x, rb.y),
                      quant8_sr(v1.x, v1.y, v1.z, v1.w, enc, rb.z, rb.w));
  }
  return make_uint2(quant8(v0.x, v0.y, v0.z, v0.w, enc), quant8(v1.x, v1.y, v1.z, v1.w, enc));
}
__device__ __forceinline__ void sts_u16(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(static_cast<uint16_t>(v)) : "memory");
}
__device__ __forceinline__ void sts_u8_if(uint32_t addr, uint32_t v, bool pred) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p st.shared.u8 [%0], %1;\n\t}" ::"r"(addr),
               "r"(v), "r"(static_cast<uint32_t>(pred))
               : "memory");
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t saddr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(saddr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, "
      "{%8, %9}, {%10, %10, %10, %10};"
      : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.0f));
}

// Shared-memory image of one tile: two TMA boxes of [128 rows][128 B], 128B-swizzled.
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk16) {
  return base + (chunk16 >> 3) * 16384 + row * 128 + (((chunk16 & 7) ^ (row & 7)) << 4);
}

// B fragments of the 16x16 transform for the two n-tiles, columns permuted so that a thread's
// four results of one block are the consecutive transformed rows 4t .. 4t+3.
__device__ __forceinline__ uint32_t transform_entry(const uint16_t *m, uint32_t mask, int i, int j) {
  if (m != nullptr) return m[i * 16 + j];
  if (mask & 0x10000u) return i == j ? 0x3f80u : 0u;                      // identity
  const uint32_t neg = (__popc(i & j) + ((mask >> i) & 1u)) & 1u;         // diag(s) . H16 . 0.25
  return 0x3e80u | (neg << 15);
}
__device__ __forceinline__ void load_transform(const uint16_t *m, uint32_t mask, int lane,
                                               uint32_t (&b)[2][2]) {
  const int t = lane & 3, gid = lane >> 2;
#pragma unroll
  for (int tt = 0; tt < 2; ++tt) {
    const int j = 4 * (gid >> 1) + 2 * tt + (gid & 1);
#pragma unroll
    for (int hk = 0; hk < 2; ++hk) {
      const int i0 = 8 * hk + 2 * t;
      b[tt][hk] = transform_entry(m, mask, i0, j) | (transform_entry(m, mask, i0 + 1, j) << 16);
    }
  }
}

#ifndef TEB_FP4_STAGES
#define TEB_FP4_STAGES 3
#endif
#ifndef TEB_FP4_CTAS
#define TEB_FP4_CTAS 2
#endif
#ifndef TEB_FP4_FORWARD
#define TEB_FP4_FORWARD 1
#endif
constexpr int kStages = TEB_FP4_STAGES;
constexpr int kCtasPerSm = TEB_FP4_CTAS;
constexpr int kStageBytes = 32768;
constexpr int kSmemBytes = kStages * kStageBytes + 8192 + 1024 + 1024 + 1024 + 64 + 1024 /*align*/;

// MODE 0: quantize; MODE 1: amax(x) and amax(bf16(transform(x^T))).
// Persistent: CTA b processes tiles b, b + grid, ... through a 3-stage TMA ring, so the loads of
// the next two tiles are in flight while a tile is being converted (2 CTAs x 96 KB per SM).
// Quantize mode: every thread owns whole 1x16 blocks in both directions (no shuffles, one scale
// computation per block); warp w owns rows 16w .. 16w+15 of the tile:
//   rowwise   : 2 x LDS.128 = one block; results kept in registers until the warp's rows are dead,
//               then staged over them and stored by the warp's own TMA store (box 16 rows x 64 B);
//   columnwise: ldmatrix.trans + mma (transform) -> bf16 -> scratch image over the warp's own rows
//               (32 B per column) -> 2 x LDS.128 = one block -> staged [128 cols][64 B] -> one TMA
//               store per tile;
//   scales    : staged in tcgen05 atom order (SWZ) and stored by two 1 KB bulk copies.
// No global store instruction is issued on the SWZ path: ncu showed the LSU data pipe at 73 % with
// 24 points of it spent on 32/64-byte store fragments.
template <int MODE, bool ROW, bool COL, bool SWZ, bool SR>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
    nvfp4_tc_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_row,
                    const __grid_constant__ CUtensorMap tmap_col, const TcArgs a) {
  using namespace sm100;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t *col_s = smem + kStages * kStageBytes;                   // [128 cols][64 B], TMA 64B swizzle
  uint8_t *scr_s = col_s + 8192;                                   // rowwise scales of the tile, 1 KB
  uint8_t *sct_s = scr_s + 1024;                                   // columnwise scales of the tile, 1 KB
  float *lut_r = reinterpret_cast<float *>(sct_s + 1024);          // [128]
  float *lut_c = lut_r + 128;                                      // [128]
  uint64_t *bar = reinterpret_cast<uint64_t *>(lut_c + 128);       // [kStages]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_x = static_cast<int>(a.cols / kTile);
  const int num_tiles = tiles_x * static_cast<int>(a.rows / kTile);
  // Programmatic dependent launch: barrier setup and descriptor prefetch overlap the tail of the
  // previous kernel in the stream (usually our own amax pass); every global access comes after.
  if (tid == 0) {
    tma_prefetch_desc(&tmap);
    if (MODE == 0 && ROW) tma_prefetch_desc(&tmap_row);
    if (MODE == 0 && COL) tma_prefetch_desc(&tmap_col);
    for (int s = 0; s < kStages; ++s) mbar_init(&bar[s], 1);
    fence_barrier_init();
  }
  grid_dependency_wait();
  grid_launch_dependents();
  if (a.noop != nullptr && a.noop[0] == 1.0f) return;
  SrState sr{0, 0};
  if (SR) {
    sr.seed = a.rng[0];
    sr.offset = a.rng[1];
  }

  // Tried (round 2): walking the cast pass BACKWARDS so that it starts on the tail of the amax pass still in L2
  // (-DTEB_FP4_FORWARD=0). No gain on the 256 MiB config-3 tensor (76.7 vs 74.1 us: the streamed TMA loads do not
  // stay resident) and the reversed walk produced wrong data bytes, so the forward order stays.
  auto tile_of = [&](int k) {
    const int lin = blockIdx.x + k * gridDim.x;
    return (MODE == 0 && !TEB_FP4_FORWARD) ? num_tiles - 1 - lin : (lin < num_tiles ? lin : -1);
  };
  auto issue = [&](int k) {   // thread 0: TMA of this CTA's k-th tile into stage k % kStages
    const int tile_id = tile_of(k);
    if (tile_id < 0) return;
    const int s = k % kStages;
    const int x = (tile_id % tiles_x) * kTile * 2, y = (tile_id / tiles_x) * kTile;
    mbar_expect_tx(&bar[s], kStageBytes);
    tma_load_2d(smem + s * kStageBytes, &tmap, &bar[s], x, y);
    tma_load_2d(smem + s * kStageBytes + 16384, &tmap, &bar[s], x + 128, y);
  };
  if (tid == 0)
    for (int k = 0; k < kStages; ++k) issue(k);

  float k_r = 0.0f, k_c = 0.0f;
  if (MODE == 0) {
    // encode-multiplier tables: entry b = min(1 / (e4m3(b) * S_dec), FLT_MAX)
    if (ROW) {
      const float S_enc = global_encode_scale(*a.amax_row);
      const float S_dec = 1.0f / S_enc;
      k_r = S_enc * (1.0f / 6.0f);
      if (tid < 128) lut_r[tid] = fminf(1.0f / (ptx::fp8x2_to_float2<false>(tid).x * S_dec), FLT_MAX);
    }
    if (COL) {
      const float S_enc = global_encode_scale(a.amax_col ? *a.amax_col : *a.amax_row);
      const float S_dec = 1.0f / S_enc;
      k_c = S_enc * (1.0f / 6.0f);
      if (tid < 128) lut_c[tid] = fminf(1.0f / (ptx::fp8x2_to_float2<false>(tid).x * S_dec), FLT_MAX);
    }
  }
  uint32_t bfrag[2][2];
  if (COL) load_transform(a.matrix, a.sign_mask, lane, bfrag);
  __syncthreads();           // barrier init + tables visible

  uint32_t racc = 0u, cacc = 0u;   // amax mode accumulators (packed)
  const uint32_t cols_base = smem_u32(col_s), scr_base = smem_u32(scr_s), sct_base = smem_u32(sct_s);
  const uint32_t lutr_base = smem_u32(lut_r), lutc_base = smem_u32(lut_c);

  for (int k = 0;; ++k) {
    const int tile_id = tile_of(k);
    if (tile_id < 0) break;
    const int s = k % kStages;
    const size_t r0 = static_cast<size_t>(tile_id / tiles_x) * kTile;
    const size_t c0 = static_cast<size_t>(tile_id % tiles_x) * kTile;
    mbar_wait(&bar[s], (k / kStages) & 1);
    const uint32_t tile = smem_u32(smem + s * kStageBytes);

    if (MODE == 0) {
      // ---------------- rowwise: 4 blocks per thread, loaded first ----------------
      const int rl = warp * 16 + (lane & 7);          // + 8 * (it & 1)
      const int bq = lane >> 3;                       // block (bq + 4 * (it >> 1)) of the row
      uint4 rv[4][2];
      if (ROW) {
        const uint32_t a0 = tile + rl * 128 + (((2 * bq) ^ (lane & 7)) << 4);
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const uint32_t ad = a0 + (it >> 1) * 16384 + (it & 1) * 1024;
          rv[it][0] = lds_v4(ad);
          rv[it][1] = lds_v4(ad ^ 16u);
        }
      }
      // The previous tile's staged bytes must have left shared memory before they are overwritten,
      // and its stage may be refilled once every warp's row store has read it.
      if (k > 0) {
        if (lane == 0) bulk_wait_read_all();
        __syncthreads();
        if (tid == 0) issue(k - 1 + kStages);
      }
      uint2 rq[4];
      if (ROW) {
        // scale staging: SWZ: atom order (kb >> 2) * 512 + (m & 31) * 16 + (m >> 5) * 4 + (kb & 3); else [m][8]
        const int m0 = rl;
        const uint32_t sbase = SWZ ? scr_base + (m0 & 31) * 16 + (m0 >> 5) * 4 + bq : scr_base + m0 * 8 + bq;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          uint32_t sf;
          rq[it] = quant_block16<SR>(rv[it][0], rv[it][1], k_r, lutr_base, sf, sr,
                                     (r0 + rl + 8 * (it & 1)) * (a.cols >> 4) + (c0 >> 4) + bq + 4 * (it >> 1), 0u);
          // row m0 + 8 * (it & 1), block bq + 4 * (it >> 1)
          sts_u8(sbase + (SWZ ? (it & 1) * 128 + (it >> 1) * 512 : (it & 1) * 64 + (it >> 1) * 4), sf);
        }
      }
      // ---------------- columnwise: transform on the tensor cores, then 4 blocks per thread -------
      if (COL) {
        const int t = lane & 3, gid = lane >> 2;
        const int mi = lane >> 3, rr = lane & 7;
        const int row_l = warp * 16 + (mi >> 1) * 8 + rr;   // my ldmatrix row address
        const uint32_t lbase = tile + row_l * 128;
        const uint32_t lx = static_cast<uint32_t>(((mi & 1) ^ (row_l & 7)) << 4);
        uint32_t pq[8][4];
#pragma unroll
        for (int cg = 0; cg < 8; ++cg) {
          uint32_t af[4];
          ldmatrix_x4_trans(lbase + (cg >> 2) * 16384 + (lx ^ (((2 * cg) & 6) << 4)), af);
          float d0[4], d1[4];
          mma_bf16_16816(d0, af, bfrag[0][0], bfrag[0][1]);
          mma_bf16_16816(d1, af, bfrag[1][0], bfrag[1][1]);
          // column 16cg + gid: transformed rows 4t..4t+3 = d0[0], d0[1], d1[0], d1[1]; column + 8: d0[2..3], d1[2..3]
          pq[cg][0] = pack_bf16x2_rn(d0[0], d0[1]);
          pq[cg][1] = pack_bf16x2_rn(d1[0], d1[1]);
          pq[cg][2] = pack_bf16x2_rn(d0[2], d0[3]);
          pq[cg][3] = pack_bf16x2_rn(d1[2], d1[3]);
        }
        __syncwarp();   // the warp's 16 rows have been read for the last time: reuse them as scratch
        // scratch image: column c of the tile -> 32 bytes (16 transformed bf16) at
        //   (c >> 6) * 16384 + warp * 2048 + (c & 63) * 32, the two 16-byte halves swapped when (c >> 2) & 1
        const uint32_t scratch = tile + warp * 2048;
        const uint32_t wb = scratch + gid * 32 + ((((t >> 1) ^ (gid >> 2)) & 1) << 4) + (t & 1) * 8;
#pragma unroll
        for (int cg = 0; cg < 8; ++cg) {
          const uint32_t ad = wb + (cg >> 2) * 16384 + (cg & 3) * 512;
          sts_v2(ad, pq[cg][0], pq[cg][1]);
          sts_v2(ad + 256, pq[cg][2], pq[cg][3]);
        }
        __syncwarp();
        const uint32_t rb = scratch + lane * 32 + (((lane >> 2) & 1) << 4);
        // staged packed FP4: column c, bytes 8 warp .. 8 warp + 7 of its 64, TMA 64B swizzle
        // (16-byte chunk index ^= (c >> 1) & 3)
        const uint32_t stg = cols_base + lane * 64 + ((((warp >> 1) ^ (lane >> 1)) & 3) << 4) + (warp & 1) * 8;
        // scale staging: SWZ: atom order with m = column, kb = warp; else [c][8]
        const uint32_t tsc = SWZ ? sct_base + (warp >> 2) * 512 + lane * 16 + (warp & 3) : sct_base + lane * 8 + warp;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const uint32_t ad = rb + (it >> 1) * 16384 + (it & 1) * 1024;
          const uint4 v0 = lds_v4(ad), v1 = lds_v4(ad ^ 16u);
          uint32_t sf;
          const uint2 q = quant_block16<SR>(v0, v1, k_c, lutc_base, sf, sr,
                                            (c0 + lane + 32 * it) * (a.rows >> 4) + (r0 >> 4) + warp, 1u);
          sts_v2(stg + it * 2048, q.x, q.y);
          sts_u8(tsc + (SWZ ? it * 4 : it * 256), sf);
        }
      }
      if (ROW) {
        // the warp's rows (input, then scratch) are dead: stage the rowwise bytes over them as
        // [16 rows][64 B] with the TMA 64B swizzle and let the warp's lane 0 store the box
        __syncwarp();
        const int r = lane & 7;   // + 8 * (it & 1)
        const uint32_t rs = tile + warp * 2048 + r * 64 + (bq & 1) * 8;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int rr2 = r + 8 * (it & 1), b = bq + 4 * (it >> 1);
          sts_v2(rs + (it & 1) * 512 + ((((b >> 1) ^ (rr2 >> 1)) & 3) << 4), rq[it].x, rq[it].y);
        }
      }
      fence_async_smem();
      if (ROW) {
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmap_row, reinterpret_cast<const void *>(smem + s * kStageBytes + warp * 2048),
                       static_cast<int>(c0 / 2), static_cast<int>(r0) + warp * 16);
          bulk_commit();
        }
      }
      __syncthreads();   // staged columnwise bytes and scales complete
      if (tid == 0) {
        if (COL) tma_store_2d(&tmap_col, col_s, static_cast<int>(r0 / 2), static_cast<int>(c0));
        if (SWZ) {
          if (ROW) bulk_store_1d(a.sc_row + ((r0 >> 7) * (a.ld_row >> 2) + (c0 >> 6)) * 512, scr_s, 1024);
          if (COL) bulk_store_1d(a.sc_col + ((c0 >> 7) * (a.ld_col >> 2) + (r0 >> 6)) * 512, sct_s, 1024);
        }
        bulk_commit();
      }
      if (!SWZ) {
        if (ROW && tid < 128) {
          const uint2 w2 = *reinterpret_cast<const uint2 *>(scr_s + tid * 8);
          *reinterpret_cast<uint2 *>(a.sc_row + (r0 + tid) * a.ld_row + c0 / 16) = w2;
        }
        if (COL && tid >= 128) {
          const int cl = tid - 128;
          const uint2 w2 = *reinterpret_cast<const uint2 *>(sct_s + cl * 8);
          *reinterpret_cast<uint2 *>(a.sc_col + (c0 + cl) * a.ld_col + r0 / 16) = w2;
        }
      }
    } else {
      // ---------------- amax mode ----------------
      if (ROW) {
        const int h = lane >> 4, j = lane & 15;
        const int rl0 = warp * 16 + h * 8;
        const uint32_t rbase = tile + (j >> 3) * 16384 + rl0 * 128;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint4 v = lds_v4(rbase + i * 128 + (((j & 7) ^ i) << 4));
          racc = amax2(racc, amax2(amax2(v.x, v.y), amax2(v.z, v.w)));
        }
      }
      if (COL) {
        const int mi = lane >> 3, rr = lane & 7;
        const int row_l = warp * 16 + (mi >> 1) * 8 + rr;
        const uint32_t lbase = tile + row_l * 128;
        const uint32_t lx = static_cast<uint32_t>(((mi & 1) ^ (row_l & 7)) << 4);
#pragma unroll
        for (int cg = 0; cg < 8; ++cg) {
          uint32_t af[4];
          ldmatrix_x4_trans(lbase + (cg >> 2) * 16384 + (lx ^ (((2 * cg) & 6) << 4)), af);
          float d0[4], d1[4];
          mma_bf16_16816(d0, af, bfrag[0][0], bfrag[0][1]);
          mma_bf16_16816(d1, af, bfrag[1][0], bfrag[1][1]);
          const uint32_t p0 = pack_bf16x2_rn(d0[0], d0[1]), p1 = pack_bf16x2_rn(d1[0], d1[1]);
          const uint32_t q0 = pack_bf16x2_rn(d0[2], d0[3]), q1 = pack_bf16x2_rn(d1[2], d1[3]);
          cacc = amax2(cacc, amax2(amax2(p0, p1), amax2(q0, q1)));
        }
      }
      __syncthreads();   // stage s fully read
      if (tid == 0) issue(k + kStages);
    }
  }

  if (MODE == 0) {
    if (lane == 0) bulk_wait_read_all();   // shared memory must outlive the stores' reads
  } else {
    float ra = ptx::warp_max(amax_of_pair_abs(racc));
    float ca = ptx::warp_max(amax_of_pair_abs(cacc));
    if (lane == 0) {
      if (ROW && ra > 0.0f) ptx::atomic_max_nonneg(a.amax_out_row, ra);
      if (COL && ca > 0.0f) ptx::atomic_max_nonneg(a.amax_out_col, ca);
    }
  }
}

template <int MODE, bool ROW, bool COL, bool SWZ, bool SR = false>
void launch_one(const CUtensorMap &tm, const CUtensorMap &tm_row, const CUtensorMap &tm_col, const TcArgs &a,
                cudaStream_t stream) {
  static bool configured = [] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(nvfp4_tc_kernel<MODE, ROW, COL, SWZ, SR>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    return true;
  }();
  (void)configured;
  const size_t tiles = (a.cols / kTile) * (a.rows / kTile);
  const size_t cap = static_cast<size_t>(sm_count()) * kCtasPerSm;
  const unsigned grid = static_cast<unsigned>(tiles < cap ? tiles : cap);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, nvfp4_tc_kernel<MODE, ROW, COL, SWZ, SR>, tm, tm_row, tm_col, a));
}

}  // namespace

namespace {
bool eligible(const Tensor &in, size_t rows, size_t cols) {
  static const bool disabled = [] { const char *e = getenv("NVTE_B200_NVFP4_NO_TC"); return e && e[0] == '1'; }();
  return !disabled && in.data.dtype == kNVTEBFloat16 && rows % kTile == 0 && cols % kTile == 0 &&
         rows > 0 && cols > 0 && reinterpret_cast<uintptr_t>(in.data.dptr) % 16 == 0 &&
         rows < (1ull << 31) && cols * 2 < (1ull << 31);
}
}  // namespace

// Returns false when the shape / dtype is not eligible (the caller then uses the generic kernel).
// `a_generic` carries the already validated pointers and strides of quantize_nvfp4.cu.
bool quantize_tc(const Tensor &in, size_t rows, size_t cols, uint8_t *out_row, uint8_t *out_col,
                 uint8_t *sc_row, uint8_t *sc_col, size_t ld_row, size_t ld_col,
                 const float *amax_row, const float *amax_col, const uint16_t *matrix,
                 uint32_t sign_mask, bool swz, const float *noop, const uint64_t *rng, cudaStream_t stream) {
  if (!eligible(in, rows, cols)) return false;
  const bool row = out_row != nullptr, col = out_col != nullptr;
  // TMA / bulk stores need 16-byte aligned bases; compact scales leave as 8-byte words
  if (row && (reinterpret_cast<uintptr_t>(out_row) % 16 != 0 || reinterpret_cast<uintptr_t>(sc_row) % 16 != 0 ||
              ld_row % 8 != 0))
    return false;
  if (col && (reinterpret_cast<uintptr_t>(out_col) % 16 != 0 || reinterpret_cast<uintptr_t>(sc_col) % 16 != 0 ||
              ld_col % 8 != 0))
    return false;
  TcArgs a{};
  a.out_row = out_row;
  a.out_col = out_col;
  a.sc_row = sc_row;
  a.sc_col = sc_col;
  a.rows = rows;
  a.cols = cols;
  a.ld_row = ld_row;
  a.ld_col = ld_col;
  a.amax_row = amax_row;
  a.amax_col = amax_col;
  a.matrix = matrix;
  a.sign_mask = sign_mask;
  a.noop = noop;
  a.rng = rng;
  const CUtensorMap tm = gemm::tensor_map_bytes_2d(in.data.dptr, cols * 2, rows, 128, kTile);
  const CUtensorMap tr = row ? gemm::tensor_map_bytes_2d_sw64(out_row, cols / 2, rows, 64, 16) : tm;
  const CUtensorMap tc = col ? gemm::tensor_map_bytes_2d_sw64(out_col, rows / 2, cols, 64, kTile) : tm;
#define TEB_TC_DISPATCH(SR_)                                                                              \
  do {                                                                                                   \
    if (row && col) { if (swz) launch_one<0, true, true, true, SR_>(tm, tr, tc, a, stream); else launch_one<0, true, true, false, SR_>(tm, tr, tc, a, stream); } \
    else if (row) { if (swz) launch_one<0, true, false, true, SR_>(tm, tr, tc, a, stream); else launch_one<0, true, false, false, SR_>(tm, tr, tc, a, stream); } \
    else { if (swz) launch_one<0, false, true, true, SR_>(tm, tr, tc, a, stream); else launch_one<0, false, true, false, SR_>(tm, tr, tc, a, stream); } \
  } while (0)
  if (rng != nullptr) TEB_TC_DISPATCH(true);
  else TEB_TC_DISPATCH(false);
#undef TEB_TC_DISPATCH
  return true;
}

// amax(x) -> amax_row (optional), amax(bf16(transform(x^T))) -> amax_col (optional); both must
// have been zeroed by the caller. matrix == nullptr: transform built from sign_mask (0x10000 = identity).
bool amax_tc(const Tensor &in, size_t rows, size_t cols, float *amax_row, float *amax_col,
             const uint16_t *matrix, uint32_t sign_mask, cudaStream_t stream) {
  if (!eligible(in, rows, cols)) return false;
  TcArgs a{};
  a.sign_mask = sign_mask;
  a.rows = rows;
  a.cols = cols;
  a.amax_out_row = amax_row;
  a.amax_out_col = amax_col;
  a.matrix = matrix;
  const CUtensorMap tm = gemm::tensor_map_bytes_2d(in.data.dptr, cols * 2, rows, 128, kTile);
  if (amax_row && amax_col) launch_one<1, true, true, false>(tm, tm, tm, a, stream);
  else if (amax_row) launch_one<1, true, false, false>(tm, tm, tm, a, stream);
  else launch_one<1, false, true, false>(tm, tm, tm, a, stream);
  return true;
}

}  // namespace nvfp4
}  // namespace teb
