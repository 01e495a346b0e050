// MXFP8 quantization, TMA variant for 16-bit inputs whose rows and columns are multiples of 128
// (every tensor the Linear path produces). Same bytes as quantize_mxfp8.cu; replaces reference
// cast/mxfp8/quantize_mxfp8.cuh:558-845 / specialized/quantize_mxfp8.cuh for eligible shapes.
//
// Why a second kernel: ncu of the register-tile kernel showed 12.2 instructions per element
// (4 lanes share a 1x32 block: shuffles plus the scale chain repeated per lane) at 49 % issue
// utilisation, i.e. instruction-bound below the HBM roofline for mid-size tensors. Here
//   * TMA brings 128 x 128 tiles into 128B-swizzled shared memory through a 3-stage ring
//     (persistent CTAs, 2 per SM), so loads of the next tiles overlap the conversion;
//   * warp w owns the 32-row x 64-column sub-tile (box w >> 2, rows 32 (w & 3) ..) in BOTH
//     directions and every thread owns whole blocks: two 1x32 row blocks (4 x LDS.128 each) and
//     two 32x1 column blocks (one 32-bit word = 2 adjacent columns down 32 rows, conflict-free);
//     no shuffles, one scale chain per block;
//   * the FP8 bytes are staged IN PLACE over the warp's own (dead) input rows in the TMA 64B
//     swizzle and leave through the warp's own TMA stores (box 32 rows x 64 B per direction);
//     the scales are staged as one 512-byte tcgen05 atom per direction and leave by bulk copy.
//     No global load/store instruction is issued on the swizzled-scale path (compact scales are
//     written from registers);
//   * a producer warp drives the loads and the scale stores; all hand-offs are mbarriers, the
//     tile loop contains no CTA-wide barrier.
#include <cfloat>

#include "core.h"
#include "ptx.cuh"
#include "sm100.cuh"
#include "tensormap.h"

namespace teb {
namespace mxfp8 {

namespace {

constexpr int kTile = 128;
constexpr int kConsumerWarps = 8;
constexpr int kThreads = 32 * (kConsumerWarps + 1);   // + one producer warp
constexpr int kStages = 3;
constexpr int kCtasPerSm = 2;
constexpr int kStageBytes = kTile * kTile * 2;
constexpr int kSmemBytes = kStages * kStageBytes + 2048 + 128 + 1024 /*align*/;

struct TmaArgs {
  uint8_t *sc_row, *sc_col;
  size_t rows, cols;
  size_t ld_row, ld_col;
  float *amax;
  const float *noop;
  float *dbias_ws;   // optional [rows / 32][cols] per-slab column sums (nvte_quantize_dbias)
};

template <typename IT>
__device__ __forceinline__ uint32_t amax2(uint32_t a, uint32_t b);   // |max| per half, garbage signs
template <>
__device__ __forceinline__ uint32_t amax2<__nv_bfloat16>(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.xorsign.abs.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
template <>
__device__ __forceinline__ uint32_t amax2<__half>(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.xorsign.abs.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

// two packed 16-bit words (4 values) times (s0, s1, s0, s1) -> four FP8 bytes, value 0 lowest
template <typename IT, bool E5M2>
__device__ __forceinline__ uint32_t cvt4(uint32_t w0, uint32_t w1, float s0, float s1) {
  using W = ptx::Word<IT>;
  uint32_t r;
  if constexpr (E5M2) {
    asm("{\n\t.reg .b16 h0, h1;\n\t.reg .b64 p0, p1, q;\n\t.reg .f32 a0, a1, a2, a3;\n\t"
        "mov.b64 p0, {%1, %2};\n\tmov.b64 p1, {%3, %4};\n\tmov.b64 q, {%5, %6};\n\t"
        "mul.f32x2 p0, p0, q;\n\tmul.f32x2 p1, p1, q;\n\t"
        "mov.b64 {a0, a1}, p0;\n\tmov.b64 {a2, a3}, p1;\n\t"
        "cvt.rn.satfinite.e5m2x2.f32 h0, a1, a0;\n\tcvt.rn.satfinite.e5m2x2.f32 h1, a3, a2;\n\t"
        "mov.b32 %0, {h0, h1};\n\t}"
        : "=r"(r)
        : "f"(W::lo(w0)), "f"(W::hi(w0)), "f"(W::lo(w1)), "f"(W::hi(w1)), "f"(s0), "f"(s1));
  } else {
    asm("{\n\t.reg .b16 h0, h1;\n\t.reg .b64 p0, p1, q;\n\t.reg .f32 a0, a1, a2, a3;\n\t"
        "mov.b64 p0, {%1, %2};\n\tmov.b64 p1, {%3, %4};\n\tmov.b64 q, {%5, %6};\n\t"
        "mul.f32x2 p0, p0, q;\n\tmul.f32x2 p1, p1, q;\n\t"
        "mov.b64 {a0, a1}, p0;\n\tmov.b64 {a2, a3}, p1;\n\t"
        "cvt.rn.satfinite.e4m3x2.f32 h0, a1, a0;\n\tcvt.rn.satfinite.e4m3x2.f32 h1, a3, a2;\n\t"
        "mov.b32 %0, {h0, h1};\n\t}"
        : "=r"(r)
        : "f"(W::lo(w0)), "f"(W::hi(w0)), "f"(W::lo(w1)), "f"(W::hi(w1)), "f"(s0), "f"(s1));
  }
  return r;
}
// one packed word (2 values) times (s0, s1) -> two FP8 bytes in the low half
template <typename IT, bool E5M2>
__device__ __forceinline__ uint32_t cvt2(uint32_t w, float s0, float s1) {
  using W = ptx::Word<IT>;
  return ptx::mul2_cvt_fp8x2<E5M2>(W::lo(w), W::hi(w), s0, s1);
}

__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(x), "r"(y), "r"(z), "r"(w)
               : "memory");
}
__device__ __forceinline__ void sts_u16(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(static_cast<uint16_t>(v)) : "memory");
}
__device__ __forceinline__ void named_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void sts_u8(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// Warp roles: warps 0..7 convert (consumers), warp 8 lane 0 drives the TMA loads and the scale
// stores (producer). Hand-offs are mbarriers only -- there is no CTA-wide barrier in the loop, so
// the consumer warps drift apart and their load / convert / store phases overlap:
//   full[s]     producer -> consumers : tile bytes have landed in stage s (TMA complete_tx)
//   empty[s]    consumers -> producer : 8 arrivals, each after that warp's TMA stores out of stage
//                                       s have finished reading it
//   sc_full[b]  consumers -> producer : 8 arrivals, the tile's scale atoms are staged in buffer b
//   sc_empty[b] producer -> consumers : the bulk store of buffer b has finished reading it
template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
    quantize_tma_kernel(const __grid_constant__ CUtensorMap tmap_in,
                        const __grid_constant__ CUtensorMap tmap_row,
                        const __grid_constant__ CUtensorMap tmap_col, const TmaArgs a) {
  using namespace sm100;
  using W = ptx::Word<IT>;
  constexpr float kMaxRcp = E5M2 ? (1.0f / 57344.0f) : (1.0f / 448.0f);
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t *sc_s = smem + kStages * kStageBytes;                  // [2 buffers][row 512 B | col 512 B]
  uint64_t *full = reinterpret_cast<uint64_t *>(sc_s + 2048);    // [kStages]
  uint64_t *empty = full + kStages;                              // [kStages]
  uint64_t *sc_full = empty + kStages;                           // [2]
  uint64_t *sc_empty = sc_full + 2;                              // [2]

  // Programmatic dependent launch: everything up to the first global access (barrier setup,
  // descriptor prefetch) overlaps the tail of the previous kernel in the stream.
  if (a.noop != nullptr) {
    grid_dependency_wait();
    if (a.noop[0] == 1.0f) return;
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_x = static_cast<int>(a.cols / kTile);
  const int num_tiles = tiles_x * static_cast<int>(a.rows / kTile);
  const int my_tiles = (num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                       static_cast<int>(gridDim.x);

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kConsumerWarps);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sc_full[i], kConsumerWarps);
      mbar_init(&sc_empty[i], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  grid_launch_dependents();

  if (warp == kConsumerWarps) {
    // ------------------------------- producer -------------------------------
    if (lane != 0) return;
    tma_prefetch_desc(&tmap_in);
    // all global traffic of this CTA is ordered after the first load (consumers wait on full[])
    grid_dependency_wait();
    auto store_scales = [&](int j) {   // tile j of this CTA: staged scales -> global
      const int tile_id = blockIdx.x + j * gridDim.x;
      const int r0 = (tile_id / tiles_x) * kTile, c0 = (tile_id % tiles_x) * kTile;
      const uint8_t *buf = sc_s + (j & 1) * 1024;
      mbar_wait(&sc_full[j & 1], (j >> 1) & 1);
      if (ROW)
        bulk_store_1d(a.sc_row + (static_cast<size_t>(r0 >> 7) * (a.ld_row >> 2) + (c0 >> 7)) * 512, buf, 512);
      if (COL)
        bulk_store_1d(a.sc_col + (static_cast<size_t>(c0 >> 7) * (a.ld_col >> 2) + (r0 >> 7)) * 512, buf + 512, 512);
      bulk_commit();
      bulk_wait_read_all();
      mbar_arrive(&sc_empty[j & 1]);
    };
    for (int k = 0; k < my_tiles; ++k) {
      const int s = k % kStages;
      if (k >= kStages) {
        if (SWZ) store_scales(k - kStages);
        mbar_wait(&empty[s], ((k / kStages) - 1) & 1);
      }
      const int tile_id = blockIdx.x + k * gridDim.x;
      const int x = (tile_id % tiles_x) * kTile * 2, y = (tile_id / tiles_x) * kTile;
      mbar_expect_tx(&full[s], kStageBytes);
      tma_load_2d(smem + s * kStageBytes, &tmap_in, &full[s], x, y);
      tma_load_2d(smem + s * kStageBytes + 16384, &tmap_in, &full[s], x + 128, y);
    }
    if (SWZ)
      for (int j = my_tiles > kStages ? my_tiles - kStages : 0; j < my_tiles; ++j) store_scales(j);
    return;
  }

  // ------------------------------- consumers -------------------------------
  if (tid == 0) {
    if (ROW) tma_prefetch_desc(&tmap_row);
    if (COL) tma_prefetch_desc(&tmap_col);
  }
  const int bx = warp >> 2, rg = warp & 3;          // my sub-tile: box bx, rows 32 rg .. 32 rg + 31
  const uint32_t sub_off = bx * 16384 + rg * 4096;
  // row pass: lane -> row r8 + 8 rh (+ 16 it), block b of the sub-tile row
  const int r8 = lane & 7, b = (lane >> 3) & 1, rh = lane >> 4;
  const uint32_t row_ld = sub_off + (r8 + 8 * rh) * 128 + (((4 * b) ^ r8) << 4);
  const uint32_t row_st = sub_off + (r8 + 8 * rh) * 64 + (((2 * b) ^ ((r8 >> 1) & 3)) << 4);
  // col pass: lane -> columns 2 lane, 2 lane + 1 of the sub-tile
  const uint32_t col_ld = sub_off + (lane & 3) * 4;          // + p * 128 + ((chunk ^ p) << 4) + q * 1024
  const uint32_t col_st = sub_off + 2048 + (lane & 7) * 2;   // + r * 64 + ((chunk16 ^ ((r >> 1) & 3)) << 4)
  const uint32_t sc_base = smem_u32(sc_s);
  uint32_t tmax = 0u;   // packed |max| over everything this thread has seen (tensor amax)

  for (int k = 0; k < my_tiles; ++k) {
    const int tile_id = blockIdx.x + k * gridDim.x;
    const int s = k % kStages;
    const int r0 = (tile_id / tiles_x) * kTile, c0 = (tile_id % tiles_x) * kTile;
    mbar_wait(&full[s], (k / kStages) & 1);
    const uint32_t tile = smem_u32(smem + s * kStageBytes);

    uint4 rv[2][4];
    uint32_t cv[32];
    if (ROW) {
#pragma unroll
      for (int it = 0; it < 2; ++it)
#pragma unroll
        for (int j = 0; j < 4; ++j) rv[it][j] = lds_v4((tile + row_ld + it * 2048) ^ (j << 4));
    }
    if (COL) {
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        const uint32_t ad = tile + col_ld + p * 128 + ((((lane >> 2) ^ p) & 7) << 4);
#pragma unroll
        for (int q = 0; q < 4; ++q) cv[8 * q + p] = lds_u32(ad + q * 1024);
      }
    }
    // My stores out of the previous stage have had a whole load phase to drain: hand it back.
    if (k > 0 && lane == 0) {
      bulk_wait_read_all();
      mbar_arrive(&empty[(k - 1) % kStages]);
    }
    __syncwarp();   // the warp's 4 KB of input are in registers: outputs are staged over them
    const uint32_t scr_base = sc_base + (k & 1) * 1024, sct_base = scr_base + 512;
    if (SWZ && k >= 2) mbar_wait(&sc_empty[k & 1], ((k >> 1) - 1) & 1);

    if (ROW) {
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        uint32_t m = amax2<IT>(amax2<IT>(amax2<IT>(rv[it][0].x, rv[it][0].y), amax2<IT>(rv[it][0].z, rv[it][0].w)),
                               amax2<IT>(amax2<IT>(rv[it][1].x, rv[it][1].y), amax2<IT>(rv[it][1].z, rv[it][1].w)));
        m = amax2<IT>(m, amax2<IT>(amax2<IT>(amax2<IT>(rv[it][2].x, rv[it][2].y), amax2<IT>(rv[it][2].z, rv[it][2].w)),
                                   amax2<IT>(amax2<IT>(rv[it][3].x, rv[it][3].y), amax2<IT>(rv[it][3].z, rv[it][3].w))));
        if (!COL) tmax = amax2<IT>(tmax, m);
        const uint32_t e = ptx::float_to_e8m0(W::hmax(W::abs(m)) * kMaxRcp);
        const float sc = ptx::e8m0_rcp(e);
        const uint32_t st = tile + row_st + it * 1024;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint4 &v0 = rv[it][2 * h], &v1 = rv[it][2 * h + 1];
          sts_v4(st ^ (h << 4), cvt4<IT, E5M2>(v0.x, v0.y, sc, sc), cvt4<IT, E5M2>(v0.z, v0.w, sc, sc),
                 cvt4<IT, E5M2>(v1.x, v1.y, sc, sc), cvt4<IT, E5M2>(v1.z, v1.w, sc, sc));
        }
        // scale of tile row m = 32 rg + r8 + 8 rh + 16 it, k-block 2 bx + b
        const int ml = r8 + 8 * rh + 16 * it;
        if (SWZ) {
          sts_u8(scr_base + ml * 16 + rg * 4 + 2 * bx + b, e);
        } else {
          a.sc_row[static_cast<size_t>(r0 + 32 * rg + ml) * a.ld_row + (c0 >> 5) + 2 * bx + b] = static_cast<uint8_t>(e);
        }
      }
    }
    if (COL) {
      uint32_t m = 0u;
#pragma unroll
      for (int i = 0; i < 32; ++i) m = amax2<IT>(m, cv[i]);
      tmax = amax2<IT>(tmax, m);
      if (a.dbias_ws != nullptr) {   // column sums of my 2 columns over this 32-row slab
        float d0 = 0.0f, d1 = 0.0f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          d0 += W::lo(cv[i]);
          d1 += W::hi(cv[i]);
        }
        *reinterpret_cast<float2 *>(a.dbias_ws + static_cast<size_t>((r0 >> 5) + rg) * a.cols + c0 + 64 * bx +
                                    2 * lane) = make_float2(d0, d1);
      }
      const uint32_t ma = W::abs(m);
      const uint32_t e0 = ptx::float_to_e8m0(W::lo(ma) * kMaxRcp), e1 = ptx::float_to_e8m0(W::hi(ma) * kMaxRcp);
      const float s0 = ptx::e8m0_rcp(e0), s1 = ptx::e8m0_rcp(e1);
#pragma unroll
      for (int r = 0; r < 32; ++r)
        sts_u16(tile + col_st + r * 64 + ((((lane >> 3) ^ (r >> 1)) & 3) << 4), cvt2<IT, E5M2>(cv[r], s0, s1));
      // scales of tile columns 64 bx + 2 lane (+1), k-block rg
      const int cl = 64 * bx + 2 * lane;
      if (SWZ) {
        const uint32_t ad = sct_base + (cl & 31) * 16 + (cl >> 5) * 4 + rg;
        sts_u8(ad, e0);
        sts_u8(ad + 16, e1);
      } else {
        *reinterpret_cast<uint16_t *>(a.sc_col + static_cast<size_t>((r0 >> 5) + rg) * a.ld_col + c0 + cl) =
            static_cast<uint16_t>(e0 | (e1 << 8));
      }
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      const uint8_t *sub = smem + s * kStageBytes + sub_off;
      if (ROW) tma_store_2d(&tmap_row, sub, c0 + 64 * bx, r0 + 32 * rg);
      if (COL) tma_store_2d(&tmap_col, sub + 2048, c0 + 64 * bx, r0 + 32 * rg);
      bulk_commit();
      if (SWZ) mbar_arrive(&sc_full[k & 1]);
    }
  }

  if (lane == 0) bulk_wait_read_all();   // shared memory must outlive the stores' reads
  if (a.amax != nullptr) {
    const float t = ptx::warp_max(W::hmax(W::abs(tmax)));
    if (lane == 0 && t > 0.0f) ptx::atomic_max_nonneg(a.amax, t);
  }
}

template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ>
void launch_one(const CUtensorMap &ti, const CUtensorMap &tr, const CUtensorMap &tc, const TmaArgs &a,
                cudaStream_t stream) {
  static bool configured = [] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(quantize_tma_kernel<IT, ROW, COL, E5M2, SWZ>,
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
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, quantize_tma_kernel<IT, ROW, COL, E5M2, SWZ>, ti, tr, tc, a));
}

template <typename IT, bool E5M2, bool SWZ>
void launch_dir(bool row, bool col, const CUtensorMap &ti, const CUtensorMap &tr, const CUtensorMap &tc,
                const TmaArgs &a, cudaStream_t stream) {
  if (row && col) launch_one<IT, true, true, E5M2, SWZ>(ti, tr, tc, a, stream);
  else if (row) launch_one<IT, true, false, E5M2, SWZ>(ti, tr, tc, a, stream);
  else launch_one<IT, false, true, E5M2, SWZ>(ti, tr, tc, a, stream);
}

template <typename IT>
void launch_type(bool row, bool col, bool e5m2, bool swz, const CUtensorMap &ti, const CUtensorMap &tr,
                 const CUtensorMap &tc, const TmaArgs &a, cudaStream_t stream) {
  if (e5m2) {
    if (swz) launch_dir<IT, true, true>(row, col, ti, tr, tc, a, stream);
    else launch_dir<IT, true, false>(row, col, ti, tr, tc, a, stream);
  } else {
    if (swz) launch_dir<IT, false, true>(row, col, ti, tr, tc, a, stream);
    else launch_dir<IT, false, false>(row, col, ti, tr, tc, a, stream);
  }
}

}  // namespace

// Returns false when the shape / dtype / alignment is not eligible (the caller then uses the
// register-tile kernel of quantize_mxfp8.cu). Pointers and strides are already validated.
bool quantize_tma(NVTEDType in_dtype, const void *in, size_t rows, size_t cols, uint8_t *out_row,
                  uint8_t *out_col, uint8_t *sc_row, uint8_t *sc_col, size_t ld_row, size_t ld_col,
                  float *amax, bool e5m2, bool swz, const float *noop, float *dbias_ws, cudaStream_t stream) {
  static const bool disabled = [] { const char *e = getenv("NVTE_B200_MXFP8_NO_TMA"); return e && e[0] == '1'; }();
  if (disabled || (in_dtype != kNVTEBFloat16 && in_dtype != kNVTEFloat16)) return false;
  if (rows == 0 || cols == 0 || rows % kTile != 0 || cols % kTile != 0) return false;
  if (rows >= (1ull << 31) || cols * 2 >= (1ull << 31)) return false;
  auto aligned = [](const void *p, size_t n) { return reinterpret_cast<uintptr_t>(p) % n == 0; };
  const bool row = out_row != nullptr, col = out_col != nullptr;
  if (!aligned(in, 16)) return false;
  if (row && (!aligned(out_row, 16) || !aligned(sc_row, swz ? 16 : 4) || ld_row % 4 != 0)) return false;
  if (col && (!aligned(out_col, 16) || !aligned(sc_col, swz ? 16 : 4) || ld_col % 4 != 0)) return false;
  TmaArgs a{};
  a.sc_row = sc_row;
  a.sc_col = sc_col;
  a.rows = rows;
  a.cols = cols;
  a.ld_row = ld_row;
  a.ld_col = ld_col;
  a.amax = amax;
  a.noop = noop;
  a.dbias_ws = dbias_ws;
  const CUtensorMap ti = gemm::tensor_map_bytes_2d(in, cols * 2, rows, 128, kTile);
  const CUtensorMap tr = row ? gemm::tensor_map_bytes_2d_sw64(out_row, cols, rows, 64, 32) : ti;
  const CUtensorMap tc = col ? gemm::tensor_map_bytes_2d_sw64(out_col, cols, rows, 64, 32) : ti;
  if (in_dtype == kNVTEBFloat16) launch_type<__nv_bfloat16>(row, col, e5m2, swz, ti, tr, tc, a, stream);
  else launch_type<__half>(row, col, e5m2, swz, ti, tr, tc, a, stream);
  return true;
}

}  // namespace mxfp8
}  // namespace teb
