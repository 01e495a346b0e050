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
// Stage count: measured, not guessed (profiles/r02_mxfp8_stage_sweep.txt). 2 CTAs x 2 stages (128 KB of tile
// loads in flight per SM) runs the config-4 dy tensor at 6.24 TB/s; 2 x 3 or 3 x 2 (192 KB) at 5.3 TB/s, 1 x 2 at
// 4.7 TB/s: beyond ~4 tiles per SM the extra outstanding loads only lengthen the DRAM queues of the store stream.
#ifndef TEB_MX_STAGES
#define TEB_MX_STAGES 2
#endif
#ifndef TEB_MX_CTAS
#define TEB_MX_CTAS 2
#endif
#ifndef TEB_MX_MAX_MEMBERS
#define TEB_MX_MAX_MEMBERS 16
#endif
constexpr int kStages = TEB_MX_STAGES;
constexpr int kCtasPerSm = TEB_MX_CTAS;
constexpr int kStageBytes = kTile * kTile * 2;
constexpr int kSmemBytes = kStages * kStageBytes + 2048 + 128 + 1024 /*align*/;

// One tensor of a launch. A launch walks the CONCATENATED tile list of up to NM tensors
// (nvte_multi_tensor_quantize: x, W and dy of a Linear step, or the experts of a GroupedLinear, in ONE
// persistent launch -- no ramp / tail / launch gap between the tensors). The table travels as a
// __grid_constant__ kernel parameter: no workspace, no copy, CUDA-graph capturable as it is.
struct alignas(64) Member {
  CUtensorMap in, row, col;
  uint8_t *sc_row, *sc_col;
  float *amax;
  float *dbias_ws;        // optional [rows / 32][cols] per-slab column sums (nvte_quantize_dbias)
  uint32_t ld_row, ld_col;
  uint32_t cols;
  int tiles_x;
  int tile_end;           // exclusive end of this member's tiles in the concatenated list
};
constexpr int kMaxMembers = TEB_MX_MAX_MEMBERS;
template <int NM>
struct Group {
  Member m[NM];
  const float *noop;
  int n;
  int total_tiles;
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
// two packed 16-bit words (4 values) times (s0, s1, s2, s3) -> four FP8 bytes, value 0 lowest
template <typename IT, bool E5M2>
__device__ __forceinline__ uint32_t cvt4s(uint32_t w0, uint32_t w1, float s0, float s1, float s2, float s3) {
  using W = ptx::Word<IT>;
  uint32_t r;
  if constexpr (E5M2) {
    asm("{\n\t.reg .b16 h0, h1;\n\t.reg .b64 p0, p1, q0, q1;\n\t.reg .f32 a0, a1, a2, a3;\n\t"
        "mov.b64 p0, {%1, %2};\n\tmov.b64 p1, {%3, %4};\n\tmov.b64 q0, {%5, %6};\n\tmov.b64 q1, {%7, %8};\n\t"
        "mul.f32x2 p0, p0, q0;\n\tmul.f32x2 p1, p1, q1;\n\t"
        "mov.b64 {a0, a1}, p0;\n\tmov.b64 {a2, a3}, p1;\n\t"
        "cvt.rn.satfinite.e5m2x2.f32 h0, a1, a0;\n\tcvt.rn.satfinite.e5m2x2.f32 h1, a3, a2;\n\t"
        "mov.b32 %0, {h0, h1};\n\t}"
        : "=r"(r)
        : "f"(W::lo(w0)), "f"(W::hi(w0)), "f"(W::lo(w1)), "f"(W::hi(w1)), "f"(s0), "f"(s1), "f"(s2), "f"(s3));
  } else {
    asm("{\n\t.reg .b16 h0, h1;\n\t.reg .b64 p0, p1, q0, q1;\n\t.reg .f32 a0, a1, a2, a3;\n\t"
        "mov.b64 p0, {%1, %2};\n\tmov.b64 p1, {%3, %4};\n\tmov.b64 q0, {%5, %6};\n\tmov.b64 q1, {%7, %8};\n\t"
        "mul.f32x2 p0, p0, q0;\n\tmul.f32x2 p1, p1, q1;\n\t"
        "mov.b64 {a0, a1}, p0;\n\tmov.b64 {a2, a3}, p1;\n\t"
        "cvt.rn.satfinite.e4m3x2.f32 h0, a1, a0;\n\tcvt.rn.satfinite.e4m3x2.f32 h1, a3, a2;\n\t"
        "mov.b32 %0, {h0, h1};\n\t}"
        : "=r"(r)
        : "f"(W::lo(w0)), "f"(W::hi(w0)), "f"(W::lo(w1)), "f"(W::hi(w1)), "f"(s0), "f"(s1), "f"(s2), "f"(s3));
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
__device__ __forceinline__ uint2 lds_v2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
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
template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ, int NM>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
    quantize_tma_kernel(const __grid_constant__ Group<NM> g) {
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
  if (g.noop != nullptr) {
    grid_dependency_wait();
    if (g.noop[0] == 1.0f) return;
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int num_tiles = g.total_tiles;
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

  // Tiles of the concatenated list are visited in increasing order, so the member index only moves forward.
  int mi = 0, mbase = 0;
  auto locate = [&](int tile_id) {
    if (NM > 1) {
      while (tile_id >= g.m[mi].tile_end) {
        mbase = g.m[mi].tile_end;
        ++mi;
      }
    }
    return tile_id - mbase;
  };

  if (warp == kConsumerWarps) {
    // ------------------------------- producer -------------------------------
    if (lane != 0) return;
    tma_prefetch_desc(&g.m[0].in);
    // all global traffic of this CTA is ordered after the first load (consumers wait on full[])
    grid_dependency_wait();
    uint8_t *rec_row[kStages], *rec_col[kStages];   // where the scales of the tile in stage s go
    auto store_scales = [&](int j) {   // tile j of this CTA: staged scales -> global
      const uint8_t *buf = sc_s + (j & 1) * 1024;
      mbar_wait(&sc_full[j & 1], (j >> 1) & 1);
      if (ROW) bulk_store_1d(rec_row[j % kStages], buf, 512);
      if (COL) bulk_store_1d(rec_col[j % kStages], buf + 512, 512);
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
      const int local = locate(blockIdx.x + k * gridDim.x);
      const Member &M = g.m[mi];
      const int tx = local % M.tiles_x, ty = local / M.tiles_x;
      if (SWZ) {
        if (ROW) rec_row[s] = M.sc_row + (static_cast<size_t>(ty) * (M.ld_row >> 2) + tx) * 512;
        if (COL) rec_col[s] = M.sc_col + (static_cast<size_t>(tx) * (M.ld_col >> 2) + ty) * 512;
      }
      mbar_expect_tx(&full[s], kStageBytes);
      tma_load_2d(smem + s * kStageBytes, &M.in, &full[s], tx * kTile * 2, ty * kTile);
      tma_load_2d(smem + s * kStageBytes + 16384, &M.in, &full[s], tx * kTile * 2 + 128, ty * kTile);
    }
    if (SWZ)
      for (int j = my_tiles > kStages ? my_tiles - kStages : 0; j < my_tiles; ++j) store_scales(j);
    return;
  }

  // ------------------------------- consumers -------------------------------
  if (tid == 0) {
    if (ROW) tma_prefetch_desc(&g.m[0].row);
    if (COL) tma_prefetch_desc(&g.m[0].col);
  }
  const int bx = warp >> 2, rg = warp & 3;          // my sub-tile: box bx, rows 32 rg .. 32 rg + 31
  const uint32_t sub_off = bx * 16384 + rg * 4096;
  // row pass: lane -> row r8 + 8 rh (+ 16 it), block b of the sub-tile row
  const int r8 = lane & 7, b = (lane >> 3) & 1, rh = lane >> 4;
  const uint32_t row_ld = sub_off + (r8 + 8 * rh) * 128 + (((4 * b) ^ r8) << 4);
  const uint32_t row_st = sub_off + (r8 + 8 * rh) * 64 + (((2 * b) ^ ((r8 >> 1) & 3)) << 4);
  // col pass: lane -> the 4 columns 4 cg .. 4 cg + 3 of the sub-tile, rows of parity hf. The lane pair
  // (cg, hf = 0 / 1) owns the four 32x1 blocks together: 64-bit loads (a half-warp reads one whole 128-byte
  // row), 32-bit stores (the two halves write rows 2i and 2i + 1 = 128 contiguous bytes), one shuffle pair
  // for the block maxima. (The 2-column-per-lane version stored 16 bits per lane: two lanes per bank word
  // cost a second wavefront per store -- 256 bank-conflict cycles per tile in ncu.)
  const int cg = lane & 15, hf = lane >> 4;
  const uint32_t col_ld = sub_off + hf * 128 + (cg & 1) * 8;     // + i * 256 + ((cg >> 1) ^ ((2 i + hf) & 7)) << 4
  const uint32_t col_st = sub_off + 2048 + hf * 64 + (cg & 3) * 4;   // + i * 128 + ((cg >> 2) ^ (i & 3)) << 4
  const uint32_t sc_base = smem_u32(sc_s);
  uint32_t tmax = 0u;   // packed |max| over everything this thread has seen of the current member
  int amax_member = 0;
  auto flush_amax = [&](int member) {
    float *dst = g.m[member].amax;
    if (dst != nullptr) {
      const float t = ptx::warp_max(W::hmax(W::abs(tmax)));
      if (lane == 0 && t > 0.0f) ptx::atomic_max_nonneg(dst, t);
    }
    tmax = 0u;
  };

  for (int k = 0; k < my_tiles; ++k) {
    const int local = locate(blockIdx.x + k * gridDim.x);
    if (NM > 1 && mi != amax_member) {
      flush_amax(amax_member);
      amax_member = mi;
    }
    const Member &M = g.m[mi];
    const int s = k % kStages;
    const int r0 = (local / M.tiles_x) * kTile, c0 = (local % M.tiles_x) * kTile;
    mbar_wait(&full[s], (k / kStages) & 1);
    const uint32_t tile = smem_u32(smem + s * kStageBytes);

    uint4 rv[2][4];
    uint2 cv[16];
    if (ROW) {
#pragma unroll
      for (int it = 0; it < 2; ++it)
#pragma unroll
        for (int j = 0; j < 4; ++j) rv[it][j] = lds_v4((tile + row_ld + it * 2048) ^ (j << 4));
    }
    if (COL) {
#pragma unroll
      for (int i = 0; i < 16; ++i)
        cv[i] = lds_v2(tile + col_ld + i * 256 + ((((cg >> 1) ^ (2 * i + hf)) & 7) << 4));
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
          M.sc_row[static_cast<size_t>(r0 + 32 * rg + ml) * M.ld_row + (c0 >> 5) + 2 * bx + b] = static_cast<uint8_t>(e);
        }
      }
    }
    if (COL) {
      uint32_t m0 = 0u, m1 = 0u;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        m0 = amax2<IT>(m0, cv[i].x);
        m1 = amax2<IT>(m1, cv[i].y);
      }
      m0 = amax2<IT>(m0, __shfl_xor_sync(0xffffffffu, m0, 16));
      m1 = amax2<IT>(m1, __shfl_xor_sync(0xffffffffu, m1, 16));
      tmax = amax2<IT>(tmax, amax2<IT>(m0, m1));
      if (M.dbias_ws != nullptr) {   // column sums of my 4 columns over this 32-row slab
        float d0 = 0.0f, d1 = 0.0f, d2 = 0.0f, d3 = 0.0f;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          d0 += W::lo(cv[i].x);
          d1 += W::hi(cv[i].x);
          d2 += W::lo(cv[i].y);
          d3 += W::hi(cv[i].y);
        }
        d0 += __shfl_xor_sync(0xffffffffu, d0, 16);
        d1 += __shfl_xor_sync(0xffffffffu, d1, 16);
        d2 += __shfl_xor_sync(0xffffffffu, d2, 16);
        d3 += __shfl_xor_sync(0xffffffffu, d3, 16);
        if (hf == 0)
          *reinterpret_cast<float4 *>(M.dbias_ws + static_cast<size_t>((r0 >> 5) + rg) * M.cols + c0 + 64 * bx +
                                      4 * cg) = make_float4(d0, d1, d2, d3);
      }
      const uint32_t a0 = W::abs(m0), a1 = W::abs(m1);
      const uint32_t e0 = ptx::float_to_e8m0(W::lo(a0) * kMaxRcp), e1 = ptx::float_to_e8m0(W::hi(a0) * kMaxRcp);
      const uint32_t e2 = ptx::float_to_e8m0(W::lo(a1) * kMaxRcp), e3 = ptx::float_to_e8m0(W::hi(a1) * kMaxRcp);
      const float s0 = ptx::e8m0_rcp(e0), s1 = ptx::e8m0_rcp(e1), s2 = ptx::e8m0_rcp(e2), s3 = ptx::e8m0_rcp(e3);
#pragma unroll
      for (int i = 0; i < 16; ++i)
        sts_u32(tile + col_st + i * 128 + ((((cg >> 2) ^ i) & 3) << 4), cvt4s<IT, E5M2>(cv[i].x, cv[i].y, s0, s1, s2, s3));
      // scales of tile columns 64 bx + 4 cg + (0..3), k-block rg: half 0 writes the first two, half 1 the others
      const int cl = 64 * bx + 4 * cg + 2 * hf;
      const uint32_t ea = hf ? e2 : e0, eb = hf ? e3 : e1;
      if (SWZ) {
        const uint32_t ad = sct_base + (cl & 31) * 16 + (cl >> 5) * 4 + rg;
        sts_u8(ad, ea);
        sts_u8(ad + 16, eb);
      } else {
        *reinterpret_cast<uint16_t *>(M.sc_col + static_cast<size_t>((r0 >> 5) + rg) * M.ld_col + c0 + cl) =
            static_cast<uint16_t>(ea | (eb << 8));
      }
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      const uint8_t *sub = smem + s * kStageBytes + sub_off;
      if (ROW) tma_store_2d(&M.row, sub, c0 + 64 * bx, r0 + 32 * rg);
      if (COL) tma_store_2d(&M.col, sub + 2048, c0 + 64 * bx, r0 + 32 * rg);
      bulk_commit();
      if (SWZ) mbar_arrive(&sc_full[k & 1]);
    }
  }

  if (lane == 0) bulk_wait_read_all();   // shared memory must outlive the stores' reads
  if (my_tiles > 0) flush_amax(amax_member);
}

template <typename IT, bool ROW, bool COL, bool E5M2, bool SWZ, int NM>
void launch_one(const Group<NM> &g, cudaStream_t stream) {
  static bool configured = [] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(quantize_tma_kernel<IT, ROW, COL, E5M2, SWZ, NM>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    return true;
  }();
  (void)configured;
  const size_t tiles = static_cast<size_t>(g.total_tiles);
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
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, quantize_tma_kernel<IT, ROW, COL, E5M2, SWZ, NM>, g));
}

template <typename IT, bool E5M2, bool SWZ, int NM>
void launch_dir(bool row, bool col, const Group<NM> &g, cudaStream_t stream) {
  if (row && col) launch_one<IT, true, true, E5M2, SWZ, NM>(g, stream);
  else if (row) launch_one<IT, true, false, E5M2, SWZ, NM>(g, stream);
  else launch_one<IT, false, true, E5M2, SWZ, NM>(g, stream);
}

template <typename IT, int NM>
void launch_type(bool row, bool col, bool e5m2, bool swz, const Group<NM> &g, cudaStream_t stream) {
  if (e5m2) {
    if (swz) launch_dir<IT, true, true, NM>(row, col, g, stream);
    else launch_dir<IT, true, false, NM>(row, col, g, stream);
  } else {
    if (swz) launch_dir<IT, false, true, NM>(row, col, g, stream);
    else launch_dir<IT, false, false, NM>(row, col, g, stream);
  }
}

bool eligible(const TmaTensor &t, bool swz) {
  if (t.in_dtype != kNVTEBFloat16 && t.in_dtype != kNVTEFloat16) return false;
  if (t.rows == 0 || t.cols == 0 || t.rows % kTile != 0 || t.cols % kTile != 0) return false;
  if (t.rows >= (1ull << 31) || t.cols * 2 >= (1ull << 31)) return false;
  auto aligned = [](const void *p, size_t n) { return reinterpret_cast<uintptr_t>(p) % n == 0; };
  const bool row = t.out_row != nullptr, col = t.out_col != nullptr;
  if (!row && !col) return false;
  if (!aligned(t.in, 16)) return false;
  if (row && (!aligned(t.out_row, 16) || !aligned(t.sc_row, swz ? 16 : 4) || t.ld_row % 4 != 0)) return false;
  if (col && (!aligned(t.out_col, 16) || !aligned(t.sc_col, swz ? 16 : 4) || t.ld_col % 4 != 0)) return false;
  return true;
}

Member make_member(const TmaTensor &t, int tile_end) {
  Member m;
  std::memset(static_cast<void *>(&m), 0, sizeof(m));
  const bool row = t.out_row != nullptr, col = t.out_col != nullptr;
  m.in = gemm::tensor_map_bytes_2d(t.in, t.cols * 2, t.rows, 128, kTile);
  m.row = row ? gemm::tensor_map_bytes_2d_sw64(t.out_row, t.cols, t.rows, 64, 32) : m.in;
  m.col = col ? gemm::tensor_map_bytes_2d_sw64(t.out_col, t.cols, t.rows, 64, 32) : m.in;
  m.sc_row = t.sc_row;
  m.sc_col = t.sc_col;
  m.amax = t.amax;
  m.dbias_ws = t.dbias_ws;
  m.ld_row = static_cast<uint32_t>(t.ld_row);
  m.ld_col = static_cast<uint32_t>(t.ld_col);
  m.cols = static_cast<uint32_t>(t.cols);
  m.tiles_x = static_cast<int>(t.cols / kTile);
  m.tile_end = tile_end;
  return m;
}

bool tma_disabled() {
  static const bool disabled = [] { const char *e = getenv("NVTE_B200_MXFP8_NO_TMA"); return e && e[0] == '1'; }();
  return disabled;
}

}  // namespace

// Returns false when the shape / dtype / alignment is not eligible (the caller then uses the
// register-tile kernel of quantize_mxfp8.cu). Pointers and strides are already validated.
bool quantize_tma(const TmaTensor &t, bool e5m2, bool swz, const float *noop, cudaStream_t stream) {
  if (tma_disabled() || !eligible(t, swz)) return false;
#ifdef TEB_MX_FORCE_GROUP
  return quantize_tma_multi(&t, 1, e5m2, swz, noop, stream);
#endif
  Group<1> g;
  g.m[0] = make_member(t, static_cast<int>((t.rows / kTile) * (t.cols / kTile)));
  g.noop = noop;
  g.n = 1;
  g.total_tiles = g.m[0].tile_end;
  const bool row = t.out_row != nullptr, col = t.out_col != nullptr;
  if (t.in_dtype == kNVTEBFloat16) launch_type<__nv_bfloat16, 1>(row, col, e5m2, swz, g, stream);
  else launch_type<__half, 1>(row, col, e5m2, swz, g, stream);
  return true;
}

// nvte_multi_tensor_quantize: every tensor of the list in ONE persistent launch (lists longer than
// kMaxMembers in several). All members must agree on input dtype, directions, FP8 type and scale layout,
// and each must be eligible for the TMA kernel; otherwise false and nothing has been launched.
bool quantize_tma_multi(const TmaTensor *ts, size_t n, bool e5m2, bool swz, const float *noop, cudaStream_t stream) {
  if (tma_disabled() || n == 0) return false;
  const bool row = ts[0].out_row != nullptr, col = ts[0].out_col != nullptr;
  size_t total = 0;
  for (size_t i = 0; i < n; ++i) {
    if (!eligible(ts[i], swz) || ts[i].in_dtype != ts[0].in_dtype || ts[i].dbias_ws != nullptr ||
        (ts[i].out_row != nullptr) != row || (ts[i].out_col != nullptr) != col)
      return false;
    total += (ts[i].rows / kTile) * (ts[i].cols / kTile);
  }
  if (total >= (1ull << 31)) return false;
  for (size_t first = 0; first < n; first += kMaxMembers) {
    const size_t cnt = n - first < static_cast<size_t>(kMaxMembers) ? n - first : kMaxMembers;
    Group<kMaxMembers> g;
    std::memset(static_cast<void *>(&g), 0, sizeof(g));
    int end = 0;
    for (size_t i = 0; i < cnt; ++i) {
      const TmaTensor &t = ts[first + i];
      end += static_cast<int>((t.rows / kTile) * (t.cols / kTile));
      g.m[i] = make_member(t, end);
    }
    for (size_t i = cnt; i < static_cast<size_t>(kMaxMembers); ++i) g.m[i].tile_end = 0x7fffffff;
    g.noop = noop;
    g.n = static_cast<int>(cnt);
    g.total_tiles = end;
    if (ts[0].in_dtype == kNVTEBFloat16) launch_type<__nv_bfloat16, kMaxMembers>(row, col, e5m2, swz, g, stream);
    else launch_type<__half, kMaxMembers>(row, col, e5m2, swz, g, stream);
  }
  return true;
}

}  // namespace mxfp8
}  // namespace teb
