// Persistent warp-specialised tcgen05 GEMM for the low-precision Linear path:
//   Out[R, C] (row-major) = alpha' * X[R, K] . Y[C, K]^T (+ bias[C]) (+ beta * Cin[R, C])
// X = activation-side operand (TE's "B"), Y = weight-side operand (TE's "A").
// KIND 0: per-tensor-scaled FP8   (tcgen05.mma.kind::f8f6f4,  alpha' = alpha * sinv_x * sinv_y)
// KIND 1: MXFP8, E8M0 1x32 scales (kind::mxf8f6f4.block_scale, scales in TMEM)
// KIND 2: NVFP4, E4M3 1x16 scales (kind::mxf4nvf4.block_scale.scale_vec::4X,
//                                  alpha' = alpha * amax_x * amax_y / (6*6*448*448))
//
// One CTA per SM, 10 warps: warp 0 = TMA producer, warp 1 = MMA issuer (one elected
// lane) + TMEM owner, warps 2-9 = epilogue. An epilogue warp drains its 32 lanes x BN/2
// columns of the accumulator into registers in one go (tcgen05.ld x BN/64), releases the
// TMEM buffer to the MMA warp at once, and only then scales / converts / stores -- so the
// tensor pipe idles for a TMEM drain (a few hundred cycles), not for the global stores,
// even when BN = 256 leaves no TMEM for a second accumulator buffer.
// Pipelines: smem full/empty ring (TMA <-> MMA), TMEM accumulator full/empty
// (MMA <-> epilogue, double-buffered when BN <= 128), static persistent tile loop.
// Per k-block of 128 bytes of K per row: X tile 128 rows, Y tile BN rows, both
// 128B-swizzled; block scales arrive as 512-byte pre-swizzled 128x4 atoms by 1-D bulk
// copy and go smem -> TMEM with tcgen05.cp.32x128b.warpx4.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../../include/transformerengine_b200.h"
#include "ptx.cuh"
#include "sm100.cuh"

namespace teb {
namespace gemm {

constexpr int kBM = 128;         // rows of X per tile (UMMA M)
constexpr int kBKBytes = 128;    // bytes of K per row per k-block (= one 128B swizzle row)
constexpr int kNumEpilogueWarps = 8;
constexpr int kNumThreads = 64 + 32 * kNumEpilogueWarps;
constexpr int kTmemCols = 512;

struct Operand {
  uint32_t lbo, sbo;    // UMMA descriptor offsets (bytes)
  uint32_t kstep;       // descriptor start-address advance per MMA (bytes)
  int mn_major;         // 1: global tensor is [K, MN] (MN contiguous), loaded as 128x128 boxes
};

struct Params {
  int R, C, K;                 // K in elements
  int kblocks;                 // ceil(K bytes / 128)
  Operand x, y;
  uint32_t idesc;
  const uint8_t *sfx, *sfy;    // swizzled scale atoms (512 B each)
  int sfx_ld, sfy_ld;          // atoms per 128-row block (= padded scale columns / 4)
  const float *s0, *s1;        // device scalars folded into alpha (scale_inv / amax), may be null
  float alpha, beta, scalar_const;
  void *out;
  const void *cin;             // accumulate source (same dtype as out), may be null
  const void *bias;            // [C], same dtype as out, may be null
  int out_dtype;               // NVTEDType of out: 4 fp32, 5 fp16, 6 bf16
  int tiles_r, tiles_c;
  int rb_size, rb_stride, rb_off;   // output row blocks (0 = off): row r -> (r / size) * stride + off + r % size
  // scatter output (sc_rows != 0): row r goes to sc_out[r / sc_rows] + (r % sc_rows) * C (peer-mapped buffers)
  int sc_rows, sc_rot, sc_tiles;    // rows per rank; tile rows the walk is rotated by, modulo all tile rows
  int sc_staged;                    // pair kernel, per-tensor FP8, 16-bit output: rows leave through shared memory
  void *sc_out[8];
  int tile_r0;                 // first tile row of this launch (the tile space can be split over 2 launches)
  int raster_g;                // tile rows per super-row of the L2-aware tile walk (pair kernel; <= 1: row-major)
  int bn;                      // pair kernel: tile width in columns, 256 (all kinds) or 224 / 192 (FP8, K-major Y):
                               // narrower tiles when that saves a wave (8192 x 3584: 448 tiles = 6.05 waves of 74
                               // pairs at 256, 512 tiles = 6.92 waves at 224 -> 7 x 224 instead of 7 x 256)
  // arrival flags (kNVTEMatmulConfigB200ArrivalFlags): X's rows are the gather buffer of a symmetric exchange
  const uint32_t *arr_flags;   // [world * arr_fpr], null = off
  const uint32_t *arr_step;
  uint32_t *arr_error;
  int arr_world, arr_rank, arr_fpr, arr_tpf;   // ranks, my rank, flags per rank, 256-row tiles per flag
  // push (arr_push_unit != 0): an idle warp of the pair's second CTA streams this rank's rows to the peers
  int arr_push_unit, arr_push_upc;   // bytes per unit, units per flag chunk
  const uint8_t *arr_push_src;       // my rows (inside operand X)
  uint8_t *arr_push_dst[8];          // peer r's address of my row window
  uint32_t *arr_push_cnt[8];         // peer r's flag words
  unsigned int *tile_counter;        // dynamic tile queue: {next tile, pairs done}, both zero between launches
  // second pushed region per chunk (MXFP8 block scales) and the step protocol of an exchange without an amax step
  int arr_push_unit2, arr_push_upc2;
  const uint8_t *arr_push_src2;
  uint8_t *arr_push_mc_dst2;
  int arr_advance;                   // this launch is step *arr_step + 1; its last pair advances the counter
  const uint32_t *arr_ready;         // local word: GPUs that finished reading, world per step (null: not needed)
  uint32_t *arr_ready_mc;            // its multicast address
  uint8_t *arr_push_mc_dst;          // NVSwitch multicast address of my row window (null: unicast bulk copies)
  uint32_t *arr_push_mc_cnt;         // multicast address of the flag words
  int debug_flags;             // dev aid: 1 = skip scale copies, 2 = skip MMAs
  unsigned long long *debug;   // optional per-CTA cycle counters of the MMA thread (dev aid)
};

// L2-aware tile order of the persistent pair kernel. The pairs of one "wave" (74 of them) work on
// consecutive tile indices; row-major order makes a wave cover ~1.3 tile rows x all tile columns, i.e. every
// wave re-reads nearly all of Y through L2 (ncu r1: 350 MB of DRAM reads for 95 MB of operands on the
// config-4 fprop). Walking super-rows of G tile rows column-first makes a wave a G x (74 / G) patch: it
// touches G row blocks of X and 74 / G column blocks of Y (G = 8: 17 operand blocks instead of ~58).
__device__ __forceinline__ void tile_coords(const Params &p, int tile, int &tr, int &tc) {
  const int g = p.raster_g;
  if (g <= 1) {
    tr = tile / p.tiles_c;
    tc = tile - tr * p.tiles_c;
  } else {
    const int per_super = g * p.tiles_c;
    const int sr = tile / per_super;
    const int rem = tile - sr * per_super;
    const int left = p.tiles_r - sr * g;
    const int rows = left < g ? left : g;   // the last super-row may be shorter
    tc = rem / rows;
    tr = sr * g + (rem - tc * rows);
  }
  tr += p.tile_r0;
  if (p.sc_rot != 0) {   // scatter output: start behind my own rows
    tr += p.sc_rot;
    if (tr >= p.sc_tiles) tr -= p.sc_tiles;
  }
  if (p.arr_flags != nullptr) {
    // arrival order: my own rows first, then chunk 0 of every peer (nearest first), chunk 1, ...
    const int cn = p.arr_fpr * p.arr_tpf;   // tile rows per rank
    if (tr < cn) {
      tr += p.arr_rank * cn;
    } else {
      const int j = tr - cn, t = j % p.arr_tpf, q = j / p.arr_tpf;
      const int peer = q % (p.arr_world - 1), c = q / (p.arr_world - 1);
      tr = ((p.arr_rank + 1 + peer) % p.arr_world) * cn + c * p.arr_tpf + t;
    }
  }
}

template <int BN, int KIND>
struct Config {
  static constexpr int kSfAtoms = KIND == 2 ? 4 : (KIND == 1 ? 1 : 0);  // 512B atoms / 128 rows / k-block
  static constexpr int kXBytes = kBM * kBKBytes;
  static constexpr int kYBytes = BN * kBKBytes;
  static constexpr int kSfxBytes = kSfAtoms * 512;
  static constexpr int kSfyBytes = kSfAtoms * 512 * (BN / 128);
  static constexpr int kStageBytes = kXBytes + kYBytes + kSfxBytes + kSfyBytes;
  static constexpr int kStages = (200 * 1024) / kStageBytes > 8 ? 8 : (200 * 1024) / kStageBytes;
  static constexpr int kAccStages = BN <= 128 ? 2 : 1;
  static constexpr int kSfColsX = kSfAtoms * 4;
  static constexpr int kSfColsY = kSfAtoms * 4 * (BN / 128);
  static constexpr int kSfColBase = kAccStages * BN;
  static_assert(kSfColBase + kStages * (kSfColsX + kSfColsY) <= kTmemCols, "TMEM overflow");
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
};

// 32-byte global store (sm_100: st.global.v8.b32): one full sector per lane and half the store
// instructions of the 16-byte form -- the epilogue's lanes each own a row, so every store instruction
// touches 32 different lines and its cost is per line, not per byte.
__device__ __forceinline__ void st_global_v8(void *dst, const uint32_t (&w)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "r"(w[0]), "r"(w[1]),
               "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
               : "memory");
}

template <typename OT>
__device__ __forceinline__ void store_chunk(OT *dst, const float (&v)[32], int valid, bool vec_ok);

template <>
__device__ __forceinline__ void store_chunk<float>(float *dst, const float (&v)[32], int valid,
                                                   bool vec_ok) {
  if (vec_ok && valid == 32 && (reinterpret_cast<uintptr_t>(dst) & 31u) == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = __float_as_uint(v[8 * i + j]);
      st_global_v8(dst + 8 * i, w);
    }
  } else if (vec_ok && valid == 32) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      reinterpret_cast<float4 *>(dst)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < valid) dst[i] = v[i];
  }
}
template <>
__device__ __forceinline__ void store_chunk<__nv_bfloat16>(__nv_bfloat16 *dst, const float (&v)[32],
                                                           int valid, bool vec_ok) {
  if (vec_ok && valid == 32 && (reinterpret_cast<uintptr_t>(dst) & 31u) == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      uint32_t w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[16 * i + 2 * j], v[16 * i + 2 * j + 1]);
        w[j] = *reinterpret_cast<uint32_t *>(&h);
      }
      st_global_v8(dst + 16 * i, w);
    }
  } else if (vec_ok && valid == 32) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 w;
      uint32_t *pw = reinterpret_cast<uint32_t *>(&w);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[8 * i + 2 * j], v[8 * i + 2 * j + 1]);
        pw[j] = *reinterpret_cast<uint32_t *>(&h);
      }
      reinterpret_cast<uint4 *>(dst)[i] = w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < valid) dst[i] = __float2bfloat16_rn(v[i]);
  }
}
template <>
__device__ __forceinline__ void store_chunk<__half>(__half *dst, const float (&v)[32], int valid,
                                                    bool vec_ok) {
  if (vec_ok && valid == 32) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 w;
      uint32_t *pw = reinterpret_cast<uint32_t *>(&w);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        __half2 h = __floats2half2_rn(v[8 * i + 2 * j], v[8 * i + 2 * j + 1]);
        pw[j] = *reinterpret_cast<uint32_t *>(&h);
      }
      reinterpret_cast<uint4 *>(dst)[i] = w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < valid) dst[i] = __float2half_rn(v[i]);
  }
}

template <typename OT>
__device__ __forceinline__ void epilogue_chunk(const Params &p, const uint32_t (&acc)[32],
                                               float alpha, size_t row, int col0, bool row_ok) {
  const int valid = p.C - col0 < 32 ? p.C - col0 : 32;
  if (!row_ok || valid <= 0) return;
  if (p.rb_size > 0) row = (row / p.rb_size) * static_cast<size_t>(p.rb_stride) + p.rb_off + row % p.rb_size;
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]) * alpha;
  if (p.bias != nullptr) {
    const OT *b = reinterpret_cast<const OT *>(p.bias) + col0;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < valid) v[i] += static_cast<float>(b[i]);
  }
  OT *dst = reinterpret_cast<OT *>(p.out) + row * static_cast<size_t>(p.C) + col0;
  if (p.sc_rows > 0) {
    const size_t q = row / static_cast<size_t>(p.sc_rows);
    dst = reinterpret_cast<OT *>(p.sc_out[q]) + (row - q * p.sc_rows) * static_cast<size_t>(p.C) + col0;
  }
  if (p.cin != nullptr && p.beta != 0.0f) {
    const OT *c = reinterpret_cast<const OT *>(p.cin) + row * static_cast<size_t>(p.C) + col0;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < valid) v[i] += p.beta * static_cast<float>(c[i]);
  }
  const bool vec_ok = (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
  store_chunk<OT>(dst, v, valid, vec_ok);
}

template <int BN, int KIND>
__global__ void __launch_bounds__(kNumThreads, 1)
    gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y,
                const Params p) {
  using Cfg = Config<BN, KIND>;
  using namespace sm100;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t *empty_bar = full_bar + Cfg::kStages;
  uint64_t *acc_full = empty_bar + Cfg::kStages;
  uint64_t *acc_empty = acc_full + Cfg::kAccStages;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_empty + Cfg::kAccStages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_r * p.tiles_c;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_y);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < Cfg::kAccStages; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], kNumEpilogueWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) overlaps the tail of the
  // previous kernel in the stream (usually the cast that produced an operand); every thread that touches global
  // memory waits for that kernel here, and the next kernel may start its own prologue from now on.
  grid_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      grid_dependency_wait();
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int tr = tile / p.tiles_c, tc = tile % p.tiles_c;
        for (int kb = 0; kb < p.kblocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t *sx = smem + stage * Cfg::kStageBytes;
          uint8_t *sy = sx + Cfg::kXBytes;
          // Y scale row blocks past ceil(C / 128) do not exist in the operand's scale buffer (BN = 256 with an
          // odd number of 128-row blocks): skip their copy; the data rows they would scale are TMA zero fill
          // and the output columns are never stored
          const int sfy_valid = KIND == 0 ? 0 : min(BN / 128, (p.C + 127) / 128 - tc * (BN / 128));
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes - (BN / 128 - sfy_valid) * Cfg::kSfxBytes);
          if (p.x.mn_major) tma_load_2d(sx, &tmap_x, &full_bar[stage], tr * kBM, kb * kBKBytes);
          else tma_load_2d(sx, &tmap_x, &full_bar[stage], kb * kBKBytes, tr * kBM);
          if (p.y.mn_major) {
#pragma unroll
            for (int b = 0; b < BN / 128; ++b)
              tma_load_2d(sy + b * 16384, &tmap_y, &full_bar[stage], tc * BN + b * 128, kb * kBKBytes);
          } else {
            tma_load_2d(sy, &tmap_y, &full_bar[stage], kb * kBKBytes, tc * BN);
          }
          if (KIND != 0) {
            uint8_t *ssx = sy + Cfg::kYBytes;
            uint8_t *ssy = ssx + Cfg::kSfxBytes;
            bulk_load_1d(ssx, p.sfx + (static_cast<size_t>(tr) * p.sfx_ld + static_cast<size_t>(kb) * Cfg::kSfAtoms) * 512,
                         Cfg::kSfxBytes, &full_bar[stage]);
#pragma unroll
            for (int b = 0; b < BN / 128; ++b)
              if (b < sfy_valid)
                bulk_load_1d(ssy + b * Cfg::kSfxBytes,
                           p.sfy + (static_cast<size_t>(tc * (BN / 128) + b) * p.sfy_ld +
                                    static_cast<size_t>(kb) * Cfg::kSfAtoms) * 512,
                           Cfg::kSfxBytes, &full_bar[stage]);
          }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // Whole warp runs the loop (warp-uniform arithmetic -> uniform registers feed UTCCP /
    // UTCQMMA without R2UR round trips); only the elected lane issues tcgen05 instructions.
    {
      const bool issuer = elect_one();
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      const uint64_t dxc = make_smem_desc(0, p.x.lbo, p.x.sbo, kSwizzle128B);
      const uint64_t dyc = make_smem_desc(0, p.y.lbo, p.y.sbo, kSwizzle128B);
      const uint64_t dsc = make_smem_desc(0, 0, 128, kSwizzleNone);
      const uint32_t xstep = p.x.kstep >> 4, ystep = p.y.kstep >> 4;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&acc_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_acc = tmem_base + as * BN;
        for (int kb = 0; kb < p.kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sx = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sy = sx + Cfg::kXBytes;
          const uint32_t tsfx = tmem_base + Cfg::kSfColBase + stage * (Cfg::kSfColsX + Cfg::kSfColsY);
          const uint32_t tsfy = tsfx + Cfg::kSfColsX;
          const uint64_t dx0 = dxc | ((sx & 0x3ffffu) >> 4);
          const uint64_t dy0 = dyc | ((sy & 0x3ffffu) >> 4);
          if (issuer) {
            if (KIND != 0) {
              // smem order: X atoms [a], then Y atoms [b][a]; TMEM order for Y is [a][b] so that
              // the scales of one MMA (k-atom a) cover all BN columns contiguously.
              const uint64_t dsx = dsc | (((sy + Cfg::kYBytes) & 0x3ffffu) >> 4);
              const uint64_t dsy = dsx + (Cfg::kSfxBytes >> 4);
#pragma unroll
              for (int a = 0; a < Cfg::kSfAtoms; ++a) {
                tmem_cp_sf(tsfx + 4 * a, dsx + 32 * a);
#pragma unroll
                for (int b = 0; b < BN / 128; ++b)
                  tmem_cp_sf(tsfy + 4 * (a * (BN / 128) + b), dsy + 32 * (b * Cfg::kSfAtoms + a));
              }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint64_t dx = dx0 + j * xstep;
              const uint64_t dy = dy0 + j * ystep;
              const uint32_t accumulate = (kb | j) != 0;
              if (KIND == 0) mma_f8(tmem_acc, dx, dy, p.idesc, accumulate);
              else if (KIND == 1) mma_mxf8(tmem_acc, dx, dy, with_sf_ids(p.idesc, j, j), accumulate, tsfx, tsfy);
              else mma_nvf4(tmem_acc, dx, dy, p.idesc, accumulate, tsfx + 4 * j, tsfy + 4 * j * (BN / 128));
            }
            tc_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        if (issuer) tc_commit(&acc_full[as]);
        __syncwarp();
        if (++as == Cfg::kAccStages) { as = 0; aphase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue =====================
    const int q = warp & 3;                  // TMEM lane quadrant this warp may access
    const int half = (warp - 2) >> 2;        // which BN/2 column half of the tile
    constexpr int kChunks = BN / 64;         // 32-column chunks per warp
    grid_dependency_wait();
    float alpha = p.alpha * p.scalar_const;
    if (p.s0 != nullptr) alpha *= *p.s0;
    if (p.s1 != nullptr) alpha *= *p.s1;
    int as = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int tr = tile / p.tiles_c, tc = tile % p.tiles_c;
      mbar_wait(&acc_full[as], aphase);
      tc_fence_after();
      uint32_t acc[kChunks][32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + half * (BN / 2);
#pragma unroll
      for (int c = 0; c < kChunks; ++c) tmem_ld_32x32(taddr + c * 32, acc[c]);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_relaxed(&acc_empty[as]);   // accumulator is in registers (wait::ld): MMA may go on
      const size_t row = static_cast<size_t>(tr) * kBM + q * 32 + lane;
      const bool row_ok = row < static_cast<size_t>(p.R);
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int col0 = tc * BN + half * (BN / 2) + c * 32;
        if (p.out_dtype == kNVTEBFloat16) epilogue_chunk<__nv_bfloat16>(p, acc[c], alpha, row, col0, row_ok);
        else if (p.out_dtype == kNVTEFloat32) epilogue_chunk<float>(p, acc[c], alpha, row, col0, row_ok);
        else epilogue_chunk<__half>(p, acc[c], alpha, row, col0, row_ok);
      }
      if (++as == Cfg::kAccStages) { as = 0; aphase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace gemm
}  // namespace teb
