// CTA-pair (tcgen05 cta_group::2) variant of the block-scaled GEMM: one 256 x 256 output tile
// per pair of SMs. Each CTA of the pair stages its own 128 rows of X and its own 128 of the
// 256 Y rows, so a k-block costs 32 KB of L2->SM traffic per SM instead of 48 KB for the
// single-CTA 128 x 256 tile (less operand traffic per flop = less energy under the 1000 W power cap that
// bounds these GEMMs). Tiles are walked in an L2-aware order (tile_coords).
//
// Roles per CTA (10 warps): warp 0 = TMA producer (both CTAs; every load signals the LEADER's
// full barrier), warp 1 = TMEM owner, and in the leader CTA the single-thread issuer of
// tcgen05.cp / tcgen05.mma.cta_group::2 for the pair, warps 2-9 = epilogue of the CTA's own
// 128 accumulator rows (drain to registers, release TMEM, then store).
// Barriers: full[s] (leader only; tx bytes of both CTAs), empty[s] (each CTA; multicast
// commit), acc_full[a] (each CTA; multicast commit), acc_empty[a] (leader; 16 epilogue warps).
#pragma once

#include "gemm_kernel.cuh"

namespace teb {
namespace gemm {

constexpr int kBN2 = 256;          // tile columns per CTA pair
constexpr int kSmemBudget2 = 220 * 1024;

template <int KIND>
struct Config2 {
  static constexpr int kSfAtoms = KIND == 2 ? 4 : (KIND == 1 ? 1 : 0);
  static constexpr int kXBytes = kBM * kBKBytes;          // own 128 rows of X
  static constexpr int kYBytes = 128 * kBKBytes;          // own 128 of the 256 Y rows
  static constexpr int kSfxBytes = kSfAtoms * 512;        // scales of the own X rows
  static constexpr int kSfyBytes = 2 * kSfAtoms * 512;    // scales of all 256 Y rows
  static constexpr int kStageBytes = kXBytes + kYBytes + kSfxBytes + kSfyBytes;
  static constexpr int kStagesRaw = (kSmemBudget2 - 2048) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kAccStages = 2;
  static constexpr int kSfColsX = kSfAtoms * 4;
  static constexpr int kSfColsY = kSfAtoms * 8;
  // TMEM: two 256-column fp32 accumulators fill all 512 columns, so the block-scaled kinds make the
  // two accumulators OVERLAP by kOverlap columns and park the scales in the columns that frees:
  //   acc0 = [0, 256), acc1 = [256 - kOverlap, 512 - kOverlap), scales = [512 - kOverlap, 512).
  // The epilogue drains the shared columns first (the LAST columns of acc0, the FIRST of acc1) and
  // signals acc_ovl; the next tile's MMAs (other accumulator) start as soon as that arrives while
  // the rest of the drain proceeds. tcgen05.cp and tcgen05.mma execute in issue order, so the scale
  // slots need no barrier of their own and do not have to match the shared-memory stage count.
  static constexpr int kSfSlots = KIND == 1 ? 2 : 1;
  static constexpr int kSfCols = kSfSlots * (kSfColsX + kSfColsY);
  static constexpr int kOverlap = (kSfCols + 31) / 32 * 32;
  static constexpr int kOvlChunks = kOverlap / 32;      // 32-column chunks of the owner warps
  static constexpr int kSfColBase = kTmemCols - kOverlap;
  static_assert(kOverlap <= 128, "the shared columns must belong to one half of the epilogue warps");
  __host__ __device__ static constexpr int acc_col(int as) { return as == 0 ? 0 : kBN2 - kOverlap; }
  // push ring (arrival flags with push, per-tensor FP8 only): two units behind the barriers
  static constexpr int kPushUnit = 16384;
  static constexpr int kPushBytes = KIND == 0 ? 2 * kPushUnit : 0;
  // control region behind the stages: [0, 256) pipeline barriers, [256, 384) tile-queue barriers, [384, 448) tile ring
  static constexpr int kCtrlBytes = 512;
  static constexpr int kTileRing = 16;     // dynamic tile queue: slots (see gemm2_kernel, DYN)
  static constexpr int kPushOff = kStages * kStageBytes + kCtrlBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + kCtrlBytes + kPushBytes;
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

// Scatter output through shared memory (16-bit output types). A lane owns a row of the accumulator, so a direct store
// touches 32 rows with 32 bytes each -- fine for local HBM, but over NVLink every 32-byte fragment is its own packet
// (measured at tp = 8: the scattering GEMM took ~0.4 ms for 117 MB of remote rows). Two 32-column chunks of a warp's 32
// rows (128 bytes per row) are therefore staged in the warp's 4 KB of shared memory (16-byte pieces xor-swizzled by row)
// and written back with 8 lanes per row: every store instruction covers 4 rows x 128 contiguous bytes.
template <typename OT>
__device__ __forceinline__ void stage_chunk16(uint32_t stg, int lane, int piece0, const uint32_t (&acc)[32], float alpha,
                                              const OT *bias) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float a = __uint_as_float(acc[8 * j + 2 * k]) * alpha, b = __uint_as_float(acc[8 * j + 2 * k + 1]) * alpha;
      if (bias != nullptr) {
        a += static_cast<float>(bias[8 * j + 2 * k]);
        b += static_cast<float>(bias[8 * j + 2 * k + 1]);
      }
      if constexpr (sizeof(OT) == 2 && !std::is_same<OT, __half>::value) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        w[k] = *reinterpret_cast<uint32_t *>(&h);
      } else {
        __half2 h = __floats2half2_rn(a, b);
        w[k] = *reinterpret_cast<uint32_t *>(&h);
      }
    }
    const uint32_t addr = stg + lane * 128 + (((piece0 + j) ^ (lane & 7)) << 4);
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
  }
}

// One member of a grouped launch (nvte_multi_tensor_gemm): its own tensor maps and parameters; the
// members' tiles are concatenated into one tile list that the persistent pairs walk, so a list of
// expert GEMMs costs one launch and no SM idles between members. The table lives in device memory
// (written by the host before the launch; TMA reads the maps through their generic address).
struct alignas(128) GroupEntry {
  CUtensorMap mx, my, msx, msy;
  Params p;
  int tile_end;      // exclusive end of this member's tiles in the concatenated list
};

// CSZ = cluster size (set at launch): 2 = one CTA pair; 4 = two pairs working on two horizontally
// adjacent tiles (same X rows). With CSZ == 4 each CTA loads only half of its 128 X rows and
// multicasts them to the CTA of the same parity in the other pair (tmap_x then has a 64-row box),
// which cuts the L2 -> SM traffic per k-block from 32 KB to 24 KB per CTA; the two pairs run in
// lockstep through the empty barriers (one arrival per pair, commits multicast to all four CTAs).
//
// DYN (CSZ == 2, not grouped): tiles come from a global counter instead of the static round-robin pair + k * pairs.
// The leader CTA's producer thread takes the next tile index with one atomicAdd and hands it to every consumer of both
// CTAs -- its own MMA and epilogue warps, the peer's producer and epilogue warps -- through a 16-slot ring in shared
// memory: the value travels by st.async, which completes the 4 transaction bytes the slot's mbarrier was armed with
// (no fence, no second signal). The ring needs no "empty" barriers: the operand-stage and accumulator barriers keep the
// producer at most kStages k-blocks + 2 accumulators (< 16 tiles) ahead of the slowest consumer. Used with arrival
// flags, where a pair whose next tile's rows have not arrived yet must not keep a static share of the work waiting
// behind it. The last pair to run dry resets the counter for the next launch.
template <int KIND, int CSZ, bool GROUPED = false, bool DYN = false>
__global__ void __launch_bounds__(kNumThreads, 1)
    gemm2_kernel(const __grid_constant__ CUtensorMap tmap_x0, const __grid_constant__ CUtensorMap tmap_y0,
                 const __grid_constant__ CUtensorMap tmap_sfx0,
                 const __grid_constant__ CUtensorMap tmap_sfy0, const __grid_constant__ Params p0,
                 const GroupEntry *__restrict__ groups, const int num_groups) {
  static_assert(!GROUPED || CSZ == 2, "grouped launches use plain CTA pairs");
  static_assert(!DYN || (CSZ == 2 && !GROUPED), "the dynamic tile queue serves plain CTA pairs");
  using Cfg = Config2<KIND>;
  using namespace sm100;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t *empty_bar = full_bar + Cfg::kStages;
  uint64_t *acc_full = empty_bar + Cfg::kStages;
  uint64_t *acc_empty = acc_full + Cfg::kAccStages;
  uint64_t *acc_ovl = acc_empty + Cfg::kAccStages;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_ovl + Cfg::kAccStages);
  uint64_t *push_bar = reinterpret_cast<uint64_t *>(smem + Cfg::kStages * Cfg::kStageBytes + 240);   // [2]
  uint64_t *tq_full = reinterpret_cast<uint64_t *>(smem + Cfg::kStages * Cfg::kStageBytes + 256);    // [kTileRing]
  int *tq_ring = reinterpret_cast<int *>(smem + Cfg::kStages * Cfg::kStageBytes + 384);              // [kTileRing]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const uint32_t rank = crank & 1;            // rank inside the CTA pair, 0 = leader
  const uint32_t pp = crank >> 1;             // pair inside the cluster
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int num_tiles = GROUPED ? groups[num_groups - 1].tile_end : p0.tiles_r * p0.tiles_c;
  // per-tile view: which member the tile belongs to (members are walked in order, so the scan
  // resumes where it stopped), its tensor maps, parameters and tile index inside the member
  int grp = 0;
  const CUtensorMap *tmx = &tmap_x0, *tmy = &tmap_y0, *tmsx = &tmap_sfx0, *tmsy = &tmap_sfy0;
  const Params *prm = &p0;
  auto locate = [&](int tile) -> int {
    if constexpr (GROUPED) {
      while (tile >= groups[grp].tile_end) ++grp;
      const GroupEntry *g = groups + grp;
      tmx = &g->mx; tmy = &g->my; tmsx = &g->msx; tmsy = &g->msy;
      prm = &g->p;
      return tile - (grp > 0 ? groups[grp - 1].tile_end : 0);
    } else {
      return tile;
    }
  };
  // k-th tile of this pair, or >= num_tiles when there is none. Static: pair + k * pairs. DYN: slot k % 16 of the
  // ring; `arm` = the caller is the CTA's producer thread, which arms the slot's barrier (arrive + expect 4 bytes)
  // before waiting on it, and on the leader CTA also fetches the tile and sends it to both CTAs.
  auto tile_at = [&](int k, bool arm) -> int {
    if constexpr (!DYN) {
      (void)arm;
      return pair + k * num_pairs;
    } else {
      const int slot = k & (Cfg::kTileRing - 1);
      if (arm) {
        mbar_expect_tx(&tq_full[slot], 4);
        if (rank == 0) {
          unsigned int t = atomicAdd(p0.tile_counter, 1u);
          if (t > static_cast<unsigned int>(num_tiles)) t = static_cast<unsigned int>(num_tiles);
          const uint32_t la = smem_u32(&tq_ring[slot]), lb = smem_u32(&tq_full[slot]);
          for (uint32_t dst = crank & ~1u; dst < (crank & ~1u) + 2u; ++dst) {
            uint32_t ra, rb;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(dst));
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rb) : "r"(lb), "r"(dst));
            asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(ra),
                         "r"(t), "r"(rb)
                         : "memory");
          }
        }
      }
      mbar_wait(&tq_full[slot], static_cast<uint32_t>(k / Cfg::kTileRing) & 1u);
      return *reinterpret_cast<volatile int *>(&tq_ring[slot]);
    }
  };
#define tmap_x (*tmx)
#define tmap_y (*tmy)
#define tmap_sfx (*tmsx)
#define tmap_sfy (*tmsy)
#define p (*prm)

  if (warp == 0 && lane == 0) {
    if constexpr (!GROUPED) {
      tma_prefetch_desc(&tmap_x);
      tma_prefetch_desc(&tmap_y);
      if (KIND != 0) {
        tma_prefetch_desc(&tmap_sfx);
        tma_prefetch_desc(&tmap_sfy);
      }
    }
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], CSZ / 2);
    }
    for (int s = 0; s < Cfg::kAccStages; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 2 * kNumEpilogueWarps);
      mbar_init(&acc_ovl[s], kNumEpilogueWarps);   // the owner half of both CTAs
    }
    if (Cfg::kPushBytes != 0) {
      mbar_init(&push_bar[0], 1);
      mbar_init(&push_bar[1], 1);
    }
    if (DYN)
      for (int s = 0; s < Cfg::kTileRing; ++s) mbar_init(&tq_full[s], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_cg2(tmem_slot, kTmemCols);
  tc_fence_before();
  cluster_sync_all();   // barriers of both CTAs are initialised before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch (see gemm_kernel.cuh): the prologue above overlaps the previous kernel's tail
  grid_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      grid_dependency_wait();
      int stage = 0;
      uint32_t phase = 0;
      long long t_empty = 0;
      unsigned long long t_flag = 0;   // ns spent waiting for arrival flags
      const bool timing = p0.debug != nullptr;   // dev aid (scripts/gemm_debug.py): off in production
      // (arr_advance: nobody advanced the step before this launch -- no amax exchange, MXFP8 -- so this launch is
      // step *counter + 1 and its last pair advances the counter at the very end of the kernel)
      const uint32_t arr_step = p0.arr_flags != nullptr
                                    ? *reinterpret_cast<const volatile uint32_t *>(p0.arr_step) + (p0.arr_advance ? 1u : 0u)
                                    : 0u;
      for (int it = 0;; ++it) {
        const int gtile = tile_at(it, true);
        if (gtile >= num_tiles) break;
        const int tile = locate(gtile);
        int tr, tc;
        tile_coords(p, tile, tr, tc);
        if (p.arr_flags != nullptr) {
          // a remote row chunk of X is loaded only after its sender raised the flag of this step
          const int cn = p.arr_fpr * p.arr_tpf, owner = tr / cn;
          if (owner != p.arr_rank) {
            // (with push the flag word counts landed units: arr_step * units per chunk, modulo 2^32)
            const uint32_t want =
                p.arr_push_unit != 0 ? arr_step * static_cast<uint32_t>(p.arr_push_upc + p.arr_push_upc2) : arr_step;
            const unsigned long long g0 = timing ? ptx::globaltimer_ns() : 0ull;
            ptx::wait_count_u32(p.arr_flags + owner * p.arr_fpr + (tr % cn) / p.arr_tpf, want, p.arr_error);
            if (timing) t_flag += ptx::globaltimer_ns() - g0;
            asm volatile("fence.proxy.async.global;" ::: "memory");   // the TMA (async proxy) reads what the flag covers
          }
        }
        const int bn = p.bn;
        const int xrow = tr * 256 + static_cast<int>(rank) * 128;         // my X rows
        const int yrow = tc * bn + static_cast<int>(rank) * (bn >> 1);    // my Y rows (box = bn / 2 rows)
        const uint32_t stage_tx = 2u * static_cast<uint32_t>(Cfg::kStageBytes - (kBN2 - bn) * (kBKBytes / 2));
        const int kblocks = p.kblocks, sfx_ld = p.sfx_ld, sfy_ld = p.sfy_ld;
        const bool x_mn = p.x.mn_major != 0, y_mn = p.y.mn_major != 0;
        for (int kb = 0; kb < kblocks; ++kb) {
          const long long t0 = timing ? clock64() : 0;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (timing) t_empty += clock64() - t0;
          uint8_t *sx = smem + stage * Cfg::kStageBytes;
          uint8_t *sy = sx + Cfg::kXBytes;
          // the leader's barrier collects the bytes of both CTAs
          if (rank == 0) mbar_expect_tx(&full_bar[stage], stage_tx);
          if (CSZ == 4) {
            // my half of the 128 X rows (K-major: 64 rows; MN-major: 64 of the 128 k-rows), to
            // both CTAs of my parity
            const uint16_t mask = static_cast<uint16_t>((1u << rank) | (1u << (2 + rank)));
            uint8_t *dst = sx + pp * (Cfg::kXBytes / 2);
            if (x_mn) tma_load_2d_cg2_mc(dst, &tmap_x, &full_bar[stage], xrow, kb * kBKBytes + static_cast<int>(pp) * 64, mask);
            else tma_load_2d_cg2_mc(dst, &tmap_x, &full_bar[stage], kb * kBKBytes, xrow + static_cast<int>(pp) * 64, mask);
          } else if (x_mn) {
            tma_load_2d_cg2(sx, &tmap_x, &full_bar[stage], xrow, kb * kBKBytes);
          } else {
            tma_load_2d_cg2(sx, &tmap_x, &full_bar[stage], kb * kBKBytes, xrow);
          }
          if (y_mn) tma_load_2d_cg2(sy, &tmap_y, &full_bar[stage], yrow, kb * kBKBytes);
          else tma_load_2d_cg2(sy, &tmap_y, &full_bar[stage], kb * kBKBytes, yrow);
          if (KIND != 0) {
            uint8_t *ssx = sy + Cfg::kYBytes;
            uint8_t *ssy = ssx + Cfg::kSfxBytes;
            // scale atoms: [row block][k atom][512 B]; tensor map = [atoms, 128 x u32]
            tma_load_2d_cg2(ssx, &tmap_sfx, &full_bar[stage], 0,
                            (tr * 2 + static_cast<int>(rank)) * sfx_ld + kb * Cfg::kSfAtoms);
#pragma unroll
            for (int b = 0; b < 2; ++b)
              tma_load_2d_cg2(ssy + b * Cfg::kSfxBytes, &tmap_sfy, &full_bar[stage], 0,
                              (tc * 2 + b) * sfy_ld + kb * Cfg::kSfAtoms);
          }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
      if (timing && rank == 0) {
        p0.debug[pair * 4 + 3] = t_empty;
        p0.debug[600 + pair] = t_flag;
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    // The whole warp runs the loop so that stage / phase / descriptor arithmetic stays
    // warp-uniform (uniform registers feed UTCCP / UTCQMMA directly: with a single divergent
    // lane every operand went through R2UR and one tcgen05 issue cost ~70 cycles, which made
    // the issuing thread -- not the tensor pipe -- the bottleneck of the block-scaled kinds).
    // Only the elected lane executes the tcgen05 instructions.
    if (rank == 0) {
      const bool issuer = elect_one();
      int stage = 0, as = 0, sfs = 0;
      uint32_t phase = 0, aphase = 0;
      long long t_full = 0, t_acc = 0;
      const bool timing = p0.debug != nullptr;
      const long long t_begin = timing ? clock64() : 0;
      const uint64_t dsc = make_smem_desc(0, 0, 128, kSwizzleNone);
      for (int it = 0;; ++it) {
        const int gtile = tile_at(it, false);
        if (gtile >= num_tiles) break;
        locate(gtile);
        // constant upper halves of the shared-memory descriptors
        const uint64_t dxc = make_smem_desc(0, p.x.lbo, p.x.sbo, kSwizzle128B);
        const uint64_t dyc = make_smem_desc(0, p.y.lbo, p.y.sbo, kSwizzle128B);
        const uint32_t xstep = p.x.kstep >> 4, ystep = p.y.kstep >> 4;
        const int kblocks = p.kblocks;
        const uint32_t idesc = p.idesc;
        const int dflags = p0.debug_flags;
        long long t0 = timing ? clock64() : 0;
        mbar_wait(&acc_empty[as], aphase ^ 1);
        // the previous tile (other accumulator) has left the columns the two accumulators share
        if (Cfg::kOverlap != 0) mbar_wait(&acc_ovl[as ^ 1], as == 0 ? aphase ^ 1 : aphase);
        if (timing) t_acc += clock64() - t0;
        if (timing && issuer && pair == 0 && gtile == pair + 2 * num_pairs) p0.debug[513] = clock64();
        tc_fence_after();
        const uint32_t tmem_acc = tmem_base + Cfg::acc_col(as);
        for (int kb = 0; kb < kblocks; ++kb) {
          if (timing) t0 = clock64();
          mbar_wait(&full_bar[stage], phase);
          if (timing) t_full += clock64() - t0;
          tc_fence_after();
          const uint32_t sx = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t sy = sx + Cfg::kXBytes;
          const uint32_t tsfx = tmem_base + Cfg::kSfColBase + sfs * (Cfg::kSfColsX + Cfg::kSfColsY);
          const uint32_t tsfy = tsfx + Cfg::kSfColsX;
          const uint64_t dx0 = dxc | ((sx & 0x3ffffu) >> 4);
          const uint64_t dy0 = dyc | ((sy & 0x3ffffu) >> 4);
          if (issuer) {
            if (KIND != 0 && !(dflags & 1)) {
              const uint64_t dsx = dsc | (((sy + Cfg::kYBytes) & 0x3ffffu) >> 4);
              const uint64_t dsy = dsx + (Cfg::kSfxBytes >> 4);
#pragma unroll
              for (int a = 0; a < Cfg::kSfAtoms; ++a) {
                tmem_cp_sf_cg2(tsfx + 4 * a, dsx + 32 * a);
#pragma unroll
                for (int b = 0; b < 2; ++b)
                  tmem_cp_sf_cg2(tsfy + 4 * (a * 2 + b), dsy + 32 * (b * Cfg::kSfAtoms + a));
              }
            }
            if (!(dflags & 2)) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint64_t dx = dx0 + j * xstep;
                const uint64_t dy = dy0 + j * ystep;
                const uint32_t accumulate = (kb | j) != 0;
                if (KIND == 0) mma_f8_cg2(tmem_acc, dx, dy, idesc, accumulate);
                else if (KIND == 1) mma_mxf8_cg2(tmem_acc, dx, dy, with_sf_ids(idesc, j, j), accumulate, tsfx, tsfy);
                else mma_nvf4_cg2(tmem_acc, dx, dy, idesc, accumulate, tsfx + 4 * j, tsfy + 8 * j);
              }
            }
            tc_commit_cg2(&empty_bar[stage], CSZ == 4 ? 0xF : 3);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
          if (++sfs == Cfg::kSfSlots) sfs = 0;
        }
        if (issuer) tc_commit_cg2(&acc_full[as], static_cast<uint16_t>(3u << (2 * pp)));
        __syncwarp();
        // dev aid: hand-off trace of pair 0, tile #1 -> #2 (raw clock64 of the leader's SM)
        if (timing && issuer && pair == 0 && gtile == pair + num_pairs) p0.debug[512] = clock64();
        if (++as == Cfg::kAccStages) { as = 0; aphase ^= 1; }
      }
      if (timing && issuer) {
        p0.debug[pair * 4 + 0] = clock64() - t_begin;
        p0.debug[pair * 4 + 1] = t_full;
        p0.debug[pair * 4 + 2] = t_acc;
      }
    } else if (!GROUPED && p0.arr_push_unit != 0 && p0.arr_push_mc_dst != nullptr) {
      // ===================== pusher, NVSwitch multicast (non-leader CTA's otherwise idle warp) =====================
      // One multimem.st reaches every rank's gather buffer, so the rank's rows leave this GPU once. Units of a
      // chunk are dealt round-robin to the pushers; after its units of chunk c a pusher fences at system scope and
      // adds their number to the chunk's counter on every rank with one multimem.red (release).
      grid_dependency_wait();   // the local cast wrote the rows
      const bool ptiming = p0.debug != nullptr;
      const unsigned long long pg0 = ptiming ? ptx::globaltimer_ns() : 0ull;
      unsigned long long t_fence = 0;
      const int upc = p0.arr_push_upc, fpr = p0.arr_fpr, me = p0.arr_rank;
      const size_t unit = static_cast<size_t>(p0.arr_push_unit);
      if (p0.arr_ready != nullptr) {
        // single buffer set, no amax exchange to order the steps: the peers must have finished READING the previous
        // step's rows before they are overwritten (each GPU says so at the end of its GEMM, see the kernel's tail)
        const uint32_t e = *reinterpret_cast<const volatile uint32_t *>(p0.arr_step) + 1u;
        if (lane == 0) ptx::wait_count_u32(p0.arr_ready, (e - 1u) * static_cast<uint32_t>(p0.arr_world), p0.arr_error);
        __syncwarp();
      }
      auto copy_unit = [&](const uint8_t *s, uint8_t *d, size_t bytes) {
        const uint4 *src = reinterpret_cast<const uint4 *>(s) + lane;
        uint8_t *dst = d + lane * 16;
        for (int i = 0; i < static_cast<int>(bytes / 512); i += 8) {   // 8 x 512 B in flight per warp
          uint4 v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = ptx::ld_stream_v4(src + (i + j) * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) ptx::multimem_st_v4(dst + static_cast<size_t>(i + j) * 512, v[j]);
        }
      };
      const int upc2 = p0.arr_push_upc2;
      const size_t unit2 = static_cast<size_t>(p0.arr_push_unit2);
      for (int c = 0; c < fpr; ++c) {
        const int lo = c * upc, hi = lo + upc;
        int u = lo + ((pair - lo) % num_pairs + num_pairs) % num_pairs;
        uint32_t mine = 0;
        for (; u < hi; u += num_pairs, ++mine) copy_unit(p0.arr_push_src + u * unit, p0.arr_push_mc_dst + u * unit, unit);
        // second region of the chunk (block scales of an MXFP8 shard), dealt the same way
        const int lo2 = c * upc2, hi2 = lo2 + upc2;
        for (u = lo2 + ((pair - lo2) % num_pairs + num_pairs) % num_pairs; u < hi2; u += num_pairs, ++mine)
          copy_unit(p0.arr_push_src2 + u * unit2, p0.arr_push_mc_dst2 + u * unit2, unit2);
        if (mine != 0) {
          const unsigned long long f0 = ptiming ? ptx::globaltimer_ns() : 0ull;
          __threadfence_system();
          __syncwarp();
          if (lane == 0) ptx::multimem_red_release_add_u32(p0.arr_push_mc_cnt + me * fpr + c, mine);
          if (ptiming) t_fence += ptx::globaltimer_ns() - f0;
        }
      }
      if (ptiming && lane == 0) {
        p0.debug[700 + pair] = ptx::globaltimer_ns() - pg0;
        p0.debug[800 + pair] = t_fence;
      }
    } else if (KIND == 0 && !GROUPED && p0.arr_push_unit != 0 && lane == 0) {
      // ===================== pusher (non-leader CTA; its MMA warp has nothing else to do) =====================
      // The all-gather of the sequence-parallel input, inside the GEMM that consumes it: this rank's rows of X go
      // to every peer's gather buffer in 16 KB units -- bulk copy global -> shared, then one bulk store per peer
      // over NVLink -- chunk-major, so that everybody's chunk 0 lands first. Two units in flight; when a unit's
      // stores have completed the peers' counters of its chunk go up by one (release, system scope).
      if constexpr (Cfg::kPushBytes != 0) {
        grid_dependency_wait();   // the local cast wrote the rows
        const int unit = p0.arr_push_unit, upc = p0.arr_push_upc;
        const int total = upc * p0.arr_fpr, world = p0.arr_world, me = p0.arr_rank;
        uint8_t *ring = smem + Cfg::kPushOff;
        auto signal = [&](int k) {
          const int c = (pair + k * num_pairs) / upc;
          for (int j = 1; j < world; ++j) {
            const int r = (me + j) % world;
            ptx::red_release_sys_add_u32(p0.arr_push_cnt[r] + me * p0.arr_fpr + c, 1u);
          }
        };
        int n = 0;
        for (int u = pair; u < total; u += num_pairs, ++n) {
          const int b = n & 1;
          if (n >= 2) {
            bulk_wait_group<1>();   // unit n - 2 has landed everywhere (and its buffer is free)
            signal(n - 2);
          }
          mbar_expect_tx(&push_bar[b], static_cast<uint32_t>(unit));
          bulk_load_1d(ring + b * unit, p0.arr_push_src + static_cast<size_t>(u) * unit, static_cast<uint32_t>(unit),
                       &push_bar[b]);
          mbar_wait(&push_bar[b], static_cast<uint32_t>(n >> 1) & 1u);
          for (int j = 1; j < world; ++j) {
            const int r = (me + j) % world;
            bulk_store_1d(p0.arr_push_dst[r] + static_cast<size_t>(u) * unit, ring + b * unit, static_cast<uint32_t>(unit));
          }
          bulk_commit();
        }
        if (n >= 2) {
          bulk_wait_group<1>();
          signal(n - 2);
        }
        if (n >= 1) {
          bulk_wait_group<0>();
          signal(n - 1);
        }
      }
    }
  } else {
    // ===================== epilogue (both CTAs, own 128 rows) =====================
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    constexpr int kChunks = kBN2 / 64;
    grid_dependency_wait();
    float alpha = 0.f;
    int as = 0, alpha_grp = -1;
    uint32_t aphase = 0;
    for (int it = 0;; ++it) {
      const int gtile = tile_at(it, false);
      if (gtile >= num_tiles) break;
      const int tile = locate(gtile);
      if (alpha_grp != grp) {   // once per member (once per kernel when not grouped)
        alpha_grp = grp;
        alpha = p.alpha * p.scalar_const;
        if (p.s0 != nullptr) alpha *= *p.s0;
        if (p.s1 != nullptr) alpha *= *p.s1;
      }
      int tr, tc;
      tile_coords(p, tile, tr, tc);
      mbar_wait(&acc_full[as], aphase);
      const bool trace = p0.debug != nullptr && pair == 0 && rank == 0 && lane == 0 && gtile == pair + num_pairs;
      if (trace) p0.debug[520 + 4 * (warp - 2)] = clock64();
      tc_fence_after();
      uint32_t acc[kChunks][32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + Cfg::acc_col(as) + half * (kBN2 / 2);
      // acc[c] holds column chunk pc(c): acc0 is drained back to front, acc1 front to back, so the
      // columns the accumulators share leave first; their owners (half 1 of acc0, half 0 of acc1)
      // signal acc_ovl as soon as those chunks are in registers
      const bool rev = Cfg::kOverlap != 0 && as == 0;
      const bool owner = Cfg::kOverlap != 0 && half == (as == 0 ? 1 : 0);
      // TMEM -> register bandwidth is ~64-90 B/clk per SM (measured: 128 KB drain in ~2.2 K cycles), so
      // the shared columns must not queue behind the other half's loads: the four non-owner warps hold
      // their loads back (named barrier 1) until the owners have the shared chunks in registers
      if (Cfg::kOverlap != 0 && !owner) asm volatile("bar.sync 1, 256;" ::: "memory");
      const int bn = p.bn;
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        if (half * (kBN2 / 2) + c * 32 < bn || Cfg::kOverlap != 0)   // narrower tiles (FP8 only) end inside half 1
          tmem_ld_32x32(taddr + (rev ? kChunks - 1 - c : c) * 32, acc[c]);
        if (Cfg::kOverlap != 0 && c == Cfg::kOvlChunks - 1 && owner) {
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster_relaxed(&acc_ovl[as], crank & ~1u);
          asm volatile("bar.arrive 1, 256;" ::: "memory");
          if (trace) p0.debug[521 + 4 * (warp - 2)] = clock64();
        }
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster_relaxed(&acc_empty[as], crank & ~1u);
      if (trace) p0.debug[522 + 4 * (warp - 2)] = clock64();
      const size_t row = static_cast<size_t>(tr) * 256 + rank * 128 + q * 32 + lane;
      const bool row_ok = row < static_cast<size_t>(p.R);
      if (KIND == 0 && p.sc_staged != 0) {
        if constexpr (Cfg::kPushBytes != 0) {
          // scatter output, staged (see stage_chunk16): chunk pairs (c, c + 1) are 64 adjacent columns
          const uint32_t stg = smem_u32(smem + Cfg::kPushOff + (warp - 2) * 4096);
          const size_t row0 = static_cast<size_t>(tr) * 256 + rank * 128 + q * 32;   // the warp's first row
          const size_t own = row0 / static_cast<size_t>(p.sc_rows);
          uint8_t *dbase = static_cast<uint8_t *>(p.sc_out[own]) + (row0 - own * p.sc_rows) * static_cast<size_t>(p.C) * 2;
#pragma unroll
          for (int g = 0; g < kChunks / 2; ++g) {
            const int ca = 2 * g, cb = 2 * g + 1;
            const int offa = half * (kBN2 / 2) + (rev ? kChunks - 1 - ca : ca) * 32;
            const int offb = half * (kBN2 / 2) + (rev ? kChunks - 1 - cb : cb) * 32;
            const int lo = offa < offb ? offa : offb;
            if (lo >= bn) continue;
            const int col0 = tc * bn + lo;
            if (p.out_dtype == kNVTEBFloat16) {
              const __nv_bfloat16 *b = p.bias != nullptr ? reinterpret_cast<const __nv_bfloat16 *>(p.bias) + tc * bn : nullptr;
              stage_chunk16<__nv_bfloat16>(stg, lane, offa < offb ? 0 : 4, acc[ca], alpha, b ? b + offa : nullptr);
              stage_chunk16<__nv_bfloat16>(stg, lane, offa < offb ? 4 : 0, acc[cb], alpha, b ? b + offb : nullptr);
            } else {
              const __half *b = p.bias != nullptr ? reinterpret_cast<const __half *>(p.bias) + tc * bn : nullptr;
              stage_chunk16<__half>(stg, lane, offa < offb ? 0 : 4, acc[ca], alpha, b ? b + offa : nullptr);
              stage_chunk16<__half>(stg, lane, offa < offb ? 4 : 0, acc[cb], alpha, b ? b + offb : nullptr);
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int r = 4 * i + (lane >> 3), j = lane & 7;
              uint4 v;
              asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                           : "r"(stg + r * 128 + ((j ^ (r & 7)) << 4)));
              if (row0 + r < static_cast<size_t>(p.R))
                *reinterpret_cast<uint4 *>(dbase + (static_cast<size_t>(r) * p.C + col0) * 2 + j * 16) = v;
            }
            __syncwarp();
          }
        }
      } else
#pragma unroll
      for (int c = 0; c < kChunks; ++c) {
        const int coff = half * (kBN2 / 2) + (rev ? kChunks - 1 - c : c) * 32;
        if (coff >= bn) continue;
        const int col0 = tc * bn + coff;
        if (p.out_dtype == kNVTEBFloat16) epilogue_chunk<__nv_bfloat16>(p, acc[c], alpha, row, col0, row_ok);
        else if (p.out_dtype == kNVTEFloat32) epilogue_chunk<float>(p, acc[c], alpha, row, col0, row_ok);
        else epilogue_chunk<__half>(p, acc[c], alpha, row, col0, row_ok);
      }
      if (trace) p0.debug[523 + 4 * (warp - 2)] = clock64();
      if (++as == Cfg::kAccStages) { as = 0; aphase ^= 1; }
    }
  }

#undef tmap_x
#undef tmap_y
#undef tmap_sfx
#undef tmap_sfy
#undef p
  tc_fence_before();
  cluster_sync_all();   // the peer's smem / TMEM stay alive until every MMA of the pair retired
  if (warp == 1) tmem_dealloc_cg2(tmem_base, kTmemCols);
  if (DYN && rank == 0 && threadIdx.x == 0) {
    // Every pair has fetched exactly once past the end and (after the cluster barrier) consumed all it loaded. The
    // last pair to get here leaves the counters at zero for the next launch, and for an exchange without an amax
    // step (arr_advance) it also advances the step and tells every rank that this GPU has finished READING its
    // gather buffer of this step (the peers' next push waits for world such messages per step).
    __threadfence();
    if (atomicAdd(p0.tile_counter + 1, 1u) == static_cast<unsigned int>(num_pairs) - 1u) {
      p0.tile_counter[0] = 0u;
      p0.tile_counter[1] = 0u;
      if (p0.arr_advance) {
        const uint32_t e = *reinterpret_cast<const volatile uint32_t *>(p0.arr_step) + 1u;
        *reinterpret_cast<volatile uint32_t *>(const_cast<uint32_t *>(p0.arr_step)) = e;
        if (p0.arr_ready_mc != nullptr) ptx::multimem_red_release_add_u32(p0.arr_ready_mc, 1u);
      }
    }
  }
}

}  // namespace gemm
}  // namespace teb
