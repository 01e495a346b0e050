// Host side of the block-scaled GEMM: operand canonicalisation (which buffer of
// each quantized tensor feeds the tensor cores for TN / NN / NT), TMA tensor maps,
// launch. Replaces reference gemm/cublaslt_gemm.cu:102-893 (cuBLASLt) with the
// hand-written tcgen05 kernel in gemm_kernel.cuh. No descriptor heuristics per
// call: tensor maps are cached, the launch is a single kernel.
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <vector>
#include <mutex>
#include <unordered_map>

#include <algorithm>
#include <cstdio>

#include "core.h"
#include "gemm2_kernel.cuh"
#include "gemm_kernel.cuh"

namespace teb {
namespace gemm {

namespace {

using EncodeTiledFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                   const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                   const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    TEB_CHECK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    TEB_CHECK(p != nullptr && q == cudaDriverEntryPointSuccess,
              "cuTensorMapEncodeTiled is not available in this driver");
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

struct MapKey {
  const void *ptr;
  uint64_t inner, outer;
  uint32_t box_inner, box_outer;
  uint32_t kind;  // 0: bytes, 128B swizzle; 1: 32-bit words, no swizzle (scale atoms)
  bool operator==(const MapKey &o) const {
    return ptr == o.ptr && inner == o.inner && outer == o.outer && box_inner == o.box_inner &&
           box_outer == o.box_outer && kind == o.kind;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey &k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    h = h * 1000003u ^ k.inner;
    h = h * 1000003u ^ k.outer;
    h = h * 1000003u ^ (static_cast<size_t>(k.box_inner) << 16 | k.box_outer | static_cast<size_t>(k.kind) << 40);
    return h;
  }
};

// kind 0: byte tensor [outer, inner] (inner contiguous), box [box_outer, box_inner], 128B swizzle.
// kind 1: the same over 32-bit words without swizzle (512-byte scale atoms: inner = 128 words).
CUtensorMap make_map(const void *ptr, uint64_t inner, uint64_t outer, uint32_t box_inner,
                     uint32_t box_outer, uint32_t kind = 0) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  const MapKey key{ptr, inner, outer, box_inner, box_outer, kind};
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
  }
  CUtensorMap m;
  const cuuint64_t dims[2] = {inner, outer};
  const cuuint64_t strides[1] = {inner * (kind == 1 ? 4u : 1u)};
  const cuuint32_t box[2] = {box_inner, box_outer};
  const cuuint32_t elem[2] = {1, 1};
  const CUresult r = encode_fn()(
      &m, kind == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2,
      const_cast<void *>(ptr), dims, strides, box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE,
      kind == 1 ? CU_TENSOR_MAP_SWIZZLE_NONE : (kind == 2 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B),
      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TEB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (", static_cast<int>(r),
            ") for a [", outer, ", ", inner, "] byte tensor at ", ptr,
            "; rows must be 16-byte multiples and the base 16-byte aligned");
  std::lock_guard<std::mutex> lock(mu);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, m);
  return m;
}

// One GEMM operand after canonicalisation.
struct Side {
  const void *data = nullptr;
  NVTEDType dtype = kNVTEByte;
  bool mn_major = false;       // global layout [K, MN] instead of [MN, K]
  size_t mn = 0, k = 0;        // logical extents (elements)
  const uint8_t *sf = nullptr; // swizzled block scales
  size_t sf_cols = 0;          // padded scale columns (K direction)
  const float *scalar = nullptr;  // scale_inv (FP8) or amax (NVFP4)
};

// Pick the buffer that feeds the contraction (reference CanonicalizeGemmInput,
// gemm/cublaslt_gemm.cu:102-307). `k_is_last` = the contraction runs along the
// tensor's last logical dimension (TE "T" for A, "N" for B).
Side canonicalize(const Tensor &t, bool k_is_last, const char *name) {
  Side s;
  const size_t rows = t.rows(), cols = t.cols();
  s.mn = k_is_last ? rows : cols;
  s.k = k_is_last ? cols : rows;
  switch (t.scaling_mode) {
    case NVTE_DELAYED_TENSOR_SCALING: {
      TEB_CHECK(t.data.dptr != nullptr, name, ": per-tensor FP8 GEMM operand needs rowwise data "
                "(sm_100 consumes MN-major FP8 directly, no transpose buffer is used)");
      TEB_CHECK(is_fp8(t.data.dtype), name, ": only FP8 operands are implemented by this GEMM, got ",
                dtype_name(t.data.dtype));
      s.data = t.data.dptr;
      s.dtype = t.data.dtype;
      s.mn_major = !k_is_last;
      s.scalar = t.scale_inv.as<float>();
      TEB_CHECK(s.scalar != nullptr, name, ": FP8 operand without scale_inv");
      break;
    }
    case NVTE_MXFP8_1D_SCALING: {
      TEB_CHECK(t.with_gemm_swizzled_scales, name,
                ": MXFP8 scaling factors must be in the GEMM-swizzled layout");
      const Buffer &d = k_is_last ? t.data : t.columnwise_data;
      const Buffer &sc = k_is_last ? t.scale_inv : t.columnwise_scale_inv;
      TEB_CHECK(d.dptr != nullptr && sc.dptr != nullptr, name, ": missing ",
                k_is_last ? "rowwise" : "columnwise", " MXFP8 data/scales");
      s.data = d.dptr;
      s.dtype = d.dtype;
      s.mn_major = !k_is_last;  // columnwise MXFP8 data keeps the [rows, cols] layout
      s.sf = sc.as<uint8_t>();
      s.sf_cols = round_up(ceil_div(s.k, 32), 4);
      TEB_CHECK(sc.numel() >= round_up(s.mn, 128) * s.sf_cols, name, ": scale buffer too small");
      break;
    }
    case NVTE_NVFP4_1D_SCALING: {
      TEB_CHECK(t.with_gemm_swizzled_scales, name,
                ": NVFP4 scaling factors must be in the GEMM-swizzled layout");
      const Buffer &d = k_is_last ? t.data : t.columnwise_data;  // columnwise is stored transposed
      const Buffer &sc = k_is_last ? t.scale_inv : t.columnwise_scale_inv;
      const Buffer &am = k_is_last ? t.amax : t.columnwise_amax;
      TEB_CHECK(d.dptr != nullptr && sc.dptr != nullptr && am.dptr != nullptr, name, ": missing ",
                k_is_last ? "rowwise" : "columnwise", " NVFP4 data/scales/amax");
      s.data = d.dptr;
      s.dtype = kNVTEFloat4E2M1;
      s.mn_major = false;
      s.sf = sc.as<uint8_t>();
      s.sf_cols = round_up(ceil_div(s.k, 16), 4);
      s.scalar = am.as<float>();
      TEB_CHECK(sc.numel() >= round_up(s.mn, 128) * s.sf_cols, name, ": scale buffer too small");
      break;
    }
    default:
      TEB_ERROR(name, ": unsupported scaling mode ", static_cast<int>(t.scaling_mode));
  }
  return s;
}

bool pdl_enabled() {
  static const bool on = [] { const char *e = getenv("NVTE_B200_GEMM_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

template <int BN, int KIND>
void launch(const CUtensorMap &mx, const CUtensorMap &my, const Params &p, int sms,
            cudaStream_t stream) {
  using Cfg = Config<BN, KIND>;
  static bool configured = [] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(gemm_kernel<BN, KIND>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    return true;
  }();
  (void)configured;
  const int tiles = p.tiles_r * p.tiles_c;
  const int grid = tiles < sms ? tiles : sms;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm_kernel<BN, KIND>, mx, my, p));
}

}  // namespace

// Cached 2-D tensor map over raw bytes with 128B swizzle (also used by the NVFP4 quantizer).
CUtensorMap tensor_map_bytes_2d(const void *ptr, uint64_t inner_bytes, uint64_t outer,
                                uint32_t box_inner_bytes, uint32_t box_outer) {
  return make_map(ptr, inner_bytes, outer, box_inner_bytes, box_outer, 0);
}
// Same with the 64-byte swizzle (64-byte-wide boxes of packed FP4 leaving the NVFP4 quantizer).
CUtensorMap tensor_map_bytes_2d_sw64(const void *ptr, uint64_t inner_bytes, uint64_t outer,
                                     uint32_t box_inner_bytes, uint32_t box_outer) {
  return make_map(ptr, inner_bytes, outer, box_inner_bytes, box_outer, 2);
}

// Development aid: with NVTE_B200_GEMM_DEBUG=1 the MMA / producer threads of the CTA-pair kernel
// leave cycle counters here ([pair][total, wait_full, wait_acc_empty, producer_wait_empty]).
unsigned long long *debug_counters() {
  static unsigned long long *buf = [] {
    unsigned long long *d = nullptr;
    const char *e = getenv("NVTE_B200_GEMM_DEBUG");
    if (e && e[0] == '1') cudaMallocManaged(&d, 1024 * sizeof(unsigned long long));
    return d;
  }();
  return buf;
}
namespace {

template <int KIND, int CSZ, bool DYN = false>
void configure2() {
  using Cfg = Config2<KIND>;
  static bool configured = [] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(gemm2_kernel<KIND, CSZ, false, DYN>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    return true;
  }();
  (void)configured;
}

// How many clusters of 4 CTAs (1 CTA per SM) the device can hold at once: clusters do not
// straddle GPCs, so this is less than sm_count / 4 (33 on a 148-SM B200).
template <int KIND>
int resident_clusters4() {
  static int n = [] {
    using Cfg = Config2<KIND>;
    configure2<KIND, 4>();
    cudaLaunchConfig_t q{};
    q.gridDim = dim3(4 * (sm_count() / 4));
    q.blockDim = dim3(kNumThreads);
    q.dynamicSmemBytes = Cfg::kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 4;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    q.attrs = attr;
    q.numAttrs = 1;
    int r = 0;
    if (cudaOccupancyMaxActiveClusters(&r, gemm2_kernel<KIND, 4>, &q) != cudaSuccess || r <= 0) {
      (void)cudaGetLastError();
      r = sm_count() / 4;
    }
    if (const char *e = getenv("NVTE_B200_GEMM_VERBOSE"); e && e[0] == '1')
      fprintf(stderr, "[teb] gemm2: %d clusters of 4 can be resident (%d SMs)\n", r, sm_count());
    return r;
  }();
  return n;
}

// `clusters` = number of CSZ-CTA clusters to launch (persistent: each walks the tile list)
template <int KIND, int CSZ, bool DYN = false>
void launch2(const CUtensorMap &mx, const CUtensorMap &my, const CUtensorMap &msx,
             const CUtensorMap &msy, const Params &p, int clusters, cudaStream_t stream) {
  using Cfg = Config2<KIND>;
  configure2<KIND, CSZ, DYN>();
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CSZ;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  const int tiles = p.tiles_r * p.tiles_c;
  const int need = static_cast<int>(ceil_div(static_cast<size_t>(tiles), static_cast<size_t>(CSZ / 2)));
  if (clusters > need) clusters = need;
  cfg.gridDim = dim3(CSZ * clusters);
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm2_kernel<KIND, CSZ, false, DYN>, mx, my, msx, msy, p,
                                    static_cast<const GroupEntry *>(nullptr), 0));
}

// Grouped launch: one persistent kernel walks the concatenated tile list of all members.
template <int KIND>
void launch_grouped(const GroupEntry *table_dev, int num_groups, int total_tiles, const Params &p0,
                    const CUtensorMap &any_map, int pairs, cudaStream_t stream) {
  using Cfg = Config2<KIND>;
  static bool configured = [] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(gemm2_kernel<KIND, 2, true>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    return true;
  }();
  (void)configured;
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (pairs > total_tiles) pairs = total_tiles;
  cfg.gridDim = dim3(2 * pairs);
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gemm2_kernel<KIND, 2, true>, any_map, any_map, any_map, any_map, p0,
                                    table_dev, num_groups));
}

// Side stream + events used to run the two halves of a split GEMM concurrently (fork / join on
// the caller's stream; capture-safe).
struct SplitStreams {
  cudaStream_t side = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
SplitStreams &split_streams() {
  static thread_local SplitStreams s = [] {
    SplitStreams r;
    TEB_CHECK_CUDA(cudaStreamCreateWithFlags(&r.side, cudaStreamNonBlocking));
    TEB_CHECK_CUDA(cudaEventCreateWithFlags(&r.fork, cudaEventDisableTiming));
    TEB_CHECK_CUDA(cudaEventCreateWithFlags(&r.join, cudaEventDisableTiming));
    return r;
  }();
  return s;
}

// Large problems: 33 clusters of 4 (two pairs sharing their X tile by TMA multicast) cover 132 of
// the 148 SMs; the 16 SMs that no cluster of 4 can use run plain CTA pairs on the last tile rows,
// concurrently, from a side stream. The split is by measured throughput (a multicast pair moves
// ~12 % more than a plain one because it pulls 24 KB instead of 32 KB per k-block through L2).
template <int KIND>
void launch_split(const CUtensorMap &mx64, const CUtensorMap &mx128, const CUtensorMap &my,
                  const CUtensorMap &msx, const CUtensorMap &msy, const Params &p, int sms,
                  cudaStream_t stream) {
  const int c4 = std::min(resident_clusters4<KIND>(), sms / 4);
  const int left_pairs = (sms - 4 * c4) / 2;
  Params pa = p;
  if (left_pairs <= 0 || p.tiles_r < 8) {
    launch2<KIND, 4>(mx64, my, msx, msy, pa, c4, stream);
    return;
  }
  const double w4 = 2.0 * c4 * 1.12, w2 = static_cast<double>(left_pairs);
  int rows4 = static_cast<int>(p.tiles_r * w4 / (w4 + w2) + 0.5);
  rows4 = std::max(1, std::min(p.tiles_r - 1, rows4));
  Params pb = p;
  pa.tiles_r = rows4;
  pb.tiles_r = p.tiles_r - rows4;
  pb.tile_r0 = p.tile_r0 + rows4;
  SplitStreams &ss = split_streams();
  TEB_CHECK_CUDA(cudaEventRecord(ss.fork, stream));
  launch2<KIND, 4>(mx64, my, msx, msy, pa, c4, stream);
  TEB_CHECK_CUDA(cudaStreamWaitEvent(ss.side, ss.fork, 0));
  launch2<KIND, 2>(mx128, my, msx, msy, pb, left_pairs, ss.side);
  TEB_CHECK_CUDA(cudaEventRecord(ss.join, ss.side));
  TEB_CHECK_CUDA(cudaStreamWaitEvent(stream, ss.join, 0));
}

}  // namespace

namespace {
// Everything about one GEMM that is decided on the host: canonical operands, kernel parameters,
// which kernel family serves it. `force_pair` = plan for the CTA-pair kernel whatever the shape
// (grouped launches use one kernel for all members).
struct Plan {
  Params p{};
  Side x, y;
  int kind = 0;
  size_t k_bytes = 0;
  bool pair = false;
  int BN = 256;
  bool empty = false;   // zero-sized output: nothing to launch
};

Plan make_plan(int transa, int transb, float alpha, const Tensor &A, const Tensor &B, float beta,
               const Tensor *C, Tensor &D, const MatmulConfig &cfg, bool force_pair) {
  Plan pl;
  TEB_CHECK(!cfg.with_gelu_epilogue && !cfg.with_dgelu_epilogue,
            "GELU epilogues are outside the Linear path and not implemented");
  TEB_CHECK(cfg.dbias_tensor == nullptr || !lookup(cfg.dbias_tensor) ||
                lookup(cfg.dbias_tensor)->data.dptr == nullptr,
            "The bias-gradient epilogue is not implemented");
  TEB_CHECK(A.scaling_mode == B.scaling_mode, "Inputs A and B to GEMM need to have the same "
            "scaling mode (got ", static_cast<int>(A.scaling_mode), " and ",
            static_cast<int>(B.scaling_mode), ")");
  // Y = weight-side operand (TE "A"), X = activation-side operand (TE "B").
  const Side y = canonicalize(A, transa != 0, "A");
  const Side x = canonicalize(B, transb == 0, "B");
  TEB_CHECK(x.k == y.k, "GEMM inner dimensions do not match: A has k=", y.k, ", B has k=", x.k);
  const size_t R = x.mn, Cc = y.mn, K = x.k;
  TEB_CHECK(D.data.dptr != nullptr, "GEMM output D has no data");
  const int32_t *rb = cfg.out_row_blocks;
  if (rb[0] > 0) {
    TEB_CHECK(R % rb[0] == 0 && rb[1] >= rb[0] && rb[2] >= 0, "Bad output row blocks {", rb[0], ", ", rb[1], ", ",
              rb[2], "} for ", R, " output rows");
    TEB_CHECK(D.data.numel() >= ((R / rb[0] - 1) * static_cast<size_t>(rb[1]) + rb[2] + rb[0]) * Cc,
              "GEMM output is too small for the requested row blocks");
    TEB_CHECK(beta == 0.0f, "Output row blocks cannot be combined with accumulation");
  } else {
    TEB_CHECK(D.data.numel() == R * Cc, "GEMM output has ", D.data.numel(), " elements, expected ",
              R, " x ", Cc);
  }
  if (cfg.scatter.world > 1) {
    TEB_CHECK(rb[0] <= 0 && beta == 0.0f && cfg.arrival.world <= 1,
              "Scatter output cannot be combined with output row blocks, accumulation or arrival flags");
    TEB_CHECK(R == static_cast<size_t>(cfg.scatter.world) * cfg.scatter.rows_per_rank,
              "Scatter output: ", R, " output rows != world ", cfg.scatter.world, " x rows_per_rank ",
              cfg.scatter.rows_per_rank);
  }
  TEB_CHECK(D.data.dtype == kNVTEBFloat16 || D.data.dtype == kNVTEFloat16 ||
                D.data.dtype == kNVTEFloat32,
            "GEMM output dtype ", dtype_name(D.data.dtype), " is not supported");
  if (R == 0 || Cc == 0) { pl.empty = true; return pl; }

  int kind;
  size_t k_bytes;
  if (A.scaling_mode == NVTE_DELAYED_TENSOR_SCALING) { kind = 0; k_bytes = K; }
  else if (A.scaling_mode == NVTE_MXFP8_1D_SCALING) { kind = 1; k_bytes = K; }
  else { kind = 2; k_bytes = K / 2; TEB_CHECK(K % 32 == 0, "NVFP4 GEMM needs K % 32 == 0"); }
  TEB_CHECK(K > 0, "GEMM with K = 0 is not supported");

  Params &p = pl.p;
  p.R = static_cast<int>(R);
  p.C = static_cast<int>(Cc);
  p.K = static_cast<int>(K);
  p.kblocks = static_cast<int>(ceil_div(k_bytes, kBKBytes));
  auto operand = [](const Side &s) {
    Operand o{};
    o.mn_major = s.mn_major;
    if (s.mn_major) { o.lbo = 16384; o.sbo = 1024; o.kstep = 32 * 128; }
    else { o.lbo = 16; o.sbo = 1024; o.kstep = 32; }
    return o;
  };
  p.x = operand(x);
  p.y = operand(y);
  // CTA pairs (256 x 256 tiles) whenever the problem has more than one 128-row block; the
  // single-CTA kernel (128 x BN) serves small or skinny outputs.
  static const bool force_1cta = [] { const char *e = getenv("NVTE_B200_GEMM_1CTA"); return e && e[0] == '1'; }();
  // NVFP4 with K % 256 != 0: the last k-block's 4 scale atoms reach past the row block's own atoms. The
  // pair kernel fetches scales through a tensor map (next block's finite scales or zero fill, times
  // zero-filled data = 0); the single-CTA kernel's linear bulk copy would run off the buffer, so such
  // problems always take the pair kernel
  const bool ragged_nvfp4 = kind == 2 && (x.sf_cols / 4) % 4 != 0;
  bool pair = force_pair || ragged_nvfp4 || (R > 128 && Cc > 128 && !force_1cta);
  // Small problems: when the 256 x 256 pair tiles make at most two waves, the single-CTA kernel's 128 x 256 or
  // 128 x 128 tiles spread the work over all SMs in smaller units (config 2: 2048 x 3072 x 768 = 96 pair tiles on
  // 74 pairs = two waves of 256 x 256, but 384 tiles of 128 x 128 on 148 SMs). Cost model: waves x tile area /
  // relative efficiency of the tile shape (pair 1.0, 128 x 256 0.85, 128 x 128 0.7); NVTE_B200_GEMM_SMALL=0 disables.
  static const int one_bn = [] { const char *e = getenv("NVTE_B200_GEMM_1CTA_BN"); return e ? atoi(e) : 0; }();
  static const bool small_on = [] { const char *e = getenv("NVTE_B200_GEMM_SMALL"); return !(e && e[0] == '0'); }();
  int small_bn = 0;
  if (pair && !force_pair && !ragged_nvfp4 && small_on && cfg.arrival.world <= 1) {
    const size_t sms = static_cast<size_t>(cfg.sm_count > 0 ? std::min(cfg.sm_count, sm_count()) : sm_count());
    const size_t pt = ceil_div(R, 256) * ceil_div(Cc, 256);
    if (pt <= 2 * (sms / 2)) {
      const double c_pair = static_cast<double>(ceil_div(pt, sms / 2)) * 65536.0;
      const double c_256 = static_cast<double>(ceil_div(ceil_div(R, 128) * ceil_div(Cc, 256), sms)) * 32768.0 / 0.85;
      const double c_128 = static_cast<double>(ceil_div(ceil_div(R, 128) * ceil_div(Cc, 128), sms)) * 16384.0 / 0.70;
      if (c_256 < c_pair && c_256 <= c_128) small_bn = 256;
      else if (c_128 < c_pair) small_bn = 128;
    }
  }
  if (one_bn == 128 || one_bn == 256) small_bn = (force_1cta && !force_pair && !ragged_nvfp4) ? one_bn : small_bn;
  if (small_bn != 0) pair = false;
  int BN = small_bn != 0 ? small_bn : ((pair || Cc > 128) ? 256 : 128);
  if (pair && !force_pair && kind == 0 && !y.mn_major) {
    // per-tensor FP8, K-major weights: 224- or 192-column tiles when they save more than 5 % of wave x width
    static const int fixed = [] { const char *e = getenv("NVTE_B200_GEMM_BN"); return e ? atoi(e) : 0; }();
    const size_t pairs = static_cast<size_t>(std::max(1, (cfg.sm_count > 0 ? std::min(cfg.sm_count, sm_count()) : sm_count()) / 2));
    const size_t tr = ceil_div(R, 256);
    auto cost = [&](int bn) { return ceil_div(tr * ceil_div(Cc, static_cast<size_t>(bn)), pairs) * bn; };
    for (int bn : {224, 192}) {
      if (bn % 64 != 0 && cfg.scatter.world > 1) continue;   // the staged scatter epilogue moves 64-column pieces
      if (cost(bn) * 100 < cost(BN) * 95) BN = bn;
    }
    if (fixed == 256 || fixed == 224 || fixed == 192) BN = fixed;
  }
  const int BMt = pair ? 256 : kBM;
  const int fx = (kind == 2) ? 1 : (x.dtype == kNVTEFloat8E5M2 ? 1 : 0);
  const int fy = (kind == 2) ? 1 : (y.dtype == kNVTEFloat8E5M2 ? 1 : 0);
  p.idesc = kind == 0 ? sm100::idesc_f8(fx, fy, x.mn_major, y.mn_major, BMt, BN)
                      : sm100::idesc_block_scaled(fx, fy, x.mn_major, y.mn_major, BMt, BN,
                                                  /*ue8m0=*/kind == 1);
  p.sfx = x.sf;
  p.sfy = y.sf;
  p.sfx_ld = static_cast<int>(x.sf_cols / 4);
  p.sfy_ld = static_cast<int>(y.sf_cols / 4);
  p.s0 = x.scalar;
  p.s1 = y.scalar;
  p.alpha = alpha;
  p.beta = beta;
  // NVFP4: alpha * amax_x * amax_y / (6*6*448*448) (reference recipe/nvfp4.cu:72-78)
  p.scalar_const = kind == 2 ? 1.0f / (6.0f * 6.0f * 448.0f * 448.0f) : 1.0f;
  p.out = D.data.dptr;
  p.out_dtype = D.data.dtype;
  if (cfg.scatter.world > 1) {
    p.sc_rows = static_cast<int>(cfg.scatter.rows_per_rank);
    for (uint32_t q = 0; q < cfg.scatter.world; ++q) p.sc_out[q] = reinterpret_cast<void *>(cfg.scatter.dst[q]);
  }
  p.rb_size = rb[0] > 0 ? rb[0] : 0;
  p.rb_stride = rb[1];
  p.rb_off = rb[2];
  if (beta != 0.0f) {
    TEB_CHECK(C != nullptr && C->data.dptr != nullptr, "beta != 0 needs a C tensor");
    TEB_CHECK(C->data.dtype == D.data.dtype && C->data.numel() == D.data.numel(),
              "C must match D in dtype and shape");
    p.cin = C->data.dptr;
  }
  if (const Tensor *bias = cfg.bias_tensor ? lookup(cfg.bias_tensor) : nullptr) {
    if (bias->data.dptr != nullptr) {
      TEB_CHECK(bias->data.dtype == D.data.dtype, "bias dtype must match the output dtype");
      TEB_CHECK(bias->data.numel() == Cc, "bias must have ", Cc, " elements");
      p.bias = bias->data.dptr;
    }
  }
  p.debug = debug_counters();
  static const int dbg_flags = [] { const char *e = getenv("NVTE_B200_GEMM_DEBUG_FLAGS"); return e ? atoi(e) : 0; }();
  p.debug_flags = dbg_flags;
  p.tiles_r = static_cast<int>(ceil_div(R, BMt));
  p.tiles_c = static_cast<int>(ceil_div(Cc, BN));
  // super-rows of 16 tile rows walked column-first (tile_coords); NVTE_B200_GEMM_RASTER=<g> overrides (1 = row-major)
  static const int raster = [] { const char *e = getenv("NVTE_B200_GEMM_RASTER"); return e ? atoi(e) : 16; }();
  p.raster_g = pair ? raster : 1;
  p.bn = BN;
  p.sc_tiles = p.tiles_r;
  // 128-byte row pieces through shared memory (gemm2_kernel.cuh, stage_chunk16) instead of 32-byte fragments
  static const bool stage_on = [] { const char *e = getenv("NVTE_B200_GEMM_SCATTER_STAGED"); return !(e && e[0] == '0'); }();
  if (p.sc_rows > 0 && pair && kind == 0 && stage_on && (D.data.dtype == kNVTEBFloat16 || D.data.dtype == kNVTEFloat16) &&
      Cc % 64 == 0 && BN % 64 == 0 && cfg.scatter.rows_per_rank % 128 == 0 && beta == 0.0f)
    p.sc_staged = 1;
  if (p.sc_rows > 0 && pair && cfg.scatter.rows_per_rank % 256 == 0)
    p.sc_rot = static_cast<int>(((cfg.scatter.rank + 1) % cfg.scatter.world) * (cfg.scatter.rows_per_rank / 256));
  if (cfg.arrival.world > 1) {
    const NVTEB200ArrivalFlags &a = cfg.arrival;
    TEB_CHECK(pair, "Arrival flags need the CTA-pair GEMM (more than 128 output rows and columns)");
    TEB_CHECK(!x.mn_major && R == static_cast<size_t>(a.world) * a.flags_per_rank * a.rows_per_flag,
              "Arrival flags: operand B must be the K-major gather buffer of ", a.world, " x ", a.flags_per_rank,
              " x ", a.rows_per_flag, " rows, got ", R);
    p.arr_flags = reinterpret_cast<const uint32_t *>(a.flags);
    p.arr_step = reinterpret_cast<const uint32_t *>(a.step);
    p.arr_error = reinterpret_cast<uint32_t *>(a.error);
    p.arr_world = static_cast<int>(a.world);
    p.arr_rank = static_cast<int>(a.rank);
    p.arr_fpr = static_cast<int>(a.flags_per_rank);
    p.arr_tpf = static_cast<int>(a.rows_per_flag / 256);
    p.tile_counter = reinterpret_cast<unsigned int *>(a.tile_counter);
    if (a.push_unit_bytes != 0) {
      const size_t chunk_bytes = static_cast<size_t>(a.rows_per_flag) * k_bytes;
      TEB_CHECK((kind == 0 || (kind == 1 && a.push_mc_dst != 0)) && a.push_unit_bytes == Config2<0>::kPushUnit &&
                    chunk_bytes % a.push_unit_bytes == 0,
                "Arrival flags with push: per-tensor FP8 operands (MXFP8 with multicast addresses only), "
                "push_unit_bytes = ", Config2<0>::kPushUnit, " dividing rows_per_flag * K bytes");
      p.arr_push_unit = static_cast<int>(a.push_unit_bytes);
      p.arr_push_upc = static_cast<int>(chunk_bytes / a.push_unit_bytes);
      p.arr_push_src = static_cast<const uint8_t *>(x.data) + static_cast<size_t>(a.rank) * a.flags_per_rank * chunk_bytes;
      TEB_CHECK((a.push_mc_dst == 0) == (a.push_mc_flags == 0) && a.push_mc_dst % 16 == 0,
                "Arrival flags with push: multicast addresses come as a (16-byte aligned) pair");
      p.arr_push_mc_dst = reinterpret_cast<uint8_t *>(a.push_mc_dst);
      p.arr_push_mc_cnt = reinterpret_cast<uint32_t *>(a.push_mc_flags);
      if (a.push_src2 != 0) {
        TEB_CHECK(a.push_mc_dst != 0 && a.push_mc_dst2 != 0 && a.push_unit2_bytes != 0 && a.push_unit2_bytes % 4096 == 0 &&
                      a.push_bytes2_per_flag % a.push_unit2_bytes == 0 && a.push_src2 % 16 == 0 && a.push_mc_dst2 % 16 == 0,
                  "Arrival flags with push: the second region needs multicast addresses, a unit that is a multiple "
                  "of 4096 bytes and divides the bytes per flag, 16-byte aligned windows");
        p.arr_push_unit2 = static_cast<int>(a.push_unit2_bytes);
        p.arr_push_upc2 = static_cast<int>(a.push_bytes2_per_flag / a.push_unit2_bytes);
        p.arr_push_src2 = reinterpret_cast<const uint8_t *>(a.push_src2);
        p.arr_push_mc_dst2 = reinterpret_cast<uint8_t *>(a.push_mc_dst2);
      }
      TEB_CHECK(kind == 0 || a.push_src2 != 0, "Arrival flags with push: MXFP8 operands also push their block scales");
      if (a.advance_step != 0) {
        TEB_CHECK(a.tile_counter != 0 && (a.ready == 0) == (a.ready_mc == 0),
                  "Arrival flags: advance_step needs tile_counter, and ready / ready_mc come as a pair");
        p.arr_advance = 1;
        p.arr_ready = reinterpret_cast<const uint32_t *>(a.ready);
        p.arr_ready_mc = reinterpret_cast<uint32_t *>(a.ready_mc);
      }
      for (uint32_t r = 0; r < a.world && a.push_mc_dst == 0; ++r) {
        TEB_CHECK(r == a.rank || (a.push_dst[r] % 16 == 0 && a.push_dst[r] != 0 && a.push_flags[r] != 0),
                  "Arrival flags with push: peer ", r, " has no (16-byte aligned) destination / flag address");
        p.arr_push_dst[r] = reinterpret_cast<uint8_t *>(a.push_dst[r]);
        p.arr_push_cnt[r] = reinterpret_cast<uint32_t *>(a.push_flags[r]);
      }
    }
    // one super-row = one rank's tile rows: the first tiles of the walk are all local
    p.raster_g = std::min(16, p.arr_fpr * p.arr_tpf);
  }
  static const bool force_dyn = [] { const char *e = getenv("NVTE_B200_GEMM_DYNAMIC"); return e && e[0] == '1'; }();
  if (force_dyn && pair && kind == 0 && p.tile_counter == nullptr && !force_pair) {
    static unsigned int *counter = [] {
      unsigned int *d = nullptr;
      if (cudaMalloc(&d, 2 * sizeof(unsigned int)) != cudaSuccess || cudaMemset(d, 0, 2 * sizeof(unsigned int)) != cudaSuccess)
        d = nullptr;
      return d;
    }();
    p.tile_counter = counter;
  }
  pl.x = x;
  pl.y = y;
  pl.kind = kind;
  pl.k_bytes = k_bytes;
  pl.pair = pair;
  pl.BN = BN;
  return pl;
}

// Tensor maps of a CTA-pair launch: operands (128-row boxes) and scale atoms.
struct PairMaps {
  CUtensorMap mx, my, msx, msy;
};
CUtensorMap operand_map(const Side &s, size_t k_bytes, uint32_t tile_rows) {
  // K-major: [MN, Kbytes], box [rows, 128]; MN-major (FP8 only): [K, MN], box [128 k-rows, 128 mn-bytes]
  if (s.mn_major) return make_map(s.data, s.mn, s.k, 128, 128);
  return make_map(s.data, k_bytes, s.mn, kBKBytes, tile_rows);
}
PairMaps pair_maps(const Plan &pl) {
  PairMaps m;
  m.mx = operand_map(pl.x, pl.k_bytes, 128);
  m.my = operand_map(pl.y, pl.k_bytes, pl.pair ? pl.BN / 2 : 128);
  m.msx = m.mx;
  m.msy = m.my;
  if (pl.kind != 0) {
    const uint32_t atoms = pl.kind == 2 ? 4 : 1;
    m.msx = make_map(pl.x.sf, 128, round_up(pl.x.mn, 128) / 128 * (pl.x.sf_cols / 4), 128, atoms, 1);
    m.msy = make_map(pl.y.sf, 128, round_up(pl.y.mn, 128) / 128 * (pl.y.sf_cols / 4), 128, atoms, 1);
  }
  return m;
}
}  // namespace

void run(int transa, int transb, float alpha, const Tensor &A, const Tensor &B, float beta,
         const Tensor *C, Tensor &D, Tensor *workspace, const MatmulConfig &cfg,
         cudaStream_t stream) {
  (void)workspace;
  require_blackwell("The block-scaled GEMM");
  const Plan pl = make_plan(transa, transb, alpha, A, B, beta, C, D, cfg, false);
  if (pl.empty) return;
  const Params &p = pl.p;
  const Side &x = pl.x, &y = pl.y;
  const int kind = pl.kind, BN = pl.BN;
  const size_t k_bytes = pl.k_bytes;
  const bool pair = pl.pair;

  // Tensor maps over the raw bytes. K-major: [MN, Kbytes], box [rows, 128];
  // MN-major (FP8 only): [K, MN], box [128 k-rows, 128 mn-bytes].
  auto map_of = [&](const Side &s, uint32_t tile_rows) {
    if (s.mn_major) return make_map(s.data, s.mn, s.k, 128, 128);
    return make_map(s.data, k_bytes, s.mn, kBKBytes, tile_rows);
  };
  int sms = sm_count();
  if (cfg.sm_count > 0 && cfg.sm_count < sms) sms = cfg.sm_count;
  if (pair) {
    // Opt-in experiment (NVTE_B200_GEMM_MULTICAST=1): clusters of two pairs sharing their X tile
    // by TMA multicast + plain pairs on the SMs no cluster of 4 can use. Measured this round
    // (profiles/r01_gemm_analysis.md): a multicast pair moves ~12 % more per SM, but only 33
    // clusters of 4 fit (132 of 148 SMs) and the coarser wave quantisation eats the rest:
    // +5 % on wgrad, -7 % on dgrad, 0 on fprop. Off by default.
    static const bool use_mc = [] { const char *e = getenv("NVTE_B200_GEMM_MULTICAST"); return e && e[0] == '1'; }();
    const bool mc = use_mc && p.tiles_c % 2 == 0 && p.tiles_r * p.tiles_c >= 2 * sms;
    const PairMaps pm = pair_maps(pl);
    const CUtensorMap &mx = pm.mx, &my = pm.my, &msx = pm.msx, &msy = pm.msy;
    if (mc) {
      const CUtensorMap mx64 = x.mn_major ? make_map(x.data, x.mn, x.k, 128, 64) : map_of(x, 64);
      if (kind == 0) launch_split<0>(mx64, mx, my, msx, msy, p, sms, stream);
      else if (kind == 1) launch_split<1>(mx64, mx, my, msx, msy, p, sms, stream);
      else launch_split<2>(mx64, mx, my, msx, msy, p, sms, stream);
    } else {
      // dynamic tile queue (per-tensor FP8 pairs): with arrival flags when the caller gave a counter, or for every
      // FP8 pair GEMM with NVTE_B200_GEMM_DYNAMIC=1 (test aid: one library-owned counter, one GEMM at a time)
      if (kind == 0 && p.tile_counter != nullptr) launch2<0, 2, true>(mx, my, msx, msy, p, sms / 2, stream);
      else if (kind == 1 && p.tile_counter != nullptr) launch2<1, 2, true>(mx, my, msx, msy, p, sms / 2, stream);
      else if (kind == 0) launch2<0, 2>(mx, my, msx, msy, p, sms / 2, stream);
      else if (kind == 1) launch2<1, 2>(mx, my, msx, msy, p, sms / 2, stream);
      else launch2<2, 2>(mx, my, msx, msy, p, sms / 2, stream);
    }
    return;
  }
  const CUtensorMap mx = map_of(x, kBM);
  const CUtensorMap my = map_of(y, BN);
#define TEB_GEMM_LAUNCH(BN_, KIND_) launch<BN_, KIND_>(mx, my, p, sms, stream)
  if (BN == 256) {
    if (kind == 0) TEB_GEMM_LAUNCH(256, 0); else if (kind == 1) TEB_GEMM_LAUNCH(256, 1); else TEB_GEMM_LAUNCH(256, 2);
  } else {
    if (kind == 0) TEB_GEMM_LAUNCH(128, 0); else if (kind == 1) TEB_GEMM_LAUNCH(128, 1); else TEB_GEMM_LAUNCH(128, 2);
  }
#undef TEB_GEMM_LAUNCH
}

// nvte_multi_tensor_gemm (reference gemm/cublaslt_gemm.cu:970-1030 loops cuBLASLt calls over its
// pool of compute streams): here the members share ONE launch. Each member is planned like a single
// GEMM (same checks, same operand selection); its tensor maps and parameters go into a table in the
// caller's workspace and the persistent CTA pairs walk the concatenated tile list. Members with no
// output rows / columns are skipped. Returns false when the list cannot share a launch (mixed
// scaling modes, no usable workspace); the caller then runs the members one by one.
namespace {
// Pinned, never-modified copy of a member table, keyed by its bytes (FNV-1a + full compare).
const GroupEntry *pinned_table(const std::vector<GroupEntry> &table) {
  static std::mutex mu;
  static std::unordered_map<uint64_t, std::vector<std::pair<size_t, GroupEntry *>>> cache;
  static size_t cached_bytes = 0;
  const size_t bytes = table.size() * sizeof(GroupEntry);
  const uint8_t *src = reinterpret_cast<const uint8_t *>(table.data());
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < bytes; ++i) h = (h ^ src[i]) * 1099511628211ull;
  std::lock_guard<std::mutex> lock(mu);
  for (auto &e : cache[h])
    if (e.first == bytes && std::memcmp(e.second, src, bytes) == 0) return e.second;
  TEB_CHECK(cached_bytes + bytes <= (256u << 20), "Too many distinct grouped-GEMM member tables are alive");
  GroupEntry *p = nullptr;
  // called under stream capture: a pinned allocation is not a stream operation, but the default
  // (global) capture mode forbids it unless this thread says it knows what it is doing
  cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
  TEB_CHECK_CUDA(cudaThreadExchangeStreamCaptureMode(&mode));
  const cudaError_t err = cudaMallocHost(reinterpret_cast<void **>(&p), bytes);
  TEB_CHECK_CUDA(cudaThreadExchangeStreamCaptureMode(&mode));
  TEB_CHECK_CUDA(err);
  std::memcpy(p, src, bytes);
  cache[h].emplace_back(bytes, p);
  cached_bytes += bytes;
  return p;
}
}  // namespace

bool run_grouped(int transa, int transb, const std::vector<GroupedMember> &members, Tensor *workspace,
                 const std::vector<MatmulConfig> &cfgs, cudaStream_t stream) {
  require_blackwell("The block-scaled GEMM");
  static const bool off = [] { const char *e = getenv("NVTE_B200_GROUPED_GEMM"); return e && e[0] == '0'; }();
  if (off || members.size() < 2) return false;
  if (workspace == nullptr || workspace->data.dptr == nullptr) return false;
  const size_t need = members.size() * sizeof(GroupEntry) + 128;
  if (workspace->data.bytes() < need) return false;
  const int mode = members[0].A->scaling_mode;
  for (const GroupedMember &m : members)
    if (m.A->scaling_mode != mode || m.B->scaling_mode != mode) return false;

  std::vector<GroupEntry> table;
  table.reserve(members.size());
  int tiles = 0, kind = 0;
  for (size_t i = 0; i < members.size(); ++i) {
    const GroupedMember &m = members[i];
    const Plan pl = make_plan(transa, transb, m.alpha, *m.A, *m.B, m.beta, m.D, *m.D, cfgs[i], true);
    if (pl.empty) continue;
    const PairMaps pm = pair_maps(pl);
    GroupEntry e;
    std::memset(static_cast<void *>(&e), 0, sizeof(e));   // padding too: the table is hashed by bytes
    e.mx = pm.mx; e.my = pm.my; e.msx = pm.msx; e.msy = pm.msy;
    e.p = pl.p;
    tiles += pl.p.tiles_r * pl.p.tiles_c;
    e.tile_end = tiles;
    kind = pl.kind;
    table.push_back(e);
  }
  if (table.empty()) return true;
  uintptr_t base = (reinterpret_cast<uintptr_t>(workspace->data.dptr) + 127) & ~uintptr_t(127);
  GroupEntry *dev = reinterpret_cast<GroupEntry *>(base);
  // Eager: the source is pageable, the runtime stages the bytes before returning, so `table` may die
  // after this. Under stream capture the copy node would read the host bytes at every replay, so the
  // source is an immutable pinned copy owned by a content-addressed cache (a CUDA graph of a
  // GroupedLinear step stays valid for as long as the process lives).
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  TEB_CHECK_CUDA(cudaStreamIsCapturing(stream, &cap));
  const GroupEntry *src = cap == cudaStreamCaptureStatusNone ? table.data() : pinned_table(table);
  TEB_CHECK_CUDA(cudaMemcpyAsync(dev, src, table.size() * sizeof(GroupEntry), cudaMemcpyHostToDevice, stream));
  int sms = sm_count();
  if (cfgs[0].sm_count > 0 && cfgs[0].sm_count < sms) sms = cfgs[0].sm_count;
  Params p0 = table[0].p;
  p0.debug = debug_counters();
  const int n = static_cast<int>(table.size());
  if (kind == 0) launch_grouped<0>(dev, n, tiles, p0, table[0].mx, sms / 2, stream);
  else if (kind == 1) launch_grouped<1>(dev, n, tiles, p0, table[0].mx, sms / 2, stream);
  else launch_grouped<2>(dev, n, tiles, p0, table[0].mx, sms / 2, stream);
  return true;
}

}  // namespace gemm
}  // namespace teb

// bytes of workspace[0] that nvte_multi_tensor_gemm needs per member for the one-launch path (+ 128 for alignment)
extern "C" size_t nvte_b200_group_entry_bytes() { return sizeof(teb::gemm::GroupEntry); }
extern "C" unsigned long long *nvte_b200_gemm_debug_counters() {
  return teb::gemm::debug_counters();
}
