// COMPARATOR HARNESS. One binary that dlopen()s a `libtransformer_engine`-ABI library -- the product
// (transformerengine_b200/libtransformer_engine.so) or the reference's own sm_100a kernels built file by
// file into baseline/_ref/libte_refkernels.so -- and times the hot-path entry points through the SAME
// nvte_* C ABI calls with the same buffers, the same rotating-buffer loop (> 2x L2 of distinct buffer
// sets for the HBM-bound kernels) and CUDA events on the launching stream. It is also a drop-in proof:
// the same call sequence runs against either library.
//
//   kernel_bench <libA.so>[=name] [<libB.so>[=name] ...] [--iters N] [--rounds R] [--only substring]
//     -> one JSON object per line on stdout. With several libraries every case is timed INTERLEAVED
//     (A x iters, B x iters, A x iters, ...) in one process, so both see the same clock / power state of
//     the part -- the block-scaled GEMMs run into the 1000 W cap and a library timed after a long GEMM
//     loop of another process starts from different clocks.
//
// Kernels timed (SURVEY 2.5 / BASELINE.md 3.2 names): K8 MXFP8 quantize row+col (C1, C4 x / W / dy), K3
// per-tensor FP8 cast, K1 amax, K12 nvte_hadamard_transform_amax (C3), K14
// nvte_quantize_with_hadamard_transform (C3), K10 NVFP4 quantize+transpose without RHT, and
// nvte_cublas_gemm_v2 (the reference: cuBLASLt; ours: the tcgen05 kernel) for C4 / C3 / C5.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <array>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/transformerengine_b200.h"

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess) {                                                               \
      std::fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      std::exit(2);                                                                        \
    }                                                                                      \
  } while (0)

namespace {

struct Api {
  void *h = nullptr;
  decltype(&nvte_create_tensor) create_tensor;
  decltype(&nvte_destroy_tensor) destroy_tensor;
  decltype(&nvte_set_tensor_param_v2) set_param;
  decltype(&nvte_create_quantization_config) create_qcfg;
  decltype(&nvte_destroy_quantization_config) destroy_qcfg;
  decltype(&nvte_quantize_v2) quantize_v2;
  decltype(&nvte_multi_tensor_quantize) multi_quantize;
  decltype(&nvte_compute_amax_with_config) compute_amax;
  decltype(&nvte_hadamard_transform_amax) rht_amax;
  decltype(&nvte_quantize_with_hadamard_transform) rht_quantize;
  decltype(&nvte_create_matmul_config) create_mcfg;
  decltype(&nvte_cublas_gemm_v2) gemm_v2;
  decltype(&nvte_swizzle_scaling_factors) swizzle;
  template <typename F>
  void load(F &f, const char *name) {
    f = reinterpret_cast<F>(dlsym(h, name));
    if (f == nullptr) {
      std::fprintf(stderr, "missing symbol %s\n", name);
      std::exit(2);
    }
  }
  explicit Api(const char *path) {
    h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (h == nullptr) {
      std::fprintf(stderr, "dlopen(%s): %s\n", path, dlerror());
      std::exit(2);
    }
    load(create_tensor, "nvte_create_tensor");
    load(destroy_tensor, "nvte_destroy_tensor");
    load(set_param, "nvte_set_tensor_param_v2");
    load(create_qcfg, "nvte_create_quantization_config");
    load(destroy_qcfg, "nvte_destroy_quantization_config");
    load(quantize_v2, "nvte_quantize_v2");
    load(multi_quantize, "nvte_multi_tensor_quantize");
    load(compute_amax, "nvte_compute_amax_with_config");
    load(rht_amax, "nvte_hadamard_transform_amax");
    load(rht_quantize, "nvte_quantize_with_hadamard_transform");
    load(create_mcfg, "nvte_create_matmul_config");
    load(gemm_v2, "nvte_cublas_gemm_v2");
    load(swizzle, "nvte_swizzle_scaling_factors");
  }
};

size_t ru(size_t a, size_t b) { return (a + b - 1) / b * b; }

std::vector<void *> g_allocs;
void *dalloc(size_t bytes, int fill = -1) {
  void *p = nullptr;
  CK(cudaMalloc(&p, bytes ? bytes : 16));
  if (fill >= 0) CK(cudaMemset(p, fill, bytes));
  g_allocs.push_back(p);
  return p;
}
void free_all() {
  for (void *p : g_allocs) cudaFree(p);
  g_allocs.clear();
}

__global__ void fill_bf16_kernel(__nv_bfloat16 *p, size_t n, uint32_t seed, float scale) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    // cheap hash -> roughly normal (sum of 4 uniforms), good enough for a timing input
    uint32_t x = static_cast<uint32_t>(i) * 2654435761u ^ seed;
    float acc = 0.f;
    for (int j = 0; j < 4; ++j) {
      x ^= x << 13; x ^= x >> 17; x ^= x << 5;
      acc += (x & 0xffff) * (1.0f / 65536.0f) - 0.5f;
    }
    p[i] = __float2bfloat16(acc * 1.7320508f * scale);
  }
}
__nv_bfloat16 *alloc_bf16(size_t n, uint32_t seed, float scale = 1.0f) {
  auto *p = static_cast<__nv_bfloat16 *>(dalloc(n * 2));
  fill_bf16_kernel<<<1184, 256>>>(p, n, seed, scale);
  CK(cudaGetLastError());
  return p;
}

struct T {   // RAII-less handle helper (handles are released at process exit)
  const Api &api;
  NVTETensor t;
  T(const Api &a, NVTEScalingMode mode) : api(a), t(a.create_tensor(mode)) {}
  T &set(NVTETensorParam p, void *ptr, NVTEDType dt, std::vector<size_t> shape) {
    NVTEBasicTensor b{};
    b.data_ptr = ptr;
    b.dtype = dt;
    b.shape.ndim = shape.size();
    for (size_t i = 0; i < shape.size(); ++i) b.shape.data[i] = shape[i];
    api.set_param(t, p, &b, sizeof(b));
    return *this;
  }
  T &flag(NVTETensorParam p, bool v) {
    uint8_t b = v;
    api.set_param(t, p, &b, 1);
    return *this;
  }
};

struct Opts {
  int iters = 20, rounds = 5;
  std::string only;
};
Opts g_opts;
cudaStream_t g_stream;

// One timed case of one library: fn(set) launches the work on g_stream; sums() (optional) returns named
// checksums of the outputs of buffer set 0 after a run.
struct Case {
  std::string name, kernel;
  double work = 0;          // bytes, or flop / 1000 (rate = work / seconds / 1e9 -> GB/s or TFLOP/s)
  const char *unit = "GB/s";
  int nsets = 1;
  std::function<void(int)> fn;
  std::function<std::vector<std::pair<std::string, unsigned long long>>()> sums;
  std::string error;        // set when building the case threw
};

bool selected(const std::string &name) {
  return g_opts.only.empty() || name.find(g_opts.only) != std::string::npos;
}

std::string clean(std::string msg) {
  for (char &c : msg)
    if (c == '"' || c == '\n' || c == '\\') c = ' ';
  return msg.substr(0, 300);
}

// Times the same case of every library interleaved: per round, each library runs `iters` launches
// between two events; mean and best over the rounds.
void time_cases(std::vector<Case *> cs, const std::vector<std::string> &libs) {
  const int n = static_cast<int>(cs.size());
  std::vector<double> total(n, 0.0), best(n, 1e30);
  std::vector<std::string> err(n);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int l = 0; l < n; ++l) {
    err[l] = cs[l]->error;
    if (!err[l].empty()) continue;
    try {
      for (int i = 0; i < 3; ++i) cs[l]->fn(i % cs[l]->nsets);
      CK(cudaStreamSynchronize(g_stream));
    } catch (const std::exception &e) {
      err[l] = clean(e.what());
      cudaGetLastError();
    }
  }
  for (int r = 0; r < g_opts.rounds; ++r)
    for (int l = 0; l < n; ++l) {
      if (!err[l].empty()) continue;
      try {
        CK(cudaEventRecord(e0, g_stream));
        for (int i = 0; i < g_opts.iters; ++i) cs[l]->fn(i % cs[l]->nsets);
        CK(cudaEventRecord(e1, g_stream));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ms /= g_opts.iters;
        total[l] += ms;
        best[l] = std::fmin(best[l], static_cast<double>(ms));
      } catch (const std::exception &e) {
        err[l] = clean(e.what());
        cudaGetLastError();
      }
    }
  for (int l = 0; l < n; ++l) {
    const Case &c = *cs[l];
    if (!err[l].empty()) {
      std::printf("{\"lib\": \"%s\", \"case\": \"%s\", \"kernel\": \"%s\", \"error\": \"%s\"}\n", libs[l].c_str(),
                  c.name.c_str(), c.kernel.c_str(), err[l].c_str());
      continue;
    }
    const double mean = total[l] / g_opts.rounds;
    std::printf("{\"lib\": \"%s\", \"case\": \"%s\", \"kernel\": \"%s\", \"us\": %.2f, \"us_best\": %.2f, "
                "\"rate\": %.1f, \"unit\": \"%s\", \"sets\": %d, \"iters\": %d, \"rounds\": %d, \"interleaved_with\": %d}\n",
                libs[l].c_str(), c.name.c_str(), c.kernel.c_str(), mean * 1e3, best[l] * 1e3, c.work / (mean * 1e-3) / 1e9,
                c.unit, c.nsets, g_opts.iters, g_opts.rounds, n - 1);
    if (c.sums) {
      std::printf("{\"lib\": \"%s\", \"case\": \"%s\", \"checksums\": {", libs[l].c_str(), c.name.c_str());
      bool first = true;
      for (auto &kv : c.sums()) {
        std::printf("%s\"%s\": \"%016llx\"", first ? "" : ", ", kv.first.c_str(), kv.second);
        first = false;
      }
      std::printf("}}\n");
    }
  }
  std::fflush(stdout);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
}

int sets_for(size_t bytes_per_set) {   // distinct buffer sets so that a pass never hits L2-resident data
  const size_t want = 320ull << 20;    // > 2x the 126 MB L2
  size_t n = (want + bytes_per_set - 1) / bytes_per_set;
  return static_cast<int>(n < 2 ? 2 : (n > 12 ? 12 : n));
}

// position-dependent 64-bit checksum of a device buffer: equal checksums from the two libraries on the same
// input = bit-identical output bytes (parity of ours with the reference's GPU kernels on this box)
__global__ void checksum_kernel(const uint8_t *p, size_t n, unsigned long long *out) {
  unsigned long long acc = 0;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    acc += (static_cast<unsigned long long>(p[i]) + 1ull) * ((i * 0x9E3779B97F4A7C15ull) >> 17 | 1ull);
  atomicAdd(out, acc);
}
unsigned long long checksum(const void *p, size_t n) {
  static unsigned long long *d = nullptr;
  if (d == nullptr) CK(cudaMalloc(&d, 8));
  CK(cudaMemsetAsync(d, 0, 8, g_stream));
  checksum_kernel<<<592, 256, 0, g_stream>>>(static_cast<const uint8_t *>(p), n, d);
  unsigned long long h = 0;
  CK(cudaMemcpyAsync(&h, d, 8, cudaMemcpyDeviceToHost, g_stream));
  CK(cudaStreamSynchronize(g_stream));
  return h;
}
NVTETensor plain(const Api &api, void *p, NVTEDType dt, std::vector<size_t> shape) {
  return T(api, NVTE_DELAYED_TENSOR_SCALING).set(kNVTERowwiseData, p, dt, shape).t;
}

// ---------------------------------------------------------------- MXFP8 quantize (K8)
struct MxOut {
  uint8_t *row, *col, *sr, *sc;
  size_t sr_bytes, sc_bytes;
  NVTETensor t;
};
MxOut mx_out(const Api &api, size_t rows, size_t cols, bool rowwise, bool colwise, bool swizzled) {
  MxOut o{};
  const size_t srr = ru(rows, 128), src = ru(cols / 32, 4), scr = ru(rows / 32, 4), scc = ru(cols, 128);
  o.sr_bytes = srr * src;
  o.sc_bytes = scr * scc;
  T t(api, NVTE_MXFP8_1D_SCALING);
  if (rowwise) {
    o.row = static_cast<uint8_t *>(dalloc(rows * cols, 0));
    o.sr = static_cast<uint8_t *>(dalloc(o.sr_bytes, 0));
    t.set(kNVTERowwiseData, o.row, kNVTEFloat8E4M3, {rows, cols});
    t.set(kNVTERowwiseScaleInv, o.sr, kNVTEFloat8E8M0, {srr, src});
  }
  if (colwise) {
    o.col = static_cast<uint8_t *>(dalloc(rows * cols, 0));
    o.sc = static_cast<uint8_t *>(dalloc(o.sc_bytes, 0));
    t.set(kNVTEColumnwiseData, o.col, kNVTEFloat8E4M3, {rows, cols});
    t.set(kNVTEColumnwiseScaleInv, o.sc, kNVTEFloat8E8M0, {scr, scc});
  }
  t.flag(kNVTEWithGEMMSwizzledScales, swizzled);
  o.t = t.t;
  return o;
}

using Sums = std::vector<std::pair<std::string, unsigned long long>>;

std::vector<Case> bench_mxfp8_quantize(const Api &api, const char *cfg, size_t rows, size_t cols, bool swizzled) {
  Case c;
  c.name = std::string("mxfp8_quantize_") + cfg + (swizzled ? "_swizzled" : "_compact");
  if (!selected(c.name)) return {};
  c.kernel = "K8 mxfp8 row+col";
  c.work = rows * cols * 4.0 + 2.0 * rows * cols / 32;
  c.nsets = sets_for(static_cast<size_t>(c.work));
  std::vector<NVTETensor> in;
  std::vector<MxOut> out;
  for (int s = 0; s < c.nsets; ++s) {
    in.push_back(plain(api, alloc_bf16(rows * cols, 17 + s), kNVTEBFloat16, {rows, cols}));
    out.push_back(mx_out(api, rows, cols, true, true, swizzled));
  }
  NVTEQuantizationConfig qc = api.create_qcfg();
  const Api *a = &api;
  c.fn = [=](int s) { a->quantize_v2(in[s], out[s].t, qc, g_stream); };
  c.sums = [=]() {
    return Sums{{"row", checksum(out[0].row, rows * cols)}, {"col", checksum(out[0].col, rows * cols)},
                {"row_scale", checksum(out[0].sr, out[0].sr_bytes)}, {"col_scale", checksum(out[0].sc, out[0].sc_bytes)}};
  };
  return {c};
}

// The bench's headline step: x, W and dy of the config-4 Linear, (a) one nvte_quantize_v2 per tensor and
// (b) the list through nvte_multi_tensor_quantize (reference: casts spread over its compute streams,
// cast/cast.cu:57-95; ours: one persistent launch over the concatenated tile list).
std::vector<Case> bench_mxfp8_step(const Api &api, bool swizzled) {
  const std::string base = std::string("mxfp8_quantize_c4_step_") + (swizzled ? "swizzled" : "compact");
  if (!selected(base)) return {};
  const std::vector<std::array<size_t, 2>> shp = {{8192, 4096}, {14336, 4096}, {8192, 14336}};
  double bytes = 0;
  for (auto &s : shp) bytes += s[0] * s[1] * 4.0 + 2.0 * s[0] * s[1] / 32;
  const int nsets = 2;
  std::vector<std::vector<NVTETensor>> in(nsets), out(nsets);
  std::vector<MxOut> first;
  for (int s = 0; s < nsets; ++s)
    for (int j = 0; j < 3; ++j) {
      in[s].push_back(plain(api, alloc_bf16(shp[j][0] * shp[j][1], 77 + 3 * s + j), kNVTEBFloat16, {shp[j][0], shp[j][1]}));
      MxOut o = mx_out(api, shp[j][0], shp[j][1], true, true, swizzled);
      if (s == 0) first.push_back(o);
      out[s].push_back(o.t);
    }
  NVTEQuantizationConfig qc = api.create_qcfg();
  const Api *a = &api;
  auto all_sums = [=]() {
    unsigned long long h = 0;
    for (int j = 0; j < 3; ++j)
      h += checksum(first[j].row, shp[j][0] * shp[j][1]) + checksum(first[j].col, shp[j][0] * shp[j][1]) +
           checksum(first[j].sr, first[j].sr_bytes) + checksum(first[j].sc, first[j].sc_bytes);
    return Sums{{"all_outputs", h}};
  };
  Case c3, c1;
  c3.name = base + "_3_calls";
  c3.kernel = "K8 x3 (nvte_quantize_v2 each)";
  c1.name = base + "_multi_tensor";
  c1.kernel = "nvte_multi_tensor_quantize(x, W, dy)";
  c3.work = c1.work = bytes;
  c3.nsets = c1.nsets = nsets;
  c3.fn = [=](int s) { for (int j = 0; j < 3; ++j) a->quantize_v2(in[s][j], out[s][j], qc, g_stream); };
  c1.fn = [=](int s) {
    std::vector<NVTETensor> i3 = in[s], o3 = out[s];
    a->multi_quantize(i3.data(), o3.data(), qc, 3, g_stream);
  };
  c3.sums = c1.sums = all_sums;
  return {c3, c1};
}

// ---------------------------------------------------------------- per-tensor FP8 cast + amax (K3, K1)
std::vector<Case> bench_fp8(const Api &api, const char *cfg, size_t rows, size_t cols) {
  const std::string base = std::string("fp8_") + cfg, aname = std::string("amax_") + cfg;
  if (!selected(base) && !selected(aname)) return {};
  const int nsets = sets_for(rows * cols * 3);
  std::vector<NVTETensor> in, out, aout;
  std::vector<uint8_t *> data;
  auto *scal = static_cast<float *>(dalloc(16 * sizeof(float), 0));
  const float three = 3.0f;
  CK(cudaMemcpy(scal, &three, 4, cudaMemcpyHostToDevice));   // scale = 3
  for (int s = 0; s < nsets; ++s) {
    in.push_back(plain(api, alloc_bf16(rows * cols, 29 + s), kNVTEBFloat16, {rows, cols}));
    data.push_back(static_cast<uint8_t *>(dalloc(rows * cols, 0)));
    out.push_back(T(api, NVTE_DELAYED_TENSOR_SCALING)
                      .set(kNVTERowwiseData, data[s], kNVTEFloat8E4M3, {rows, cols})
                      .set(kNVTEScale, scal, kNVTEFloat32, {1})
                      .set(kNVTEAmax, scal + 1, kNVTEFloat32, {1})
                      .set(kNVTERowwiseScaleInv, scal + 2, kNVTEFloat32, {1}).t);
    aout.push_back(T(api, NVTE_DELAYED_TENSOR_SCALING)
                       .set(kNVTERowwiseData, data[s], kNVTEFloat8E4M3, {rows, cols})
                       .set(kNVTEScale, scal, kNVTEFloat32, {1})
                       .set(kNVTERowwiseScaleInv, scal + 5, kNVTEFloat32, {1})
                       .set(kNVTEAmax, scal + 4, kNVTEFloat32, {1}).t);
  }
  NVTEQuantizationConfig qc = api.create_qcfg();
  const Api *a = &api;
  Case cc, ca;
  cc.name = base + "_cast";
  cc.kernel = "K3 fp8 cast (delayed: +amax)";
  cc.work = rows * cols * 3.0;
  ca.name = aname;
  ca.kernel = "K1 amax";
  ca.work = rows * cols * 2.0;
  cc.nsets = ca.nsets = nsets;
  cc.fn = [=](int s) { a->quantize_v2(in[s], out[s], qc, g_stream); };
  uint8_t *d0 = data[0];
  cc.sums = [=]() { return Sums{{"data", checksum(d0, rows * cols)}, {"amax_scale_inv", checksum(scal + 1, 8)}}; };
  ca.fn = [=](int s) { a->compute_amax(in[s], aout[s], qc, g_stream); };
  ca.sums = [=]() { return Sums{{"amax", checksum(scal + 4, 4)}}; };
  std::vector<Case> r;
  if (selected(cc.name)) r.push_back(cc);
  if (selected(ca.name)) r.push_back(ca);
  return r;
}

// ---------------------------------------------------------------- NVFP4 (K12 amax, K14 RHT cast, K10 plain)
struct Fp4Out {
  uint8_t *row = nullptr, *col = nullptr, *sr = nullptr, *sc = nullptr;
  float *amax;   // [0] rowwise, [1] columnwise
  size_t sr_bytes, sc_bytes;
  NVTETensor t;
};
Fp4Out fp4_out(const Api &api, size_t rows, size_t cols, bool rowwise, bool colwise, bool swizzled) {
  Fp4Out o{};
  const size_t srr = ru(rows, 128), src = ru(cols / 16, 4), scr = ru(cols, 128), scc = ru(rows / 16, 4);
  o.sr_bytes = srr * src;
  o.sc_bytes = scr * scc;
  o.amax = static_cast<float *>(dalloc(16, 0));
  T t(api, NVTE_NVFP4_1D_SCALING);
  if (rowwise) {
    o.row = static_cast<uint8_t *>(dalloc(rows * cols / 2, 0));
    o.sr = static_cast<uint8_t *>(dalloc(o.sr_bytes, 0));
    t.set(kNVTERowwiseData, o.row, kNVTEFloat4E2M1, {rows, cols});
    t.set(kNVTERowwiseScaleInv, o.sr, kNVTEFloat8E4M3, {srr, src});
    t.set(kNVTEAmax, o.amax, kNVTEFloat32, {1});
  }
  if (colwise) {
    o.col = static_cast<uint8_t *>(dalloc(rows * cols / 2, 0));
    o.sc = static_cast<uint8_t *>(dalloc(o.sc_bytes, 0));
    t.set(kNVTEColumnwiseData, o.col, kNVTEFloat4E2M1, {cols, rows});
    t.set(kNVTEColumnwiseScaleInv, o.sc, kNVTEFloat8E4M3, {scr, scc});
    t.set(kNVTEColumnwiseAmax, o.amax + 1, kNVTEFloat32, {1});
  }
  t.flag(kNVTEWithGEMMSwizzledScales, swizzled);
  o.t = t.t;
  return o;
}

__nv_bfloat16 *rht_matrix_device() {   // diag(signs) . H16 (Sylvester) . 0.25, bf16 (reference nvfp4_tensor.py:48-110)
  static const int signs[16] = {1, 1, 1, -1, 1, -1, -1, -1, -1, -1, -1, 1, -1, 1, -1, -1};
  __nv_bfloat16 h[256];
  for (int i = 0; i < 16; ++i)
    for (int j = 0; j < 16; ++j)
      h[i * 16 + j] = __float2bfloat16(((__builtin_popcount(i & j) & 1) ? -0.25f : 0.25f) * signs[i]);
  auto *d = static_cast<__nv_bfloat16 *>(dalloc(512));
  CK(cudaMemcpy(d, h, 512, cudaMemcpyHostToDevice));
  return d;
}

std::vector<Case> bench_nvfp4(const Api &api, const char *cfg, size_t rows, size_t cols) {
  const std::string base = std::string("nvfp4_") + cfg;
  if (!selected(base)) return {};
  const double cast_bytes = rows * cols * 3.0 + 2.0 * rows * cols / 16;
  const int nsets = sets_for(static_cast<size_t>(rows * cols * 2));
  std::vector<NVTETensor> in, amax_only;
  std::vector<Fp4Out> out, plain_out;
  for (int s = 0; s < nsets; ++s) {
    in.push_back(plain(api, alloc_bf16(rows * cols, 41 + s), kNVTEBFloat16, {rows, cols}));
    out.push_back(fp4_out(api, rows, cols, true, true, true));
    amax_only.push_back(T(api, NVTE_NVFP4_1D_SCALING).set(kNVTEAmax, out[s].amax, kNVTEFloat32, {1})
                            .set(kNVTEColumnwiseAmax, out[s].amax + 1, kNVTEFloat32, {1}).t);
    // K10: plain NVFP4 quantize + transposed copy (weights), compact scales, amax given
    plain_out.push_back(fp4_out(api, rows, cols, true, true, false));
    const float am[2] = {3.5f, 3.5f};
    CK(cudaMemcpy(plain_out[s].amax, am, 8, cudaMemcpyHostToDevice));
  }
  NVTETensor hm = plain(api, rht_matrix_device(), kNVTEBFloat16, {16, 16});
  NVTEQuantizationConfig qc = api.create_qcfg();
  const int mask_t = 0xD7E8;
  const Api *a = &api;
  auto sums_of = [=](const Fp4Out &o, bool with_amax) {
    Sums r{{"row", checksum(o.row, rows * cols / 2)}, {"col", checksum(o.col, rows * cols / 2)},
           {"row_scale", checksum(o.sr, o.sr_bytes)}, {"col_scale", checksum(o.sc, o.sc_bytes)}};
    if (with_amax) r.push_back({"amax", checksum(o.amax, 8)});
    return r;
  };
  Case c_amax, c_cast, c_both, c_plain;
  c_amax.name = base + "_rht_amax";
  c_amax.kernel = "K12 hadamard_transform_amax";
  c_amax.work = rows * cols * 2.0;
  c_amax.fn = [=](int s) { a->rht_amax(in[s], amax_only[s], 0, mask_t, g_stream); };
  c_cast.name = base + "_rht_cast";
  c_cast.kernel = "K14 quantize_with_hadamard_transform";
  c_cast.work = cast_bytes;
  c_cast.fn = [=](int s) { a->rht_quantize(in[s], out[s].t, hm, qc, g_stream); };
  Fp4Out o0 = out[0], p0 = plain_out[0];
  c_cast.sums = [=]() { return sums_of(o0, true); };
  c_both.name = base + "_rht_amax_cast";
  c_both.kernel = "K12+K14";
  c_both.work = cast_bytes + rows * cols * 2.0;
  c_both.fn = [=](int s) {
    a->rht_amax(in[s], amax_only[s], 0, mask_t, g_stream);
    a->rht_quantize(in[s], out[s].t, hm, qc, g_stream);
  };
  c_plain.name = base + "_plain_cast";
  c_plain.kernel = "K10 nvfp4 quantize_transpose";
  c_plain.work = cast_bytes;
  c_plain.fn = [=](int s) { a->quantize_v2(in[s], plain_out[s].t, qc, g_stream); };
  c_plain.sums = [=]() { return sums_of(p0, false); };
  std::vector<Case> r;
  for (Case *c : {&c_amax, &c_cast, &c_both, &c_plain}) {
    c->nsets = nsets;
    if (selected(c->name)) r.push_back(*c);
  }
  return r;
}

// ---------------------------------------------------------------- GEMM (nvte_cublas_gemm_v2)
struct GemmOperand {
  NVTETensor t;
  void *data;
  size_t data_bytes;
};
// kind 0 = per-tensor FP8, 1 = MXFP8, 2 = NVFP4. `shape` = the tensor's own [d0, d1]; colwise = the GEMM will
// read the columnwise copy. Quantized by the library under test (bf16 -> kind), scales swizzled by it too.
GemmOperand make_operand(const Api &api, int kind, size_t d0, size_t d1, bool colwise, uint32_t seed, float scale) {
  __nv_bfloat16 *src = alloc_bf16(d0 * d1, seed, scale);
  NVTETensor in = plain(api, src, kNVTEBFloat16, {d0, d1});
  NVTEQuantizationConfig qc = api.create_qcfg();
  GemmOperand op{};
  const auto swizzle = api.swizzle;
  if (kind == 0) {
    auto *scal = static_cast<float *>(dalloc(16, 0));
    const float s = 8.0f;
    CK(cudaMemcpy(scal, &s, 4, cudaMemcpyHostToDevice));
    op.data_bytes = d0 * d1;
    op.data = dalloc(op.data_bytes, 0);
    op.t = T(api, NVTE_DELAYED_TENSOR_SCALING).set(kNVTERowwiseData, op.data, kNVTEFloat8E4M3, {d0, d1})
               .set(kNVTEScale, scal, kNVTEFloat32, {1}).set(kNVTERowwiseScaleInv, scal + 1, kNVTEFloat32, {1}).t;
    api.quantize_v2(in, op.t, qc, g_stream);
  } else if (kind == 1) {
    MxOut q = mx_out(api, d0, d1, !colwise, colwise, false);
    api.quantize_v2(in, q.t, qc, g_stream);
    const size_t sb = colwise ? q.sc_bytes : q.sr_bytes;
    void *sw = dalloc(sb, 0);
    const std::vector<size_t> ss = colwise ? std::vector<size_t>{ru(d0 / 32, 4), ru(d1, 128)}
                                           : std::vector<size_t>{ru(d0, 128), ru(d1 / 32, 4)};
    op.data = colwise ? q.col : q.row;
    op.data_bytes = d0 * d1;
    op.t = T(api, NVTE_MXFP8_1D_SCALING)
               .set(colwise ? kNVTEColumnwiseData : kNVTERowwiseData, op.data, kNVTEFloat8E4M3, {d0, d1})
               .set(colwise ? kNVTEColumnwiseScaleInv : kNVTERowwiseScaleInv, sw, kNVTEFloat8E8M0, ss)
               .flag(kNVTEWithGEMMSwizzledScales, true).t;
    swizzle(q.t, op.t, g_stream);
  } else {
    Fp4Out q = fp4_out(api, d0, d1, !colwise, colwise, false);
    const float am[2] = {3.0f * scale, 3.0f * scale};
    CK(cudaMemcpy(q.amax, am, 8, cudaMemcpyHostToDevice));
    api.quantize_v2(in, q.t, qc, g_stream);
    const size_t sb = colwise ? q.sc_bytes : q.sr_bytes;
    void *sw = dalloc(sb, 0);
    op.data = colwise ? q.col : q.row;
    op.data_bytes = d0 * d1 / 2;
    T t(api, NVTE_NVFP4_1D_SCALING);
    if (colwise)
      t.set(kNVTEColumnwiseData, op.data, kNVTEFloat4E2M1, {d1, d0})
          .set(kNVTEColumnwiseScaleInv, sw, kNVTEFloat8E4M3, {ru(d1, 128), ru(d0 / 16, 4)})
          .set(kNVTEColumnwiseAmax, q.amax + 1, kNVTEFloat32, {1});
    else
      t.set(kNVTERowwiseData, op.data, kNVTEFloat4E2M1, {d0, d1})
          .set(kNVTERowwiseScaleInv, sw, kNVTEFloat8E4M3, {ru(d0, 128), ru(d1 / 16, 4)})
          .set(kNVTEAmax, q.amax, kNVTEFloat32, {1});
    t.flag(kNVTEWithGEMMSwizzledScales, true);
    op.t = t.t;
    swizzle(q.t, op.t, g_stream);
  }
  CK(cudaStreamSynchronize(g_stream));
  api.destroy_qcfg(qc);
  return op;
}

// TE convention: D[n, m]; TN: A[m,k] B[n,k]; NN: A[k,m] B[n,k]; NT: A[k,m] B[k,n]
std::vector<Case> bench_gemm(const Api &api, const char *cfg, int kind, const char *layout, size_t n, size_t m, size_t k) {
  static const char *kinds[3] = {"fp8", "mxfp8", "nvfp4"};
  Case c;
  c.name = std::string("gemm_") + kinds[kind] + "_" + cfg + "_" + layout;
  if (!selected(c.name)) return {};
  c.kernel = kind == 0 ? "fp8 gemm" : (kind == 1 ? "mxfp8 block-scaled gemm" : "nvfp4 block-scaled gemm");
  c.work = 2.0 * n * m * k / 1000.0;
  c.unit = "TFLOP/s";
  const bool ta = layout[0] == 'T', tb = layout[1] == 'T';
  try {
    GemmOperand A = ta ? make_operand(api, kind, m, k, false, 101, 0.05f) : make_operand(api, kind, k, m, true, 101, 0.05f);
    GemmOperand B = tb ? make_operand(api, kind, k, n, true, 202, 1.0f) : make_operand(api, kind, n, k, false, 202, 1.0f);
    void *d = dalloc(n * m * 2, 0);
    NVTETensor D = plain(api, d, kNVTEBFloat16, {n, m});
    const size_t ws_bytes = (32u << 20) + 1024;
    NVTETensor W = plain(api, dalloc(ws_bytes, 0), kNVTEByte, {ws_bytes});
    NVTEMatmulConfig mc = api.create_mcfg();
    const Api *a = &api;
    c.fn = [=](int) {
      const float alpha = 1.0f, beta = 0.0f;
      a->gemm_v2(ta, tb, &alpha, A.t, B.t, &beta, D, D, W, mc, g_stream);
    };
    c.sums = [=]() {
      return Sums{{"A", checksum(A.data, A.data_bytes)}, {"B", checksum(B.data, B.data_bytes)}, {"D_bf16", checksum(d, n * m * 2)}};
    };
  } catch (const std::exception &e) {
    c.error = clean(e.what());
    cudaGetLastError();
  }
  return {c};
}

}  // namespace

int main(int argc, char **argv) {
  std::vector<std::string> paths, names;
  for (int i = 1; i < argc; ++i) {
    if (!std::strcmp(argv[i], "--iters") && i + 1 < argc) g_opts.iters = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--rounds") && i + 1 < argc) g_opts.rounds = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--only") && i + 1 < argc) g_opts.only = argv[++i];
    else if (!std::strcmp(argv[i], "--name") && i + 1 < argc) names.back() = argv[++i];
    else {
      std::string a = argv[i];
      const size_t eq = a.find('=');
      paths.push_back(eq == std::string::npos ? a : a.substr(0, eq));
      names.push_back(eq == std::string::npos ? a : a.substr(eq + 1));
    }
  }
  if (paths.empty()) {
    std::fprintf(stderr, "usage: %s <lib.so>[=name] [<lib.so>[=name] ...] [--iters N] [--rounds R] [--only substr]\n", argv[0]);
    return 2;
  }
  CK(cudaSetDevice(0));
  CK(cudaStreamCreate(&g_stream));
  std::vector<Api *> apis;
  for (auto &p : paths) apis.push_back(new Api(p.c_str()));
  using Builder = std::function<std::vector<Case>(const Api &)>;
  std::vector<Builder> groups;
  // config 1 and config 4 tensors (MXFP8), compact and GEMM-swizzled scale output
  for (int sw = 0; sw < 2; ++sw) {
    groups.push_back([=](const Api &a) { return bench_mxfp8_quantize(a, "c1_4096x4096", 4096, 4096, sw); });
    groups.push_back([=](const Api &a) { return bench_mxfp8_quantize(a, "c4_x_8192x4096", 8192, 4096, sw); });
    groups.push_back([=](const Api &a) { return bench_mxfp8_quantize(a, "c4_w_14336x4096", 14336, 4096, sw); });
    groups.push_back([=](const Api &a) { return bench_mxfp8_quantize(a, "c4_dy_8192x14336", 8192, 14336, sw); });
  }
  groups.push_back([](const Api &a) { return bench_mxfp8_step(a, false); });
  groups.push_back([](const Api &a) { return bench_mxfp8_step(a, true); });
  groups.push_back([](const Api &a) { return bench_fp8(a, "c5_shard_8192x8192", 8192, 8192); });
  groups.push_back([](const Api &a) { return bench_fp8(a, "c2_dy_2048x3072", 2048, 3072); });
  groups.push_back([](const Api &a) { return bench_nvfp4(a, "c3_16384x8192", 16384, 8192); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c4_fprop", 1, "TN", 8192, 14336, 4096); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c4_dgrad", 1, "NN", 8192, 4096, 14336); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c4_wgrad", 1, "NT", 14336, 4096, 8192); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c3", 2, "TN", 16384, 8192, 8192); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c5", 0, "TN", 8192, 28672, 8192); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c5_tp8", 0, "TN", 8192, 3584, 8192); });
  groups.push_back([](const Api &a) { return bench_gemm(a, "c2", 0, "TN", 2048, 3072, 768); });
  for (auto &build : groups) {
    std::vector<std::vector<Case>> per_lib;
    for (Api *a : apis) {
      try {
        per_lib.push_back(build(*a));
      } catch (const std::exception &e) {
        std::printf("{\"lib\": \"%s\", \"error\": \"building a case failed: %s\"}\n", names[per_lib.size()].c_str(),
                    clean(e.what()).c_str());
        cudaGetLastError();
        per_lib.push_back({});
      }
    }
    size_t ncase = 0;
    for (auto &v : per_lib) ncase = v.size() > ncase ? v.size() : ncase;
    for (size_t ci = 0; ci < ncase; ++ci) {
      std::vector<Case *> cs;
      std::vector<std::string> ln;
      for (size_t l = 0; l < per_lib.size(); ++l)
        if (ci < per_lib[l].size()) {
          cs.push_back(&per_lib[l][ci]);
          ln.push_back(names[l]);
        }
      time_cases(cs, ln);
    }
    free_all();
  }
  return 0;
}
