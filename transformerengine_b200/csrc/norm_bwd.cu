// LayerNorm / RMSNorm backward (reference normalization/layernorm/ln_api.cpp:153-260,
// rmsnorm/rmsnorm_api.cpp:132-300; kernels ln_bwd_kernels.cuh, rmsnorm_bwd_kernels.cuh).
//   y = (x - mu) * rsigma,  dy = g * dz
//   dx = rsigma * (dy - mean(dy * y) * y - mean(dy)) [+ add]     (mean(dy) term: LayerNorm only)
//   dgamma = sum_rows y * dz,  dbeta = sum_rows dz
// HBM-bound: 4 B read + 2 B written per element. Persistent CTAs, one row at a time held in
// registers as packed 16-bit words (the next row's loads are issued before the block reduction of
// the current one); each thread keeps fp32 dgamma / dbeta accumulators for its own columns across
// all of its CTA's rows, so the column reductions cost one [grid, H] fp32 workspace write at the
// end and a small second kernel. Row order per CTA and CTA order in the second kernel are fixed:
// the result is deterministic for a given shape and device.
#include "core.h"
#include "ptx.cuh"
#include "sm100.cuh"

namespace teb {
namespace norm {

namespace {

constexpr int kMaxThreads = 512;

struct BwdArgs {
  const void *dz, *x, *add, *gamma;
  const float *mu, *rsigma;
  void *dx;
  float *ws_dgamma, *ws_dbeta;   // [grid][H]
  int wtype;
  size_t rows, cols;
  float inv_h;
  bool zero_centered;
};

__device__ __forceinline__ float load_weight(const void *p, int wtype, size_t i) {
  switch (wtype) {
    case kNVTEFloat32: return static_cast<const float *>(p)[i];
    case kNVTEFloat16: return __half2float(static_cast<const __half *>(p)[i]);
    default: return __bfloat162float(static_cast<const __nv_bfloat16 *>(p)[i]);
  }
}
template <typename IT>
__device__ __forceinline__ uint32_t pack16(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack16<__nv_bfloat16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <>
__device__ __forceinline__ uint32_t pack16<__half>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

template <typename IT, bool LN, bool ADD, int VPT>
__global__ void __launch_bounds__(kMaxThreads) norm_bwd_kernel(const BwdArgs a) {
  using W = ptx::Word<IT>;
  extern __shared__ float smem_f[];
  float *gam = smem_f;                       // [H] fp32 gamma (+1)
  float *red = smem_f + a.cols;              // [2][warps][2]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const size_t nvec = a.cols >> 3;

  sm100::grid_dependency_wait();
  for (size_t i = tid; i < a.cols; i += blockDim.x) {
    float g = load_weight(a.gamma, a.wtype, i);
    if (a.zero_centered) g += 1.0f;
    gam[i] = g;
  }
  __syncthreads();
  sm100::grid_launch_dependents();

  float dg[VPT][8], db[VPT][8];
#pragma unroll
  for (int k = 0; k < VPT; ++k)
#pragma unroll
    for (int i = 0; i < 8; ++i) dg[k][i] = db[k][i] = 0.0f;

  uint4 xv[VPT], zv[VPT], nx[VPT], nz[VPT];
  auto load_row = [&](size_t row, uint4 (&xo)[VPT], uint4 (&zo)[VPT]) {
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const size_t v = tid + static_cast<size_t>(k) * blockDim.x;
      if (v < nvec) {
        xo[k] = ptx::ld_stream_v4(static_cast<const uint4 *>(a.x) + row * nvec + v);
        zo[k] = ptx::ld_stream_v4(static_cast<const uint4 *>(a.dz) + row * nvec + v);
      } else {
        xo[k] = zo[k] = make_uint4(0, 0, 0, 0);
      }
    }
  };
  size_t row = blockIdx.x;
  if (row < a.rows) load_row(row, nx, nz);
  int parity = 0;
  for (; row < a.rows; row += gridDim.x, parity ^= 1) {
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      xv[k] = nx[k];
      zv[k] = nz[k];
    }
    const float mu = LN ? a.mu[row] : 0.0f, rs = a.rsigma[row];
    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const size_t v = tid + static_cast<size_t>(k) * blockDim.x;
      if (v >= nvec) continue;
      const uint32_t xw[4] = {xv[k].x, xv[k].y, xv[k].z, xv[k].w}, zw[4] = {zv[k].x, zv[k].y, zv[k].z, zv[k].w};
      const float4 g0 = *reinterpret_cast<const float4 *>(gam + v * 8), g1 = *reinterpret_cast<const float4 *>(gam + v * 8 + 4);
      const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float xf = (i & 1) ? W::hi(xw[i >> 1]) : W::lo(xw[i >> 1]);
        const float zf = (i & 1) ? W::hi(zw[i >> 1]) : W::lo(zw[i >> 1]);
        const float y = (xf - mu) * rs, dy = g[i] * zf;
        dg[k][i] = fmaf(y, zf, dg[k][i]);
        if (LN) {
          db[k][i] += zf;
          s1 += dy;
        }
        s2 = fmaf(dy, y, s2);
      }
    }
    if (row + gridDim.x < a.rows) load_row(row + gridDim.x, nx, nz);   // in flight during the reduction
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      if (LN) s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    float *rbuf = red + parity * 2 * (kMaxThreads / 32);
    if (lane == 0) {
      rbuf[2 * warp] = s1;
      rbuf[2 * warp + 1] = s2;
    }
    __syncthreads();
    float t1 = 0.0f, t2 = 0.0f;
    for (int w = 0; w < nwarps; ++w) {
      t1 += rbuf[2 * w];
      t2 += rbuf[2 * w + 1];
    }
    const float mdy = LN ? t1 * a.inv_h : 0.0f, mdyy = t2 * a.inv_h;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const size_t v = tid + static_cast<size_t>(k) * blockDim.x;
      if (v >= nvec) continue;
      const uint32_t xw[4] = {xv[k].x, xv[k].y, xv[k].z, xv[k].w}, zw[4] = {zv[k].x, zv[k].y, zv[k].z, zv[k].w};
      const float4 g0 = *reinterpret_cast<const float4 *>(gam + v * 8), g1 = *reinterpret_cast<const float4 *>(gam + v * 8 + 4);
      const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      uint4 av = make_uint4(0, 0, 0, 0);
      if (ADD) av = ptx::ld_stream_v4(static_cast<const uint4 *>(a.add) + row * nvec + v);
      const uint32_t aw[4] = {av.x, av.y, av.z, av.w};
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float xf = (i & 1) ? W::hi(xw[i >> 1]) : W::lo(xw[i >> 1]);
        const float zf = (i & 1) ? W::hi(zw[i >> 1]) : W::lo(zw[i >> 1]);
        const float y = (xf - mu) * rs, dy = g[i] * zf;
        o[i] = rs * (dy - (mdyy * y + mdy));
        if (ADD) o[i] += (i & 1) ? W::hi(aw[i >> 1]) : W::lo(aw[i >> 1]);
      }
      *(static_cast<uint4 *>(a.dx) + row * nvec + v) =
          make_uint4(pack16<IT>(o[0], o[1]), pack16<IT>(o[2], o[3]), pack16<IT>(o[4], o[5]), pack16<IT>(o[6], o[7]));
    }
  }
  // per-CTA column partial sums
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const size_t v = tid + static_cast<size_t>(k) * blockDim.x;
    if (v >= nvec) continue;
    float4 *pg = reinterpret_cast<float4 *>(a.ws_dgamma + blockIdx.x * a.cols + v * 8);
    pg[0] = make_float4(dg[k][0], dg[k][1], dg[k][2], dg[k][3]);
    pg[1] = make_float4(dg[k][4], dg[k][5], dg[k][6], dg[k][7]);
    if (LN) {
      float4 *pb = reinterpret_cast<float4 *>(a.ws_dbeta + blockIdx.x * a.cols + v * 8);
      pb[0] = make_float4(db[k][0], db[k][1], db[k][2], db[k][3]);
      pb[1] = make_float4(db[k][4], db[k][5], db[k][6], db[k][7]);
    }
  }
}

// dgamma[c] = sum_p ws[p][c] (CTA order), cast to the weight dtype; y = 0: dgamma, y = 1: dbeta
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float *ws_dgamma, const float *ws_dbeta, int parts,
                                                              size_t cols, void *dgamma, void *dbeta, int wtype) {
  sm100::grid_dependency_wait();
  const size_t c = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (c >= cols) return;
  const float *ws = blockIdx.y == 0 ? ws_dgamma : ws_dbeta;
  void *out = blockIdx.y == 0 ? dgamma : dbeta;
  float s = 0.0f;
  for (int p = 0; p < parts; ++p) s += ws[p * cols + c];
  switch (wtype) {
    case kNVTEFloat32: static_cast<float *>(out)[c] = s; break;
    case kNVTEFloat16: static_cast<__half *>(out)[c] = __float2half_rn(s); break;
    default: static_cast<__nv_bfloat16 *>(out)[c] = __float2bfloat16_rn(s); break;
  }
}

template <typename IT, bool LN, bool ADD, int VPT>
void launch(const BwdArgs &a, int threads, int *grid_out, bool query, cudaStream_t stream) {
  auto *kern = norm_bwd_kernel<IT, LN, ADD, VPT>;
  const size_t smem = a.cols * 4 + 2 * 2 * (kMaxThreads / 32) * 4;
  static bool configured = [kern] {
    TEB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 4 + 1024));
    return true;
  }();
  (void)configured;
  int per_sm = 0;
  TEB_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem));
  if (per_sm < 1) per_sm = 1;
  const size_t cap = static_cast<size_t>(sm_count()) * per_sm;
  *grid_out = static_cast<int>(a.rows < cap ? (a.rows ? a.rows : 1) : cap);
  if (query) return;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(*grid_out);
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
}

template <typename IT, bool LN, bool ADD>
void launch_vpt(int vpt, const BwdArgs &a, int threads, int *grid, bool query, cudaStream_t stream) {
  switch (vpt) {
    case 1: launch<IT, LN, ADD, 1>(a, threads, grid, query, stream); break;
    case 2: launch<IT, LN, ADD, 2>(a, threads, grid, query, stream); break;
    default: launch<IT, LN, ADD, 4>(a, threads, grid, query, stream); break;
  }
}

template <typename IT>
void launch_mode(bool ln, bool add, int vpt, const BwdArgs &a, int threads, int *grid, bool query,
                 cudaStream_t stream) {
  if (ln) launch_vpt<IT, true, false>(vpt, a, threads, grid, query, stream);
  else if (add) launch_vpt<IT, false, true>(vpt, a, threads, grid, query, stream);
  else launch_vpt<IT, false, false>(vpt, a, threads, grid, query, stream);
}

}  // namespace

// norm_generic.cu
int generic_bwd_grid(size_t N);
void backward_generic(bool layernorm, const Tensor &dz, const Tensor &x, const Tensor *add, const Tensor *mu,
                      const Tensor &rsigma, const Tensor &gamma, Tensor &dx, float *ws_dgamma, float *ws_dbeta, int grid,
                      bool zero_centered_gamma, cudaStream_t stream);

// Shared front end of nvte_layernorm_bwd / nvte_rmsnorm_bwd / nvte_rmsnorm_bwd_add.
void backward(bool layernorm, const Tensor &dz, const Tensor &x, const Tensor *add, const Tensor *mu,
              const Tensor &rsigma, const Tensor &gamma, Tensor &dx, Tensor &dgamma, Tensor *dbeta,
              Tensor &workspace, bool zero_centered_gamma, cudaStream_t stream) {
  require_blackwell("Normalization");
  const char *name = layernorm ? "LayerNorm backward" : "RMSNorm backward";
  // reference ln_api.cpp:160-185
  TEB_CHECK(x.data.shape.ndim == 2, "x must be 2D tensor.");
  const size_t N = x.data.shape.data[0], H = x.data.shape.data[1];
  TEB_CHECK(dz.data.shape.ndim == 2 && dz.data.shape.data[0] == N && dz.data.shape.data[1] == H,
            "dz must have the shape of x.");
  TEB_CHECK(rsigma.data.dtype == kNVTEFloat32 && rsigma.data.numel() == N, "rsigma must be float32 [x.shape[0]].");
  TEB_CHECK(gamma.data.numel() == H, "Gamma must have the same hidden size.");
  TEB_CHECK(dx.data.shape.ndim == 2 && dx.data.shape.data[0] == N && dx.data.shape.data[1] == H,
            "dx must have the shape of x.");
  TEB_CHECK(dx.data.dtype == x.data.dtype, "dx must have the dtype of x.");
  TEB_CHECK(dgamma.data.numel() == H && dgamma.data.dtype == gamma.data.dtype, "dgamma must match gamma.");
  if (layernorm) {
    TEB_CHECK(mu != nullptr && mu->data.dtype == kNVTEFloat32 && mu->data.numel() == N, "mu must be float32 [x.shape[0]].");
    TEB_CHECK(dbeta != nullptr && dbeta->data.numel() == H && dbeta->data.dtype == gamma.data.dtype,
              "dbeta must match gamma.");
  }
  if (add != nullptr)
    TEB_CHECK(add->data.shape.ndim == 2 && add->data.shape.data[0] == N && add->data.shape.data[1] == H &&
                  add->data.dtype == x.data.dtype,
              "add must have the shape and dtype of x.");
  // The fast kernel takes 16-bit tensors of one dtype (gamma may be fp32) with H % 8 == 0, H <= 16384 and 16-byte
  // aligned buffers; everything else goes to the generic kernel of norm_generic.cu.
  auto aligned16 = [](const void *p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  const bool fast = (x.data.dtype == kNVTEBFloat16 || x.data.dtype == kNVTEFloat16) && dz.data.dtype == x.data.dtype &&
                    (gamma.data.dtype == x.data.dtype || gamma.data.dtype == kNVTEFloat32) && H % 8 == 0 && H <= 16384 &&
                    aligned16(dz.data.dptr) && aligned16(x.data.dptr) && aligned16(dx.data.dptr) &&
                    (add == nullptr || aligned16(add->data.dptr));
  if (!fast) {
    const int ggrid = generic_bwd_grid(N);
    if (workspace.data.numel() == 0) {
      workspace.data.shape.ndim = 1;
      workspace.data.shape.data[0] = static_cast<size_t>(ggrid) * H * 4 * (layernorm ? 2 : 1) + 16;
      workspace.data.dtype = kNVTEByte;
      return;
    }
    TEB_CHECK(dz.data.dptr && x.data.dptr && rsigma.data.dptr && gamma.data.dptr && dx.data.dptr && dgamma.data.dptr &&
                  workspace.data.dptr,
              name, ": input and output tensors must be allocated");
    const size_t need = static_cast<size_t>(ggrid) * H * 4 * (layernorm ? 2 : 1);
    TEB_CHECK(workspace.data.bytes() >= need, name, ": workspace too small (", workspace.data.bytes(), " < ", need, ")");
    if (N == 0 || H == 0) {
      TEB_CHECK_CUDA(cudaMemsetAsync(dgamma.data.dptr, 0, dgamma.data.bytes(), stream));
      if (layernorm) TEB_CHECK_CUDA(cudaMemsetAsync(dbeta->data.dptr, 0, dbeta->data.bytes(), stream));
      return;
    }
    float *wsg = reinterpret_cast<float *>((reinterpret_cast<uintptr_t>(workspace.data.dptr) + 15) & ~uintptr_t(15));
    float *wsb = wsg + static_cast<size_t>(ggrid) * H;
    backward_generic(layernorm, dz, x, add, mu, rsigma, gamma, dx, wsg, wsb, ggrid, zero_centered_gamma, stream);
    const dim3 rgrid(static_cast<unsigned>((H + 255) / 256), layernorm ? 2 : 1);
    reduce_partials_kernel<<<rgrid, 256, 0, stream>>>(wsg, wsb, ggrid, H, dgamma.data.dptr,
                                                      layernorm ? dbeta->data.dptr : nullptr,
                                                      static_cast<int>(gamma.data.dtype));
    TEB_CHECK_CUDA(cudaGetLastError());
    return;
  }

  const size_t nvec = H / 8;
  int threads = 128;
  while (threads < kMaxThreads && static_cast<size_t>(threads) * 2 < nvec) threads *= 2;
  const int vpt_raw = static_cast<int>((nvec + threads - 1) / threads);
  const int vpt = vpt_raw <= 1 ? 1 : vpt_raw <= 2 ? 2 : 4;

  BwdArgs a{};
  a.dz = dz.data.dptr;
  a.x = x.data.dptr;
  a.add = add ? add->data.dptr : nullptr;
  a.gamma = gamma.data.dptr;
  a.mu = layernorm ? mu->data.as<float>() : nullptr;
  a.rsigma = rsigma.data.as<float>();
  a.dx = dx.data.dptr;
  a.wtype = gamma.data.dtype;
  a.rows = N;
  a.cols = H;
  a.inv_h = 1.0f / static_cast<float>(H);
  a.zero_centered = zero_centered_gamma;

  const bool query = workspace.data.numel() == 0;
  int grid = 1;
  if (query) {
    if (x.data.dtype == kNVTEBFloat16) launch_mode<__nv_bfloat16>(layernorm, add != nullptr, vpt, a, threads, &grid, true, stream);
    else launch_mode<__half>(layernorm, add != nullptr, vpt, a, threads, &grid, true, stream);
    workspace.data.shape.ndim = 1;
    workspace.data.shape.data[0] = static_cast<size_t>(grid) * H * 4 * (layernorm ? 2 : 1) + 16;
    workspace.data.dtype = kNVTEByte;
    return;
  }
  TEB_CHECK(dz.data.dptr && x.data.dptr && rsigma.data.dptr && gamma.data.dptr && dx.data.dptr && dgamma.data.dptr &&
                workspace.data.dptr,
            name, ": input and output tensors must be allocated");
  TEB_CHECK(aligned16(workspace.data.dptr), name, ": the workspace must be 16-byte aligned");
  if (N == 0) {
    TEB_CHECK_CUDA(cudaMemsetAsync(dgamma.data.dptr, 0, dgamma.data.bytes(), stream));
    if (layernorm) TEB_CHECK_CUDA(cudaMemsetAsync(dbeta->data.dptr, 0, dbeta->data.bytes(), stream));
    return;
  }
  // the partial-sum layout depends on the grid: recompute it and check the workspace holds it
  if (x.data.dtype == kNVTEBFloat16) launch_mode<__nv_bfloat16>(layernorm, add != nullptr, vpt, a, threads, &grid, true, stream);
  else launch_mode<__half>(layernorm, add != nullptr, vpt, a, threads, &grid, true, stream);
  const size_t need = static_cast<size_t>(grid) * H * 4 * (layernorm ? 2 : 1);
  TEB_CHECK(workspace.data.bytes() >= need, name, ": workspace too small (", workspace.data.bytes(), " < ", need, ")");
  a.ws_dgamma = workspace.data.as<float>();
  a.ws_dbeta = a.ws_dgamma + static_cast<size_t>(grid) * H;
  if (x.data.dtype == kNVTEBFloat16) launch_mode<__nv_bfloat16>(layernorm, add != nullptr, vpt, a, threads, &grid, false, stream);
  else launch_mode<__half>(layernorm, add != nullptr, vpt, a, threads, &grid, false, stream);

  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(static_cast<unsigned>((H + 255) / 256), layernorm ? 2 : 1);
  cfg.blockDim = dim3(256);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TEB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, reduce_partials_kernel, static_cast<const float *>(a.ws_dgamma),
                                    static_cast<const float *>(a.ws_dbeta), grid, H, dgamma.data.dptr,
                                    layernorm ? dbeta->data.dptr : nullptr, static_cast<int>(a.wtype)));
}

}  // namespace norm
}  // namespace teb
