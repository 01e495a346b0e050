// LayerNorm / RMSNorm forward for every configuration the fused cluster kernel (norm_fwd.cu) does not take:
// fp32 inputs, hidden sizes that are not a multiple of 64 or do not split over a cluster, E5M2 outputs, output
// dtypes that differ from the input's, MXFP8 columnwise outputs with ragged row counts. Correctness path, not a
// hot path (the Linear path's activations are bf16 with H % 64 == 0): one CTA per row, three passes over the row
// in fp32 (mean, centred variance, output), every dtype read through a runtime switch. MXFP8 outputs go through
// an fp32 image of y in the caller's workspace and the regular MXFP8 quantizer, so they carry the quantizer's
// bytes. Replaces, for those configurations, reference normalization/layernorm/ln_api.cpp:27-150 and
// rmsnorm/rmsnorm_api.cpp:25-130 (whose general path is a cuDNN graph).
#include <cfloat>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "core.h"
#include "ptx.cuh"

namespace teb {
namespace norm {

namespace {

__device__ __forceinline__ float load_f(const void *p, NVTEDType t, size_t i) {
  switch (t) {
    case kNVTEFloat32: return reinterpret_cast<const float *>(p)[i];
    case kNVTEBFloat16: return __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(p)[i]);
    default: return __half2float(reinterpret_cast<const __half *>(p)[i]);
  }
}
__device__ __forceinline__ void store_f(void *p, NVTEDType t, size_t i, float v) {
  switch (t) {
    case kNVTEFloat32: reinterpret_cast<float *>(p)[i] = v; break;
    case kNVTEBFloat16: reinterpret_cast<__nv_bfloat16 *>(p)[i] = __float2bfloat16_rn(v); break;
    case kNVTEFloat16: reinterpret_cast<__half *>(p)[i] = __float2half_rn(v); break;
    case kNVTEFloat8E4M3: reinterpret_cast<uint8_t *>(p)[i] = static_cast<uint8_t>(ptx::cvt_fp8x2<false>(v, 0.0f)); break;
    default: reinterpret_cast<uint8_t *>(p)[i] = static_cast<uint8_t>(ptx::cvt_fp8x2<true>(v, 0.0f)); break;
  }
}

__device__ __forceinline__ float block_sum(float v, float *red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.0f;
  for (int w = 0; w < static_cast<int>(blockDim.x >> 5); ++w) t += red[w];
  return t;
}

struct GenArgs {
  const void *x, *gamma, *beta;
  void *z;
  NVTEDType xt, wt, zt;
  float *mu, *rsigma;
  const float *scale;
  float *scale_inv, *amax;
  size_t rows, cols;
  float eps;
  bool layernorm, zero_centered;
};

__global__ void __launch_bounds__(256) norm_generic_kernel(const GenArgs a) {
  __shared__ float red[8];
  const bool fp8 = a.zt == kNVTEFloat8E4M3 || a.zt == kNVTEFloat8E5M2;
  const float s = (fp8 && a.scale != nullptr) ? *a.scale : 1.0f;
  if (blockIdx.x == 0 && threadIdx.x == 0 && fp8 && a.scale_inv != nullptr) *a.scale_inv = __frcp_rn(s);
  float m = 0.0f;
  for (size_t r = blockIdx.x; r < a.rows; r += gridDim.x) {
    const size_t base = r * a.cols;
    float mu = 0.0f;
    if (a.layernorm) {
      float t = 0.0f;
      for (size_t c = threadIdx.x; c < a.cols; c += blockDim.x) t += load_f(a.x, a.xt, base + c);
      mu = block_sum(t, red) / static_cast<float>(a.cols);
    }
    float q = 0.0f;
    for (size_t c = threadIdx.x; c < a.cols; c += blockDim.x) {
      const float d = load_f(a.x, a.xt, base + c) - mu;
      q += d * d;
    }
    const float rs = rsqrtf(block_sum(q, red) / static_cast<float>(a.cols) + a.eps);
    if (threadIdx.x == 0) {
      if (a.layernorm) a.mu[r] = mu;
      a.rsigma[r] = rs;
    }
    for (size_t c = threadIdx.x; c < a.cols; c += blockDim.x) {
      float g = load_f(a.gamma, a.wt, c);
      if (a.zero_centered) g += 1.0f;
      float y = (load_f(a.x, a.xt, base + c) - mu) * rs * g;
      if (a.layernorm) y += load_f(a.beta, a.wt, c);
      m = fmaxf(m, fabsf(y));
      store_f(a.z, a.zt, base + c, y * s);
    }
  }
  if (a.amax != nullptr) {
    m = ptx::warp_max(m);
    if ((threadIdx.x & 31) == 0 && m > 0.0f) ptx::atomic_max_nonneg(a.amax, m);
  }
}

}  // namespace

// bytes of workspace the generic path needs for this output (an fp32 image of y for MXFP8 outputs)
size_t generic_workspace_bytes(const Tensor &z, size_t N, size_t H) {
  return z.scaling_mode == NVTE_MXFP8_1D_SCALING ? N * H * sizeof(float) + 256 : 16;
}

void forward_generic(bool layernorm, const Tensor &x, const Tensor &gamma, const Tensor *beta, float epsilon, Tensor &z,
                     Tensor *mu, Tensor &rsigma, Tensor &workspace, bool zero_centered_gamma, cudaStream_t stream) {
  const size_t N = x.data.shape.data[0], H = x.data.shape.data[1];
  TEB_CHECK(is_high_precision(x.data.dtype) && is_high_precision(gamma.data.dtype), "norm: x and gamma must be fp32 / bf16 / fp16");
  if (N == 0 || H == 0) return;
  GenArgs a{};
  a.x = x.data.dptr;
  a.gamma = gamma.data.dptr;
  a.beta = layernorm ? beta->data.dptr : nullptr;
  a.xt = x.data.dtype;
  a.wt = gamma.data.dtype;
  a.mu = layernorm ? mu->data.as<float>() : nullptr;
  a.rsigma = rsigma.data.as<float>();
  a.rows = N;
  a.cols = H;
  a.eps = epsilon;
  a.layernorm = layernorm;
  a.zero_centered = zero_centered_gamma;
  const int grid = static_cast<int>(std::min<size_t>(N, static_cast<size_t>(sm_count()) * 8));
  if (z.scaling_mode == NVTE_MXFP8_1D_SCALING) {
    TEB_CHECK(workspace.data.dptr != nullptr && workspace.data.bytes() >= N * H * sizeof(float),
              "norm: workspace too small for the MXFP8 output (query it with an empty workspace first)");
    float *tmp = reinterpret_cast<float *>((reinterpret_cast<uintptr_t>(workspace.data.dptr) + 15) & ~uintptr_t(15));
    a.z = tmp;
    a.zt = kNVTEFloat32;
    norm_generic_kernel<<<grid, 256, 0, stream>>>(a);
    TEB_CHECK_CUDA(cudaGetLastError());
    Tensor y{};
    y.scaling_mode = NVTE_DELAYED_TENSOR_SCALING;
    y.data.dptr = tmp;
    y.data.dtype = kNVTEFloat32;
    y.data.shape.ndim = 2;
    y.data.shape.data[0] = N;
    y.data.shape.data[1] = H;
    mxfp8::quantize(y, z, nullptr, stream);
    return;
  }
  TEB_CHECK(z.columnwise_data.dptr == nullptr, "norm: a transposed per-tensor copy is not produced");
  a.z = z.data.dptr;
  a.zt = z.data.dtype;
  if (is_fp8(z.data.dtype)) {
    TEB_CHECK(z.scale.dptr != nullptr, "FP8 output needs a scale");
    a.scale = z.scale.as<float>();
    a.scale_inv = z.scale_inv.as<float>();
  }
  a.amax = z.amax.as<float>();
  norm_generic_kernel<<<grid, 256, 0, stream>>>(a);
  TEB_CHECK_CUDA(cudaGetLastError());
}

// ---- backward for what the fast kernel (norm_bwd.cu) does not take: fp32 tensors, mixed dtypes, hidden sizes that
// are not a multiple of 8 or exceed 16384 (the reference's tests/cpp/operator/test_normalization.cu sweeps them).
// One CTA walks rows; per row two block reductions (mean(dy), mean(dy * xhat)), then dx; the CTA's dgamma / dbeta
// partial sums live in its own [H] slab of the workspace (a thread owns its columns, so plain read-modify-write), and
// the fast path's reduce kernel sums the slabs in CTA order. Correctness path: ~3 passes over the row.
namespace {

struct GenBwdArgs {
  const void *dz, *x, *add, *gamma;
  void *dx;
  NVTEDType xt, dzt, wt;
  const float *mu, *rsigma;
  float *ws_dgamma, *ws_dbeta;
  size_t rows, cols;
  bool layernorm, zero_centered;
};

__global__ void __launch_bounds__(256) norm_bwd_generic_kernel(const GenBwdArgs a) {
  __shared__ float red[8];
  float *wg = a.ws_dgamma + static_cast<size_t>(blockIdx.x) * a.cols;
  float *wb = a.layernorm ? a.ws_dbeta + static_cast<size_t>(blockIdx.x) * a.cols : nullptr;
  for (size_t c = threadIdx.x; c < a.cols; c += blockDim.x) {
    wg[c] = 0.0f;
    if (wb != nullptr) wb[c] = 0.0f;
  }
  const float inv_h = 1.0f / static_cast<float>(a.cols);
  for (size_t r = blockIdx.x; r < a.rows; r += gridDim.x) {
    const size_t base = r * a.cols;
    const float mu = a.layernorm ? a.mu[r] : 0.0f, rs = a.rsigma[r];
    float s1 = 0.0f, s2 = 0.0f;
    for (size_t c = threadIdx.x; c < a.cols; c += blockDim.x) {
      float g = load_f(a.gamma, a.wt, c);
      if (a.zero_centered) g += 1.0f;
      const float dzv = load_f(a.dz, a.dzt, base + c);
      const float xhat = (load_f(a.x, a.xt, base + c) - mu) * rs;
      const float dy = dzv * g;
      s1 += dy;
      s2 += dy * xhat;
      wg[c] += dzv * xhat;
      if (wb != nullptr) wb[c] += dzv;
    }
    const float m1 = a.layernorm ? block_sum(s1, red) * inv_h : 0.0f;
    const float m2 = block_sum(s2, red) * inv_h;
    for (size_t c = threadIdx.x; c < a.cols; c += blockDim.x) {
      float g = load_f(a.gamma, a.wt, c);
      if (a.zero_centered) g += 1.0f;
      const float xhat = (load_f(a.x, a.xt, base + c) - mu) * rs;
      float dxv = rs * (load_f(a.dz, a.dzt, base + c) * g - m1 - xhat * m2);
      if (a.add != nullptr) dxv += load_f(a.add, a.xt, base + c);
      store_f(a.dx, a.xt, base + c, dxv);
    }
  }
}

}  // namespace

int generic_bwd_grid(size_t N) { return static_cast<int>(std::min<size_t>(N ? N : 1, static_cast<size_t>(sm_count()) * 4)); }

void backward_generic(bool layernorm, const Tensor &dz, const Tensor &x, const Tensor *add, const Tensor *mu,
                      const Tensor &rsigma, const Tensor &gamma, Tensor &dx, float *ws_dgamma, float *ws_dbeta, int grid,
                      bool zero_centered_gamma, cudaStream_t stream) {
  TEB_CHECK(is_high_precision(x.data.dtype) && is_high_precision(dz.data.dtype) && is_high_precision(gamma.data.dtype),
            "norm backward: tensors must be fp32 / bf16 / fp16");
  GenBwdArgs a{};
  a.dz = dz.data.dptr;
  a.x = x.data.dptr;
  a.add = add ? add->data.dptr : nullptr;
  a.gamma = gamma.data.dptr;
  a.dx = dx.data.dptr;
  a.xt = x.data.dtype;
  a.dzt = dz.data.dtype;
  a.wt = gamma.data.dtype;
  a.mu = layernorm ? mu->data.as<float>() : nullptr;
  a.rsigma = rsigma.data.as<float>();
  a.ws_dgamma = ws_dgamma;
  a.ws_dbeta = ws_dbeta;
  a.rows = x.data.shape.data[0];
  a.cols = x.data.shape.data[1];
  a.layernorm = layernorm;
  a.zero_centered = zero_centered_gamma;
  norm_bwd_generic_kernel<<<grid, 256, 0, stream>>>(a);
  TEB_CHECK_CUDA(cudaGetLastError());
}

}  // namespace norm
}  // namespace teb
