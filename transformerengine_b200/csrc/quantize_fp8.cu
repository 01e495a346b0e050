// Per-tensor FP8: amax, scale-from-amax, cast (+ optional transposed copy),
// dequantize, and the delayed-scaling history/scale update.
// Replaces, for the Linear path, the reference's
//   recipe/current_scaling.cu:24-263      (amax, scale from amax)
//   cast/fp8/quantize_fp8.cuh:252-573     (cast, amax, scale_inv)
//   transpose/cast_transpose.cu:215-340   (cast + transposed copy)
//   recipe/delayed_scaling.cu:77-420      (history roll + scale update)
// Numerics per SURVEY.md section 3A. All kernels are streaming, HBM-bound:
// 16-byte loads, 4 independent vectors in flight per thread, grid sized to the
// SM count (persistent grid-stride), warp-shuffle + one atomic per warp for amax.
#include <cfloat>

#include "core.h"
#include "ptx.cuh"

namespace teb {
namespace fp8 {

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

inline int persistent_grid(size_t work_items, int ctas_per_sm) {
  const size_t cap = static_cast<size_t>(sm_count()) * ctas_per_sm;
  return static_cast<int>(work_items < cap ? (work_items ? work_items : 1) : cap);
}

// ---- amax ------------------------------------------------------------------
__global__ void zero_amax_kernel(float *amax, const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  *amax = 0.0f;
}

// n16 = number of 16-byte vectors; the scalar tail [tail_begin, n) is handled by block 0.
template <typename IT>
__global__ void __launch_bounds__(kThreads)
    amax_kernel(const IT *__restrict__ in, size_t n, size_t n16, float *amax, const float *noop) {
  using W = ptx::Word<IT>;
  if (noop != nullptr && noop[0] == 1.0f) return;
  uint32_t m0 = 0u, m1 = 0u, m2 = 0u, m3 = 0u;
  const uint4 *p = reinterpret_cast<const uint4 *>(in);
  const size_t stride = static_cast<size_t>(gridDim.x) * kThreads;
  size_t i = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
  for (; i + (kUnroll - 1) * stride < n16; i += kUnroll * stride) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) v[u] = ptx::ld_stream_v4(p + i + u * stride);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      m0 = W::max(m0, W::abs(v[u].x));
      m1 = W::max(m1, W::abs(v[u].y));
      m2 = W::max(m2, W::abs(v[u].z));
      m3 = W::max(m3, W::abs(v[u].w));
    }
  }
  for (; i < n16; i += stride) {
    const uint4 v = ptx::ld_stream_v4(p + i);
    m0 = W::max(m0, W::abs(v.x));
    m1 = W::max(m1, W::abs(v.y));
    m2 = W::max(m2, W::abs(v.z));
    m3 = W::max(m3, W::abs(v.w));
  }
  float m = W::hmax(W::max(W::max(m0, m1), W::max(m2, m3)));
  if (blockIdx.x == 0) {
    const size_t per = 16 / sizeof(IT);
    for (size_t j = n16 * per + threadIdx.x; j < n; j += kThreads)
      m = fmaxf(m, fabsf(static_cast<float>(in[j])));
  }
  m = ptx::warp_max(m);
  __shared__ float wmax[kThreads / 32];
  if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < kThreads / 32 ? wmax[threadIdx.x] : 0.0f;
    m = ptx::warp_max(m);
    if (threadIdx.x == 0 && m > 0.0f) ptx::atomic_max_nonneg(amax, m);
  }
}

// ---- scale from amax (reference recipe/recipe_common.cuh:14-50) --------------
__global__ void scale_from_amax_kernel(const float *amax, float *scale, float max_fp8,
                                       bool force_pow2, float eps, const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  float a = *amax;
  if (a < eps) a = eps;
  float s = 1.0f;
  if (!(isinf(a) || a == 0.0f || isnan(a))) {
    s = __fdiv_rn(max_fp8, a);
    if (isinf(s)) s = FLT_MAX;
    if (force_pow2) s = __uint_as_float(__float_as_uint(s) & 0xFF800000u);
  }
  *scale = s;
}

// ---- cast --------------------------------------------------------------------
template <typename IT, bool E5M2>
__device__ __forceinline__ void cast_vec(const uint4 &v, float s, uint32_t (&q)[2], int &nwords,
                                         uint32_t &m0) {
  using W = ptx::Word<IT>;
  if (W::kElems == 2) {
    q[0] = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v.x), W::hi(v.x), s) |
           (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v.y), W::hi(v.y), s) << 16);
    q[1] = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v.z), W::hi(v.z), s) |
           (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v.w), W::hi(v.w), s) << 16);
    nwords = 2;
  } else {
    q[0] = ptx::mul_cvt_fp8x2<E5M2>(W::lo(v.x), W::lo(v.y), s) |
           (ptx::mul_cvt_fp8x2<E5M2>(W::lo(v.z), W::lo(v.w), s) << 16);
    nwords = 1;
  }
  m0 = W::max(m0, W::max(W::max(W::abs(v.x), W::abs(v.y)), W::max(W::abs(v.z), W::abs(v.w))));
}

template <typename IT, bool E5M2>
__global__ void __launch_bounds__(kThreads)
    cast_kernel(const IT *__restrict__ in, uint8_t *__restrict__ out, size_t n, size_t n16,
                const float *scale_ptr, float *amax, float *scale_inv, const float *noop) {
  using W = ptx::Word<IT>;
  if (noop != nullptr && noop[0] == 1.0f) return;
  const float s = scale_ptr ? *scale_ptr : 1.0f;
  if (blockIdx.x == 0 && threadIdx.x == 0 && scale_inv != nullptr) *scale_inv = __frcp_rn(s);
  constexpr int kOutBytes = 16 / sizeof(IT);  // fp8 bytes per 16-byte input vector
  uint32_t m0 = 0u;
  const uint4 *p = reinterpret_cast<const uint4 *>(in);
  const size_t stride = static_cast<size_t>(gridDim.x) * kThreads;
  size_t i = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x;
  for (; i + (kUnroll - 1) * stride < n16; i += kUnroll * stride) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) v[u] = ptx::ld_stream_v4(p + i + u * stride);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      uint32_t q[2];
      int nw;
      cast_vec<IT, E5M2>(v[u], s, q, nw, m0);
      uint8_t *dst = out + (i + u * stride) * kOutBytes;
      if (kOutBytes == 8) *reinterpret_cast<uint2 *>(dst) = make_uint2(q[0], q[1]);
      else *reinterpret_cast<uint32_t *>(dst) = q[0];
    }
  }
  for (; i < n16; i += stride) {
    uint32_t q[2];
    int nw;
    cast_vec<IT, E5M2>(ptx::ld_stream_v4(p + i), s, q, nw, m0);
    uint8_t *dst = out + i * kOutBytes;
    if (kOutBytes == 8) *reinterpret_cast<uint2 *>(dst) = make_uint2(q[0], q[1]);
    else *reinterpret_cast<uint32_t *>(dst) = q[0];
  }
  float m = W::hmax(m0);
  if (blockIdx.x == 0) {
    for (size_t j = n16 * kOutBytes + threadIdx.x; j < n; j += kThreads) {
      const float x = static_cast<float>(in[j]);
      m = fmaxf(m, fabsf(x));
      out[j] = static_cast<uint8_t>(ptx::cvt_fp8x2<E5M2>(x * s, 0.0f));
    }
  }
  if (amax != nullptr) {
    m = ptx::warp_max(m);
    if ((threadIdx.x & 31) == 0 && m > 0.0f) ptx::atomic_max_nonneg(amax, m);
  }
}

// Cast + push: every 16-byte output vector also goes to up to 8 peer-mapped destinations (the other
// ranks' all-gather buffers over NVLink). One thread = two adjacent 16-byte input vectors = one
// 16-byte store per destination; a warp writes 512 contiguous bytes per destination.
struct PeerDsts {
  uint8_t *p[8];
  int n;
};
template <typename IT, bool E5M2>
__global__ void __launch_bounds__(kThreads)
    cast_push_kernel(const IT *__restrict__ in, uint8_t *__restrict__ out, const PeerDsts peers, size_t n32,
                     const float *scale_ptr, float *amax, float *scale_inv, const float *noop) {
  static_assert(sizeof(IT) == 2, "16-bit inputs");
  using W = ptx::Word<IT>;
  if (noop != nullptr && noop[0] == 1.0f) return;
  const float s = scale_ptr ? *scale_ptr : 1.0f;
  if (blockIdx.x == 0 && threadIdx.x == 0 && scale_inv != nullptr) *scale_inv = __frcp_rn(s);
  uint32_t m0 = 0u;
  const uint4 *p = reinterpret_cast<const uint4 *>(in);
  const size_t stride = static_cast<size_t>(gridDim.x) * kThreads;
  for (size_t i = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x; i < n32; i += stride) {
    const uint4 v0 = ptx::ld_stream_v4(p + 2 * i), v1 = ptx::ld_stream_v4(p + 2 * i + 1);
    uint32_t q0[2], q1[2];
    int nw;
    cast_vec<IT, E5M2>(v0, s, q0, nw, m0);
    cast_vec<IT, E5M2>(v1, s, q1, nw, m0);
    const uint4 w = make_uint4(q0[0], q0[1], q1[0], q1[1]);
    reinterpret_cast<uint4 *>(out)[i] = w;
#pragma unroll
    for (int d = 0; d < 8; ++d)
      if (d < peers.n) reinterpret_cast<uint4 *>(peers.p[d])[i] = w;
  }
  if (amax != nullptr) {
    float m = ptx::warp_max(W::hmax(m0));
    if ((threadIdx.x & 31) == 0 && m > 0.0f) ptx::atomic_max_nonneg(amax, m);
  }
}

// Cast + per-slab column sums (nvte_quantize_dbias, per-tensor FP8): one thread = one 16-byte
// vector of adjacent columns walked down a 32-row slab; the slab's partial sums go to ws[slab][col].
template <typename IT, bool E5M2>
__global__ void __launch_bounds__(256)
    cast_dbias_kernel(const IT *__restrict__ in, uint8_t *__restrict__ out, float *__restrict__ ws,
                      size_t rows, size_t cols, const float *scale_ptr, float *amax, float *scale_inv) {
  using W = ptx::Word<IT>;
  constexpr int VEC = 16 / sizeof(IT);
  const float s = scale_ptr ? *scale_ptr : 1.0f;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && scale_inv != nullptr) *scale_inv = __frcp_rn(s);
  const size_t c = (static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x) * VEC;
  uint32_t m0 = 0u;
  if (c < cols) {
    const size_t r0 = static_cast<size_t>(blockIdx.y) * 32;
    const size_t r1 = r0 + 32 < rows ? r0 + 32 : rows;
    float sum[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) sum[k] = 0.0f;
    for (size_t r = r0; r < r1; ++r) {
      const uint4 v = ptx::ld_stream_v4(in + r * cols + c);
      uint32_t q[2];
      int nw;
      cast_vec<IT, E5M2>(v, s, q, nw, m0);
      uint8_t *dst = out + r * cols + c;
      if (VEC == 8) *reinterpret_cast<uint2 *>(dst) = make_uint2(q[0], q[1]);
      else *reinterpret_cast<uint32_t *>(dst) = q[0];
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (W::kElems == 2) {
          sum[2 * k] += W::lo(w[k]);
          sum[2 * k + 1] += W::hi(w[k]);
        } else {
          sum[k] += W::lo(w[k]);
        }
      }
    }
    float *dst = ws + static_cast<size_t>(blockIdx.y) * cols + c;
#pragma unroll
    for (int k = 0; k < VEC; ++k) dst[k] = sum[k];
  }
  if (amax != nullptr) {
    const float m = ptx::warp_max(W::hmax(m0));
    if ((threadIdx.x & 31) == 0 && m > 0.0f) ptx::atomic_max_nonneg(amax, m);
  }
}

// Cast with transposed copy (and optional rowwise copy): 64x64 tile through
// padded shared memory. Only reached when a caller hands a columnwise buffer
// to a per-tensor FP8 tensor (pre-Blackwell layout); B200 GEMMs do not need it.
template <typename IT, bool E5M2>
__global__ void __launch_bounds__(256)
    cast_transpose_kernel(const IT *__restrict__ in, uint8_t *out, uint8_t *out_t, size_t rows,
                          size_t cols, const float *scale_ptr, float *amax, float *scale_inv,
                          const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  __shared__ uint8_t tile[64][65];
  const float s = scale_ptr ? *scale_ptr : 1.0f;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && scale_inv != nullptr)
    *scale_inv = __frcp_rn(s);
  const size_t r0 = static_cast<size_t>(blockIdx.y) * 64, c0 = static_cast<size_t>(blockIdx.x) * 64;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  float m = 0.0f;
  for (int r = ty; r < 64; r += 4) {
    const size_t rr = r0 + r, cc = c0 + tx;
    uint8_t q = 0;
    if (rr < rows && cc < cols) {
      const float x = static_cast<float>(in[rr * cols + cc]);
      m = fmaxf(m, fabsf(x));
      q = static_cast<uint8_t>(ptx::cvt_fp8x2<E5M2>(x * s, 0.0f));
      if (out != nullptr) out[rr * cols + cc] = q;
    }
    tile[r][tx] = q;
  }
  __syncthreads();
  for (int c = ty; c < 64; c += 4) {
    const size_t cc = c0 + c, rr = r0 + tx;
    if (rr < rows && cc < cols) out_t[cc * rows + rr] = tile[tx][c];
  }
  if (amax != nullptr) {
    m = ptx::warp_max(m);
    if ((threadIdx.x & 31) == 0 && m > 0.0f) ptx::atomic_max_nonneg(amax, m);
  }
}

// nvte_quantize with a HIGH-PRECISION output type (reference transpose/cast_transpose.cu accepts any output
// type: its test_cast_transpose.cu runs bf16 / fp16 / fp32 outputs): plain cast into the rowwise buffer and a
// transposed copy into the columnwise one, scale = 1, no amax. Not a hot path: 64 x 64 tiles through shared memory.
template <typename IT, typename OT>
__global__ void __launch_bounds__(256)
    hp_cast_transpose_kernel(const IT *__restrict__ in, OT *out, OT *out_t, size_t rows, size_t cols,
                             const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  __shared__ float tile[64][65];
  const size_t r0 = static_cast<size_t>(blockIdx.y) * 64, c0 = static_cast<size_t>(blockIdx.x) * 64;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  for (int r = ty; r < 64; r += 4) {
    const size_t rr = r0 + r, cc = c0 + tx;
    float x = 0.0f;
    if (rr < rows && cc < cols) {
      x = static_cast<float>(in[rr * cols + cc]);
      if (out != nullptr) out[rr * cols + cc] = static_cast<OT>(x);
    }
    tile[r][tx] = x;
  }
  if (out_t == nullptr) return;
  __syncthreads();
  for (int c = ty; c < 64; c += 4) {
    const size_t cc = c0 + c, rr = r0 + tx;
    if (rr < rows && cc < cols) out_t[cc * rows + rr] = static_cast<OT>(tile[tx][c]);
  }
}

// ---- dequantize: out = fp8 * scale_inv ----------------------------------------
template <typename OT, bool E5M2>
__global__ void __launch_bounds__(kThreads)
    dequantize_kernel(const uint8_t *__restrict__ in, OT *__restrict__ out, size_t n,
                      const float *scale_inv) {
  const float s = *scale_inv;
  const size_t stride = static_cast<size_t>(gridDim.x) * kThreads;
  const size_t n2 = n / 2;
  for (size_t i = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x; i < n2; i += stride) {
    const float2 f = ptx::fp8x2_to_float2<E5M2>(reinterpret_cast<const uint16_t *>(in)[i]);
    out[2 * i] = static_cast<OT>(f.x * s);
    out[2 * i + 1] = static_cast<OT>(f.y * s);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && (n & 1)) {
    const float2 f = ptx::fp8x2_to_float2<E5M2>(in[n - 1]);
    out[n - 1] = static_cast<OT>(f.x * s);
  }
}

// ---- delayed scaling (reference recipe/delayed_scaling.cu:77-253) ------------
struct BulkEntry {
  float *history;
  float *scale;
  int num_scale;
  int offset;
};
constexpr int kBulkMax = 1024;
struct BulkParams {
  BulkEntry e[kBulkMax];
};

// One warp per scale. history: [len, stride]; in-place roll allowed.
__global__ void __launch_bounds__(kThreads)
    delayed_update_kernel(const float *hist_in, const float *scale_in, float *hist_out,
                          float *scale_out, size_t len, size_t stride, size_t num_scales,
                          bool most_recent, float scaled_max) {
  const size_t sidx = static_cast<size_t>(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
  if (sidx >= num_scales) return;
  const int lane = threadIdx.x & 31;
  const float last = hist_in[sidx];
  float amax = 0.0f;
  for (size_t base = 0; base < len; base += 32) {
    const size_t i = base + lane;
    float a = 0.0f;
    if (i < len) {
      a = (i + 1 < len) ? hist_in[(i + 1) * stride + sidx] : last;
      amax = fmaxf(amax, a);
    }
    __syncwarp();  // in-place roll: all reads of this chunk happen before its writes
    if (i < len) hist_out[i * stride + sidx] = (i > 0) ? a : 0.0f;
  }
  amax = most_recent ? last : ptx::warp_max(amax);
  if (lane == 0) {
    float s = (isfinite(amax) && amax > 0.0f) ? scaled_max / amax : scale_in[sidx];
    if (isinf(s)) s = FLT_MAX;
    scale_out[sidx] = s;
  }
}

__global__ void __launch_bounds__(kThreads)
    delayed_update_bulk_kernel(const float *reduction_buffer, BulkParams p, int num_tensors,
                               size_t len, bool most_recent, float scaled_max) {
  const int t = blockIdx.x;
  if (t >= num_tensors) return;
  const BulkEntry e = p.e[t];
  const int lane = threadIdx.x & 31;
  for (int sidx = threadIdx.x >> 5; sidx < e.num_scale; sidx += kThreads / 32) {
    float *hist = e.history + sidx;
    const size_t stride = e.num_scale;
    float last = hist[0];
    if (reduction_buffer != nullptr && reduction_buffer[e.offset + sidx] != 0.0f)
      last = reduction_buffer[e.offset + sidx];
    float amax = 0.0f;
    if (last != 0.0f) {
      for (size_t base = 0; base < len; base += 32) {
        const size_t i = base + lane;
        float a = 0.0f;
        if (i < len) {
          a = (i + 1 < len) ? hist[(i + 1) * stride] : last;
          amax = fmaxf(amax, a);
        }
        __syncwarp();
        if (i < len) hist[i * stride] = (i > 0) ? a : 0.0f;
      }
    }
    amax = most_recent ? last : ptx::warp_max(amax);
    if (lane == 0) {
      float s = (isfinite(amax) && amax > 0.0f) ? scaled_max / amax : e.scale[sidx];
      if (isinf(s)) s = FLT_MAX;
      e.scale[sidx] = s;
    }
  }
}

template <typename F>
void dispatch_in(NVTEDType t, F &&f) {
  switch (t) {
    case kNVTEBFloat16: f(__nv_bfloat16{}); break;
    case kNVTEFloat16: f(__half{}); break;
    case kNVTEFloat32: f(float{}); break;
    default: TEB_ERROR("Unsupported high-precision dtype ", dtype_name(t));
  }
}

}  // namespace

void compute_amax(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream) {
  TEB_CHECK(in.scaling_mode == NVTE_DELAYED_TENSOR_SCALING && is_high_precision(in.data.dtype),
            "Input tensor for amax computation must be unquantized, but got dtype=",
            dtype_name(in.data.dtype));
  TEB_CHECK(in.data.dptr != nullptr || in.numel() == 0, "Input tensor for amax has no data");
  TEB_CHECK(out.scaling_mode == NVTE_DELAYED_TENSOR_SCALING ||
                out.scaling_mode == NVTE_NVFP4_1D_SCALING,
            "Output tensor for amax computation must be FP8 tensor with per-tensor scaling or "
            "NVFP4 1D scaling");
  // NVFP4 without rowwise usage keeps its amax in the columnwise slot
  // (reference recipe/current_scaling.cu:118-176).
  float *amax = out.amax.as<float>();
  if (amax == nullptr) amax = out.columnwise_amax.as<float>();
  TEB_CHECK(amax != nullptr, "Output tensor for amax computation has no amax buffer");
  const float *np = noop ? noop->data.as<float>() : nullptr;
  zero_amax_kernel<<<1, 1, 0, stream>>>(amax, np);
  const size_t n = in.numel();
  if (n > 0) {
    dispatch_in(in.data.dtype, [&](auto tag) {
      using IT = decltype(tag);
      const bool aligned = reinterpret_cast<uintptr_t>(in.data.dptr) % 16 == 0;
      const size_t n16 = aligned ? n * sizeof(IT) / 16 : 0;
      const int grid = persistent_grid(ceil_div(n16 ? n16 : 1, kThreads * kUnroll), 8);
      amax_kernel<IT><<<grid, kThreads, 0, stream>>>(in.data.as<IT>(), n, n16, amax, np);
    });
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}

// amax all-reduce without a collective library, step 1 of 2: every rank stores its amax into its own
// slot of every rank's slot array (peer-mapped addresses); after a barrier each rank takes the max
// of the slots locally.
__global__ void push_scalar_kernel(const float *src, const PeerDsts peers, const float *noop) {
  if (noop != nullptr && noop[0] == 1.0f) return;
  const float v = *src;
  if (static_cast<int>(threadIdx.x) < peers.n) *reinterpret_cast<float *>(peers.p[threadIdx.x]) = v;
}
void push_amax(const Tensor &out, const uint64_t *peer_slots, const Tensor *noop, cudaStream_t stream) {
  TEB_CHECK(out.amax.dptr != nullptr, "Output needs an amax buffer");
  PeerDsts peers{};
  peers.n = static_cast<int>(peer_slots[0]);
  for (int d = 0; d < peers.n; ++d) {
    peers.p[d] = reinterpret_cast<uint8_t *>(peer_slots[1 + d]);
    TEB_CHECK(peers.p[d] != nullptr && peer_slots[1 + d] % 4 == 0, "Peer amax slot ", d, " is null or misaligned");
  }
  push_scalar_kernel<<<1, 32, 0, stream>>>(out.amax.as<float>(), peers, noop ? noop->data.as<float>() : nullptr);
  TEB_CHECK_CUDA(cudaGetLastError());
}

// ---- symmetric exchange (NVTEB200SymmetricExchange): the amax all-reduce and the all-gather of the
// sequence-parallel column-parallel Linear as flag-carrying stores into peer-mapped memory. No collective, no
// barrier kernel: step e of every rank pairs up through (value, e) words.
__global__ void push_amax_step_kernel(const float *amax, const NVTEB200SymmetricExchange x) {
  const uint32_t e = *reinterpret_cast<const volatile uint32_t *>(x.step) + 1u;
  const uint64_t word = (static_cast<uint64_t>(e) << 32) | __float_as_uint(*amax);
  if (threadIdx.x < x.world)
    ptx::st_release_sys_u64(reinterpret_cast<uint64_t *>(x.amax_slots[threadIdx.x]) + (e & 1u) * x.world + x.rank, word);
}
void push_amax_step(const Tensor &out, const NVTEB200SymmetricExchange &x, cudaStream_t stream) {
  TEB_CHECK(out.amax.dptr != nullptr, "Output needs an amax buffer");
  push_amax_step_kernel<<<1, 32, 0, stream>>>(out.amax.as<float>(), x);
  TEB_CHECK_CUDA(cudaGetLastError());
}

// waits for the amax of step e from every rank of the exchange, max over them -> amax, then the scale (and its
// reciprocal when the output has a scale_inv). flags_per_rank == 0 (the GEMM pushes the rows itself, the cast is
// an ordinary local cast): this kernel also advances the step counter.
__global__ void scale_from_slots_kernel(float *amax, float *scale, float *scale_inv, float max_fp8, bool force_pow2,
                                        float eps, const NVTEB200SymmetricExchange x) {
  const uint32_t e = *reinterpret_cast<const volatile uint32_t *>(x.step) + 1u;
  const uint64_t *slots = reinterpret_cast<const uint64_t *>(x.amax_slots[x.rank]) + (e & 1u) * x.world;
  float a = 0.0f;
  if (threadIdx.x < x.world) {
    uint64_t v = ptx::ld_acquire_sys_u64(slots + threadIdx.x);
    const uint64_t t0 = ptx::globaltimer_ns();
    while (static_cast<uint32_t>(v >> 32) != e) {
      __nanosleep(64);
      v = ptx::ld_acquire_sys_u64(slots + threadIdx.x);
      if (ptx::globaltimer_ns() - t0 > 5000000000ull) {
        *reinterpret_cast<uint32_t *>(x.error) = 1u;
        break;
      }
    }
    a = __uint_as_float(static_cast<uint32_t>(v));
  }
  a = ptx::warp_max(a);
  if (threadIdx.x == 0) {
    *amax = a;
    if (a < eps) a = eps;
    float s = 1.0f;
    if (!(isinf(a) || a == 0.0f || isnan(a))) {
      s = __fdiv_rn(max_fp8, a);
      if (isinf(s)) s = FLT_MAX;
      if (force_pow2) s = __uint_as_float(__float_as_uint(s) & 0xFF800000u);
    }
    *scale = s;
    if (scale_inv != nullptr) *scale_inv = __frcp_rn(s);
    if (x.flags_per_rank == 0u) *reinterpret_cast<volatile uint32_t *>(x.step) = e;
  }
}

// The cast of this rank's shard [flags_per_rank * rows_per_flag, cols] into EVERY rank's gather buffer (x.gather[r],
// r = rank included), chunk by chunk; when the last CTA has finished chunk c (its stores fenced at system scope) it
// raises flag [rank * flags_per_rank + c] = e on every rank, and after the last chunk advances the step counter.
// (A GEMM cannot run next to this kernel: its 10 warps of 168 registers fill three of the four register-file
// partitions of an SM. The overlapped form of the exchange is the GEMM that pushes the rows itself,
// NVTEB200ArrivalFlags.push_unit_bytes.)
constexpr int kXchWorkers = 256;             // casting threads of a CTA
constexpr int kXchThreads = kXchWorkers + 32;  // + the publisher warp

// Two roles per CTA. The 8 worker warps cast and store chunk after chunk WITHOUT ever waiting for a flag to go
// out (a system-scope fence behind a stream of NVLink stores takes microseconds): after a chunk each warp bumps a
// shared counter (release, CTA scope). Lane 0 of the publisher warp trails them: once the CTA's warps are past
// chunk c it fences at system scope (cumulative over the workers' stores it observed through the counter), counts
// the CTA in, and the last CTA of the grid raises the flags.
template <typename IT, bool E5M2>
__global__ void __launch_bounds__(kXchThreads)
    cast_exchange_kernel(const IT *__restrict__ in, size_t vec_per_chunk, const NVTEB200SymmetricExchange x,
                         const float *scale_ptr, float *scale_inv) {
  static_assert(sizeof(IT) == 2, "16-bit inputs");
  __shared__ uint32_t s_warps_done;   // worker warps x chunks finished so far
  if (threadIdx.x == 0) s_warps_done = 0u;
  __syncthreads();
  const uint32_t e = *reinterpret_cast<const volatile uint32_t *>(x.step) + 1u;
  const uint32_t s_addr = static_cast<uint32_t>(__cvta_generic_to_shared(&s_warps_done));
  if (threadIdx.x >= kXchWorkers) {
    if (threadIdx.x != kXchWorkers) return;
    uint32_t *done = reinterpret_cast<uint32_t *>(x.done);
    for (uint32_t c = 0; c < x.flags_per_rank; ++c) {
      const uint32_t want = (kXchWorkers / 32) * (c + 1u);
      uint32_t have;
      do {
        asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(have) : "r"(s_addr) : "memory");
        if (have < want) __nanosleep(100);
      } while (have < want);
      __threadfence_system();
      if (atomicAdd(done + c, 1u) == gridDim.x - 1) {
        done[c] = 0u;
        __threadfence_system();
        for (uint32_t r = 0; r < x.world; ++r)
          ptx::st_release_sys_u32(reinterpret_cast<uint32_t *>(x.flags[r]) + x.rank * x.flags_per_rank + c, e);
        if (c + 1 == x.flags_per_rank) *reinterpret_cast<volatile uint32_t *>(x.step) = e;
      }
    }
    return;
  }
  const float s = scale_ptr ? *scale_ptr : 1.0f;
  if (blockIdx.x == 0 && threadIdx.x == 0 && scale_inv != nullptr) *scale_inv = __frcp_rn(s);
  const uint4 *p = reinterpret_cast<const uint4 *>(in);
  const size_t stride = static_cast<size_t>(gridDim.x) * kXchWorkers;
  auto one = [&](size_t v, const uint4 &v0, const uint4 &v1) {
    uint32_t q0[2], q1[2], m0 = 0u;
    int nw;
    cast_vec<IT, E5M2>(v0, s, q0, nw, m0);
    cast_vec<IT, E5M2>(v1, s, q1, nw, m0);
    const uint4 w = make_uint4(q0[0], q0[1], q1[0], q1[1]);
#pragma unroll
    for (int r = 0; r < 8; ++r)
      if (r < static_cast<int>(x.world)) reinterpret_cast<uint4 *>(x.gather[r])[v] = w;
  };
  for (uint32_t c = 0; c < x.flags_per_rank; ++c) {
    const size_t base = static_cast<size_t>(c) * vec_per_chunk;
    size_t i = static_cast<size_t>(blockIdx.x) * kXchWorkers + threadIdx.x;
    for (; i + stride < vec_per_chunk; i += 2 * stride) {   // two 32-byte loads in flight per thread
      const size_t va = base + i, vb = va + stride;
      const uint4 a0 = ptx::ld_stream_v4(p + 2 * va), a1 = ptx::ld_stream_v4(p + 2 * va + 1);
      const uint4 b0 = ptx::ld_stream_v4(p + 2 * vb), b1 = ptx::ld_stream_v4(p + 2 * vb + 1);
      one(va, a0, a1);
      one(vb, b0, b1);
    }
    if (i < vec_per_chunk) {
      const size_t va = base + i;
      one(va, ptx::ld_stream_v4(p + 2 * va), ptx::ld_stream_v4(p + 2 * va + 1));
    }
    __syncwarp();
    if ((threadIdx.x & 31u) == 0u)
      asm volatile("red.release.cta.shared.add.u32 [%0], 1;" ::"r"(s_addr) : "memory");
  }
}

void quantize_exchange(const Tensor &in, Tensor &out, const NVTEB200SymmetricExchange &x, cudaStream_t stream) {
  TEB_CHECK(in.has_data() && (in.data.dtype == kNVTEBFloat16 || in.data.dtype == kNVTEFloat16),
            "The symmetric-exchange cast needs a bf16 / fp16 input");
  TEB_CHECK(is_fp8(out.dtype()) && out.columnwise_data.dptr == nullptr, "The symmetric-exchange cast writes rowwise FP8 only");
  TEB_CHECK(out.scale.dptr != nullptr, "Per-tensor FP8 output needs a scale tensor");
  const size_t rows = in.rows(), cols = in.cols();
  TEB_CHECK(rows == static_cast<size_t>(x.flags_per_rank) * x.rows_per_flag, "Shard has ", rows, " rows, the exchange expects ",
            static_cast<size_t>(x.flags_per_rank) * x.rows_per_flag);
  TEB_CHECK(cols % 16 == 0 && reinterpret_cast<uintptr_t>(in.data.dptr) % 16 == 0,
            "The symmetric-exchange cast needs 16-element rows and a 16-byte aligned input");
  TEB_CHECK(out.data.dptr == reinterpret_cast<void *>(x.gather[x.rank]) && out.data.numel() == rows * cols,
            "The symmetric-exchange cast writes the rank's own rows of its gather buffer: output data must be that window");
  const size_t vec_per_chunk = static_cast<size_t>(x.rows_per_flag) * cols / 16;
  const bool e5m2 = out.dtype() == kNVTEFloat8E5M2;
  const int grid = static_cast<int>(std::min<size_t>(ceil_div(vec_per_chunk, static_cast<size_t>(kXchWorkers)),
                                                     static_cast<size_t>(2 * sm_count())));
  dispatch_in(in.data.dtype, [&](auto tag) {
    using IT = decltype(tag);
    if constexpr (sizeof(IT) == 2) {
      auto go = [&](auto kernel) {
        kernel<<<grid, kXchThreads, 0, stream>>>(in.data.as<IT>(), vec_per_chunk, x, out.scale.as<float>(),
                                                out.scale_inv.as<float>());
      };
      if (e5m2) go(cast_exchange_kernel<IT, true>);
      else go(cast_exchange_kernel<IT, false>);
    }
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}

void compute_scale_from_amax(Tensor &out, const QuantConfig &cfg, cudaStream_t stream) {
  TEB_CHECK(out.scaling_mode == NVTE_DELAYED_TENSOR_SCALING,
            "Tensor must be FP8 tensor with per-tensor scaling");
  TEB_CHECK(is_fp8(out.data.dtype), "Tensor must be FP8, but got dtype=",
            dtype_name(out.data.dtype));
  TEB_CHECK(out.amax.numel() == 1 && out.amax.dptr != nullptr,
            "Tensor has invalid amax tensor (expected 1 entry)");
  TEB_CHECK(out.scale.numel() == 1 && out.scale.dptr != nullptr,
            "Tensor has invalid scale tensor (expected 1 entry)");
  const float max_fp8 = out.data.dtype == kNVTEFloat8E4M3 ? 448.0f : 57344.0f;
  const Tensor *noop = cfg.noop_tensor ? lookup(cfg.noop_tensor) : nullptr;
  if (cfg.symm.world > 1) {   // the amax is the max over every rank's slot of this step (no collective)
    scale_from_slots_kernel<<<1, 32, 0, stream>>>(out.amax.as<float>(), out.scale.as<float>(), out.scale_inv.as<float>(),
                                                  max_fp8, cfg.force_pow_2_scales, cfg.amax_epsilon, cfg.symm);
    TEB_CHECK_CUDA(cudaGetLastError());
    return;
  }
  scale_from_amax_kernel<<<1, 1, 0, stream>>>(out.amax.as<float>(), out.scale.as<float>(), max_fp8,
                                              cfg.force_pow_2_scales, cfg.amax_epsilon,
                                              noop ? noop->data.as<float>() : nullptr);
  TEB_CHECK_CUDA(cudaGetLastError());
}

// Per-tensor FP8 cast whose rowwise bytes also go to peer-mapped buffers (QuantConfig extension
// kNVTEQuantizationConfigB200PeerRowwiseData): the cast of a sequence-parallel shard is its own
// all-gather -- stores travel over NVLink while the local copy is written, no separate collective.
void quantize_push(const Tensor &in, Tensor &out, const uint64_t *peer_rowwise, const Tensor *noop,
                   cudaStream_t stream) {
  TEB_CHECK(in.has_data() && (in.data.dtype == kNVTEBFloat16 || in.data.dtype == kNVTEFloat16),
            "Cast with peer destinations needs a bf16 / fp16 input");
  TEB_CHECK(out.data.dptr != nullptr && out.columnwise_data.dptr == nullptr && is_fp8(out.data.dtype),
            "Cast with peer destinations writes rowwise FP8 only");
  TEB_CHECK(out.scale.dptr != nullptr, "Per-tensor FP8 output needs a scale tensor");
  const size_t n = in.numel();
  TEB_CHECK(out.data.numel() == n, "Input and output shapes need to match.");
  TEB_CHECK(n % 16 == 0 && reinterpret_cast<uintptr_t>(in.data.dptr) % 16 == 0 &&
                reinterpret_cast<uintptr_t>(out.data.dptr) % 16 == 0,
            "Cast with peer destinations needs 16-element multiples and 16-byte aligned buffers");
  PeerDsts peers{};
  peers.n = static_cast<int>(peer_rowwise[0]);
  for (int d = 0; d < peers.n; ++d) {
    peers.p[d] = reinterpret_cast<uint8_t *>(peer_rowwise[1 + d]);
    TEB_CHECK(peers.p[d] != nullptr && reinterpret_cast<uintptr_t>(peers.p[d]) % 16 == 0,
              "Peer destination ", d, " is null or not 16-byte aligned");
  }
  if (n == 0) return;
  const bool e5m2 = out.data.dtype == kNVTEFloat8E5M2;
  const float *np = noop ? noop->data.as<float>() : nullptr;
  const size_t n32 = n / 16;
  const int grid = persistent_grid(ceil_div(n32, static_cast<size_t>(kThreads)), 8);
  dispatch_in(in.data.dtype, [&](auto tag) {
    using IT = decltype(tag);
    if constexpr (sizeof(IT) == 2) {
      if (e5m2)
        cast_push_kernel<IT, true><<<grid, kThreads, 0, stream>>>(in.data.as<IT>(), out.data.as<uint8_t>(), peers,
                                                                 n32, out.scale.as<float>(), out.amax.as<float>(),
                                                                 out.scale_inv.as<float>(), np);
      else
        cast_push_kernel<IT, false><<<grid, kThreads, 0, stream>>>(in.data.as<IT>(), out.data.as<uint8_t>(), peers,
                                                                  n32, out.scale.as<float>(), out.amax.as<float>(),
                                                                  out.scale_inv.as<float>(), np);
    }
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}

void quantize(const Tensor &in, Tensor &out, const Tensor *noop, cudaStream_t stream) {
  TEB_CHECK(in.has_data(), "Cannot quantize tensor without rowwise data.");
  TEB_CHECK(is_high_precision(in.data.dtype), "Input must be in higher precision.");
  const bool row = out.data.dptr != nullptr, col = out.columnwise_data.dptr != nullptr;
  TEB_CHECK(row || col, "Either rowwise or columnwise output data need to be allocated.");
  const NVTEDType otype = out.dtype();
  const size_t n = in.numel();
  if (row) TEB_CHECK(out.data.numel() == n, "Input and output shapes need to match.");
  if (col) TEB_CHECK(out.columnwise_data.numel() == n, "Input and output shapes need to match.");
  if (is_high_precision(otype)) {   // cast (+ transposed copy) between high-precision types
    if (n == 0) return;
    const size_t rows = in.rows(), cols = in.cols();
    const dim3 grid(ceil_div(cols, 64), ceil_div(rows, 64));
    const float *np = noop ? noop->data.as<float>() : nullptr;
    dispatch_in(in.data.dtype, [&](auto itag) {
      using IT = decltype(itag);
      auto go = [&](auto otag) {
        using OT = decltype(otag);
        hp_cast_transpose_kernel<IT, OT><<<grid, 256, 0, stream>>>(in.data.as<IT>(), out.data.as<OT>(),
                                                                   out.columnwise_data.as<OT>(), rows, cols, np);
      };
      if (otype == kNVTEFloat32) go(float{});
      else if (otype == kNVTEBFloat16) go(__nv_bfloat16{});
      else go(__half{});
    });
    TEB_CHECK_CUDA(cudaGetLastError());
    return;
  }
  TEB_CHECK(is_fp8(otype), "Output must have FP8 type.");
  TEB_CHECK(out.scale.dptr != nullptr, "Per-tensor FP8 output needs a scale tensor");
  if (n == 0) return;
  const bool e5m2 = otype == kNVTEFloat8E5M2;
  const float *np = noop ? noop->data.as<float>() : nullptr;
  const float *scale = out.scale.as<float>();
  float *amax = out.amax.as<float>();
  float *sinv = out.scale_inv.as<float>();
  dispatch_in(in.data.dtype, [&](auto tag) {
    using IT = decltype(tag);
    const IT *src = in.data.as<IT>();
    if (col) {
      const size_t rows = in.rows(), cols = in.cols();
      const dim3 grid(ceil_div(cols, 64), ceil_div(rows, 64));
      if (e5m2)
        cast_transpose_kernel<IT, true><<<grid, 256, 0, stream>>>(
            src, out.data.as<uint8_t>(), out.columnwise_data.as<uint8_t>(), rows, cols, scale, amax,
            sinv, np);
      else
        cast_transpose_kernel<IT, false><<<grid, 256, 0, stream>>>(
            src, out.data.as<uint8_t>(), out.columnwise_data.as<uint8_t>(), rows, cols, scale, amax,
            sinv, np);
      return;
    }
    const bool aligned = reinterpret_cast<uintptr_t>(src) % 16 == 0 &&
                         reinterpret_cast<uintptr_t>(out.data.dptr) % 8 == 0;
    const size_t n16 = aligned ? n * sizeof(IT) / 16 : 0;
    const int grid = persistent_grid(ceil_div(n16 ? n16 : 1, kThreads * kUnroll), 8);
    if (e5m2)
      cast_kernel<IT, true><<<grid, kThreads, 0, stream>>>(src, out.data.as<uint8_t>(), n, n16,
                                                          scale, amax, sinv, np);
    else
      cast_kernel<IT, false><<<grid, kThreads, 0, stream>>>(src, out.data.as<uint8_t>(), n, n16,
                                                           scale, amax, sinv, np);
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}

// nvte_quantize_dbias, per-tensor FP8, rowwise output: cast + slab partial sums in one pass.
void quantize_dbias_slabs(const Tensor &in, Tensor &out, float *ws, cudaStream_t stream) {
  TEB_CHECK(out.data.dptr != nullptr && is_fp8(out.data.dtype), "Output must have FP8 rowwise data.");
  TEB_CHECK(out.scale.dptr != nullptr, "Per-tensor FP8 output needs a scale tensor");
  const size_t rows = in.rows(), cols = in.cols();
  TEB_CHECK(out.data.numel() == rows * cols, "Input and output shapes need to match.");
  const bool e5m2 = out.data.dtype == kNVTEFloat8E5M2;
  dispatch_in(in.data.dtype, [&](auto tag) {
    using IT = decltype(tag);
    constexpr int VEC = 16 / sizeof(IT);
    const dim3 grid(static_cast<unsigned>(ceil_div(cols, static_cast<size_t>(256 * VEC))),
                    static_cast<unsigned>(ceil_div(rows, static_cast<size_t>(32))));
    if (e5m2)
      cast_dbias_kernel<IT, true><<<grid, 256, 0, stream>>>(in.data.as<IT>(), out.data.as<uint8_t>(), ws, rows, cols,
                                                           out.scale.as<float>(), out.amax.as<float>(),
                                                           out.scale_inv.as<float>());
    else
      cast_dbias_kernel<IT, false><<<grid, 256, 0, stream>>>(in.data.as<IT>(), out.data.as<uint8_t>(), ws, rows, cols,
                                                            out.scale.as<float>(), out.amax.as<float>(),
                                                            out.scale_inv.as<float>());
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}

void dequantize(const Tensor &in, Tensor &out, cudaStream_t stream) {
  TEB_CHECK(is_fp8(in.data.dtype), "Input must have FP8 type.");
  TEB_CHECK(is_high_precision(out.data.dtype), "Output must be in higher precision.");
  TEB_CHECK(in.scale_inv.dptr != nullptr, "Input needs scale_inv");
  const size_t n = in.data.numel();
  TEB_CHECK(out.data.numel() == n, "Input and output shapes need to match.");
  if (n == 0) return;
  const int grid = persistent_grid(ceil_div(n / 2 + 1, kThreads), 8);
  const bool e5m2 = in.data.dtype == kNVTEFloat8E5M2;
  dispatch_in(out.data.dtype, [&](auto tag) {
    using OT = decltype(tag);
    if (e5m2)
      dequantize_kernel<OT, true><<<grid, kThreads, 0, stream>>>(
          in.data.as<uint8_t>(), out.data.as<OT>(), n, in.scale_inv.as<float>());
    else
      dequantize_kernel<OT, false><<<grid, kThreads, 0, stream>>>(
          in.data.as<uint8_t>(), out.data.as<OT>(), n, in.scale_inv.as<float>());
  });
  TEB_CHECK_CUDA(cudaGetLastError());
}

static float fp8_max_of(NVTEDType t) {
  TEB_CHECK(is_fp8(t), "Expected FP8 dtype, but got ", dtype_name(t));
  return t == kNVTEFloat8E4M3 ? 448.0f : 57344.0f;
}
static bool parse_algo(const char *algo) {
  const std::string a = algo ? algo : "";
  if (a == "max") return false;
  if (a == "most_recent") return true;
  TEB_ERROR("Unsupported amax compute algorithm (", a, ")");
}

void delayed_update(const Tensor &hist, const Tensor &scale, Tensor &hist_out, Tensor &scale_out,
                    const char *algo, NVTEDType fp8_dtype, float margin, cudaStream_t stream) {
  const bool most_recent = parse_algo(algo);
  const float scaled_max = fp8_max_of(fp8_dtype) * std::pow(2.0f, -margin);
  TEB_CHECK(hist.data.shape.ndim == 2, "Found ", hist.data.shape.ndim, " dims");
  const size_t len = hist.data.shape.data[0], num = hist.data.shape.data[1];
  TEB_CHECK(hist.data.dtype == kNVTEFloat32 && scale.data.dtype == kNVTEFloat32,
            "amax history and scale must be Float32");
  TEB_CHECK(scale.data.numel() == num && scale_out.data.numel() == num &&
                hist_out.data.numel() == len * num,
            "Mismatched delayed-scaling tensor shapes");
  if (num == 0 || len == 0) return;
  const int warps = kThreads / 32;
  delayed_update_kernel<<<ceil_div(num, warps), kThreads, 0, stream>>>(
      hist.data.as<float>(), scale.data.as<float>(), hist_out.data.as<float>(),
      scale_out.data.as<float>(), len, num, num, most_recent, scaled_max);
  TEB_CHECK_CUDA(cudaGetLastError());
}

void delayed_update_after_reduction(const Tensor *reduction, const NVTETensor *hists,
                                    const NVTETensor *scales, size_t num_tensors, const char *algo,
                                    NVTEDType fp8_dtype, float margin, cudaStream_t stream) {
  const bool most_recent = parse_algo(algo);
  const float scaled_max = fp8_max_of(fp8_dtype) * std::pow(2.0f, -margin);
  if (num_tensors == 0) return;
  size_t len = 0;
  int offset = 0;
  const float *buf = reduction ? reduction->data.as<float>() : nullptr;
  BulkParams p;
  size_t filled = 0;
  int launched_base = 0;
  for (size_t i = 0; i < num_tensors; ++i) {
    const Tensor &h = lookup_checked(hists[i]);
    const Tensor &s = lookup_checked(scales[i]);
    TEB_CHECK(h.data.shape.ndim == 2, "Found ", h.data.shape.ndim, " dims");
    TEB_CHECK(h.data.dtype == kNVTEFloat32 && s.data.dtype == kNVTEFloat32,
              "amax history and scale must be Float32");
    const size_t n = h.data.shape.data[1];
    if (i == 0) len = h.data.shape.data[0];
    TEB_CHECK(h.data.shape.data[0] == len, "All amax histories must have the same length");
    TEB_CHECK(s.data.numel() == n, "Mismatched scale tensor size");
    p.e[filled++] = BulkEntry{h.data.as<float>(), s.data.as<float>(), static_cast<int>(n), offset};
    offset += static_cast<int>(n);
    if (filled == kBulkMax || i + 1 == num_tensors) {
      delayed_update_bulk_kernel<<<static_cast<int>(filled), kThreads, 0, stream>>>(
          buf, p, static_cast<int>(filled), len, most_recent, scaled_max);
      TEB_CHECK_CUDA(cudaGetLastError());
      launched_base += static_cast<int>(filled);
      filled = 0;
    }
  }
  (void)launched_base;
}

}  // namespace fp8
}  // namespace teb
