// nvte_b200_reduce_partials: out[i] = sum_s partials[s][i], fp32 accumulation in slot order, one rounding.
// Second half of the reduce-scatter whose first half is the GEMM epilogue that scatters its rows into the
// ranks' partial buffers (kNVTEMatmulConfigB200ScatterOutput; reference: the reduce-scatter after a row-parallel
// GEMM, pytorch/module/linear.py:560-600, and Userbuffers' split-GEMM reduce-scatter overlap,
// common/comm_gemm_overlap/comm_gemm_overlap.cpp:611-760). HBM-bound: world reads + one write per element.
#include "core.h"
#include "ptx.cuh"

namespace teb {
namespace rs {

namespace {

template <typename T>
__device__ __forceinline__ void unpack8(const uint4 &v, float (&f)[8]);
template <>
__device__ __forceinline__ void unpack8<__nv_bfloat16>(const uint4 &v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <>
__device__ __forceinline__ void unpack8<__half>(const uint4 &v, float (&f)[8]) {
  const __half2 *h = reinterpret_cast<const __half2 *>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// 16-bit types: one thread = 8 elements (16 bytes) of every slot
template <typename T>
__global__ void __launch_bounds__(256) reduce16_kernel(const uint4 *__restrict__ in, uint4 *__restrict__ out,
                                                       size_t vecs, uint32_t world) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < vecs;
       i += static_cast<size_t>(gridDim.x) * 256) {
    uint4 v[8];
#pragma unroll
    for (uint32_t s = 0; s < 8; ++s)
      if (s < world) v[s] = ptx::ld_stream_v4(in + s * vecs + i);
    float acc[8];
    unpack8<T>(v[0], acc);
#pragma unroll
    for (uint32_t s = 1; s < 8; ++s)
      if (s < world) {
        float f[8];
        unpack8<T>(v[s], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    uint4 o;
    T *po = reinterpret_cast<T *>(&o);
#pragma unroll
    for (int j = 0; j < 8; ++j) po[j] = static_cast<T>(acc[j]);
    out[i] = o;
  }
}

__global__ void __launch_bounds__(256) reduce32_kernel(const float4 *__restrict__ in, float4 *__restrict__ out,
                                                       size_t vecs, uint32_t world) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * 256 + threadIdx.x; i < vecs;
       i += static_cast<size_t>(gridDim.x) * 256) {
    float4 a = in[i];
    for (uint32_t s = 1; s < world; ++s) {
      const float4 b = in[s * vecs + i];
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    out[i] = a;
  }
}

}  // namespace

void reduce_partials(const void *partials, uint32_t world, size_t count, NVTEDType dtype, void *out,
                     cudaStream_t stream) {
  TEB_CHECK(world >= 1 && world <= 8, "reduce_partials: world must be 1..8, got ", world);
  TEB_CHECK(partials != nullptr && out != nullptr, "reduce_partials: null buffer");
  TEB_CHECK(reinterpret_cast<uintptr_t>(partials) % 16 == 0 && reinterpret_cast<uintptr_t>(out) % 16 == 0,
            "reduce_partials: buffers must be 16-byte aligned");
  if (count == 0) return;
  const size_t per = dtype == kNVTEFloat32 ? 4 : 8;
  TEB_CHECK(count % per == 0, "reduce_partials: the element count per slot must be a multiple of ", per);
  const size_t vecs = count / per;
  const unsigned grid = static_cast<unsigned>(std::min<size_t>(ceil_div(vecs, static_cast<size_t>(256)),
                                                               static_cast<size_t>(8 * sm_count())));
  switch (dtype) {
    case kNVTEBFloat16:
      reduce16_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const uint4 *>(partials),
                                                              static_cast<uint4 *>(out), vecs, world);
      break;
    case kNVTEFloat16:
      reduce16_kernel<__half><<<grid, 256, 0, stream>>>(static_cast<const uint4 *>(partials), static_cast<uint4 *>(out),
                                                       vecs, world);
      break;
    case kNVTEFloat32:
      reduce32_kernel<<<grid, 256, 0, stream>>>(static_cast<const float4 *>(partials), static_cast<float4 *>(out), vecs,
                                                world);
      break;
    default: TEB_ERROR("reduce_partials: dtype must be bf16, fp16 or fp32");
  }
  TEB_CHECK_CUDA(cudaGetLastError());
}

}  // namespace rs
}  // namespace teb
