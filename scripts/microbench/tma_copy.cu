// Memory-system ceiling of the quantizers' access pattern: the MXFP8 TMA kernel's pipeline with
// the conversion removed (TMA load of a bf16 tile -> per-warp TMA stores of two 1-byte outputs).
//   mode 0: tile 128 rows x 128 cols (two [128][128 B] load boxes), per-warp store boxes 32 x 64 B
//   mode 1: tile  32 rows x 512 cols (eight [32][128 B] load boxes), per-warp store boxes 32 x 64 B
//   mode 2: like 0 but only ONE output (rowwise only)
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a tma_copy.cu -o tma_copy -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("cuda error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t *b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t ph) {
  asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(s32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma_load(void *sm, const CUtensorMap *m, uint64_t *b, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(s32(sm)), "l"(m), "r"(s32(b)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tma_store(const CUtensorMap *m, const void *sm, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(m), "r"(s32(sm)), "r"(x), "r"(y) : "memory");
}

constexpr int kStages = 3, kStageBytes = 32768;

template <int MODE>
__global__ void __launch_bounds__(288, 2) copy_kernel(const __grid_constant__ CUtensorMap tin, const __grid_constant__ CUtensorMap t0,
                                                       const __grid_constant__ CUtensorMap t1, int rows, int cols) {
  extern __shared__ uint8_t raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint64_t *full = (uint64_t *)(smem + kStages * kStageBytes), *empty = full + kStages;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int TR = MODE == 1 ? 32 : 128, TC = MODE == 1 ? 512 : 128;
  const int tiles_x = cols / TC, num_tiles = tiles_x * (rows / TR);
  const int my = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 8) {
    if (lane != 0) return;
    for (int k = 0; k < my; ++k) {
      const int s = k % kStages;
      if (k >= kStages) mbar_wait(&empty[s], ((k / kStages) - 1) & 1);
      const int t = blockIdx.x + k * gridDim.x;
      const int r0 = (t / tiles_x) * TR, c0 = (t % tiles_x) * TC;
      mbar_expect(&full[s], kStageBytes);
      if (MODE == 1) {
        for (int w = 0; w < 8; ++w) tma_load(smem + s * kStageBytes + w * 4096, &tin, &full[s], (c0 + 64 * w) * 2, r0);
      } else {
        tma_load(smem + s * kStageBytes, &tin, &full[s], c0 * 2, r0);
        tma_load(smem + s * kStageBytes + 16384, &tin, &full[s], c0 * 2 + 128, r0);
      }
    }
    return;
  }
  for (int k = 0; k < my; ++k) {
    const int s = k % kStages;
    const int t = blockIdx.x + k * gridDim.x;
    const int r0 = (t / tiles_x) * TR, c0 = (t % tiles_x) * TC;
    mbar_wait(&full[s], (k / kStages) & 1);
    if (k > 0 && lane == 0) {
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      mbar_arrive(&empty[(k - 1) % kStages]);
    }
    __syncwarp();
    if (lane == 0) {
      const int bx = warp >> 2, rg = warp & 3;
      const uint8_t *sub = smem + s * kStageBytes + (MODE == 1 ? warp * 4096 : bx * 16384 + rg * 4096);
      const int x = MODE == 1 ? c0 + 64 * warp : c0 + 64 * bx, y = MODE == 1 ? r0 : r0 + 32 * rg;
      tma_store(&t0, sub, x, y);
      if (MODE != 2) tma_store(&t1, sub + 2048, x, y);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static CUtensorMap mk(EncodeFn f, void *p, uint64_t inner, uint64_t outer, uint32_t bi, uint32_t bo, CUtensorMapSwizzle sw) {
  CUtensorMap m;
  cuuint64_t d[2] = {inner, outer}, st[1] = {inner};
  cuuint32_t b[2] = {bi, bo}, e[2] = {1, 1};
  CUresult r = f(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p, d, st, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

template <int MODE>
void run(EncodeFn f, int rows, int cols, int nset) {
  size_t n = (size_t)rows * cols;
  uint8_t **in = new uint8_t *[nset], **o0 = new uint8_t *[nset], **o1 = new uint8_t *[nset];
  for (int i = 0; i < nset; ++i) { CK(cudaMalloc(&in[i], n * 2)); CK(cudaMalloc(&o0[i], n)); CK(cudaMalloc(&o1[i], n)); CK(cudaMemset(in[i], 1, n * 2)); }
  CK(cudaFuncSetAttribute(copy_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStages * kStageBytes + 2048));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 30;
  for (int it = -3; it < iters; ++it) {
    if (it == 0) CK(cudaEventRecord(e0));
    int i = (it + 3) % nset;
    CUtensorMap ti = mk(f, in[i], (uint64_t)cols * 2, rows, 128, MODE == 1 ? 32 : 128, CU_TENSOR_MAP_SWIZZLE_128B);
    CUtensorMap ta = mk(f, o0[i], cols, rows, 64, 32, CU_TENSOR_MAP_SWIZZLE_64B);
    CUtensorMap tb = mk(f, o1[i], cols, rows, 64, 32, CU_TENSOR_MAP_SWIZZLE_64B);
    copy_kernel<MODE><<<296, 288, kStages * kStageBytes + 2048>>>(ti, ta, tb, rows, cols);
  }
  CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= iters;
  double bytes = (double)n * (MODE == 2 ? 3 : 4);
  printf("mode %d %dx%d: %.1f us  %.0f GB/s\n", MODE, rows, cols, ms * 1e3, bytes / ms / 1e6);
  for (int i = 0; i < nset; ++i) { cudaFree(in[i]); cudaFree(o0[i]); cudaFree(o1[i]); }
}

int main(int argc, char **argv) {
  int rows = argc > 2 ? atoi(argv[1]) : 16384, cols = argc > 2 ? atoi(argv[2]) : 8192;
  CK(cudaFree(0));
  EncodeFn f; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&f, cudaEnableDefault, &q));
  // plain device-to-device copy of the same bytes for reference
  {
    size_t n = (size_t)rows * cols; uint8_t *a, *b; CK(cudaMalloc(&a, n * 2)); CK(cudaMalloc(&b, n * 2));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int it = -2; it < 10; ++it) { if (it == 0) cudaEventRecord(e0); CK(cudaMemcpyAsync(b, a, n * 2, cudaMemcpyDeviceToDevice)); }
    cudaEventRecord(e1); CK(cudaDeviceSynchronize()); float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
    printf("cudaMemcpy D2D %dx%d bf16: %.1f us  %.0f GB/s (read+write)\n", rows, cols, ms * 1e3, 4.0 * n / ms / 1e6);
    cudaFree(a); cudaFree(b);
  }
  run<0>(f, rows, cols, 3); run<1>(f, rows, cols, 3); run<2>(f, rows, cols, 3);
  return 0;
}
