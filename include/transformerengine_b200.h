/* C ABI of the B200-native low-precision Linear path.
 *
 * This header declares exactly the entry points the reference's own bindings
 * (its PyTorch extension and its C++ operator tests) resolve from
 * `libtransformer_engine.so` for the quantize -> swizzle -> block-scaled GEMM
 * path. Enum values, struct layouts and argument order are binary compatible
 * with the reference headers cited beside each group, so the library built
 * from this repo can be loaded in place of the reference's core library for
 * this path. Everything is `extern "C"`, plain pointers and sizes.
 *
 * Reference paths are relative to the reference checkout:
 *   TE = transformer_engine/common/include/transformer_engine
 *
 * Error convention (TE/../util/logging.h:38-56): the `nvte_*` functions throw
 * std::runtime_error through the C boundary, as the reference does. For FFI
 * callers that cannot unwind C++ exceptions (ctypes, cgo, JNI) every compute
 * entry point also has a `<name>_status` twin that returns 0 on success and
 * non-zero on failure, with the message available from nvte_b200_last_error().
 */
#ifndef TRANSFORMER_ENGINE_B200_H_
#define TRANSFORMER_ENGINE_B200_H_

#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
#include <vector> /* one entry point keeps the reference's std::vector signature */
#else
#include <stdbool.h> /* plain C callers (cgo, JNI stubs): the reference's signatures use bool */
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st *nvte_stream_t; /* == cudaStream_t */

/* ---- basic types (TE/transformer_engine.h:22-136) ----------------------- */
typedef enum {
  kNVTEByte = 0,
  kNVTEInt16 = 1,
  kNVTEInt32 = 2,
  kNVTEInt64 = 3,
  kNVTEFloat32 = 4,
  kNVTEFloat16 = 5,
  kNVTEBFloat16 = 6,
  kNVTEFloat8E4M3 = 7,
  kNVTEFloat8E5M2 = 8,
  kNVTEFloat8E8M0 = 9,
  kNVTEFloat4E2M1 = 10,
  kNVTENumTypes
} NVTEDType;

typedef struct {
  size_t data[15];
  size_t ndim;
} NVTEShape;

typedef struct {
  void *data_ptr;
  NVTEDType dtype;
  NVTEShape shape;
} NVTEBasicTensor;

typedef enum {
  kNVTERowwiseData = 0,
  kNVTEColumnwiseData = 1,
  kNVTEScale = 2,
  kNVTEAmax = 3,
  kNVTERowwiseScaleInv = 4,
  kNVTEColumnwiseScaleInv = 5,
  kNVTEColumnwiseAmax = 6,
  kNVTEWithGEMMSwizzledScales = 7, /* uint8_t */
  kNVTERowScaledNVFP4 = 8,         /* uint8_t; row scaling is not implemented: must stay 0 */
  kNVTENVFP4E4M3Max = 9,           /* int; only 448 is implemented */
  kNVTENumTensorParams
} NVTETensorParam;

typedef enum {
  NVTE_DELAYED_TENSOR_SCALING = 0, /* unquantized, or per-tensor-scaled FP8 */
  NVTE_MXFP8_1D_SCALING = 1,       /* 1x32 / 32x1 blocks, E8M0 scales */
  NVTE_BLOCK_SCALING_1D = 2,       /* not implemented (Hopper recipe) */
  NVTE_BLOCK_SCALING_2D = 3,       /* not implemented (Hopper recipe) */
  NVTE_NVFP4_1D_SCALING = 4,       /* 1x16 blocks, E4M3 scales + fp32 global amax */
  NVTE_INVALID_SCALING = 100
} NVTEScalingMode;

typedef void *NVTETensor; /* 1-based index into the handle table; 0 = invalid */

/* ---- tensor handles (TE/transformer_engine.h:148-372,483,494) ----------- */
NVTETensor nvte_create_tensor(NVTEScalingMode scaling_mode);
void nvte_create_tensors(NVTEScalingMode scaling_mode, NVTETensor *tensors, size_t N);
void nvte_destroy_tensor(NVTETensor tensor);
void nvte_destroy_tensors(NVTETensor *tensors, size_t N);
void *nvte_tensor_data(const NVTETensor tensor);
void *nvte_tensor_columnwise_data(const NVTETensor tensor);
NVTEShape nvte_make_shape(const size_t *data, size_t ndim);
NVTEShape nvte_tensor_shape(const NVTETensor tensor);
NVTEShape nvte_tensor_columnwise_shape(const NVTETensor tensor);
size_t nvte_tensor_ndims(const NVTETensor tensor);
size_t nvte_tensor_size(const NVTETensor tensor, const size_t dim);
size_t nvte_tensor_size_bytes(const NVTETensor tensor);
size_t nvte_tensor_numel(const NVTETensor tensor);
size_t nvte_tensor_element_size(const NVTETensor tensor);
size_t nvte_tensor_element_size_bits(const NVTETensor tensor);
NVTEDType nvte_tensor_type(const NVTETensor tensor);
float *nvte_tensor_amax(const NVTETensor tensor);
float *nvte_tensor_scale(const NVTETensor tensor);
float *nvte_tensor_scale_inv(const NVTETensor tensor);
NVTEShape nvte_tensor_scale_inv_shape(const NVTETensor tensor);
NVTEScalingMode nvte_tensor_scaling_mode(const NVTETensor tensor);
void nvte_zero_tensor(const NVTETensor tensor, nvte_stream_t stream);
void nvte_set_tensor_param(NVTETensor *tensor, NVTETensorParam param_name,
                           const NVTEBasicTensor *param);
NVTEBasicTensor nvte_get_tensor_param(const NVTETensor tensor, NVTETensorParam param_name);
void nvte_set_tensor_param_v2(NVTETensor tensor, NVTETensorParam param, const void *buf,
                              size_t size_in_bytes);
void nvte_get_tensor_param_v2(const NVTETensor tensor, NVTETensorParam param, void *buf,
                              size_t size_in_bytes, size_t *size_written);
int nvte_is_non_tn_fp8_gemm_supported(void);
void nvte_memset(void *ptr, int value, size_t size_in_bytes, nvte_stream_t stream);

/* ---- grouped tensor handles (TE/transformer_engine.h:536-644): host-side views; the grouped
 * quantize / GEMM kernels are a later row ------------------------------------ */
typedef void *NVTEGroupedTensor;
typedef enum {
  kNVTEGroupedRowwiseData = 0,
  kNVTEGroupedColumnwiseData = 1,
  kNVTEGroupedScale = 2,
  kNVTEGroupedAmax = 3,
  kNVTEGroupedRowwiseScaleInv = 4,
  kNVTEGroupedColumnwiseScaleInv = 5,
  kNVTEGroupedColumnwiseAmax = 6,
  kNVTEGroupedFirstDims = 7,
  kNVTEGroupedLastDims = 8,
  kNVTEGroupedTensorOffsets = 9,
  kNVTEGroupedWithGEMMSwizzledScales = 10,
  kNVTENumGroupedTensorParams
} NVTEGroupedTensorParam;
NVTEGroupedTensor nvte_create_grouped_tensor(NVTEScalingMode scaling_mode, size_t num_tensors,
                                             NVTEShape logical_shape);
void nvte_destroy_grouped_tensor(NVTEGroupedTensor tensor);
void nvte_set_grouped_tensor_param(NVTEGroupedTensor tensor, NVTEGroupedTensorParam param,
                                   const void *buf, size_t size_in_bytes);
void nvte_get_grouped_tensor_param(const NVTEGroupedTensor tensor, NVTEGroupedTensorParam param,
                                   void *buf, size_t size_in_bytes, size_t *size_written);
size_t nvte_grouped_tensor_num_tensors(const NVTEGroupedTensor tensor);
NVTEDType nvte_grouped_tensor_type(const NVTEGroupedTensor tensor);
NVTEScalingMode nvte_grouped_tensor_scaling_mode(const NVTEGroupedTensor tensor);
NVTEShape nvte_get_grouped_tensor_logical_shape(const NVTEGroupedTensor tensor);
/* TE/multi_stream.h:21 -- this path only ever uses the caller's stream */
int nvte_get_num_compute_streams(void);

/* ---- quantization config (TE/transformer_engine.h:395-477) -------------- */
typedef void *NVTEQuantizationConfig;
typedef enum {
  kNVTEQuantizationConfigForcePow2Scales = 0,               /* uint8_t */
  kNVTEQuantizationConfigAmaxEpsilon = 1,                   /* float */
  kNVTEQuantizationConfigNoopTensor = 2,                    /* NVTETensor (device flag) */
  kNVTEQuantizationConfigFloat8BlockScaleTensorFormat = 3,  /* int, deprecated, ignored */
  kNVTEQuantizationConfigRNGState = 4,                      /* NVTETensor */
  kNVTEQuantizationConfigNVFP42DQuantization = 5,           /* uint8_t */
  kNVTEQuantizationConfigStochasticRounding = 6,            /* uint8_t; NVFP4 (Philox4x32, cvt.rs) */
  kNVTEQuantizationConfigUseFastMath = 7,                   /* uint8_t; must be 0 */
  kNVTEQuantizationConfigNVFP44Over6Mode = 8,               /* uint8_t; must be 0 */
  kNVTEQuantizationConfigNVFP44Over6ErrUseFastMath = 9,     /* uint8_t */
  kNVTEQuantizationConfigNumAttributes = 10,                /* == the reference's */
  /* Extension (not in the reference; numbered from 0x1000 so that attributes the reference adds
   * later cannot collide with it; size_in_bytes must be exactly 72): uint64_t[9] = {count <= 8, peer pointers...}. The per-tensor
   * FP8 cast writes its rowwise bytes to every listed address as well as to the output tensor: with
   * peer-mapped (symmetric-memory / NVLink) addresses of the other ranks' gather buffers the cast IS
   * the all-gather of the sequence-parallel column-parallel Linear (the reference: cast, then NCCL
   * all-gather, pytorch/distributed.py:982-1073). 16-bit inputs, 16-byte aligned, no columnwise.
   * With nvte_compute_amax_with_config the addresses are float slots: the amax is stored to each
   * of them after the reduction kernel (all-reduce(MAX) = push, barrier, max over the slots). */
  kNVTEQuantizationConfigB200PeerRowwiseData = 0x1000,
  /* Extension: NVTEB200SymmetricExchange by value (size_in_bytes must be sizeof of it). The exchange of
   * the sequence-parallel column-parallel Linear (amax all-reduce + all-gather of the FP8 shard) WITHOUT a
   * collective and without barriers: arrival flags in peer-mapped memory, see the struct below. With it
   *   nvte_compute_amax_with_config  also pushes (amax, step) to every rank's slot array,
   *   nvte_compute_scale_from_amax   waits for the world's slots of this step, takes their max, then the scale,
   *   nvte_quantize_v2 (per-tensor FP8 output, 16-bit input) stores its bytes into every rank's gather buffer in
   *       row chunks and raises a flag on every rank after each chunk (the last chunk advances the step),
   * and the GEMM (kNVTEMatmulConfigB200ArrivalFlags) loads a row chunk of X only after its flag is up. */
  kNVTEQuantizationConfigB200SymmetricExchange = 0x1001
} NVTEQuantizationConfigAttribute;
/* All addresses are device addresses valid on THIS rank (peer-mapped symmetric memory for the peers). */
typedef struct {
  uint32_t world, rank;
  uint32_t rows_per_flag;   /* rows of a shard per arrival flag (a multiple of 256 = GEMM tile rows) */
  uint32_t flags_per_rank;  /* shard rows / rows_per_flag; 0 = amax exchange only: the scale kernel advances the
                             * step, the cast is an ordinary local cast and the GEMM sends the rows
                             * (NVTEB200ArrivalFlags.push_unit_bytes) */
  uint64_t gather[8];       /* [r]: where this rank's shard starts inside rank r's gather buffer */
  uint64_t flags[8];        /* [r]: rank r's flag array, uint32[world * flags_per_rank], zero-initialised */
  uint64_t amax_slots[8];   /* [r]: rank r's amax slots, uint64[2 * world] = (step << 32 | fp32 bits), zeroed */
  uint64_t step;            /* this rank's uint32 step counter (zero-initialised, advanced by the cast, or by
                             * the scale kernel when flags_per_rank == 0) */
  uint64_t done;            /* this rank's uint32[flags_per_rank + 1] completion counters (zeroed) */
  uint64_t error;           /* this rank's uint32: set to 1 when a wait for a peer timed out (5 s) */
} NVTEB200SymmetricExchange;
NVTEQuantizationConfig nvte_create_quantization_config(void);
void nvte_get_quantization_config_attribute(NVTEQuantizationConfig config,
                                            NVTEQuantizationConfigAttribute attr, void *buf,
                                            size_t size_in_bytes, size_t *size_written);
void nvte_set_quantization_config_attribute(NVTEQuantizationConfig config,
                                            NVTEQuantizationConfigAttribute attr,
                                            const void *buf, size_t size_in_bytes);
void nvte_destroy_quantization_config(NVTEQuantizationConfig config);

/* ---- quantize / dequantize (TE/cast.h:90,115,125,145) ------------------- */
void nvte_quantize(const NVTETensor input, NVTETensor output, nvte_stream_t stream);
void nvte_quantize_noop(const NVTETensor input, NVTETensor output, NVTETensor noop,
                        nvte_stream_t stream);
void nvte_quantize_v2(const NVTETensor input, NVTETensor output,
                      const NVTEQuantizationConfig quant_config, nvte_stream_t stream);
void nvte_dequantize(const NVTETensor input, NVTETensor output, nvte_stream_t stream);
/* the same cast for a list of tensors (TE/cast.h:438; MoE experts / tex.split_quantize) */
void nvte_multi_tensor_quantize(const NVTETensor *inputs, NVTETensor *outputs,
                                const NVTEQuantizationConfig quant_config, const size_t num_tensors,
                                nvte_stream_t stream);
/* cast + dbias[c] = sum_r input[r, c] (TE/cast.h:145; the backward of a Linear's bias). Called
 * with an empty workspace it only fills the workspace's shape/dtype (fp32 [ceil(rows/32), cols]). */
void nvte_quantize_dbias(const NVTETensor input, NVTETensor output, NVTETensor dbias,
                         NVTETensor workspace, nvte_stream_t stream);

/* ---- normalization (TE/normalization.h:44,76,107,138,161,170,176,186) -------
 * LayerNorm / RMSNorm of x[N, H] with z in any of: the input's 16-bit dtype (optionally with z's
 * amax filled: current scaling), per-tensor FP8 E4M3 (scale in, amax / scale_inv out), MXFP8 E4M3
 * rowwise (+ columnwise when z has columnwise data; compact or GEMM-swizzled scales). Calling with
 * an empty workspace only fills the workspace's shape/dtype. multiprocessorCount is accepted and
 * ignored. Fast kernels: bf16/fp16 x with H % 64 == 0 (forward), H % 8 == 0 and H <= 16384 (backward); everything
 * else the reference accepts (fp32 tensors, other hidden sizes, E5M2 / mixed output dtypes) runs generic kernels. */
void nvte_layernorm_fwd(const NVTETensor x, const NVTETensor gamma, const NVTETensor beta,
                        const float epsilon, NVTETensor z, NVTETensor mu, NVTETensor rsigma,
                        NVTETensor workspace, const int multiprocessorCount,
                        const bool zero_centered_gamma, nvte_stream_t stream);
void nvte_layernorm_bwd(const NVTETensor dz, const NVTETensor x, const NVTETensor mu,
                        const NVTETensor rsigma, const NVTETensor gamma, NVTETensor dx,
                        NVTETensor dgamma, NVTETensor dbeta, NVTETensor workspace,
                        const int multiprocessorCount, const bool zero_centered_gamma,
                        nvte_stream_t stream);
void nvte_rmsnorm_fwd(const NVTETensor x, const NVTETensor gamma, const float epsilon, NVTETensor z,
                      NVTETensor rsigma, NVTETensor workspace, const int multiprocessorCount,
                      const bool zero_centered_gamma, nvte_stream_t stream);
void nvte_rmsnorm_bwd(const NVTETensor dz, const NVTETensor x, const NVTETensor rsigma,
                      const NVTETensor gamma, NVTETensor dx, NVTETensor dgamma,
                      NVTETensor workspace, const int multiprocessorCount,
                      const bool zero_centered_gamma, nvte_stream_t stream);
void nvte_rmsnorm_bwd_add(const NVTETensor dz, const NVTETensor x, const NVTETensor add,
                          const NVTETensor rsigma, const NVTETensor gamma, NVTETensor dx,
                          NVTETensor dgamma, NVTETensor workspace, const int multiprocessorCount,
                          const bool zero_centered_gamma, nvte_stream_t stream);
void nvte_enable_cudnn_norm_fwd(bool enable);                       /* accepted, no effect */
void nvte_enable_cudnn_norm_bwd(bool enable);                       /* accepted, no effect */
void nvte_enable_zero_centered_gamma_in_weight_dtype(bool enable);  /* accepted, no effect */

/* ---- amax / scale / recipes (TE/recipe.h:41,85,99,111,308) -------------- */
void nvte_compute_amax(const NVTETensor input, NVTETensor output, nvte_stream_t stream);
void nvte_compute_amax_with_config(const NVTETensor input, NVTETensor output,
                                   const NVTEQuantizationConfig config, nvte_stream_t stream);
void nvte_compute_scale_from_amax(NVTETensor output, const NVTEQuantizationConfig config,
                                  nvte_stream_t stream);
void nvte_delayed_scaling_recipe_amax_and_scale_update(
    const NVTETensor amax_history, const NVTETensor scale, NVTETensor updated_amax_history,
    NVTETensor updated_scale, const char *amax_compute_algo, NVTEDType fp8_dtype, float margin,
    nvte_stream_t stream);
/* TE/recipe.h:71-74 declares `..._after_reduction` inside `extern "C"` with two by-value
 * std::vector<NVTETensor> arguments: exported here under that unmangled name with that
 * signature for C++ callers (the reference's csrc/extensions/recipe.cpp binds it). */
#ifdef __cplusplus
void nvte_delayed_scaling_recipe_amax_and_scale_update_after_reduction(
    const NVTETensor amax_reduction_buffer, std::vector<NVTETensor> amax_histories,
    std::vector<NVTETensor> scales, const char *amax_compute_algo, NVTEDType fp8_dtype,
    float margin, nvte_stream_t stream);
#endif
/* The same operation over plain arrays, for C / FFI callers. */
void nvte_b200_delayed_scaling_update_after_reduction(
    const NVTETensor amax_reduction_buffer, const NVTETensor *amax_histories,
    const NVTETensor *scales, size_t num_tensors, const char *amax_compute_algo,
    NVTEDType fp8_dtype, float margin, nvte_stream_t stream);
void nvte_nvfp4_compute_per_tensor_scale(const NVTETensor inpA, const bool use_rowwise_amax_A,
                                         const NVTETensor inpB, const bool use_rowwise_amax_B,
                                         float alpha_in, NVTETensor alpha_out,
                                         nvte_stream_t stream);

/* ---- random Hadamard transform (TE/hadamard_transform.h:30,46,59,74) ---- */
void nvte_hadamard_transform(const NVTETensor input, NVTETensor output, int random_sign_mask,
                             int random_sign_mask_t, nvte_stream_t stream);
void nvte_hadamard_transform_amax(const NVTETensor input, NVTETensor output,
                                  int random_sign_mask, int random_sign_mask_t,
                                  nvte_stream_t stream);
void nvte_hadamard_transform_cast_fusion_columnwise(const NVTETensor input, NVTETensor output,
                                                    const NVTETensor hadamard_matrix,
                                                    const NVTEQuantizationConfig quant_config,
                                                    nvte_stream_t stream);
void nvte_quantize_with_hadamard_transform(const NVTETensor input, NVTETensor output,
                                           const NVTETensor hadamard_matrix,
                                           const NVTEQuantizationConfig quant_config,
                                           nvte_stream_t stream);

/* ---- scale-factor swizzle (TE/swizzle.h:31,70) -------------------------- */
void nvte_swizzle_scaling_factors(const NVTETensor input, NVTETensor output,
                                  nvte_stream_t stream);
/* TE/swizzle.h:45,84: the same for a list of tensors */
void nvte_multi_tensor_swizzle_scaling_factors(const NVTETensor *inputs, NVTETensor *outputs,
                                               const size_t num_tensors, nvte_stream_t stream);
void nvte_multi_tensor_unswizzle_scaling_factors(const NVTETensor *inputs, NVTETensor *outputs,
                                                 const size_t num_tensors, nvte_stream_t stream);
void nvte_unswizzle_scaling_factors(const NVTETensor input, NVTETensor output,
                                    nvte_stream_t stream);

/* ---- GEMM (TE/gemm.h:31-58,93-120,178,200,231) -------------------------- */
typedef void *NVTEMatmulConfig;
typedef enum {
  kNVTEMatmulConfigBiasTensor = 0,          /* NVTETensor */
  kNVTEMatmulConfigDBiasTensor = 1,         /* NVTETensor */
  kNVTEMatmulConfigWithGELUEpilogue = 2,    /* uint8_t; not implemented */
  kNVTEMatmulConfigWithDGELUEpilogue = 3,   /* uint8_t; not implemented */
  kNVTEMatmulConfigEpilogueAuxTensor = 4,   /* NVTETensor */
  kNVTEMatmulConfigUseSplitAccumulator = 5, /* uint8_t; accepted, no effect (fp32 TMEM accum) */
  kNVTEMatmulConfigSMCount = 6,             /* int32_t */
  kNVTEMatmulConfigNumAttributes = 7,       /* == the reference's */
  /* extension (not in the reference; 0x1000 range, size_in_bytes must be exactly 12): int32_t[3] = {rows per block, block stride, first row}:
   * output row r of the GEMM is written to row (r / rows) * stride + first + r % rows of D, so a
   * GEMM over one chunk of an all-gather lands in place in the full output (comm/GEMM overlap) */
  kNVTEMatmulConfigB200OutputRowBlocks = 0x1000,
  /* extension: NVTEB200ArrivalFlags by value. Operand B's rows are the gather buffer of a symmetric exchange
   * (above): rank r's shard = rows [r * rows_per_rank, (r + 1) * rows_per_rank). The GEMM walks the tile rows
   * in arrival order -- the local rank's rows first, then chunk 0 of every peer, chunk 1, ... -- and its TMA
   * producer loads a remote chunk only once its flag word has reached this step's value. No barrier, no
   * collective.
   *   push_unit_bytes == 0: someone else fills the peers' rows and raises flags[r * flags_per_rank + chunk] to
   *     *step (the exchange cast, nvte_quantize_v2 with kNVTEQuantizationConfigB200SymmetricExchange).
   *   push_unit_bytes != 0 (16384; per-tensor FP8 operands): THE GEMM ITSELF is the all-gather. The local rows
   *     are already in place (an ordinary local cast); an otherwise idle warp of every second CTA streams them
   *     to the peers with bulk copies (global -> shared -> NVLink, push_dst[r]) in units of push_unit_bytes,
   *     chunk 0 first, and adds 1 to the peer's flag word of the chunk (push_flags[r]) when a unit has landed.
   *     The flag words are then counters: a chunk has arrived when its word equals
   *     *step * (rows_per_flag * K / push_unit_bytes) (mod 2^32). One launch hides the NVLink push behind its
   *     own first tiles.
   *     push_mc_dst / push_mc_flags != 0: NVSwitch multicast addresses of the same two windows (one
   *     multimem.st reaches every rank, so a rank sends its rows ONCE instead of world - 1 times; the counters
   *     go up by multimem.red, per chunk). The unicast addresses are then unused. */
  kNVTEMatmulConfigB200ArrivalFlags = 0x1001,
  /* extension: NVTEB200ScatterOutput by value. The reduce-scatter of a row-parallel GEMM, first half, in the
   * GEMM's epilogue: output row r belongs to rank q = r / rows_per_rank and is stored -- straight from the
   * accumulator, over NVLink for q != rank -- to dst[q] + (r % rows_per_rank) * N (elements of D's dtype), i.e.
   * into THIS rank's slot of rank q's buffer of partial results [world][rows_per_rank][N]. The tile rows are
   * walked starting behind the rank's own rows so that the senders never gang up on one receiver. D supplies
   * shape and dtype only. After every rank's GEMM has completed (a barrier), nvte_b200_reduce_partials sums
   * the slots in rank order (fp32) -- the second half. No accumulation (beta), no output row blocks. */
  kNVTEMatmulConfigB200ScatterOutput = 0x1002
} NVTEMatmulConfigAttribute;
typedef struct {
  uint64_t dst[8];          /* [q]: rank q's address of slot [rank] of its partial buffer */
  uint32_t world, rank;
  uint32_t rows_per_rank;   /* output rows / world */
  uint32_t reserved;        /* 0 */
} NVTEB200ScatterOutput;
typedef struct {
  uint64_t flags;          /* this rank's uint32[world * flags_per_rank] */
  uint64_t step;           /* this rank's uint32 step counter (already advanced by the cast of this step) */
  uint64_t error;          /* uint32, set to 1 on a wait time-out */
  uint32_t world, rank, rows_per_flag, flags_per_rank;
  uint32_t push_unit_bytes;  /* see above; 0 or 16384 */
  uint32_t reserved;         /* 0 */
  uint64_t push_dst[8];      /* push: rank r's address of THIS rank's row window in r's gather buffer */
  uint64_t push_flags[8];    /* push: rank r's address of r's flag words (uint32[world * flags_per_rank]) */
  uint64_t push_mc_dst;      /* push: multicast address of THIS rank's row window (0 = unicast bulk copies) */
  uint64_t push_mc_flags;    /* push: multicast address of the flag words */
  uint64_t tile_counter;     /* 0, or the address of two zeroed uint32 in this rank's memory: the GEMM then takes its
                              * tiles from a shared counter (in arrival order) instead of a static assignment, so a
                              * CTA pair waiting for rows that have not arrived holds up nobody else's share; the
                              * kernel leaves both words zero. One GEMM at a time per counter. */
  /* MXFP8 shards (multicast push only; needs tile_counter): a second pushed region per chunk -- the shard's
   * GEMM-swizzled block scales, whose 128-row blocks are contiguous in the gathered scale buffer -- and a step
   * protocol that needs no amax exchange: with advance_step the launch is step *step + 1, its last CTA pair
   * advances *step at the end of the kernel and adds 1 to `ready` on every rank (ready_mc); the next launch's push
   * starts once `ready` has reached world x *step, i.e. every GPU has finished reading the previous rows. */
  uint64_t push_src2;             /* this rank's scale window (local address); 0 = no second region */
  uint64_t push_mc_dst2;          /* multicast address of the same window */
  uint32_t push_unit2_bytes;      /* multiple of 4096 */
  uint32_t push_bytes2_per_flag;  /* scale bytes per flag chunk, a multiple of push_unit2_bytes */
  uint32_t advance_step;          /* see above */
  uint32_t reserved2;             /* 0 */
  uint64_t ready;                 /* this rank's uint32 `ready` word (zero-initialised); 0 = none */
  uint64_t ready_mc;              /* its multicast address */
} NVTEB200ArrivalFlags;
NVTEMatmulConfig nvte_create_matmul_config(void);
void nvte_get_matmul_config_attribute(NVTEMatmulConfig config, NVTEMatmulConfigAttribute attr,
                                      void *buf, size_t size_in_bytes, size_t *size_written);
void nvte_set_matmul_config_attribute(NVTEMatmulConfig config, NVTEMatmulConfigAttribute attr,
                                      const void *buf, size_t size_in_bytes);
void nvte_destroy_matmul_config(NVTEMatmulConfig config);
void nvte_cublas_gemm(const NVTETensor A, const NVTETensor B, NVTETensor D, const NVTETensor bias,
                      NVTETensor pre_gelu_out, bool transa, bool transb, bool grad,
                      NVTETensor workspace, bool accumulate, bool use_split_accumulator,
                      int math_sm_count, nvte_stream_t stream);
void nvte_cublas_gemm_v2(int transa, int transb, const float *alpha, const NVTETensor A,
                         const NVTETensor B, const float *beta, const NVTETensor C, NVTETensor D,
                         NVTETensor workspace, NVTEMatmulConfig config, nvte_stream_t stream);
void nvte_cublas_gemm_scaled(const NVTETensor A, const NVTETensor B, NVTETensor D,
                             const NVTETensor bias, NVTETensor pre_gelu_out, bool transa,
                             bool transb, bool grad, NVTETensor workspace, float alpha, float beta,
                             bool use_split_accumulator, int math_sm_count, nvte_stream_t stream);
/* TE/gemm.h:295 (GroupedLinear, one GEMM per expert; reference cublaslt_gemm.cu:970-1061 loops
 * cuBLASLt over its compute streams). Here: ONE persistent launch over the concatenated tile list
 * of all members when they share a scaling mode and workspace[0] can hold the member table
 * (nvte_b200_group_entry_bytes() per member + 128), else a loop. Members with an empty output are skipped. */
size_t nvte_b200_group_entry_bytes(void);
void nvte_multi_tensor_gemm(const NVTETensor *A, const NVTETensor *B, NVTETensor *D,
                            const NVTETensor *bias, NVTETensor *pre_gelu_out, const int num_gemms,
                            bool transa, bool transb, bool grad, NVTETensor *workspace,
                            bool accumulate, bool use_split_accumulator, int math_sm_count,
                            nvte_stream_t stream);

/* ---- FFI-safe twins ------------------------------------------------------ */
const char *nvte_b200_last_error(void); /* thread-local, valid until the next failing call */
int nvte_b200_sm_arch(void);            /* 10*major + minor of the current device, 0 if none */
int nvte_b200_sm_count(void);
/* FFI fast paths (no struct marshalling): set a buffer parameter / create a plain rowwise-data
 * tensor from scalars, up to 4 dimensions. 0 / non-null on success. */
int nvte_b200_set_tensor_param_flat(NVTETensor tensor, int param, void *ptr, int dtype, int ndim,
                                    size_t d0, size_t d1, size_t d2, size_t d3);
NVTETensor nvte_b200_create_plain_tensor(void *ptr, int dtype, int ndim, size_t d0, size_t d1,
                                         size_t d2, size_t d3);

int nvte_quantize_v2_status(const NVTETensor input, NVTETensor output,
                            const NVTEQuantizationConfig quant_config, nvte_stream_t stream);
int nvte_dequantize_status(const NVTETensor input, NVTETensor output, nvte_stream_t stream);
int nvte_layernorm_fwd_status(const NVTETensor x, const NVTETensor gamma, const NVTETensor beta,
                              float epsilon, NVTETensor z, NVTETensor mu, NVTETensor rsigma,
                              NVTETensor workspace, int multiprocessorCount,
                              int zero_centered_gamma, nvte_stream_t stream);
int nvte_rmsnorm_fwd_status(const NVTETensor x, const NVTETensor gamma, float epsilon, NVTETensor z,
                            NVTETensor rsigma, NVTETensor workspace, int multiprocessorCount,
                            int zero_centered_gamma, nvte_stream_t stream);
int nvte_layernorm_bwd_status(const NVTETensor dz, const NVTETensor x, const NVTETensor mu,
                              const NVTETensor rsigma, const NVTETensor gamma, NVTETensor dx,
                              NVTETensor dgamma, NVTETensor dbeta, NVTETensor workspace,
                              int multiprocessorCount, int zero_centered_gamma,
                              nvte_stream_t stream);
int nvte_rmsnorm_bwd_status(const NVTETensor dz, const NVTETensor x, const NVTETensor rsigma,
                            const NVTETensor gamma, NVTETensor dx, NVTETensor dgamma,
                            NVTETensor workspace, int multiprocessorCount, int zero_centered_gamma,
                            nvte_stream_t stream);
int nvte_rmsnorm_bwd_add_status(const NVTETensor dz, const NVTETensor x, const NVTETensor add,
                                const NVTETensor rsigma, const NVTETensor gamma, NVTETensor dx,
                                NVTETensor dgamma, NVTETensor workspace, int multiprocessorCount,
                                int zero_centered_gamma, nvte_stream_t stream);
int nvte_multi_tensor_quantize_status(const NVTETensor *inputs, NVTETensor *outputs,
                                      const NVTEQuantizationConfig quant_config, size_t num_tensors,
                                      nvte_stream_t stream);
int nvte_quantize_dbias_status(const NVTETensor input, NVTETensor output, NVTETensor dbias,
                               NVTETensor workspace, nvte_stream_t stream);
/* No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * nvte_quantize_dbias without the workspace argument: the fp32 partial sums [ceil(rows/32), cols] go to a
 * static device array inside the library, so the call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine). Same inputs, outputs and results as nvte_quantize_dbias. Limits and rules:
 *   - ceil(rows/32) * cols * 4 <= NVTE_B200_DBIAS_STATIC_SCRATCH_BYTES (8 MiB: e.g. 8192 x 8192 or
 *     16384 x 4096); a larger input is rejected with an error that says so;
 *   - every GPU that loads the library holds those 8 MiB until the process exits;
 *   - all calls on a device share the array: calls that overlap on the GPU (different streams with no
 *     ordering between them) race -- order them on one stream or with events;
 *   - returns 0, or nonzero with the message in nvte_b200_last_error() (it never throws). */
#define NVTE_B200_DBIAS_STATIC_SCRATCH_BYTES (8u << 20)
/* out[i] = sum over s < world of partials[s * count + i], accumulated in fp32 in slot order and rounded once to
 * dtype (kNVTEBFloat16 / kNVTEFloat16 / kNVTEFloat32): the second half of the reduce-scatter whose first half is
 * the GEMM epilogue above (kNVTEMatmulConfigB200ScatterOutput). Deterministic. Returns 0 or nonzero + last_error. */
int nvte_b200_reduce_partials(const void *partials, uint32_t world, size_t count, NVTEDType dtype, void *out,
                              nvte_stream_t stream);
int nvte_b200_quantize_dbias_static_scratch(const NVTETensor input, NVTETensor output,
                                            NVTETensor dbias, nvte_stream_t stream);
int nvte_compute_amax_with_config_status(const NVTETensor input, NVTETensor output,
                                         const NVTEQuantizationConfig config,
                                         nvte_stream_t stream);
int nvte_compute_scale_from_amax_status(NVTETensor output, const NVTEQuantizationConfig config,
                                        nvte_stream_t stream);
int nvte_delayed_scaling_recipe_amax_and_scale_update_status(
    const NVTETensor amax_history, const NVTETensor scale, NVTETensor updated_amax_history,
    NVTETensor updated_scale, const char *amax_compute_algo, NVTEDType fp8_dtype, float margin,
    nvte_stream_t stream);
int nvte_b200_delayed_scaling_update_after_reduction_status(
    const NVTETensor amax_reduction_buffer, const NVTETensor *amax_histories,
    const NVTETensor *scales, size_t num_tensors, const char *amax_compute_algo,
    NVTEDType fp8_dtype, float margin, nvte_stream_t stream);
int nvte_nvfp4_compute_per_tensor_scale_status(const NVTETensor inpA,
                                               const bool use_rowwise_amax_A,
                                               const NVTETensor inpB,
                                               const bool use_rowwise_amax_B, float alpha_in,
                                               NVTETensor alpha_out, nvte_stream_t stream);
int nvte_hadamard_transform_status(const NVTETensor input, NVTETensor output,
                                   int random_sign_mask, int random_sign_mask_t,
                                   nvte_stream_t stream);
int nvte_hadamard_transform_amax_status(const NVTETensor input, NVTETensor output,
                                        int random_sign_mask, int random_sign_mask_t,
                                        nvte_stream_t stream);
int nvte_quantize_with_hadamard_transform_status(const NVTETensor input, NVTETensor output,
                                                 const NVTETensor hadamard_matrix,
                                                 const NVTEQuantizationConfig quant_config,
                                                 nvte_stream_t stream);
int nvte_swizzle_scaling_factors_status(const NVTETensor input, NVTETensor output,
                                        nvte_stream_t stream);
int nvte_unswizzle_scaling_factors_status(const NVTETensor input, NVTETensor output,
                                          nvte_stream_t stream);
int nvte_multi_tensor_gemm_status(const NVTETensor *A, const NVTETensor *B, NVTETensor *D,
                                  const NVTETensor *bias, NVTETensor *pre_gelu_out, int num_gemms,
                                  int transa, int transb, int grad, NVTETensor *workspace,
                                  int accumulate, int use_split_accumulator, int math_sm_count,
                                  nvte_stream_t stream);
int nvte_multi_tensor_swizzle_scaling_factors_status(const NVTETensor *inputs, NVTETensor *outputs,
                                                     size_t num_tensors, nvte_stream_t stream);
int nvte_cublas_gemm_v2_status(int transa, int transb, const float *alpha, const NVTETensor A,
                               const NVTETensor B, const float *beta, const NVTETensor C,
                               NVTETensor D, NVTETensor workspace, NVTEMatmulConfig config,
                               nvte_stream_t stream);

#ifdef __cplusplus
} /* extern "C" */
#endif

#endif /* TRANSFORMER_ENGINE_B200_H_ */
