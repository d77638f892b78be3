/*
 * b200ops -- C ABI of the B200-native (sm_100a) decode hot path.
 *
 * This is the drop-in boundary for the operator surface of yiliu30/vllm-fork that BASELINE.json's
 * north_star names.  Every entry point replaces ONE reference operator (cited per function as
 * file:line relative to the reference tree) and keeps that operator's conventions
 * (SURVEY.md section 8b):
 *   - outputs are preallocated by the caller; nothing is allocated, nothing synchronises;
 *   - the launch is asynchronous on the `stream` handle given (a cudaStream_t); CUDA-graph capturable;
 *   - argument validation happens before any launch; a failed check returns a negative code and
 *     b200_last_error() describes it (the Python shim raises RuntimeError, like STD_TORCH_CHECK);
 *   - tensors are raw device pointers + explicit ELEMENT strides (strides are part of the ABI:
 *     K/V are views of the fused QKV output, the KV cache is a permuted view);
 *   - scales stay on the device as `const float*` and are read in-kernel.
 * No torch types appear here.  There is no CPU fallback: without a CUDA device every compute
 * entry point fails with B200_ERR_CUDA.
 */
#ifndef B200OPS_H_
#define B200OPS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

/* element types */
enum {
  B200_F32 = 0,
  B200_F16 = 1,
  B200_BF16 = 2,
  B200_FP8_E4M3 = 3,
  B200_U8 = 4,
  B200_I32 = 5,
  B200_I64 = 6
};

/* return codes */
enum {
  B200_OK = 0,
  B200_ERR_INVALID = -1,     /* argument check failed (shape / stride / alignment / dtype) */
  B200_ERR_UNSUPPORTED = -2, /* valid request this build does not cover */
  B200_ERR_CUDA = -3,        /* CUDA runtime / driver error (message holds cudaGetErrorString) */
  B200_ERR_WORKSPACE = -4    /* workspace too small */
};

typedef void* b200_stream_t; /* cudaStream_t */

/* Message of the last failure on the calling thread ("" if none). */
B200_API const char* b200_last_error(void);
/* ABI version: major*1000 + minor. */
B200_API int b200_version(void);
/* Number of SMs of the current CUDA device (148 on B200), or a negative error code. */
B200_API int b200_sm_count(void);
/* Programmatic dependent launch between consecutive b200 kernels on a stream (results are unchanged):
 *   0 = off; 1 (default) = the next kernel's prologue (barrier init, TMEM allocation, descriptor prefetch) runs
 *   under the previous kernel's tail; 2 = the GEMMs additionally request their first ring pass of WEIGHT tiles
 *   before waiting for the previous kernel, and b200_paged_attention_decode computes its tile plan from `seq_lens`
 *   before waiting. Because every b200 kernel releases its dependents at entry, such a pre-wait read can run beside
 *   ANY earlier kernel of the programmatic chain, not only the one launched immediately before. The caller's
 *   promise at level 2 is therefore: the weight operand (`b` of cutlass_scaled_mm / cutlass_scaled_fp4_mm) and
 *   `seq_lens` are not written by any kernel launched on the stream since the last non-programmatic operation
 *   (memcpy / memset, event wait, a foreign kernel, an NVLink collective in its default mode). That holds for model
 *   weights (written at load time) and for `seq_lens` (a step input copied in before the step); a caller that
 *   produces `b` on the stream must use level <= 1. The level is process-global host state: set it around your own
 *   launches and restore the returned value (b200ops.decoder.DecoderModel.forward_device does).
 *   Initial value from env B200_PDL. Returns the previous level.
 * The NVLink collectives (b200_allreduce_run, b200_allreduce_fused_add_rms_norm) are launched as plain kernels at every
 * level unless env B200_PDL_AR is 0 / 1 (measured slower: DESIGN.md section 5).
 * Not part of the reference operator surface. */
B200_API int b200_set_pdl(int level);

/* ------------------------------------------------------------------------------------------------
 * KV-cache write.
 * Replaces torch.ops._C_cache_ops.reshape_and_cache_flash
 *   (csrc/libtorch_stable/cache_kernels.cu:314-400 kernel, :795-864 host; vllm/_custom_ops.py:2614-2633).
 * key/value: [num_tokens, num_heads, head_size], last dim contiguous, token stride given.
 * key_cache/value_cache: logical [num_blocks, block_size, num_heads, head_size] views with element
 * strides (block_stride, page_stride, head_stride) -- NHD and HND are both just strides.
 * slot_mapping: int64 [num_tokens]; slot < 0 => token skipped. num_tokens = slot_mapping.size(0).
 * cache_dtype == in_dtype: plain copy ("auto"); cache_dtype == B200_FP8_E4M3:
 *   byte = e4m3_satfinite_rne(float(x) / scale), scale = k_scale[head * kv_scale_stride].
 * kv_scale_stride is 0 (per-tensor scale [1]) or 1 (per-head scale [num_heads]).
 */
B200_API int b200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                          void* value_cache, const int64_t* slot_mapping,
                                          int64_t num_tokens, int num_heads, int head_size,
                                          int block_size, int64_t key_stride, int64_t value_stride,
                                          int64_t block_stride, int64_t page_stride,
                                          int64_t head_stride, int in_dtype, int cache_dtype,
                                          const float* k_scale, const float* v_scale,
                                          int kv_scale_stride, b200_stream_t stream);

/* Replaces torch.ops._C_cache_ops.reshape_and_cache (legacy x-packed layout used by
 * paged_attention_v1/v2): csrc/libtorch_stable/cache_kernels.cu:254-312, :749-777.
 * key_cache [num_blocks, num_heads, head_size/x, block_size, x], value_cache
 * [num_blocks, num_heads, head_size, block_size], both contiguous; x = 16 / sizeof(cache elem). */
B200_API int b200_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                    void* value_cache, const int64_t* slot_mapping,
                                    int64_t num_tokens, int num_heads, int head_size, int block_size,
                                    int x, int64_t key_stride, int64_t value_stride, int in_dtype,
                                    int cache_dtype, const float* k_scale, const float* v_scale,
                                    b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * FP8 activation quantisers.  Replace torch.ops._C.{static,dynamic,dynamic_per_token}_scaled_fp8_quant
 *   (csrc/libtorch_stable/quantization/w8a8/fp8/common.cu:13-179 kernels, :183-422 hosts;
 *    conversion csrc/quantization/w8a8/fp8/common.cuh:58-79; vllm/_custom_ops.py:1832-1898).
 * input [num_tokens, hidden] with row stride in_row_stride (elements), out e4m3 with out_row_stride.
 */
/* out = e4m3(clamp(x * (1.0f / scale[gi*scale_stride_i + gj*scale_stride_j]), +-448)),
 * gi = row / group_m, gj = col / group_n  (per-tensor: both strides 0). */
B200_API int b200_fp8_quant_static(void* out, const void* input, const float* scale,
                                   int64_t num_tokens, int hidden, int64_t in_row_stride,
                                   int64_t out_row_stride, int group_m, int group_n,
                                   int64_t scale_stride_i, int64_t scale_stride_j, int in_dtype,
                                   b200_stream_t stream);
/* scale[0] = max|x| / 448 over the whole tensor (written), out = e4m3(clamp(x * (1/scale))). */
B200_API int b200_fp8_quant_dynamic_tensor(void* out, const void* input, float* scale,
                                           int64_t num_tokens, int hidden, int64_t in_row_stride,
                                           int64_t out_row_stride, int in_dtype,
                                           b200_stream_t stream);
/* scales[t] = max(min(max|x_t|, *scale_ub) / 448, 1/(448*512)); out = e4m3(clamp(x / scales[t])). */
B200_API int b200_fp8_quant_dynamic_token(void* out, const void* input, float* scales,
                                          const float* scale_ub /* nullable */, int64_t num_tokens,
                                          int hidden, int64_t in_row_stride, int64_t out_row_stride,
                                          int in_dtype, b200_stream_t stream);

/* Per-token-group quantiser (1 x group_size activation scales of block-scaled FP8 linears).
 * Replaces torch.ops._C.per_token_group_fp8_quant
 *   (csrc/libtorch_stable/quantization/w8a8/fp8/per_token_group_quant.cu:43-161,195-300;
 *    vllm/model_executor/layers/quantization/utils/fp8_utils.py:533-616).
 *   s = max(eps, max|x_group|) / fp8_max;  ue8m0: s = 2^ceil(log2(max(|s|, 1e-10)));
 *   out = e4m3(min(max(x / s, fp8_min), fp8_max));  scales[r*scale_stride_row + g*scale_stride_col] = s.
 * Row-major, column-major and TMA-aligned column-major scale tensors are stride choices (in floats).
 * hidden % group_size == 0. */
B200_API int b200_fp8_quant_token_group(void* out, float* scales, const void* input,
                                        int64_t num_tokens, int hidden, int group_size,
                                        int64_t in_row_stride, int64_t out_row_stride,
                                        int64_t scale_stride_row, int64_t scale_stride_col, float eps,
                                        float fp8_min, float fp8_max, int scale_ue8m0, int in_dtype,
                                        b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Paged-attention decode (one query token per sequence).
 * Replaces the decode branch of AttentionImpl.forward (vllm/v1/attention/backend.py:895-934) as
 * implemented by TritonAttentionImpl.forward -> unified_attention
 * (vllm/v1/attention/backends/triton_attn.py:580-730, ops/triton_unified_attention.py:179-772) and
 * FlashAttentionImpl.forward (backends/flash_attn.py:874-1110); the legacy
 * torch.ops._C.paged_attention_v1/v2 adapters below call the same kernel.
 *
 *   out[s,h,:] = softmax_t( softcap( scale*k_scale * q[s,h,:].K[s,t,h/g,:] ) + alibi[h]*(t-L_s+1) ) . V * v_scale
 *   over t in [max(0, L_s - sliding_window), L_s),  g = num_heads / num_kv_heads.
 *
 * q/out: [num_seqs, num_heads, head_size] (element strides q_stride_s/q_stride_h, last dim
 * contiguous), dtype q_dtype (BF16 or F16). lse (nullable): float [num_seqs, num_heads], natural log.
 * k_cache/v_cache: paged cache views, element (t, h, d) of physical block b lives at
 *   base + b*kv_block_stride + t*kv_token_stride + h*kv_head_stride + d   (elements of kv_dtype),
 * kv_dtype = q_dtype ("auto") or B200_FP8_E4M3 (dequant = float(byte) * *k_scale / *v_scale).
 * All byte strides and base pointers must be 16-byte aligned; head_size*sizeof(kv) >= 64.
 * block_table int32 [num_seqs, bt_stride]; seq_lens int32 [num_seqs]; block_size % 16 == 0.
 * workspace: b200_paged_attention_workspace_bytes(...) bytes of device memory owned by the caller
 * (the AttentionMetadataBuilder keeps one persistent buffer, like triton_attn.py:154-175).
 * Everything is read on the device: no host sync, CUDA-graph safe.
 */
typedef struct b200_paged_attn_args {
  void* out;
  float* lse; /* nullable */
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_table;
  const int32_t* seq_lens;
  const float* k_scale;      /* device scalar; nullable => 1.0 */
  const float* v_scale;      /* device scalar; nullable => 1.0 */
  const float* alibi_slopes; /* [num_heads] float, nullable */
  void* workspace;
  size_t workspace_bytes;
  int64_t num_blocks; /* physical blocks in the cache (bounds the TMA descriptor) */
  int64_t q_stride_s, q_stride_h;
  int64_t out_stride_s, out_stride_h;
  int64_t kv_block_stride, kv_token_stride, kv_head_stride;
  int64_t bt_stride;
  int32_t num_seqs, num_heads, num_kv_heads, head_size, block_size;
  int32_t max_seq_len; /* upper bound, used only for validation / heuristics */
  int32_t q_dtype, kv_dtype;
  int32_t sliding_window; /* <= 0: none */
  float scale;
  float softcap; /* <= 0: none */
  /* nullable. When set, `out` is e4m3 (1 byte / element, strides in elements) and holds
   * e4m3(clamp(attention * (1 / *output_scale))): the fused static FP8 output the reference kernels
   * offer through `output_scale` (backend.py:895-934, triton_unified_attention.py out_scale_inv). */
  const float* output_scale;
} b200_paged_attn_args;

B200_API size_t b200_paged_attention_workspace_bytes(int num_seqs, int num_heads, int num_kv_heads,
                                                     int head_size);
B200_API int b200_paged_attention_decode(const b200_paged_attn_args* args, b200_stream_t stream);

/* Paged-attention decode for callers that cannot allocate device memory.
 * No counterpart in the reference: vLLM always has torch's caching allocator at hand, so none of its operators
 * needs this. It exists for users of this C ABI who do not - a worker whose device memory was fully handed out
 * at start-up (weights, then every remaining byte as KV-cache blocks), for which a later cudaMalloc fails and
 * stalls the device, and bindings from runtimes without a device allocator (cgo, JNI, plain ctypes) that hold
 * only the buffers they were given.
 *
 * Same arguments, validation, kernels and results as b200_paged_attention_decode, except that
 * args->workspace / args->workspace_bytes are IGNORED (pass NULL / 0): the split-KV scratch is a fixed
 * B200_PAGED_ATTN_STATIC_WS_BYTES array inside the library's device image. Consequences a user must know:
 *   - Limit. A shape whose b200_paged_attention_workspace_bytes(...) exceeds the array is refused with
 *     B200_ERR_WORKSPACE (message names both numbers) before anything is launched. 8 MiB covers
 *     num_heads * (head_size + 2) <= 5418 at any batch up to 16384 sequences: e.g. 32 or 40 query heads of
 *     size 128 (Llama-3.1-8B unsharded, 70B from TP=2), 16 heads of size 256. Larger: use the regular entry.
 *   - Footprint. Every GPU on which the process loads this library holds the 8 MiB until exit, used or not.
 *   - Sharing. All calls on a device use the same array and it keeps the previous call's bytes. That is
 *     harmless for results (each call writes what it reads; nothing has to be zero), but two calls that can
 *     overlap on the GPU - different streams without an event between them, or different host threads - race
 *     on it and corrupt each other's split sequences. Calls on ONE stream are ordered and safe.
 *   - Set-up. The first call on each device resolves the array's device address (cudaGetSymbolAddress, with
 *     that device current), which can make the runtime load the library's device code there. Make that first
 *     call during initialisation, before any stream capture is begun. Later calls only read a table.
 * The entry point itself calls no cudaMalloc / cudaFree / allocating library routine. */
#define B200_PAGED_ATTN_STATIC_WS_BYTES (8u << 20)
B200_API int b200_paged_attention_decode_static(const b200_paged_attn_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Variable-length paged attention: prefill, chunked prefill and mixed batches (query_len >= 1).
 * Replaces the query_len > 1 branch of AttentionImpl.forward (vllm/v1/attention/backend.py:895-934)
 * as implemented by unified_attention (vllm/v1/attention/ops/triton_unified_attention.py:802-1140,
 * kernel :179-683; reference semantics tests/kernels/attention/test_triton_unified_attention.py:33-88)
 * and flash_attn_varlen_func(..., block_table=...) (vllm/v1/attention/backends/flash_attn.py:1084).
 *   q:   [num_tokens, num_heads, head_size] bf16 / f16, packed; cu_seqlens_q: int32 [num_seqs + 1]
 *   k_cache / v_cache: [num_blocks, block_size, num_kv_heads, head_size] (strided), q dtype or e4m3
 *   seq_lens: int32 [num_seqs] = context length + query length; block_table: int32 [num_seqs, bt_stride]
 *   query token i of sequence s (context c = seq_len - query_len) attends kv positions j <= c + i, and
 *   with sliding_window W > 0 only those with (c + i - j) < W (W = 1 + window_size[0] of the reference).
 *   softcap: s -> softcap * tanh(s / softcap) before masking; ALiBi adds slope * (j - c) (kernel :392-397).
 *   out: [num_tokens, num_heads, head_size] q dtype, or e4m3 when output_scale is set;
 *   lse: optional float [num_tokens, num_heads], natural log.
 * Device-only inputs, no host sync, CUDA-graph safe. head_size 64 / 128, block_size % 16 == 0.
 * Three kernels behind this entry point: tcgen05 / TMEM / TMA for head_size 128 with bf16 / f16 caches
 * (csrc/attention_prefill_tc.cuh) and with FP8 caches (csrc/attention_prefill_tc8.cuh: raw FP8 pages by TMA, converter
 * warps expand them to f16), and mma.sync for head_size 64 (csrc/attention_prefill.cu). B200_PREFILL_IMPL=mma forces
 * the last one.
 */
typedef struct b200_varlen_attn_args {
  void* out;
  float* lse; /* nullable */
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int32_t* cu_seqlens_q;
  const int32_t* seq_lens;
  const int32_t* block_table;
  const float* k_scale;      /* device scalar; nullable => 1.0 */
  const float* v_scale;      /* device scalar; nullable => 1.0 */
  const float* alibi_slopes; /* [num_heads] float, nullable */
  const float* output_scale; /* nullable: fused static FP8 output */
  int64_t q_stride_t, q_stride_h;     /* elements */
  int64_t out_stride_t, out_stride_h; /* elements of the out dtype */
  int64_t kv_block_stride, kv_token_stride, kv_head_stride; /* elements of the kv dtype */
  int64_t bt_stride;
  int64_t num_blocks; /* physical blocks in the cache (bounds the TMA descriptors of the tcgen05 kernel) */
  int32_t num_seqs, num_tokens, num_heads, num_kv_heads, head_size, block_size;
  int32_t q_dtype, kv_dtype;
  int32_t causal;         /* must be 1 */
  int32_t sliding_window; /* <= 0: none */
  float scale;
  float softcap; /* <= 0: none */
} b200_varlen_attn_args;

B200_API int b200_paged_attention_varlen(const b200_varlen_attn_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * FP8 W8A8 GEMM with scale epilogue (tcgen05 / TMEM / TMA).
 * Replaces torch.ops._C.cutlass_scaled_mm
 *   (csrc/libtorch_stable/quantization/w8a8/cutlass/scaled_mm_entry.cu:197-270; sm100 kernels
 *    c3x/scaled_mm_sm100_fp8_dispatch.cuh:18-295; vllm/_custom_ops.py:756-804).
 *   out[M,N] = a_scale (.) (b_scale (.) (a[M,K] . b[K,N])) (+ bias[N]), fp32 accumulate.
 * a: e4m3 [M,K] row-major (row stride lda); b: e4m3 [K,N] COLUMN-major, i.e. the weight stored
 * [N,K] with row stride ldb; K, lda, ldb multiples of 16. a_scale: float [1] or [M];
 * b_scale: float [1] or [N]; bias: [N] in out dtype or NULL; out bf16/f16 with row stride ldo.
 */
B200_API int b200_fp8_gemm(void* out, const void* a, const void* b, const float* a_scale,
                           const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                           int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                           int out_dtype, b200_stream_t stream);

/* Block-scaled FP8 GEMM (DeepSeek-style 1 x 128 activation scales, 128 x 128 weight scales).
 * Replaces the blockwise branch of torch.ops._C.cutlass_scaled_mm
 *   (csrc/libtorch_stable/quantization/w8a8/cutlass/c3x/scaled_mm_helper.hpp:40-60,
 *    c3x/scaled_mm_blockwise_sm100_fp8_dispatch.cuh; oracle tests/kernels/quant_utils.py:91-155):
 *   out[m,n] = sum_kb a_scales[m,kb] * b_scales[kb, n/128] * (a[m, 128 kb ..] . b[128 kb .., n]).
 * a / b / out as b200_fp8_gemm. a_scales float [M, ceil(K/128)] with element strides
 * (as_stride_m, as_stride_k) - row-major, column-major and TMA-aligned layouts all work;
 * b_scales float [ceil(K/128), ceil(N/128)] with element strides (bs_stride_k, bs_stride_n). No bias
 * (the reference rejects it). N and K need not be multiples of 128 (K % 16 == 0).
 * split: K slices over the CTAs of one thread-block cluster for batches of at most 128 tokens (partials
 * reduced through distributed shared memory in slice order: deterministic, no workspace); 0 = auto,
 * 1 = off, else an upper bound (powers of two up to 8). */
B200_API int b200_fp8_gemm_blockscaled(void* out, const void* a, const void* b, const float* a_scales,
                                       const float* b_scales, int M, int N, int K, int64_t lda,
                                       int64_t ldb, int64_t ldo, int64_t as_stride_m, int64_t as_stride_k,
                                       int64_t bs_stride_k, int64_t bs_stride_n, int out_dtype, int split,
                                       b200_stream_t stream);

/* gate_up GEMM + silu_and_mul + static FP8 quant in ONE kernel, decode batches (M <= 256).
 * Replaces the sequence cutlass_scaled_mm (gate_up_proj) -> torch.ops._C.silu_and_mul_quant
 * (csrc/libtorch_stable/quantization/activation_kernels.cu:33-104; the reference fuses the activation
 * and the quantisation, this also folds them into the GEMM epilogue) and is bit-identical to that pair.
 * w_gate_up: e4m3 [2*inter, K] row-major (gate rows then up rows, MergedColumnParallelLinear order),
 * a: e4m3 [M, K]; scales as b200_fp8_gemm (w_scale per-tensor or per-row [2*inter]); bias [2*inter] in
 * act_dtype or NULL; quant_scale: device float[1] of the next GEMM's static input scale;
 * out8: e4m3 [M, inter], row stride ldo bytes. inter % 128 == 0. */
B200_API int b200_fp8_gemm_silu_mul_fp8(void* out8, const void* a, const void* w_gate_up, const float* a_scale,
                                        const float* w_scale, const void* bias, const float* quant_scale,
                                        int M, int inter, int K, int64_t lda, int64_t ldw, int64_t ldo,
                                        int a_scale_per_row, int w_scale_per_row, int act_dtype,
                                        b200_stream_t stream);

/* Same contract as b200_fp8_gemm, third-generation kernel (persistent CTA pairs with
 * tcgen05.mma.cta_group::2, stream-K split of the (tile, k-block) space, double-buffered TMEM).
 * Needs a caller-owned, 256-byte aligned workspace of b200_fp8_gemm_v3_workspace_bytes() bytes that was
 * ZEROED once (it carries the stream-K partial tiles and their flags; the kernel restores the flags
 * to zero before it ends, so the same workspace serves every call on one stream). */
B200_API size_t b200_fp8_gemm_v3_workspace_bytes(void);
B200_API int b200_fp8_gemm_v3(void* out, const void* a, const void* b, const float* a_scale,
                              const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                              int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                              int out_dtype, void* workspace, size_t workspace_bytes,
                              b200_stream_t stream);

/* Same contract as b200_fp8_gemm for decode batches (M <= 256) whose N gives fewer 256-row weight tiles than the
 * chip has CTA pairs: the K range of each tile is split over the `split` CTA pairs of one thread-block cluster
 * (tcgen05.mma.cta_group::2 per pair), the fp32 partials are reduce-scattered through distributed shared memory
 * and summed in pair order (deterministic). split = 0 picks the largest split (<= 4) that keeps all clusters
 * co-resident. No workspace. out must be 16-byte aligned with ldo * 2 a multiple of 16. */
B200_API int b200_fp8_gemm_splitk(void* out, const void* a, const void* b, const float* a_scale,
                                  const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                                  int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                                  int out_dtype, int split, b200_stream_t stream);

/* Same contract as b200_fp8_gemm for decode batches (M <= 256): two SMs issue ONE tcgen05.mma.cta_group::2 of
 * UMMA M = 128 per k-step (swap-AB: 128 weight rows x batch_tile batch rows), each CTA staging its 64 weight rows
 * and HALF of the batch tile, which halves the bytes an SM has to pull from L2 (the bound of these kernels, ~65 B per
 * clock and SM) against the one-CTA kernel. This is the tile shape of the reference's CUTLASS config for
 * M in (64, 256] (c3x/scaled_mm_sm100_fp8_dispatch.cuh:118-134). batch_tile: 0 = auto (the smallest of 32 / 64 /
 * 128 / 256 whose grid still fits one wave). No workspace. out must be 16-byte aligned with ldo * 2 a multiple of 16. */
B200_API int b200_fp8_gemm_pair(void* out, const void* a, const void* b, const float* a_scale,
                                const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                                int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                                int out_dtype, int batch_tile, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Gated activation between the MLP / MoE GEMMs.
 * Replaces torch.ops._C.silu_and_mul (csrc/libtorch_stable/activation_kernels.cu:102-160,291-296):
 *   out[t,i] = silu(x[t,i]) * x[t,d+i], input [num_tokens, 2d], out [num_tokens, d] same dtype;
 * and torch.ops._C.silu_and_mul_quant (csrc/libtorch_stable/quantization/activation_kernels.cu:33-104):
 *   out e4m3 = quant_static(silu_and_mul(x), *scale).
 */
B200_API int b200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d,
                               int64_t in_row_stride, int64_t out_row_stride, int dtype,
                               b200_stream_t stream);
B200_API int b200_silu_and_mul_fp8(void* out, const void* input, const float* scale,
                                   int64_t num_tokens, int d, int64_t in_row_stride,
                                   int64_t out_row_stride, int in_dtype, b200_stream_t stream);

/* Replaces torch.ops._C.silu_and_mul_nvfp4_quant (csrc/libtorch_stable/quantization/fp4/
 * activation_nvfp4_quant_fusion_kernels.cu:37-175): input [num_tokens, 2d] f16/bf16,
 * out uint8 [num_tokens, d/2] = nvfp4(T(silu(gate) * up)), out_scale in the swizzled 128x4 layout
 * (see b200_nvfp4_quant), global_scale device float[1]; d % 16 == 0. */
B200_API int b200_silu_and_mul_nvfp4(void* out, void* out_scale, const void* input,
                                     const float* global_scale, int64_t num_tokens, int d,
                                     int64_t in_row_stride, int in_dtype, b200_stream_t stream);

/* rotary_embedding + reshape_and_cache_flash in ONE launch: the backend hook
 * AttentionImpl.do_rope_and_kv_cache_update (vllm/v1/attention/backend.py:950-1006) = the pair
 * torch.ops._C.rotary_embedding (pos_encoding_kernels.cu:8-99) -> _C_cache_ops::reshape_and_cache_flash
 * (cache_kernels.cu:314-400); bit-identical to running the two. query [T, num_heads, head_size] and key
 * [T, num_kv_heads, head_size] are rotated in place (token strides given, heads contiguous); value
 * [T, num_kv_heads, head_size]; cache arguments as b200_reshape_and_cache_flash; cos_sin_cache
 * [max_pos, rot_dim] in the q dtype; positions, slot_mapping int64 [T] (slot < 0: rotated, not cached). */
B200_API int b200_rope_and_cache_flash(const int64_t* positions, void* query, void* key, const void* value,
                                       const void* cos_sin_cache, void* key_cache, void* value_cache,
                                       const int64_t* slot_mapping, int64_t num_tokens, int rot_dim,
                                       int head_size, int num_heads, int num_kv_heads, int block_size,
                                       int64_t query_stride, int64_t key_stride, int64_t value_stride,
                                       int64_t block_stride, int64_t page_stride, int64_t head_stride,
                                       int is_neox, int dtype, int cache_dtype, const float* k_scale,
                                       const float* v_scale, int kv_scale_stride, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Merge of two partial attention results by log-sum-exp.
 * Replaces torch.ops._C.merge_attn_states (csrc/libtorch_stable/attention/merge_attn_states.cu:13-330;
 * vllm/v1/attention/ops/merge_attn_states.py:9-60).
 * prefix_output / suffix_output [T, H, D] f32/f16/bf16 (token strides given, one shared head stride,
 * last dim contiguous, D % (16/sizeof) == 0); prefix_lse / suffix_lse float [H, T] (strides given);
 * output [T, H, D] same dtype, or e4m3 when output_scale != NULL (static quantisation by 1/scale);
 * output_lse nullable. Tokens >= prefill_tokens_with_context copy the suffix.
 * prefill_tokens_with_context < 0 means "all tokens have context".
 */
B200_API int b200_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                                    const float* prefix_lse, const void* suffix_output,
                                    const float* suffix_lse, const float* output_scale,
                                    int64_t num_tokens, int num_heads, int head_size,
                                    int64_t prefix_tok_stride, int64_t suffix_tok_stride,
                                    int64_t in_head_stride, int64_t out_tok_stride,
                                    int64_t out_head_stride, int64_t prefix_lse_head_stride,
                                    int64_t prefix_lse_tok_stride, int64_t suffix_lse_head_stride,
                                    int64_t suffix_lse_tok_stride, int64_t out_lse_head_stride,
                                    int64_t out_lse_tok_stride, int64_t prefill_tokens_with_context,
                                    int in_dtype, int out_dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Neighbours of the hot path: RMSNorm (+ residual add) fused with the
 * static FP8 quantisation of the next GEMM input, and rotary embedding.
 * Replace torch.ops._C.rms_norm / fused_add_rms_norm (csrc/libtorch_stable/layernorm_kernels.cu),
 * rms_norm_static_fp8_quant / fused_add_rms_norm_static_fp8_quant
 * (csrc/libtorch_stable/layernorm_quant_kernels.cu:21-156) and rotary_embedding
 * (csrc/libtorch_stable/pos_encoding_kernels.cu:8-160). hidden % 8 == 0, half / bfloat16.
 */
B200_API int b200_rms_norm(void* out, const void* input, const void* weight, float eps,
                           int64_t num_tokens, int hidden, int64_t in_row_stride, int dtype,
                           b200_stream_t stream);
/* residual += input (rounded to dtype); input = rms_norm(residual) * weight  (both in place) */
B200_API int b200_fused_add_rms_norm(void* input, void* residual, const void* weight, float eps,
                                     int64_t num_tokens, int hidden, int64_t in_row_stride, int dtype,
                                     b200_stream_t stream);
B200_API int b200_rms_norm_static_fp8_quant(void* out, const void* input, const void* weight,
                                            const float* scale, float eps, int64_t num_tokens,
                                            int hidden, int64_t in_row_stride, int dtype,
                                            b200_stream_t stream);
/* residual += input; out(e4m3) = quant_static(rms_norm(residual) * weight, *scale) */
B200_API int b200_fused_add_rms_norm_static_fp8_quant(void* out, void* input, void* residual,
                                                      const void* weight, const float* scale,
                                                      float eps, int64_t num_tokens, int hidden,
                                                      int64_t in_row_stride, int dtype,
                                                      b200_stream_t stream);
/* In-place RoPE on query [T, num_heads, head] and key [T, num_kv_heads, head] (key nullable);
 * cos_sin_cache [max_pos, rot_dim] (cos | sin) in the same dtype; positions int64 [T]. */
B200_API int b200_rotary_embedding(const int64_t* positions, void* query, void* key,
                                   const void* cos_sin_cache, int64_t num_tokens, int rot_dim,
                                   int64_t query_stride, int64_t key_stride, int64_t head_stride,
                                   int num_heads, int num_kv_heads, int is_neox, int dtype,
                                   b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Fused MoE.
 * b200_moe_align_block_size replaces torch.ops._moe_C.moe_align_block_size
 *   (csrc/libtorch_stable/moe/moe_align_sum_kernels.cu:100-350,625-723): histogram of topk_ids per
 *   (mapped) expert, counts padded to block_size, exclusive scan; sorted_token_ids (pad = numel),
 *   expert_ids per block (-1 tail), num_tokens_post_pad. ids < 0, >= num_experts or mapped to -1
 *   by expert_map (int32 [num_experts], nullable) are dropped. Tokens of one expert keep ascending
 *   flat-index order (deterministic). inv_perm (nullable, int32 [numel]): sorted position of each
 *   flat (token, k) index, -1 if dropped -- consumed by b200_moe_unpermute_reduce.
 * b200_moe_topk_softmax replaces torch.ops._moe_C.topk_softmax (topk_softmax_kernels.cu:76-262).
 * b200_moe_sum replaces torch.ops._moe_C.moe_sum (moe_align_sum_kernels.cu:352-500, 757-930).
 * b200_moe_permute_fp8 / b200_moe_unpermute_reduce: moe_permute / moe_unpermute
 *   (csrc/libtorch_stable/moe/moe_permute_unpermute_op.cu) fused with the static FP8 quantisation
 *   of the GEMM input and with the topk-weighted sum respectively. inv_perm entries < 0 or
 *   >= *num_valid_rows (device int32, nullable) are skipped (dropped (token, k) pairs).
 * b200_moe_grouped_gemm_fp8: the grouped GEMM of fused_experts
 *   (vllm/model_executor/layers/fused_moe/fused_moe.py:299-612; cutlass_moe_mm,
 *   csrc/libtorch_stable/quantization/w8a8/cutlass/moe/grouped_mm_c3x_sm100.cu:16-115):
 *   out[i,:] = a_scale * w_scale[e] * (a_sorted[i,:] . w[e]^T) (* row_weight[i]) (a_scale: one float, or one per
 *   sorted row when a_scale_per_row) for every block of
 *   block_size sorted rows with expert e = expert_ids[block] >= 0. w: e4m3 [E, N, K] contiguous.
 */
B200_API int b200_moe_align_block_size(const void* topk_ids, int topk_ids_dtype, int64_t numel,
                                       int num_experts, int block_size, int32_t* sorted_token_ids,
                                       int64_t sorted_len, int32_t* expert_ids, int64_t expert_ids_len,
                                       int32_t* num_tokens_post_pad, const int32_t* expert_map,
                                       int32_t* inv_perm, b200_stream_t stream);
B200_API int b200_moe_topk_softmax(float* topk_weights, int32_t* topk_ids,
                                   int32_t* token_expert_indices, const void* gating_output,
                                   const float* bias, int64_t num_tokens, int num_experts, int topk,
                                   int renormalize, int gating_dtype, b200_stream_t stream);
B200_API int b200_moe_sum(void* out, const void* input, int64_t num_tokens, int topk, int d,
                          int64_t stride_token, int64_t stride_topk, const void* topk_ids,
                          int topk_ids_dtype, const int32_t* expert_map, int dtype,
                          b200_stream_t stream);
B200_API int b200_moe_permute_fp8(void* dst, const void* src, const int32_t* sorted_token_ids,
                                  const int32_t* num_tokens_post_pad, const float* scale,
                                  int64_t max_rows, int topk, int64_t numel, int hidden,
                                  int64_t src_stride, int src_dtype, b200_stream_t stream);
B200_API int b200_moe_unpermute_reduce(void* out, const void* src, const int32_t* inv_perm,
                                       const float* topk_weights, const int32_t* num_valid_rows,
                                       int64_t num_tokens, int topk, int hidden, int64_t out_stride,
                                       int dtype, b200_stream_t stream);
B200_API int b200_moe_grouped_gemm_fp8(void* out, const void* a_sorted, const void* w,
                                       const int32_t* expert_ids, const int32_t* num_tokens_post_pad,
                                       const float* a_scale, const float* w_scale,
                                       const float* row_weight, int64_t max_rows, int num_experts,
                                       int N, int K, int block_size, int w_scale_per_channel,
                                       int a_scale_per_row, int out_dtype, b200_stream_t stream);
/* Per-token / block-scaled activation quantisation happens in TOKEN order (scaled_fp8_quant with
 * use_per_token_if_dynamic, per_token_group_quant_fp8: fused_moe/utils.py:138-161), then
 * b200_moe_permute_rows gathers the quantised rows and their scale rows into sorted order:
 * dst[i,:] = src[sorted_token_ids[i] / topk, :], scale_dst[i, 0..scale_cols) likewise (pad rows: zeros).
 * b200_moe_grouped_gemm_fp8_blockscaled: fused_experts with block_shape = [128, 128] (fused_moe.py:478-560):
 * out[i,n] = sum_kb a_scales[i,kb] * w_scales[e, n/128, kb] * (a_sorted[i, kb-th 128 columns] . w[e, n, same]),
 * a_scales float [max_rows, ceil(K/128)] row-major, w_scales float [E, ceil(N/128), ceil(K/128)]. */
B200_API int b200_moe_permute_rows(void* dst, const void* src, const int32_t* sorted_token_ids,
                                   const int32_t* num_tokens_post_pad, float* scale_dst, const float* scale_src,
                                   int scale_cols, int64_t scale_src_stride, int64_t max_rows, int topk,
                                   int64_t numel, int64_t row_bytes, int64_t src_stride_bytes,
                                   b200_stream_t stream);
B200_API int b200_moe_grouped_gemm_fp8_blockscaled(void* out, const void* a_sorted, const void* w,
                                                   const int32_t* expert_ids,
                                                   const int32_t* num_tokens_post_pad, const float* a_scales,
                                                   const float* w_scales, int64_t max_rows, int num_experts,
                                                   int N, int K, int block_size, int out_dtype,
                                                   b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Legacy-layout paged attention. Replaces torch.ops._C.paged_attention_v1 / paged_attention_v2
 * (op absent from the reference tree, present in the wheel; semantics
 * tests/kernels/attention/test_attention.py:57-109, docs/design/paged_attention.md:45-139,
 * vllm/v1/attention/ops/paged_attn.py:15-51; schema SURVEY.md section 8b).
 * query/out [num_seqs, num_heads, head_size] (q_stride = query.stride(0), heads contiguous; out
 * contiguous); key_cache [num_blocks, num_kv_heads, head_size/x, block_size, x], value_cache
 * [num_blocks, num_kv_heads, head_size, block_size] with kv_block_stride = key_cache.stride(0),
 * kv_head_stride = key_cache.stride(1) (elements), x = 16 / sizeof(cache element);
 * block_tables int32 [num_seqs, bt_stride]; seq_lens int32; alibi_slopes float[num_heads] or NULL;
 * kv_dtype = q_dtype ("auto") or B200_FP8_E4M3 with device scalars k_scale / v_scale.
 * v2: exp_sums, max_logits float [num_seqs, num_heads, max_num_partitions] and tmp_out (q dtype)
 * [num_seqs, num_heads, max_num_partitions, head_size] are written (512-token partitions, natural-log
 * max logits, partition outputs normalised by exp_sum + 1e-6) and then reduced into out.
 * blocksparse_vert_stride > 1 (block-sparse attention) is B200_ERR_UNSUPPORTED.
 */
B200_API int b200_paged_attention_v1(void* out, const void* query, const void* key_cache,
                                     const void* value_cache, int num_seqs, int num_heads,
                                     int num_kv_heads, int head_size, float scale,
                                     const int32_t* block_tables, const int32_t* seq_lens,
                                     int block_size, int max_seq_len, const float* alibi_slopes,
                                     int q_dtype, int kv_dtype, const float* k_scale,
                                     const float* v_scale, int64_t q_stride, int64_t kv_block_stride,
                                     int64_t kv_head_stride, int64_t bt_stride,
                                     int blocksparse_vert_stride, b200_stream_t stream);
B200_API int b200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                     const void* query, const void* key_cache,
                                     const void* value_cache, int num_seqs, int num_heads,
                                     int num_kv_heads, int head_size, float scale,
                                     const int32_t* block_tables, const int32_t* seq_lens,
                                     int block_size, int max_seq_len, const float* alibi_slopes,
                                     int q_dtype, int kv_dtype, const float* k_scale,
                                     const float* v_scale, int64_t q_stride, int64_t kv_block_stride,
                                     int64_t kv_head_stride, int64_t bt_stride, int max_num_partitions,
                                     int blocksparse_vert_stride, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * NVFP4 activation quantiser.
 * Replaces torch.ops._C.scaled_fp4_quant.out (csrc/libtorch_stable/quantization/fp4/
 * nvfp4_quant_kernels.cu:39-281, nvfp4_utils.cuh:158-302; vllm/_custom_ops.py:1523-1598).
 * input [M, K] f16/bf16 (K % 16 == 0); out uint8 [M, K/2] (low nibble = even element);
 * out_scale e4m3 bytes: swizzled 128x4 layout [round_up(M,128), round_up(K/16,4)] (pads zeroed) or
 * row-major [M, K/16]; global_scale: device float[1].
 */
B200_API int b200_nvfp4_quant(void* out, void* out_scale, const void* input,
                              const float* global_scale, int64_t num_tokens, int hidden,
                              int64_t in_row_stride, int is_sf_swizzled_layout, int in_dtype,
                              b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * NVFP4 W4A4 GEMM (block-scaled tcgen05 MMA).
 * Replaces torch.ops._C.cutlass_scaled_fp4_mm (csrc/libtorch_stable/quantization/fp4/
 * nvfp4_scaled_mm_kernels.cu:76-140,279-348; vllm/_custom_ops.py:733-745).
 * out[M,N] = alpha * (a (.) block_scale_a)(b (.) block_scale_b)^T, fp32 accumulate.
 * a uint8 [M, K/2], b uint8 [N, K/2] (two e2m1 per byte, low nibble = even k; row strides in BYTES);
 * block_scale_a / block_scale_b: e4m3 bytes in the swizzled 128x4 layout of b200_nvfp4_quant,
 * [round_up(rows,128), round_up(K/16,4)]; alpha: device float[1]; out bf16/f16 [M, N], row stride ldo.
 * K is the number of e2m1 ELEMENTS (K % 32 == 0).
 */
B200_API int b200_nvfp4_gemm(void* out, const void* a, const void* b, const void* block_scale_a,
                             const void* block_scale_b, const float* alpha, int M, int N, int K,
                             int64_t lda_bytes, int64_t ldb_bytes, int64_t ldo, int out_dtype,
                             b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * NVFP4 fused MoE (packed, expert-sorted rows without block padding: the CUTLASS-style layout).
 * b200_moe_mm_data replaces torch.ops._C.get_cutlass_moe_mm_data
 *   (csrc/libtorch_stable/quantization/w8a8/cutlass/moe/moe_data.cu:17-117,236-289; schema
 *   csrc/libtorch_stable/torch_bindings.cpp get_cutlass_moe_mm_data; vllm/_custom_ops.py:847-899):
 *   topk_ids int32 [numel] -> expert_offsets [E+1], problem_sizes1 / problem_sizes2 [E,3]
 *   ((m_e, n1, k) / (m_e, k, n), n1 = is_gated ? 2n : n; (n1, m_e, k) / (k, m_e, n) when swap_ab),
 *   input_permutation [numel] (source token of each packed row), output_permutation [numel] (packed row of
 *   each flat (token, k) index; the packed row count for ids == -1; input_permutation rows past the packed
 *   row count are -1), blockscale_offsets [E+1] (nullable;
 *   running sum of m_e rounded up to 128). Rows of one expert keep ascending flat-index order.
 * b200_shuffle_rows replaces torch.ops._moe_C.shuffle_rows (csrc/libtorch_stable/moe/
 *   moe_permute_unpermute_op.cu; vllm/_custom_ops.py:919-931): dst[i,:] = src[dst2src_map[i],:], rows of
 *   row_bytes (multiple of 16) bytes; a map entry outside [0, src_rows) gives a zero row.
 * b200_nvfp4_experts_quant / b200_silu_and_mul_nvfp4_experts_quant replace
 *   torch.ops._C.scaled_fp4_experts_quant / silu_and_mul_scaled_fp4_experts_quant
 *   (csrc/libtorch_stable/quantization/fp4/nvfp4_experts_quant.cu:36-250, nvfp4_quant_entry.cu:120;
 *   vllm/_custom_ops.py:1601-1725): input [m_topk, k] (or [m_topk, 2k] gate|up) f16/bf16 sorted by expert,
 *   global_scale float [E] (nullable = 1), out uint8 [m_topk, k/2], out_scale e4m3 bytes: one swizzled
 *   128x4-tile matrix per expert starting blockscale_offsets[e] rows (x round_up(k/16,4) bytes) in.
 * b200_nvfp4_moe_gemm replaces torch.ops._C.cutlass_fp4_group_mm (csrc/libtorch_stable/quantization/fp4/
 *   nvfp4_blockwise_moe_kernel.cu:676-746; vllm/_custom_ops.py:1010-1046):
 *   out[r,:] = alphas[e] * (a[r,:] (.) sfa_e) (w[e] (.) sfw_e)^T for the problem_sizes[e][0] rows of expert e
 *   starting at expert_offsets[e]; a uint8 [max_rows, K/2], w uint8 [E, N, K/2], a_blockscale as written by
 *   b200_nvfp4_experts_quant at sf_offsets[e], w_blockscale [E, round_up(N,128), round_up(K/16,4)] swizzled
 *   per expert, out bf16/f16 [max_rows, N] contiguous. K = e2m1 ELEMENTS (K % 32 == 0), N % 8 == 0.
 */
B200_API int b200_moe_mm_data(const int32_t* topk_ids, int32_t* expert_offsets, int32_t* problem_sizes1,
                              int32_t* problem_sizes2, int32_t* input_permutation,
                              int32_t* output_permutation, int32_t* blockscale_offsets, int num_experts,
                              int64_t numel, int topk, int n, int k, int is_gated, int swap_ab,
                              b200_stream_t stream);
B200_API int b200_shuffle_rows(void* dst, const void* src, const int32_t* dst2src_map, int64_t dst_rows,
                               int64_t src_rows, int64_t row_bytes, int64_t src_stride_bytes,
                               b200_stream_t stream);
B200_API int b200_nvfp4_experts_quant(void* out, void* out_scale, const void* input,
                                      const float* global_scale, const int32_t* expert_offsets,
                                      const int32_t* blockscale_offsets, int num_experts, int64_t m_topk, int k,
                                      int64_t in_row_stride, int in_dtype, b200_stream_t stream);
B200_API int b200_silu_and_mul_nvfp4_experts_quant(void* out, void* out_scale, const void* input,
                                                   const float* global_scale, const int32_t* expert_offsets,
                                                   const int32_t* blockscale_offsets, int num_experts,
                                                   int64_t m_topk, int k, int64_t in_row_stride, int in_dtype,
                                                   b200_stream_t stream);
B200_API int b200_nvfp4_moe_gemm(void* out, const void* a, const void* w, const void* a_blockscale,
                                 const void* w_blockscale, const float* alphas, const int32_t* problem_sizes,
                                 const int32_t* expert_offsets, const int32_t* sf_offsets, int64_t max_rows,
                                 int num_experts, int N, int K, int out_dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Tensor-parallel all-reduce over NVLink peer memory (one process per GPU).
 * Replaces torch.ops._C_custom_ar.{init_custom_ar, register_buffer, all_reduce, dispose, meta_size}
 * (csrc/custom_all_reduce.cuh:7-120, csrc/custom_all_reduce.cu; python life cycle
 * vllm/distributed/device_communicators/custom_all_reduce.py:56-560) and, fused, the pair
 * all_reduce -> fused_add_rms_norm_static_fp8_quant (docs/design/fusions.md:23 fuse_allreduce_rms).
 *
 * Life cycle: every rank b200_ipc_alloc()s one symmetric buffer, the 64-byte handles are exchanged by
 * the host (torch.distributed all_gather), peers b200_ipc_open() them, and b200_allreduce_init() gets
 * the per-rank pointers (own buffer at index `rank`). Buffer layout: [signals (b200_allreduce_meta_bytes)
 * | staging (max_message_bytes) | scratch (max_message_bytes) | user heap]. Tensors placed in the user
 * heap at the same offset on every rank are reduced in place by the peers (zero copy); any other input
 * is staged by the kernel. These calls DO allocate / synchronise (init-time only); run/fused do not.
 */
B200_API int b200_ipc_alloc(size_t bytes, void** ptr, void* handle64);   /* cudaMalloc + zero + IPC handle */
B200_API int b200_ipc_open(const void* handle64, void** ptr);
B200_API int b200_ipc_close(void* ptr);
B200_API int b200_ipc_free(void* ptr);
B200_API size_t b200_allreduce_meta_bytes(void);
B200_API int b200_allreduce_init(void** ctx_out, int rank, int world, void* const* buffer_ptrs,
                                 size_t buffer_bytes, size_t max_message_bytes);
B200_API int b200_allreduce_heap(void* ctx, size_t* offset, size_t* bytes);
B200_API int b200_allreduce_register(void* ctx, void* const* buffer_ptrs, size_t bytes);
B200_API int b200_allreduce_destroy(void* ctx);
/* NVSwitch multicast (NVLS) mapping of the symmetric buffer passed to b200_allreduce_init: one virtual address bound
 * to the same allocation on every rank (cuMulticastBindMem; torch symmetric memory's multicast_ptr). With it
 * b200_allreduce_run algo 3 and b200_allreduce_fused_add_rms_norm reduce INSIDE the switch (multimem.ld_reduce,
 * fp32 accumulation) and broadcast with multimem.st - what the reference reaches through
 * vllm/distributed/device_communicators/symm_mem.py:127-150 (multimem_all_reduce_). NULL = peer loads / stores. */
B200_API int b200_allreduce_set_multicast(void* ctx, void* multicast_ptr);
B200_API int b200_allreduce_has_multicast(void* ctx);
/* out = sum over ranks of in (fp32 accumulate in rank order: bit-identical on all ranks), out-of-place
 * (in == out allowed only for staged inputs). numel * sizeof(dtype) % 16 == 0.
 * algo: 0 auto (reference thresholds), 1 one-shot, 2 two-shot, 3 NVSwitch multimem (needs b200_allreduce_set_multicast
 * and an input inside the first symmetric buffer; the switch's accumulation order is not the rank order: equal within
 * one fp32 rounding, identical on all ranks when out also lives in the symmetric buffer); +4 skips the end barrier. */
B200_API int b200_allreduce_run(void* ctx, const void* in, void* out, int64_t numel, int dtype, int algo,
                                b200_stream_t stream);
/* One kernel: z = T(T(sum_r in_r) + residual); residual = z; y = rmsnorm(z) * weight;
 * out = scale ? e4m3(clamp(T(y) / *scale)) : T(y), for all num_tokens rows on every rank.
 * Rank r computes rows [r*Mp, (r+1)*Mp), Mp = ceil(num_tokens / world), and `residual` holds ONLY those
 * rows ([Mp, hidden], NULL = no residual add). `out` must live in the symmetric heap. */
B200_API int b200_allreduce_fused_add_rms_norm(void* ctx, void* out, const void* in, void* residual,
                                               const void* weight, const float* scale, float eps,
                                               int64_t num_tokens, int hidden, int dtype,
                                               b200_stream_t stream);

/* Decode-context-parallel attention merge across ranks: cp_lse_ag_out_rs / cp_lse_ag_out_ar
 * (vllm/v1/attention/ops/common.py:185-258; correction kernel :13-93) as ONE NVLink kernel instead of
 * all_gather(lse) -> Triton correction -> reduce_scatter / all_reduce(out):
 *   lse = log(sum_r exp(lse_r)) (NaN / +inf partial LSEs count as -inf; all -inf -> lse = -inf),
 *   out = T(sum_r T(out_r * exp(lse_r - lse))) in rank order with fp32 accumulation.
 * attn_out [num_tokens, num_heads, head_size] (dtype) and attn_lse [num_tokens, num_heads] (f32) must live in the
 * symmetric heap at the same offset on every rank. mode 0: out [num_tokens, num_heads/world, head_size] and
 * lse_out [num_tokens, num_heads/world] hold heads [rank*H/world, (rank+1)*H/world); mode 1: all heads on every
 * rank (bit-identical). out / lse_out are ordinary device memory; lse_out may be NULL.
 * is_lse_base_on_e: 1 natural-log LSEs, 0 log2 LSEs. */
B200_API int b200_dcp_lse_merge(void* ctx, void* out, float* lse_out, const void* attn_out, const float* attn_lse,
                                int64_t num_tokens, int num_heads, int head_size, int dtype, int mode,
                                int is_lse_base_on_e, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Sampler-side kernels of the decode step.
 * b200_argmax: greedy sampling, out[r] = argmax_v logits[r, v] (vllm/v1/sample/sampler.py greedy_sample =
 *   logits.argmax(dim=-1)); ties -> lowest index, NaN is the maximum; logits bf16 / f16 / f32.
 * b200_top_k_top_p: apply_top_k_top_p (vllm/v1/sample/ops/topk_topp_sampler.py:349-408) on fp32 logits, in
 *   place: values below the k[r]-th largest become -inf (ties kept; k[r] >= vocab or <= 0: no top-k), then
 *   the tokens whose ascending cumulative softmax mass is <= 1 - p[r] become -inf; the largest always stays.
 *   k (int32 / int64, k_dtype) and p (float) are device arrays [rows], either may be NULL.
 * b200_apply_repetition_penalties: torch.ops._C.apply_repetition_penalties_
 *   (csrc/libtorch_stable/sampler.cu:14-45,618-657): masks are bool [rows, vocab], penalties [rows] in the
 *   logits dtype; contiguous. */
B200_API int b200_argmax(int64_t* out, const void* logits, int64_t rows, int vocab, int64_t row_stride,
                         int dtype, b200_stream_t stream);
B200_API int b200_top_k_top_p(float* logits, const void* k, int k_dtype, const float* p, int64_t rows,
                              int vocab, int64_t row_stride, b200_stream_t stream);
B200_API int b200_apply_repetition_penalties(void* logits, const void* prompt_mask, const void* output_mask,
                                             const void* penalties, int64_t rows, int vocab, int dtype,
                                             b200_stream_t stream);

/* Development probe (tools/trace_gemm.py): device buffer receiving 8 x uint64 %globaltimer stamps per CTA of the
 * v2 FP8 GEMM kernel; NULL (default) switches it off. Not part of the reference surface. */
B200_API int b200_debug_set_trace(void* dev_ptr);

#ifdef __cplusplus
}
#endif
#endif /* B200OPS_H_ */
