// b200ops -- MoE routing: moe_align_block_size, topk_softmax, moe_sum, permute(+FP8 quant) and the
// topk-weighted un-permute. Integer / bandwidth work, no tensor cores.
//
// Contracts (reference):
//   moe_align_block_size   csrc/libtorch_stable/moe/moe_align_sum_kernels.cu:85-350, 625-723
//                          (python vllm/model_executor/layers/fused_moe/moe_align_block_size.py:11-103)
//   topk_softmax           csrc/libtorch_stable/moe/topk_softmax_kernels.cu:76-262
//   moe_sum                csrc/libtorch_stable/moe/moe_align_sum_kernels.cu:352-500
//   moe_permute/unpermute  csrc/libtorch_stable/moe/moe_permute_unpermute_op.cu (semantics)
//
// moe_align_block_size here is ONE deterministic single-CTA kernel for every size: tokens inside an
// expert segment keep ascending flat-index order (the reference's large-batch path scatters with
// atomicAdd and is order-nondeterministic, :311-322; its own test only checks per-expert sets).
#include "common.cuh"

namespace b200 {
namespace moe {

constexpr int ALIGN_THREADS = 1024;
constexpr int ALIGN_WARPS = ALIGN_THREADS / 32;

template <typename idx_t>
__device__ __forceinline__ int local_expert(const idx_t* __restrict__ topk_ids, int64_t i,
                                            const int32_t* __restrict__ expert_map, int num_experts) {
  const int64_t e = (int64_t)topk_ids[i];
  if (e >= num_experts || e < 0) return -1;
  return expert_map ? expert_map[e] : (int)e;
}

// smem: hist[W][E] (per-warp strip histograms -> running offsets), cum[E+1]
template <typename idx_t>
__global__ void __launch_bounds__(ALIGN_THREADS)
    moe_align_kernel(const idx_t* __restrict__ topk_ids, int32_t* __restrict__ sorted_token_ids,
                     int32_t* __restrict__ expert_ids, int32_t* __restrict__ num_tokens_post_pad,
                     int32_t* __restrict__ inv_perm /* nullable: [numel] flat index -> sorted pos */,
                     const int32_t* __restrict__ expert_map, int num_experts, int block_size,
                     int64_t numel, int max_num_tokens_padded, int max_num_m_blocks, int W) {
  extern __shared__ int32_t sm[];
  int32_t* hist = sm;                        // [W][num_experts]
  int32_t* cum = sm + W * num_experts;       // [num_experts + 1]
  __shared__ int32_t s_warp_tot[ALIGN_WARPS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int E = num_experts;

  for (int i = tid; i < max_num_tokens_padded; i += ALIGN_THREADS) sorted_token_ids[i] = (int32_t)numel;
  if (inv_perm)
    for (int64_t i = tid; i < numel; i += ALIGN_THREADS) inv_perm[i] = -1;
  for (int i = tid; i < W * E; i += ALIGN_THREADS) hist[i] = 0;
  __syncthreads();

  // strip of warp w: [w*strip, (w+1)*strip), strip a multiple of 32
  const int64_t strip = ((numel + W - 1) / W + 31) / 32 * 32;
  const int64_t s0 = (int64_t)warp * strip, s1 = s0 + strip < numel ? s0 + strip : numel;
  if (warp < W) {
    for (int64_t i = s0 + lane; i < s1; i += 32) {
      const int e = local_expert(topk_ids, i, expert_map, E);
      if (e >= 0) atomicAdd(&hist[warp * E + e], 1);
    }
  }
  __syncthreads();

  // per expert: total, padded total -> exclusive scan over experts (E <= 1024: one thread each)
  int padded = 0;
  if (tid < E) {
    int tot = 0;
    for (int w = 0; w < W; w++) tot += hist[w * E + tid];
    padded = (tot + block_size - 1) / block_size * block_size;
  }
  int inc = padded;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) s_warp_tot[warp] = inc;
  __syncthreads();
  int woff = 0;
  for (int w = 0; w < warp; w++) woff += s_warp_tot[w];
  const int excl = woff + inc - padded;
  if (tid < E) cum[tid] = excl;
  if (tid == E - 1) {
    cum[E] = excl + padded;
    *num_tokens_post_pad = excl + padded;
  }
  __syncthreads();
  const int total = cum[E];
  if (tid < E)
    for (int i = cum[tid]; i < cum[tid + 1]; i += block_size) expert_ids[i / block_size] = tid;
  for (int i = total / block_size + tid; i < max_num_m_blocks; i += ALIGN_THREADS) expert_ids[i] = -1;
  // turn hist[w][e] into the start offset of warp w's elements of expert e
  if (tid < E) {
    int run = cum[tid];
    for (int w = 0; w < W; w++) {
      const int c = hist[w * E + tid];
      hist[w * E + tid] = run;
      run += c;
    }
  }
  __syncthreads();

  // stable scatter: each warp walks its strip in order, ranks lanes with the same expert
  if (warp < W) {
    for (int64_t i0 = s0; i0 < s1; i0 += 32) {
      const int64_t i = i0 + lane;
      const int e = i < s1 ? local_expert(topk_ids, i, expert_map, E) : -1;
      const unsigned active = __ballot_sync(0xffffffffu, e >= 0);
      if (e >= 0) {
        const unsigned peers = __match_any_sync(active, e);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        const int pos = hist[warp * E + e] + rank;
        sorted_token_ids[pos] = (int32_t)i;
        if (inv_perm) inv_perm[i] = pos;
        __syncwarp(active);
        if (rank == 0) hist[warp * E + e] += __popc(peers);
      }
      __syncwarp();
    }
  }
}

// ---------------------------------------------------------------------------------- topk_softmax
// one warp per row; softmax in fp32 exactly as moeSoftmax (expf(x - max) * (1/Z), NaN/Inf -> 0),
// then k rounds of arg-max (ties -> lower expert index), optional selection bias, renormalise.
template <typename in_t>
__global__ void __launch_bounds__(256)
    topk_softmax_kernel(float* __restrict__ topk_weights, int32_t* __restrict__ topk_ids,
                        int32_t* __restrict__ token_expert_indices, const in_t* __restrict__ gating,
                        const float* __restrict__ bias, int num_rows, int E, int k, int renormalize) {
  extern __shared__ float srow[];  // [warps][E]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + warp;
  if (row >= num_rows) return;
  float* p = srow + warp * E;
  const in_t* g = gating + (int64_t)row * E;
  float mx = -3.402823466e+38f;
  for (int e = lane; e < E; e += 32) {
    const float v = Cvt<in_t>::to_f32(g[e]);
    p[e] = v;
    mx = fmaxf(mx, v);
  }
  mx = warp_max(mx);
  float z = 0.f;
  for (int e = lane; e < E; e += 32) z += expf(p[e] - mx);
  z = warp_sum(z);
  const float nf = 1.f / z;
  for (int e = lane; e < E; e += 32) {
    float s = expf(p[e] - mx) * nf;
    if (isnan(s) || isinf(s)) s = 0.f;
    p[e] = s;
  }
  __syncwarp();
  float selected_sum = 0.f;
  for (int ki = 0; ki < k; ki++) {
    float best = -INFINITY;
    int best_e = 0;
    for (int e = lane; e < E; e += 32) {
      float v = p[e];
      if (v < -0.5f) continue;  // already selected (marked below)
      if (bias) v += bias[e];
      if (v > best) { best = v; best_e = e; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oe = __shfl_xor_sync(0xffffffffu, best_e, o);
      if (ov > best || (ov == best && oe < best_e)) { best = ov; best_e = oe; }
    }
    const float w = p[best_e] < -0.5f ? 0.f : p[best_e];
    __syncwarp();
    if (lane == 0) {
      const int idx = k * row + ki;
      topk_weights[idx] = w;
      topk_ids[idx] = best_e;
      token_expert_indices[idx] = ki * num_rows + row;
      p[best_e] = -1.f;
    }
    selected_sum += w;
    __syncwarp();
  }
  if (renormalize) {
    const float denom = selected_sum > 0.f ? selected_sum : 1.f;
    const float scale = 1.f / denom;
    for (int ki = lane; ki < k; ki += 32) topk_weights[k * row + ki] *= scale;
  }
}

// ---------------------------------------------------------------------------------- moe_sum
template <typename T, typename idx_t>
__global__ void __launch_bounds__(256)
    moe_sum_kernel(T* __restrict__ out, const T* __restrict__ in, int64_t num_tokens, int topk, int d,
                   int64_t stride_token, int64_t stride_topk, const idx_t* __restrict__ topk_ids,
                   const int32_t* __restrict__ expert_map) {
  constexpr int N = 16 / sizeof(T);
  const int64_t nvec = d / N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < num_tokens * nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = i / nvec, v = i - t * nvec;
    float acc[N];
#pragma unroll
    for (int j = 0; j < N; j++) acc[j] = 0.f;
    for (int kk = 0; kk < topk; kk++) {
      if (topk_ids) {
        const int64_t e = (int64_t)topk_ids[t * topk + kk];
        if (e < 0 || (expert_map && expert_map[e] < 0)) continue;
      }
      const uint4 x = *reinterpret_cast<const uint4*>(in + t * stride_token + kk * stride_topk + v * N);
      const T* xe = reinterpret_cast<const T*>(&x);
#pragma unroll
      for (int j = 0; j < N; j++) acc[j] += Cvt<T>::to_f32(xe[j]);
    }
    uint4 r;
    T* re = reinterpret_cast<T*>(&r);
#pragma unroll
    for (int j = 0; j < N; j++) re[j] = Cvt<T>::from_f32(acc[j]);
    *reinterpret_cast<uint4*>(out + t * d + v * N) = r;
  }
}

// ------------------------------------------------------------- permute (+ static FP8 quant) rows
// dst[i, :] = quant(src[sorted_token_ids[i] / topk, :]) for i < *num_tokens_post_pad with a real
// token; pad rows are zero-filled so the grouped GEMM reads defined bytes.
template <typename T>
__global__ void __launch_bounds__(256)
    moe_permute_quant_kernel(uint8_t* __restrict__ dst, const T* __restrict__ src,
                             const int32_t* __restrict__ sorted_token_ids,
                             const int32_t* __restrict__ num_tokens_post_pad,
                             const float* __restrict__ scale, int topk, int64_t numel, int hidden,
                             int64_t src_stride) {
  constexpr int N = 16 / sizeof(T);
  const int row = blockIdx.x;
  if (row >= *num_tokens_post_pad) return;
  const int32_t flat = sorted_token_ids[row];
  uint8_t* d = dst + (int64_t)row * hidden;
  if (flat >= numel) {
    for (int v = threadIdx.x; v < hidden / 8; v += blockDim.x)
      reinterpret_cast<uint2*>(d)[v] = make_uint2(0u, 0u);
    return;
  }
  const T* s = src + (int64_t)(flat / topk) * src_stride;
  const float inv = __fdiv_rn(1.0f, *scale);
  for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
    const uint4 x = *reinterpret_cast<const uint4*>(s + v * N);
    const T* xe = reinterpret_cast<const T*>(&x);
    uint16_t o[N / 2];
#pragma unroll
    for (int j = 0; j < N / 2; j++) {
      const float a = fmaxf(-448.f, fminf(__fmul_rn(Cvt<T>::to_f32(xe[2 * j]), inv), 448.f));
      const float b = fmaxf(-448.f, fminf(__fmul_rn(Cvt<T>::to_f32(xe[2 * j + 1]), inv), 448.f));
      o[j] = f32x2_to_e4m3x2(a, b);
    }
    if constexpr (N == 8) *reinterpret_cast<uint2*>(d + v * N) = *reinterpret_cast<const uint2*>(o);
    else *reinterpret_cast<uint32_t*>(d + v * N) = *reinterpret_cast<const uint32_t*>(o);
  }
}

// Gather of already-quantised rows (per-token / per-group activation quant happens in token order, before the
// sort): dst[i, :] = src[sorted_token_ids[i] / topk, :] (bytes) and scale_dst[i, :] = scale_src[..., :] for
// i < *num_tokens_post_pad; pad rows get zero bytes and zero scales (their GEMM output is 0 and never gathered).
__global__ void __launch_bounds__(256)
    moe_permute_rows_kernel(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src,
                            const int32_t* __restrict__ sorted_token_ids,
                            const int32_t* __restrict__ num_tokens_post_pad, float* __restrict__ scale_dst,
                            const float* __restrict__ scale_src, int scale_cols, int64_t scale_src_stride,
                            int topk, int64_t numel, int row_vecs, int64_t src_stride_bytes) {
  const int row = blockIdx.x;
  if (row >= *num_tokens_post_pad) return;
  const int32_t flat = sorted_token_ids[row];
  uint4* d = reinterpret_cast<uint4*>(dst) + (int64_t)row * row_vecs;
  if (flat >= numel || flat < 0) {
    for (int v = threadIdx.x; v < row_vecs; v += blockDim.x) d[v] = make_uint4(0u, 0u, 0u, 0u);
    for (int c = threadIdx.x; c < scale_cols; c += blockDim.x) scale_dst[(int64_t)row * scale_cols + c] = 0.f;
    return;
  }
  const int64_t t = flat / topk;
  const uint4* sp = reinterpret_cast<const uint4*>(src + t * src_stride_bytes);
  for (int v = threadIdx.x; v < row_vecs; v += blockDim.x) d[v] = ld_nc_v4(sp + v);
  for (int c = threadIdx.x; c < scale_cols; c += blockDim.x)
    scale_dst[(int64_t)row * scale_cols + c] = scale_src[t * scale_src_stride + c];
}

// out[t, :] = sum_k w[t,k] * src[inv_perm[t*topk + k], :]  (fp32 accumulate; skipped if inv < 0)
template <typename T>
__global__ void __launch_bounds__(256)
    moe_unpermute_reduce_kernel(T* __restrict__ out, const T* __restrict__ src,
                                const int32_t* __restrict__ inv_perm,
                                const float* __restrict__ topk_weights /* nullable */,
                                const int32_t* __restrict__ num_valid_rows /* nullable */, int topk,
                                int hidden, int64_t out_stride) {
  constexpr int N = 16 / sizeof(T);
  const int64_t t = blockIdx.x;
  const int valid = num_valid_rows ? *num_valid_rows : 0x7fffffff;
  for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
    float acc[N];
#pragma unroll
    for (int j = 0; j < N; j++) acc[j] = 0.f;
    for (int kk = 0; kk < topk; kk++) {
      const int pos = inv_perm[t * topk + kk];
      if (pos < 0 || pos >= valid) continue;
      const float w = topk_weights ? topk_weights[t * topk + kk] : 1.f;
      const uint4 x = *reinterpret_cast<const uint4*>(src + (int64_t)pos * hidden + v * N);
      const T* xe = reinterpret_cast<const T*>(&x);
#pragma unroll
      for (int j = 0; j < N; j++) acc[j] += w * Cvt<T>::to_f32(xe[j]);
    }
    uint4 r;
    T* re = reinterpret_cast<T*>(&r);
#pragma unroll
    for (int j = 0; j < N; j++) re[j] = Cvt<T>::from_f32(acc[j]);
    *reinterpret_cast<uint4*>(out + t * out_stride + v * N) = r;
  }
}


// ------------------------------------------------------------- get_cutlass_moe_mm_data / shuffle_rows
// torch.ops._C.get_cutlass_moe_mm_data (csrc/libtorch_stable/quantization/w8a8/cutlass/moe/moe_data.cu:17-117,
// 236-289): the tightly packed (no block padding) expert sort used by the CUTLASS-style grouped GEMMs:
//   expert_offsets[e]      first packed row of expert e (E + 1 entries),
//   problem_sizes1/2[e]    (m_e, n1, k) and (m_e, k, n) -- (n1, m_e, k) / (k, m_e, n) when swap_ab,
//   input_permutation[r]   source token of packed row r,  output_permutation[t*topk+j] packed row of (t, j)
//                          (= expert_offsets[E] for topk_ids == -1),
//   blockscale_offsets[e]  sum of the previous experts' rows rounded up to 128 (NVFP4 scale-factor rows).
// One deterministic single-CTA kernel (the reference's four kernels order an expert's rows by atomicAdd):
// rows of an expert keep ascending flat-index order.
__global__ void __launch_bounds__(ALIGN_THREADS)
    moe_mm_data_kernel(const int32_t* __restrict__ topk_ids, int32_t* __restrict__ expert_offsets,
                       int32_t* __restrict__ problem_sizes1, int32_t* __restrict__ problem_sizes2,
                       int32_t* __restrict__ input_permutation, int32_t* __restrict__ output_permutation,
                       int32_t* __restrict__ blockscale_offsets, int E, int64_t numel, int topk, int n, int k,
                       int n1, int swap_ab, int W) {
  extern __shared__ int32_t sm[];
  int32_t* hist = sm;               // [W][E]
  int32_t* cum = sm + W * E;        // [E + 1]
  __shared__ int32_t s_warp_tot[ALIGN_WARPS], s_warp_tot_r[ALIGN_WARPS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < W * E; i += ALIGN_THREADS) hist[i] = 0;
  __syncthreads();
  const int64_t strip = ((numel + W - 1) / W + 31) / 32 * 32;
  const int64_t s0 = (int64_t)warp * strip, s1 = s0 + strip < numel ? s0 + strip : numel;
  if (warp < W) {
    for (int64_t i = s0 + lane; i < s1; i += 32) {
      const int e = topk_ids[i];
      if (e >= 0 && e < E) atomicAdd(&hist[warp * E + e], 1);
    }
  }
  __syncthreads();
  int tot = 0;
  if (tid < E)
    for (int w = 0; w < W; w++) tot += hist[w * E + tid];
  const int rounded = (tot + 127) / 128 * 128;
  int inc = tot, inc_r = rounded;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int a = __shfl_up_sync(0xffffffffu, inc, o);
    const int b = __shfl_up_sync(0xffffffffu, inc_r, o);
    if (lane >= o) { inc += a; inc_r += b; }
  }
  if (lane == 31) { s_warp_tot[warp] = inc; s_warp_tot_r[warp] = inc_r; }
  __syncthreads();
  int woff = 0, woff_r = 0;
  for (int w = 0; w < warp; w++) { woff += s_warp_tot[w]; woff_r += s_warp_tot_r[w]; }
  const int excl = woff + inc - tot, excl_r = woff_r + inc_r - rounded;
  if (tid < E) {
    cum[tid] = excl;
    expert_offsets[tid] = excl;
    if (blockscale_offsets) blockscale_offsets[tid] = excl_r;
    int32_t* p1 = problem_sizes1 + tid * 3;
    int32_t* p2 = problem_sizes2 + tid * 3;
    if (swap_ab) { p1[0] = n1; p1[1] = tot; p1[2] = k; p2[0] = k; p2[1] = tot; p2[2] = n; }
    else { p1[0] = tot; p1[1] = n1; p1[2] = k; p2[0] = tot; p2[1] = k; p2[2] = n; }
    if (tid == E - 1) {
      cum[E] = excl + tot;
      expert_offsets[E] = excl + tot;
      if (blockscale_offsets) blockscale_offsets[E] = excl_r + rounded;
    }
    int run = excl;
    for (int w = 0; w < W; w++) {
      const int c = hist[w * E + tid];
      hist[w * E + tid] = run;
      run += c;
    }
  }
  __syncthreads();
  const int num_rows = cum[E];
  // packed rows past the last expert exist only when pairs were dropped: no source token
  for (int64_t i = num_rows + tid; i < numel; i += ALIGN_THREADS) input_permutation[i] = -1;
  if (warp < W) {
    for (int64_t i0 = s0; i0 < s1; i0 += 32) {
      const int64_t i = i0 + lane;
      int e = i < s1 ? topk_ids[i] : -2;
      if (e >= E) e = -2;   // out-of-range ids are dropped without an output row
      if (i < s1 && e < 0) output_permutation[i] = num_rows;
      const unsigned active = __ballot_sync(0xffffffffu, e >= 0);
      if (e >= 0) {
        const unsigned peers = __match_any_sync(active, e);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        const int pos = hist[warp * E + e] + rank;
        input_permutation[pos] = (int32_t)(i / topk);
        output_permutation[i] = pos;
        __syncwarp(active);
        if (rank == 0) hist[warp * E + e] += __popc(peers);
      }
      __syncwarp();
    }
  }
}

// torch.ops._moe_C.shuffle_rows (csrc/libtorch_stable/moe/moe_permute_unpermute_op.cu shuffle_rows):
// dst[i, :] = src[dst2src_map[i], :]; a map entry outside [0, src_rows) yields a zero row (the reference's
// output_permutation marks dropped (token, k) pairs with the packed row count).
__global__ void __launch_bounds__(256)
    shuffle_rows_kernel(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src,
                        const int32_t* __restrict__ map, int64_t dst_rows, int64_t src_rows, int row_vecs,
                        int64_t src_stride_bytes) {
  for (int64_t row = blockIdx.x; row < dst_rows; row += gridDim.x) {
    const int32_t s = map[row];
    uint4* d = reinterpret_cast<uint4*>(dst) + row * row_vecs;
    if (s < 0 || s >= src_rows) {
      for (int v = threadIdx.x; v < row_vecs; v += blockDim.x) d[v] = make_uint4(0u, 0u, 0u, 0u);
    } else {
      const uint4* sp = reinterpret_cast<const uint4*>(src + (int64_t)s * src_stride_bytes);
      for (int v = threadIdx.x; v < row_vecs; v += blockDim.x) d[v] = ld_nc_v4(sp + v);
    }
  }
}

}  // namespace moe
}  // namespace b200

using namespace b200;
using namespace b200::moe;

extern "C" int b200_moe_align_block_size(const void* topk_ids, int topk_ids_dtype, int64_t numel,
                                         int num_experts, int block_size, int32_t* sorted_token_ids,
                                         int64_t sorted_len, int32_t* expert_ids,
                                         int64_t expert_ids_len, int32_t* num_tokens_post_pad,
                                         const int32_t* expert_map, int32_t* inv_perm,
                                         b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(numel >= 0 && num_experts > 0 && block_size > 0, B200_ERR_INVALID,
             "moe_align_block_size: bad sizes");
  B200_CHECK(((num_experts + 31) / 32) * 32 < 1024, B200_ERR_INVALID,
             "padded_num_experts must be less than 1024");
  B200_CHECK(topk_ids_dtype == B200_I32 || topk_ids_dtype == B200_I64, B200_ERR_INVALID,
             "moe_align_block_size: topk_ids must be int32 or int64");
  B200_CHECK(sorted_token_ids && expert_ids && num_tokens_post_pad && (numel == 0 || topk_ids),
             B200_ERR_INVALID, "moe_align_block_size: null pointer");
  B200_CHECK(numel < (1LL << 31), B200_ERR_UNSUPPORTED, "moe_align_block_size: numel too large");
  const int smem_max = max_smem_optin();
  B200_CHECK(smem_max > 0, B200_ERR_CUDA, "moe_align_block_size: no CUDA device (no CPU fallback)");
  int W = ALIGN_WARPS;
  while (W > 1 && (size_t)(W * num_experts + num_experts + 1) * 4 > (size_t)smem_max - 2048) W >>= 1;
  const size_t smem = (size_t)(W * num_experts + num_experts + 1) * 4;
  if (topk_ids_dtype == B200_I32) {
    static size_t cfg = 48 * 1024;
    if (smem > cfg) { B200_CUDA(cudaFuncSetAttribute(moe_align_kernel<int32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); cfg = smem; }
    moe_align_kernel<int32_t><<<1, ALIGN_THREADS, smem, stream>>>(
        (const int32_t*)topk_ids, sorted_token_ids, expert_ids, num_tokens_post_pad, inv_perm, expert_map,
        num_experts, block_size, numel, (int)sorted_len, (int)expert_ids_len, W);
  } else {
    static size_t cfg = 48 * 1024;
    if (smem > cfg) { B200_CUDA(cudaFuncSetAttribute(moe_align_kernel<int64_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); cfg = smem; }
    moe_align_kernel<int64_t><<<1, ALIGN_THREADS, smem, stream>>>(
        (const int64_t*)topk_ids, sorted_token_ids, expert_ids, num_tokens_post_pad, inv_perm, expert_map,
        num_experts, block_size, numel, (int)sorted_len, (int)expert_ids_len, W);
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_moe_topk_softmax(float* topk_weights, int32_t* topk_ids,
                                     int32_t* token_expert_indices, const void* gating_output,
                                     const float* bias, int64_t num_tokens, int num_experts, int topk,
                                     int renormalize, int gating_dtype, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && num_experts > 0 && topk > 0 && topk <= num_experts, B200_ERR_INVALID,
             "topk_softmax: bad sizes (tokens %lld experts %d topk %d)", (long long)num_tokens,
             num_experts, topk);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(topk_weights && topk_ids && token_expert_indices && gating_output, B200_ERR_INVALID,
             "topk_softmax: null pointer");
  B200_CHECK(num_experts <= 4096, B200_ERR_UNSUPPORTED, "topk_softmax: more than 4096 experts");
  const int warps = num_experts <= 1024 ? 8 : 2;
  const size_t smem = (size_t)warps * num_experts * 4;
  dim3 grid((unsigned)((num_tokens + warps - 1) / warps));
#define B200_TOPK(T)                                                                            \
  topk_softmax_kernel<T><<<grid, warps * 32, smem, stream>>>(topk_weights, topk_ids,             \
                                                             token_expert_indices, (const T*)gating_output, \
                                                             bias, (int)num_tokens, num_experts, topk, renormalize)
  switch (gating_dtype) {
    case B200_F32: B200_TOPK(float); break;
    case B200_BF16: B200_TOPK(__nv_bfloat16); break;
    case B200_F16: B200_TOPK(__half); break;
    default: set_last_error("topk_softmax: unsupported gating dtype %d", gating_dtype); return B200_ERR_INVALID;
  }
#undef B200_TOPK
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_moe_sum(void* out, const void* input, int64_t num_tokens, int topk, int d,
                            int64_t stride_token, int64_t stride_topk, const void* topk_ids,
                            int topk_ids_dtype, const int32_t* expert_map, int dtype,
                            b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && topk > 0 && d > 0, B200_ERR_INVALID, "moe_sum: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(out && input, B200_ERR_INVALID, "moe_sum: null pointer");
  const int esz = dtype_size(dtype);
  B200_CHECK(dtype == B200_BF16 || dtype == B200_F16 || dtype == B200_F32, B200_ERR_INVALID,
             "moe_sum: unsupported dtype");
  B200_CHECK(d % (16 / esz) == 0 && (stride_token * esz) % 16 == 0 && (stride_topk * esz) % 16 == 0 &&
                 ((uintptr_t)input & 15) == 0 && ((uintptr_t)out & 15) == 0,
             B200_ERR_UNSUPPORTED, "moe_sum: hidden size / strides must be 16-byte aligned");
  B200_CHECK(!topk_ids || topk_ids_dtype == B200_I32 || topk_ids_dtype == B200_I64, B200_ERR_INVALID,
             "moe_sum: topk_ids must be int32 or int64");
  const int64_t work = num_tokens * (d / (16 / esz));
  int blocks = (int)((work + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
#define B200_MSUM(T, I)                                                                   \
  moe_sum_kernel<T, I><<<blocks, 256, 0, stream>>>((T*)out, (const T*)input, num_tokens, topk, d, \
                                                   stride_token, stride_topk, (const I*)topk_ids, expert_map)
#define B200_MSUM_T(T)                                      \
  if (topk_ids && topk_ids_dtype == B200_I64) B200_MSUM(T, int64_t); \
  else B200_MSUM(T, int32_t)
  if (dtype == B200_BF16) { B200_MSUM_T(__nv_bfloat16); }
  else if (dtype == B200_F16) { B200_MSUM_T(__half); }
  else { B200_MSUM_T(float); }
#undef B200_MSUM_T
#undef B200_MSUM
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_moe_permute_fp8(void* dst, const void* src, const int32_t* sorted_token_ids,
                                    const int32_t* num_tokens_post_pad, const float* scale,
                                    int64_t max_rows, int topk, int64_t numel, int hidden,
                                    int64_t src_stride, int src_dtype, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(max_rows >= 0 && hidden > 0 && topk > 0, B200_ERR_INVALID, "moe_permute: bad sizes");
  if (max_rows == 0) return B200_OK;
  B200_CHECK(dst && src && sorted_token_ids && num_tokens_post_pad && scale, B200_ERR_INVALID,
             "moe_permute: null pointer");
  B200_CHECK(hidden % 8 == 0 && (src_stride * 2) % 16 == 0 && ((uintptr_t)src & 15) == 0 &&
                 ((uintptr_t)dst & 15) == 0,
             B200_ERR_UNSUPPORTED, "moe_permute: hidden must be a multiple of 8 and pointers 16-byte aligned");
  if (src_dtype == B200_BF16)
    moe_permute_quant_kernel<__nv_bfloat16><<<(unsigned)max_rows, 256, 0, stream>>>(
        (uint8_t*)dst, (const __nv_bfloat16*)src, sorted_token_ids, num_tokens_post_pad, scale, topk, numel, hidden, src_stride);
  else if (src_dtype == B200_F16)
    moe_permute_quant_kernel<__half><<<(unsigned)max_rows, 256, 0, stream>>>(
        (uint8_t*)dst, (const __half*)src, sorted_token_ids, num_tokens_post_pad, scale, topk, numel, hidden, src_stride);
  else {
    set_last_error("moe_permute: src must be half or bfloat16");
    return B200_ERR_INVALID;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_moe_unpermute_reduce(void* out, const void* src, const int32_t* inv_perm,
                                         const float* topk_weights, const int32_t* num_valid_rows,
                                         int64_t num_tokens, int topk,
                                         int hidden, int64_t out_stride, int dtype,
                                         b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && hidden > 0 && topk > 0, B200_ERR_INVALID, "moe_unpermute: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(out && src && inv_perm, B200_ERR_INVALID, "moe_unpermute: null pointer");
  B200_CHECK(hidden % 8 == 0 && (out_stride * 2) % 16 == 0, B200_ERR_UNSUPPORTED,
             "moe_unpermute: hidden must be a multiple of 8");
  if (dtype == B200_BF16)
    moe_unpermute_reduce_kernel<__nv_bfloat16><<<(unsigned)num_tokens, 256, 0, stream>>>(
        (__nv_bfloat16*)out, (const __nv_bfloat16*)src, inv_perm, topk_weights, num_valid_rows, topk, hidden, out_stride);
  else if (dtype == B200_F16)
    moe_unpermute_reduce_kernel<__half><<<(unsigned)num_tokens, 256, 0, stream>>>(
        (__half*)out, (const __half*)src, inv_perm, topk_weights, num_valid_rows, topk, hidden, out_stride);
  else {
    set_last_error("moe_unpermute: dtype must be half or bfloat16");
    return B200_ERR_INVALID;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_moe_mm_data(const int32_t* topk_ids, int32_t* expert_offsets, int32_t* problem_sizes1,
                                int32_t* problem_sizes2, int32_t* input_permutation,
                                int32_t* output_permutation, int32_t* blockscale_offsets, int num_experts,
                                int64_t numel, int topk, int n, int k, int is_gated, int swap_ab,
                                b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(numel >= 0 && num_experts > 0 && topk > 0 && n > 0 && k > 0, B200_ERR_INVALID,
             "get_cutlass_moe_mm_data: bad sizes");
  B200_CHECK(num_experts <= 1024, B200_ERR_UNSUPPORTED, "get_cutlass_moe_mm_data: at most 1024 experts");
  B200_CHECK(expert_offsets && problem_sizes1 && problem_sizes2 && (numel == 0 || (topk_ids && input_permutation &&
                                                                                   output_permutation)),
             B200_ERR_INVALID, "get_cutlass_moe_mm_data: null pointer");
  B200_CHECK(numel < (1LL << 31), B200_ERR_UNSUPPORTED, "get_cutlass_moe_mm_data: numel too large");
  const int smem_max = max_smem_optin();
  B200_CHECK(smem_max > 0, B200_ERR_CUDA, "get_cutlass_moe_mm_data: no CUDA device (no CPU fallback)");
  int W = ALIGN_WARPS;
  while (W > 1 && (size_t)(W * num_experts + num_experts + 1) * 4 > (size_t)smem_max - 2048) W >>= 1;
  const size_t smem = (size_t)(W * num_experts + num_experts + 1) * 4;
  static size_t cfg = 48 * 1024;
  if (smem > cfg) {
    B200_CUDA(cudaFuncSetAttribute(moe_mm_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cfg = smem;
  }
  moe_mm_data_kernel<<<1, ALIGN_THREADS, smem, stream>>>(topk_ids, expert_offsets, problem_sizes1, problem_sizes2,
                                                         input_permutation, output_permutation, blockscale_offsets,
                                                         num_experts, numel, topk, n, k, is_gated ? 2 * n : n,
                                                         swap_ab, W);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_shuffle_rows(void* dst, const void* src, const int32_t* dst2src_map, int64_t dst_rows,
                                 int64_t src_rows, int64_t row_bytes, int64_t src_stride_bytes,
                                 b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(dst_rows >= 0 && src_rows >= 0 && row_bytes > 0, B200_ERR_INVALID, "shuffle_rows: bad sizes");
  if (dst_rows == 0) return B200_OK;
  B200_CHECK(dst && dst2src_map && (src || src_rows == 0), B200_ERR_INVALID, "shuffle_rows: null pointer");
  B200_CHECK(row_bytes % 16 == 0 && src_stride_bytes % 16 == 0 && ((uintptr_t)src & 15) == 0 &&
                 ((uintptr_t)dst & 15) == 0 && row_bytes / 16 < (1LL << 31),
             B200_ERR_UNSUPPORTED, "shuffle_rows: rows must be 16-byte aligned multiples of 16 bytes");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "shuffle_rows: no CUDA device (no CPU fallback)");
  const int64_t cap = (int64_t)sm_count() * 16;
  const unsigned blocks = (unsigned)(dst_rows < cap ? dst_rows : cap);
  shuffle_rows_kernel<<<blocks, 256, 0, stream>>>((uint8_t*)dst, (const uint8_t*)src, dst2src_map, dst_rows,
                                                  src_rows, (int)(row_bytes / 16), src_stride_bytes);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_moe_permute_rows(void* dst, const void* src, const int32_t* sorted_token_ids,
                                     const int32_t* num_tokens_post_pad, float* scale_dst, const float* scale_src,
                                     int scale_cols, int64_t scale_src_stride, int64_t max_rows, int topk,
                                     int64_t numel, int64_t row_bytes, int64_t src_stride_bytes,
                                     b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(max_rows >= 0 && row_bytes > 0 && topk > 0 && scale_cols >= 0, B200_ERR_INVALID,
             "moe_permute_rows: bad sizes");
  if (max_rows == 0) return B200_OK;
  B200_CHECK(dst && src && sorted_token_ids && num_tokens_post_pad && (scale_cols == 0 || (scale_dst && scale_src)),
             B200_ERR_INVALID, "moe_permute_rows: null pointer");
  B200_CHECK(row_bytes % 16 == 0 && src_stride_bytes % 16 == 0 && ((uintptr_t)src & 15) == 0 &&
                 ((uintptr_t)dst & 15) == 0 && row_bytes / 16 < (1LL << 31),
             B200_ERR_UNSUPPORTED, "moe_permute_rows: rows must be 16-byte aligned multiples of 16 bytes");
  moe_permute_rows_kernel<<<(unsigned)max_rows, 256, 0, stream>>>(
      (uint8_t*)dst, (const uint8_t*)src, sorted_token_ids, num_tokens_post_pad, scale_dst, scale_src, scale_cols,
      scale_src_stride, topk, numel, (int)(row_bytes / 16), src_stride_bytes);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
