// b200ops -- sampler-side kernels on the decode path: greedy argmax, top-k / top-p logit masking,
// repetition penalties.
//
// Contracts:
//   argmax                  : Sampler.greedy_sample = logits.argmax(dim=-1) (vllm/v1/sample/sampler.py); ties ->
//                             lowest index, NaN counts as the maximum (torch semantics).
//   top_k_top_p             : apply_top_k_top_p (vllm/v1/sample/ops/topk_topp_sampler.py:349-408): fp32 logits
//                             masked in place with -inf.  top-k keeps every value >= the k-th largest (ties kept);
//                             top-p then drops, in ascending order, the tokens whose cumulative softmax mass is
//                             <= 1 - p and always keeps the largest token.
//   apply_repetition_penalties_ : csrc/libtorch_stable/sampler.cu:14-45 (logit > 0 ? logit / pen : logit * pen on
//                             tokens present in the prompt or the output), bit-exact.
//
// B200 design. All three are HBM / L2 bandwidth kernels over [rows, vocab]. The reference sorts the whole vocabulary
// (torch.sort + softmax + cumsum + scatter: ~6 full-size temporaries); here one CTA per row finds both thresholds by
// MSB-first radix selection over order-preserving 32-bit keys - 4 counting passes for the k-th largest value,
// 4 mass passes for the top-p boundary - and rewrites the row once. The row (<= 0.5 MB) stays in L2 across passes.
// Probability mass is accumulated in 64-bit fixed point (2^-40 of the largest probability), so the histograms are
// order-independent and the result is deterministic.
#include "common.cuh"

namespace b200 {
namespace sampler {

constexpr int TPB = 1024;

__device__ __forceinline__ uint32_t f32_key(float x) {   // ascending float order == ascending unsigned order
  const uint32_t b = __float_as_uint(x);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

template <typename T>
__device__ __forceinline__ bool better(float v, int i, float bv, int bi) {
  const bool vn = v != v, bn = bv != bv;
  if (vn != bn) return vn;          // NaN beats everything
  if (vn) return i < bi;
  return v > bv || (v == bv && i < bi);
}

template <typename in_t>
__global__ void __launch_bounds__(512)
    argmax_kernel(int64_t* __restrict__ out, const in_t* __restrict__ logits, int vocab, int64_t row_stride) {
  constexpr int N = 16 / sizeof(in_t);
  const in_t* row = logits + (int64_t)blockIdx.x * row_stride;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  const bool vec = (reinterpret_cast<uintptr_t>(row) & 15) == 0;
  const int nvec = vec ? vocab / N : 0;
  // 4 independent 16-byte loads in flight per thread; a vector enters the exact (tie / NaN aware) comparison only
  // when its fmaxf-tree maximum beats the running best or its sum is NaN (a NaN element, or +inf with -inf)
  constexpr int U = 4;
  for (int v0 = threadIdx.x; v0 < nvec; v0 += blockDim.x * U) {
    uint4 raw[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int v = v0 + u * blockDim.x;
      if (v < nvec) raw[u] = ld_nc_v4(row + (int64_t)v * N);
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int v = v0 + u * blockDim.x;
      if (v >= nvec) break;
      const in_t* e = reinterpret_cast<const in_t*>(&raw[u]);
      float m = -INFINITY, sum = 0.f;
#pragma unroll
      for (int j = 0; j < N; j++) {
        const float x = Cvt<in_t>::to_f32(e[j]);
        m = fmaxf(m, x);
        sum += x;
      }
      if (m > bv || sum != sum || bi == 0x7fffffff) {
#pragma unroll
        for (int j = 0; j < N; j++) {
          const float x = Cvt<in_t>::to_f32(e[j]);
          if (better<in_t>(x, v * N + j, bv, bi)) { bv = x; bi = v * N + j; }
        }
      }
    }
  }
  for (int i = nvec * N + threadIdx.x; i < vocab; i += blockDim.x) {
    const float x = Cvt<in_t>::to_f32(row[i]);
    if (better<in_t>(x, i, bv, bi)) { bv = x; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (better<in_t>(ov, oi, bv, bi)) { bv = ov; bi = oi; }
  }
  __shared__ float sv[16];
  __shared__ int si[16];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sv[w] = bv; si[w] = bi; }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    bv = l < nw ? sv[l] : -INFINITY;
    bi = l < nw ? si[l] : 0x7fffffff;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (better<in_t>(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (l == 0) out[blockIdx.x] = bi == 0x7fffffff ? 0 : bi;
  }
}

template <typename T>
__global__ void __launch_bounds__(1024)
    repetition_penalty_kernel(T* __restrict__ logits, const uint8_t* __restrict__ prompt_mask,
                              const uint8_t* __restrict__ output_mask, const T* __restrict__ penalties,
                              int vocab, int tile) {
  const int64_t base = (int64_t)blockIdx.x * vocab;
  const int lo = blockIdx.y * tile, hi = min(lo + tile, vocab);
  const T pen = penalties[blockIdx.x];
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    if (prompt_mask[base + i] || output_mask[base + i]) {
      const T x = logits[base + i];
      // the reference's arithmetic in T (operator/ and operator* of the scalar type: fp32 op, one rounding)
      const float xf = Cvt<T>::to_f32(x), pf = Cvt<T>::to_f32(pen);
      logits[base + i] = xf > 0.f ? Cvt<T>::from_f32(__fdiv_rn(xf, pf)) : Cvt<T>::from_f32(__fmul_rn(xf, pf));
    }
  }
}

// ---- top-k / top-p thresholds by radix selection, one CTA per row
struct TopKPShared {
  unsigned int cnt[256];
  unsigned long long mass[256];
  unsigned int prefix, rank;        // selection state
  unsigned long long below;         // mass strictly below the current prefix range
  unsigned long long total;
  float red[32];
};

__device__ __forceinline__ float block_max_f(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float r = l < (int)(blockDim.x >> 5) ? red[l] : -INFINITY;
  r = warp_max(r);
  return __shfl_sync(0xffffffffu, r, 0);
}

template <typename k_t>
__global__ void __launch_bounds__(TPB)
    top_k_top_p_kernel(float* __restrict__ logits, const k_t* __restrict__ ks, const float* __restrict__ ps,
                       int vocab, int64_t row_stride) {
  __shared__ TopKPShared sh;
  float* row = logits + (int64_t)blockIdx.x * row_stride;
  const int tid = threadIdx.x;
  const long long kreq = ks ? (long long)ks[blockIdx.x] : (long long)vocab;
  const bool do_k = ks != nullptr && kreq > 0 && kreq < vocab;
  const bool do_p = ps != nullptr;
  // row maximum (also the element that is always kept)
  float mx = -INFINITY;
  for (int i = tid; i < vocab; i += TPB) mx = fmaxf(mx, row[i]);
  mx = block_max_f(mx, sh.red);
  const uint32_t key_max = f32_key(mx);
  uint32_t thr_k = 0;   // keys below are masked by top-k
  if (do_k) {
    if (tid == 0) { sh.prefix = 0; sh.rank = (unsigned)kreq; }
    for (int d = 3; d >= 0; d--) {
      const int shift = d * 8;
      for (int b = tid; b < 256; b += TPB) sh.cnt[b] = 0;
      __syncthreads();
      const uint32_t prefix = sh.prefix;
      const uint32_t himask = d == 3 ? 0u : (0xffffffffu << (shift + 8));
      for (int i = tid; i < vocab; i += TPB) {
        const uint32_t key = f32_key(row[i]);
        if ((key & himask) == prefix) atomicAdd(&sh.cnt[(key >> shift) & 255u], 1u);
      }
      __syncthreads();
      if (tid == 0) {   // walk the bins from the top: the bin holding the rank-th largest key
        unsigned r = sh.rank, b = 255;
        for (;; b--) {
          const unsigned c = sh.cnt[b];
          if (c >= r || b == 0) break;
          r -= c;
        }
        sh.rank = r;
        sh.prefix = prefix | (b << shift);
      }
      __syncthreads();
    }
    thr_k = sh.prefix;
    __syncthreads();   // everyone has read the threshold before the top-p stage reuses the selection state
  }
  uint32_t thr = thr_k;
  if (do_p) {
    const float p = ps[blockIdx.x];
    // fixed-point masses: exp(x - max) * 2^40 (largest element = 2^40), order-independent sums
    const float kScale = 1099511627776.0f;
    auto mass_of = [&](float x) -> unsigned long long {
      return (unsigned long long)(__expf(x - mx) * kScale);
    };
    if (tid == 0) { sh.total = 0; sh.prefix = 0; sh.below = 0; }
    __syncthreads();
    unsigned long long part = 0;
    for (int i = tid; i < vocab; i += TPB) {
      const float x = row[i];
      if (f32_key(x) >= thr_k) part += mass_of(x);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((tid & 31) == 0) atomicAdd(&sh.total, part);
    __syncthreads();
    // tokens whose inclusive ascending cumulative mass is <= (1 - p) * total are dropped
    const double tgt_d = (1.0 - (double)p) * (double)sh.total;
    const unsigned long long target = tgt_d <= 0.0 ? 0ull : (unsigned long long)tgt_d;
    for (int d = 3; d >= 0; d--) {
      const int shift = d * 8;
      for (int b = tid; b < 256; b += TPB) { sh.mass[b] = 0; sh.cnt[b] = 0; }
      __syncthreads();
      const uint32_t prefix = sh.prefix;
      const uint32_t himask = d == 3 ? 0u : (0xffffffffu << (shift + 8));
      for (int i = tid; i < vocab; i += TPB) {
        const float x = row[i];
        const uint32_t key = f32_key(x);
        if (key >= thr_k && (key & himask) == prefix) {
          atomicAdd(&sh.mass[(key >> shift) & 255u], mass_of(x));
          atomicAdd(&sh.cnt[(key >> shift) & 255u], 1u);
        }
      }
      __syncthreads();
      if (tid == 0) {   // first bin (ascending) where the inclusive cumulative mass exceeds the target
        unsigned long long below = sh.below;
        int b = 0, last = -1;
        for (; b < 256; b++) {
          if (sh.cnt[b] == 0) continue;
          last = b;
          if (below + sh.mass[b] > target) break;
          below += sh.mass[b];
        }
        if (b == 256) {            // rounding pushed the target past the total: keep the top bin
          b = last < 0 ? 255 : last;
          below -= sh.mass[b];
        }
        sh.below = below;
        sh.prefix = prefix | ((uint32_t)b << shift);
      }
      __syncthreads();
    }
    thr = max(thr_k, sh.prefix);
  }
  thr = min(thr, key_max);   // the largest token always survives (top_p_mask[:, -1] = False)
  if (thr == 0) return;
  for (int i = tid; i < vocab; i += TPB) {
    if (f32_key(row[i]) < thr) row[i] = -INFINITY;
  }
}

}  // namespace sampler
}  // namespace b200

using namespace b200;

extern "C" int b200_argmax(int64_t* out, const void* logits, int64_t rows, int vocab, int64_t row_stride,
                           int dtype, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(dtype == B200_BF16 || dtype == B200_F16 || dtype == B200_F32, B200_ERR_INVALID,
             "argmax: logits must be bf16 / f16 / f32 (dtype %d)", dtype);
  B200_CHECK(rows >= 0 && vocab > 0 && row_stride >= vocab, B200_ERR_INVALID, "argmax: bad shape");
  if (rows == 0) return B200_OK;
  B200_CHECK(out && logits, B200_ERR_INVALID, "argmax: null pointer");
  B200_CHECK((row_stride * dtype_size(dtype)) % 16 == 0 || row_stride == vocab, B200_ERR_INVALID,
             "argmax: row stride must keep rows 16-byte aligned");
  switch (dtype) {
    case B200_BF16:
      sampler::argmax_kernel<__nv_bfloat16><<<(unsigned)rows, 512, 0, stream>>>(out, (const __nv_bfloat16*)logits, vocab, row_stride);
      break;
    case B200_F16:
      sampler::argmax_kernel<__half><<<(unsigned)rows, 512, 0, stream>>>(out, (const __half*)logits, vocab, row_stride);
      break;
    default:
      sampler::argmax_kernel<float><<<(unsigned)rows, 512, 0, stream>>>(out, (const float*)logits, vocab, row_stride);
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_top_k_top_p(float* logits, const void* k, int k_dtype, const float* p, int64_t rows,
                                int vocab, int64_t row_stride, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(rows >= 0 && vocab > 0 && row_stride >= vocab, B200_ERR_INVALID, "apply_top_k_top_p: bad shape");
  B200_CHECK(!k || k_dtype == B200_I32 || k_dtype == B200_I64, B200_ERR_INVALID,
             "apply_top_k_top_p: k must be int32 or int64");
  if (rows == 0 || (!k && !p)) return B200_OK;
  B200_CHECK(logits, B200_ERR_INVALID, "apply_top_k_top_p: null logits");
  if (k && k_dtype == B200_I64)
    sampler::top_k_top_p_kernel<int64_t><<<(unsigned)rows, sampler::TPB, 0, stream>>>(logits, (const int64_t*)k, p, vocab, row_stride);
  else
    sampler::top_k_top_p_kernel<int32_t><<<(unsigned)rows, sampler::TPB, 0, stream>>>(logits, (const int32_t*)k, p, vocab, row_stride);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_apply_repetition_penalties(void* logits, const void* prompt_mask, const void* output_mask,
                                               const void* penalties, int64_t rows, int vocab, int dtype,
                                               b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(dtype == B200_BF16 || dtype == B200_F16 || dtype == B200_F32, B200_ERR_INVALID,
             "apply_repetition_penalties_: logits must be bf16 / f16 / f32");
  B200_CHECK(rows >= 0 && vocab > 0, B200_ERR_INVALID, "apply_repetition_penalties_: bad shape");
  if (rows == 0) return B200_OK;
  B200_CHECK(logits && prompt_mask && output_mask && penalties, B200_ERR_INVALID,
             "apply_repetition_penalties_: null pointer");
  const int sms = sm_count() > 0 ? sm_count() : 148;
  int tiles = (int)((sms + rows - 1) / rows);
  if (tiles < 1) tiles = 1;
  if (tiles > vocab) tiles = vocab;
  const int tile = (vocab + tiles - 1) / tiles;
  dim3 grid((unsigned)rows, (unsigned)tiles);
  const int threads = tile < 1024 ? ((tile + 31) / 32 * 32) : 1024;
  switch (dtype) {
    case B200_BF16:
      sampler::repetition_penalty_kernel<__nv_bfloat16><<<grid, threads, 0, stream>>>(
          (__nv_bfloat16*)logits, (const uint8_t*)prompt_mask, (const uint8_t*)output_mask, (const __nv_bfloat16*)penalties, vocab, tile);
      break;
    case B200_F16:
      sampler::repetition_penalty_kernel<__half><<<grid, threads, 0, stream>>>(
          (__half*)logits, (const uint8_t*)prompt_mask, (const uint8_t*)output_mask, (const __half*)penalties, vocab, tile);
      break;
    default:
      sampler::repetition_penalty_kernel<float><<<grid, threads, 0, stream>>>(
          (float*)logits, (const uint8_t*)prompt_mask, (const uint8_t*)output_mask, (const float*)penalties, vocab, tile);
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}
