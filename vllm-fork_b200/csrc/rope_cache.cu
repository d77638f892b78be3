// b200ops -- rotary embedding fused with the KV-cache write (one launch instead of two).
//
// Contract: the pair  torch.ops._C.rotary_embedding (csrc/libtorch_stable/pos_encoding_kernels.cu:8-99)
// -> torch.ops._C_cache_ops.reshape_and_cache_flash (cache_kernels.cu:314-400), i.e. the backend hook
// AttentionImpl.do_rope_and_kv_cache_update (vllm/v1/attention/backend.py:950-1006):
//   q, k rotated IN PLACE (neox or gpt-j pairing, cos|sin rows of cos_sin_cache at positions[t]);
//   key_cache / value_cache [slot] <- rotated k, v   (FP8: e4m3_satfinite(float(x) / scale), true division;
//   slot < 0: the token is rotated but not cached, as the unfused pair would do).
// Arithmetic is the unfused pair's, operation for operation (rope_rotate() is shared with rotary_kernel),
// so the fused launch is bit-identical to the two launches.
//
// B200 design: latency kernel (2.6 MB touched per decode layer). One CTA per token; phase 1 rotates
// (Hq + Hkv) heads with 16-byte vectors (neox: 8 x's and their 8 y's per thread), bar.sync, phase 2 streams
// the token's K and V rows into the paged cache with 16-byte loads / 16- or 8-byte stores.
#include "common.cuh"

namespace b200 {

template <typename T, bool NEOX, bool FP8>
__global__ void __launch_bounds__(256)
    rope_and_cache_kernel(const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
                          const T* __restrict__ value, const T* __restrict__ cos_sin, void* __restrict__ key_cache,
                          void* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int rot_dim,
                          int head_size, int num_heads, int num_kv_heads, int block_size, int64_t q_stride,
                          int64_t k_stride, int64_t v_stride, int64_t block_stride, int64_t page_stride,
                          int64_t head_stride, const float* __restrict__ k_scale, const float* __restrict__ v_scale,
                          int kv_scale_stride, int vec_ok) {
  using out_t = typename std::conditional<FP8, uint8_t, T>::type;
  constexpr int N = 16 / sizeof(T);
  pdl_launch_dependents();
  pdl_wait();   // q / k / v are the qkv GEMM's output
  const int64_t tok = blockIdx.x;
  const int64_t pos = positions[tok];
  const int embed = rot_dim / 2;
  const T* cosp = cos_sin + pos * rot_dim;
  const T* sinp = cosp + embed;
  T* qrow = query + tok * q_stride;
  T* krow = key + tok * k_stride;
  // ---------------- phase 1: rotate q and k in place
  if (NEOX && vec_ok) {
    const int vpe = embed / N;                       // 16-byte vectors per half
    const int items = (num_heads + num_kv_heads) * vpe;
    for (int i = threadIdx.x; i < items; i += blockDim.x) {
      const int h = i / vpe, v = i - h * vpe;
      T* arr = (h < num_heads ? qrow + (int64_t)h * head_size : krow + (int64_t)(h - num_heads) * head_size);
      uint4 xv = *reinterpret_cast<const uint4*>(arr + v * N);
      uint4 yv = *reinterpret_cast<const uint4*>(arr + embed + v * N);
      const uint4 cv = *reinterpret_cast<const uint4*>(cosp + v * N);
      const uint4 sv = *reinterpret_cast<const uint4*>(sinp + v * N);
      T* xe = reinterpret_cast<T*>(&xv);
      T* ye = reinterpret_cast<T*>(&yv);
      const T* ce = reinterpret_cast<const T*>(&cv);
      const T* se = reinterpret_cast<const T*>(&sv);
#pragma unroll
      for (int j = 0; j < N; j++) rope_rotate<T>(xe[j], ye[j], Cvt<T>::to_f32(ce[j]), Cvt<T>::to_f32(se[j]));
      *reinterpret_cast<uint4*>(arr + v * N) = xv;
      *reinterpret_cast<uint4*>(arr + embed + v * N) = yv;
    }
  } else {
    const int nq = num_heads * embed, nk = num_kv_heads * embed;
    for (int i = threadIdx.x; i < nq + nk; i += blockDim.x) {
      const bool is_k = i >= nq;
      const int j = is_k ? i - nq : i;
      const int h = j / embed, ro = j - h * embed;
      T* arr = (is_k ? krow : qrow) + (int64_t)h * head_size;
      const int xi = NEOX ? ro : 2 * ro;
      const int yi = NEOX ? embed + ro : 2 * ro + 1;
      T x = arr[xi], y = arr[yi];
      rope_rotate<T>(x, y, Cvt<T>::to_f32(cosp[ro]), Cvt<T>::to_f32(sinp[ro]));
      arr[xi] = x;
      arr[yi] = y;
    }
  }
  const int64_t slot = slot_mapping[tok];
  if (slot < 0) return;   // whole CTA: padded token, nothing to cache
  __syncthreads();        // this CTA's rotated k is visible to all of its threads
  // ---------------- phase 2: K (rotated) and V -> paged cache
  const int64_t blk = slot / block_size;
  const int64_t off = slot % block_size;
  out_t* kc = reinterpret_cast<out_t*>(key_cache) + blk * block_stride + off * page_stride;
  out_t* vc = reinterpret_cast<out_t*>(value_cache) + blk * block_stride + off * page_stride;
  const T* vrow = value + tok * v_stride;
  if (vec_ok) {
    const int vph = head_size / N;
    const int nvec = num_kv_heads * vph;
    for (int i = threadIdx.x; i < 2 * nvec; i += blockDim.x) {
      const bool is_v = i >= nvec;
      const int j = is_v ? i - nvec : i;
      const int h = j / vph, v = j - h * vph;
      const T* src = (is_v ? vrow : krow) + (int64_t)h * head_size + v * N;
      out_t* dst = (is_v ? vc : kc) + (int64_t)h * head_stride + v * N;
      const uint4 raw = *reinterpret_cast<const uint4*>(src);
      if constexpr (!FP8) {
        *reinterpret_cast<uint4*>(dst) = raw;
      } else {
        const float sc = (is_v ? v_scale : k_scale)[h * kv_scale_stride];
        const T* e = reinterpret_cast<const T*>(&raw);
        uint16_t o[N / 2];
#pragma unroll
        for (int q = 0; q < N / 2; q++)
          o[q] = f32x2_to_e4m3x2(__fdiv_rn(Cvt<T>::to_f32(e[2 * q]), sc), __fdiv_rn(Cvt<T>::to_f32(e[2 * q + 1]), sc));
        if constexpr (N == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
        else *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(o);
      }
    }
  } else {
    const int n = num_kv_heads * head_size;
    for (int i = threadIdx.x; i < 2 * n; i += blockDim.x) {
      const bool is_v = i >= n;
      const int j = is_v ? i - n : i;
      const int h = j / head_size, d = j - h * head_size;
      const T x = (is_v ? vrow : krow)[j];
      out_t* dst = (is_v ? vc : kc) + (int64_t)h * head_stride + d;
      if constexpr (FP8) {
        const float sc = (is_v ? v_scale : k_scale)[h * kv_scale_stride];
        *dst = f32_to_e4m3(__fdiv_rn(Cvt<T>::to_f32(x), sc));
      } else {
        *dst = x;
      }
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rope_and_cache_flash(const int64_t* positions, void* query, void* key, const void* value,
                                         const void* cos_sin_cache, void* key_cache, void* value_cache,
                                         const int64_t* slot_mapping, int64_t num_tokens, int rot_dim,
                                         int head_size, int num_heads, int num_kv_heads, int block_size,
                                         int64_t query_stride, int64_t key_stride, int64_t value_stride,
                                         int64_t block_stride, int64_t page_stride, int64_t head_stride,
                                         int is_neox, int dtype, int cache_dtype, const float* k_scale,
                                         const float* v_scale, int kv_scale_stride, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && head_size >= rot_dim && num_heads > 0 &&
                 num_kv_heads > 0 && block_size > 0,
             B200_ERR_INVALID, "rope_and_cache_flash: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(positions && query && key && value && cos_sin_cache && key_cache && value_cache && slot_mapping,
             B200_ERR_INVALID, "rope_and_cache_flash: null pointer");
  B200_CHECK(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_INVALID,
             "rope_and_cache_flash: dtype must be bf16 or f16");
  const bool fp8 = cache_dtype == B200_FP8_E4M3;
  B200_CHECK(fp8 || cache_dtype == dtype, B200_ERR_INVALID,
             "rope_and_cache_flash: cache dtype must equal the input dtype (\"auto\") or be fp8_e4m3");
  B200_CHECK(!fp8 || (k_scale && v_scale), B200_ERR_INVALID, "rope_and_cache_flash: fp8 cache needs k_scale and v_scale");
  B200_CHECK(kv_scale_stride == 0 || kv_scale_stride == 1, B200_ERR_INVALID,
             "rope_and_cache_flash: k_scale and v_scale must be of shape [1] or [num_heads]");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "rope_and_cache_flash: no CUDA device (no CPU fallback)");
  const int osz = fp8 ? 1 : 2;
  auto al = [](const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const int vec_ok = (rot_dim / 2) % 8 == 0 && head_size % 8 == 0 && al(query, 16) && al(key, 16) && al(value, 16) &&
                     al(cos_sin_cache, 16) && (rot_dim * 2) % 16 == 0 && (query_stride * 2) % 16 == 0 &&
                     (key_stride * 2) % 16 == 0 && (value_stride * 2) % 16 == 0 && al(key_cache, 8 * osz) &&
                     al(value_cache, 8 * osz) && (block_stride * osz) % (8 * osz) == 0 &&
                     (page_stride * osz) % (8 * osz) == 0 && (head_stride * osz) % (8 * osz) == 0;
  dim3 grid((unsigned)num_tokens);
#define B200_RC(T, NX, F8)                                                                                   \
  B200_CUDA(launch_pdl(rope_and_cache_kernel<T, NX, F8>, grid, dim3(256), 0, stream,                         \
      positions, (T*)query, (T*)key, (const T*)value, (const T*)cos_sin_cache, key_cache, value_cache,       \
      slot_mapping, rot_dim, head_size, num_heads, num_kv_heads, block_size, query_stride, key_stride,       \
      value_stride, block_stride, page_stride, head_stride, k_scale, v_scale, kv_scale_stride, vec_ok))
#define B200_RC2(T)                         \
  do {                                      \
    if (is_neox) {                          \
      if (fp8) B200_RC(T, true, true);      \
      else B200_RC(T, true, false);         \
    } else {                                \
      if (fp8) B200_RC(T, false, true);     \
      else B200_RC(T, false, false);        \
    }                                       \
  } while (0)
  if (dtype == B200_BF16) B200_RC2(__nv_bfloat16);
  else B200_RC2(__half);
#undef B200_RC2
#undef B200_RC
  B200_LAUNCH_CHECK();
  return B200_OK;
}
