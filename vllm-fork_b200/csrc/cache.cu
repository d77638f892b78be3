// b200ops -- KV-cache scatter write (reshape_and_cache_flash / reshape_and_cache).
//
// Behavioural contract: csrc/libtorch_stable/cache_kernels.cu:254-400,749-864 of the reference
// (slot = slot_mapping[t]; block = slot / block_size; offset = slot % block_size; slot < 0 skipped;
// FP8: e4m3_satfinite(float(x) / scale) with TRUE division, quant_utils.cuh:492-503).
//
// B200 design: pure bandwidth / latency kernel. One CTA per token; each thread moves 16-byte vectors
// (8 x 16-bit or 4 x f32 in, 16 B or 8 B / 4 B out) with fully coalesced accesses along head_size.
// K and V of a token are handled by the same CTA so a decode step (T = batch) is ONE launch of T
// CTAs. A scalar path covers misaligned shapes (head_size = 46 etc., test_cache.py:429-463).
#include "common.cuh"

namespace b200 {

template <typename in_t>
struct VecIn {
  static constexpr int N = 16 / sizeof(in_t);
};

template <typename in_t, bool FP8>
__device__ __forceinline__ void convert_store_vec(const in_t* __restrict__ src, void* dst,
                                                  float scale) {
  constexpr int N = VecIn<in_t>::N;
  uint4 raw = *reinterpret_cast<const uint4*>(src);
  if constexpr (!FP8) {
    *reinterpret_cast<uint4*>(dst) = raw;
  } else {
    const in_t* e = reinterpret_cast<const in_t*>(&raw);
    uint16_t o[N / 2];
#pragma unroll
    for (int i = 0; i < N / 2; i++) {
      // true division, as the reference (not multiplication by a reciprocal)
      float a = __fdiv_rn(Cvt<in_t>::to_f32(e[2 * i]), scale);
      float b = __fdiv_rn(Cvt<in_t>::to_f32(e[2 * i + 1]), scale);
      o[i] = f32x2_to_e4m3x2(a, b);
    }
    if constexpr (N == 8) {
      *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
    } else {
      *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(o);
    }
  }
}

template <typename in_t, bool FP8>
__global__ void __launch_bounds__(256)
    reshape_and_cache_flash_vec_kernel(const in_t* __restrict__ key, const in_t* __restrict__ value,
                                       void* __restrict__ key_cache, void* __restrict__ value_cache,
                                       const int64_t* __restrict__ slot_mapping, int num_heads,
                                       int head_size, int block_size, int64_t key_stride,
                                       int64_t value_stride, int64_t block_stride,
                                       int64_t page_stride, int64_t head_stride,
                                       const float* __restrict__ k_scale,
                                       const float* __restrict__ v_scale, int kv_scale_stride) {
  using out_t = typename std::conditional<FP8, uint8_t, in_t>::type;
  constexpr int N = VecIn<in_t>::N;
  const int64_t tok = blockIdx.x;
  const int64_t slot = slot_mapping[tok];
  if (slot < 0) return;  // padded token
  const int64_t blk = slot / block_size;
  const int64_t off = slot % block_size;
  const int vph = head_size / N;       // vectors per head
  const int nvec = num_heads * vph;    // per K or V
  out_t* kc = reinterpret_cast<out_t*>(key_cache) + blk * block_stride + off * page_stride;
  out_t* vc = reinterpret_cast<out_t*>(value_cache) + blk * block_stride + off * page_stride;
  const in_t* ks = key + tok * key_stride;
  const in_t* vs = value + tok * value_stride;
  for (int i = threadIdx.x; i < 2 * nvec; i += blockDim.x) {
    const bool is_v = i >= nvec;
    const int j = is_v ? i - nvec : i;
    const int h = j / vph;
    const int v = j - h * vph;
    const in_t* src = (is_v ? vs : ks) + (int64_t)h * head_size + v * N;
    out_t* dst = (is_v ? vc : kc) + (int64_t)h * head_stride + v * N;
    float sc = 1.f;
    if constexpr (FP8) sc = (is_v ? v_scale : k_scale)[h * kv_scale_stride];
    convert_store_vec<in_t, FP8>(src, dst, sc);
  }
}

template <typename in_t, bool FP8>
__global__ void __launch_bounds__(256)
    reshape_and_cache_flash_scalar_kernel(const in_t* __restrict__ key,
                                          const in_t* __restrict__ value,
                                          void* __restrict__ key_cache,
                                          void* __restrict__ value_cache,
                                          const int64_t* __restrict__ slot_mapping, int num_heads,
                                          int head_size, int block_size, int64_t key_stride,
                                          int64_t value_stride, int64_t block_stride,
                                          int64_t page_stride, int64_t head_stride,
                                          const float* __restrict__ k_scale,
                                          const float* __restrict__ v_scale, int kv_scale_stride) {
  using out_t = typename std::conditional<FP8, uint8_t, in_t>::type;
  const int64_t tok = blockIdx.x;
  const int64_t slot = slot_mapping[tok];
  if (slot < 0) return;
  const int64_t blk = slot / block_size;
  const int64_t off = slot % block_size;
  const int n = num_heads * head_size;
  out_t* kc = reinterpret_cast<out_t*>(key_cache) + blk * block_stride + off * page_stride;
  out_t* vc = reinterpret_cast<out_t*>(value_cache) + blk * block_stride + off * page_stride;
  const in_t* ks = key + tok * key_stride;
  const in_t* vs = value + tok * value_stride;
  for (int i = threadIdx.x; i < 2 * n; i += blockDim.x) {
    const bool is_v = i >= n;
    const int j = is_v ? i - n : i;
    const int h = j / head_size;
    const int d = j - h * head_size;
    const in_t x = (is_v ? vs : ks)[j];
    out_t* dst = (is_v ? vc : kc) + (int64_t)h * head_stride + d;
    if constexpr (FP8) {
      const float sc = (is_v ? v_scale : k_scale)[h * kv_scale_stride];
      *dst = f32_to_e4m3(__fdiv_rn(Cvt<in_t>::to_f32(x), sc));
    } else {
      *dst = x;
    }
  }
}

// legacy layout: K [nb, H, D/x, bs, x], V [nb, H, D, bs]  (cache_kernels.cu:254-312)
template <typename in_t, bool FP8>
__global__ void __launch_bounds__(256)
    reshape_and_cache_legacy_kernel(const in_t* __restrict__ key, const in_t* __restrict__ value,
                                    void* __restrict__ key_cache, void* __restrict__ value_cache,
                                    const int64_t* __restrict__ slot_mapping, int num_heads,
                                    int head_size, int block_size, int x, int64_t key_stride,
                                    int64_t value_stride, const float* __restrict__ k_scale,
                                    const float* __restrict__ v_scale) {
  using out_t = typename std::conditional<FP8, uint8_t, in_t>::type;
  const int64_t tok = blockIdx.x;
  const int64_t slot = slot_mapping[tok];
  if (slot < 0) return;
  const int64_t blk = slot / block_size;
  const int64_t off = slot % block_size;
  const int n = num_heads * head_size;
  out_t* kc = reinterpret_cast<out_t*>(key_cache);
  out_t* vc = reinterpret_cast<out_t*>(value_cache);
  float ksc = 1.f, vsc = 1.f;
  if constexpr (FP8) {
    ksc = *k_scale;
    vsc = *v_scale;
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int h = i / head_size;
    const int d = i - h * head_size;
    const int64_t kdst = (((blk * num_heads + h) * (head_size / x) + d / x) * block_size + off) * x +
                         (d % x);
    const int64_t vdst = ((blk * num_heads + h) * head_size + d) * block_size + off;
    const in_t kx = key[tok * key_stride + i];
    const in_t vx = value[tok * value_stride + i];
    if constexpr (FP8) {
      kc[kdst] = f32_to_e4m3(__fdiv_rn(Cvt<in_t>::to_f32(kx), ksc));
      vc[vdst] = f32_to_e4m3(__fdiv_rn(Cvt<in_t>::to_f32(vx), vsc));
    } else {
      kc[kdst] = kx;
      vc[vdst] = vx;
    }
  }
}

template <typename in_t, bool FP8>
static int launch_flash(const void* key, const void* value, void* key_cache, void* value_cache,
                        const int64_t* slot_mapping, int64_t num_tokens, int num_heads,
                        int head_size, int block_size, int64_t key_stride, int64_t value_stride,
                        int64_t block_stride, int64_t page_stride, int64_t head_stride,
                        const float* k_scale, const float* v_scale, int kv_scale_stride,
                        cudaStream_t stream) {
  constexpr int N = VecIn<in_t>::N;
  const int osz = FP8 ? 1 : (int)sizeof(in_t);
  auto al = [](const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const int oal = N * osz;  // bytes written per vector
  const bool vec_ok = head_size % N == 0 && al(key, 16) && al(value, 16) &&
                      (key_stride * (int64_t)sizeof(in_t)) % 16 == 0 &&
                      (value_stride * (int64_t)sizeof(in_t)) % 16 == 0 && al(key_cache, oal) &&
                      al(value_cache, oal) && (block_stride * osz) % oal == 0 &&
                      (page_stride * osz) % oal == 0 && (head_stride * osz) % oal == 0;
  dim3 grid((unsigned)num_tokens);
  if (vec_ok) {
    int nvec = 2 * num_heads * head_size / N;
    int threads = nvec < 256 ? ((nvec + 31) / 32) * 32 : 256;
    reshape_and_cache_flash_vec_kernel<in_t, FP8><<<grid, threads, 0, stream>>>(
        (const in_t*)key, (const in_t*)value, key_cache, value_cache, slot_mapping, num_heads,
        head_size, block_size, key_stride, value_stride, block_stride, page_stride, head_stride,
        k_scale, v_scale, kv_scale_stride);
  } else {
    reshape_and_cache_flash_scalar_kernel<in_t, FP8><<<grid, 256, 0, stream>>>(
        (const in_t*)key, (const in_t*)value, key_cache, value_cache, slot_mapping, num_heads,
        head_size, block_size, key_stride, value_stride, block_stride, page_stride, head_stride,
        k_scale, v_scale, kv_scale_stride);
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_reshape_and_cache_flash(const void* key, const void* value, void* key_cache,
                                            void* value_cache, const int64_t* slot_mapping,
                                            int64_t num_tokens, int num_heads, int head_size,
                                            int block_size, int64_t key_stride,
                                            int64_t value_stride, int64_t block_stride,
                                            int64_t page_stride, int64_t head_stride, int in_dtype,
                                            int cache_dtype, const float* k_scale,
                                            const float* v_scale, int kv_scale_stride,
                                            b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && num_heads > 0 && head_size > 0 && block_size > 0, B200_ERR_INVALID,
             "reshape_and_cache_flash: bad sizes (tokens %lld heads %d head_size %d block_size %d)",
             (long long)num_tokens, num_heads, head_size, block_size);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(key && value && key_cache && value_cache && slot_mapping, B200_ERR_INVALID,
             "reshape_and_cache_flash: null pointer");
  const bool fp8 = cache_dtype == B200_FP8_E4M3;
  B200_CHECK(fp8 || cache_dtype == in_dtype, B200_ERR_INVALID,
             "reshape_and_cache_flash: cache dtype %d must equal input dtype %d (\"auto\") or be fp8_e4m3",
             cache_dtype, in_dtype);
  B200_CHECK(!fp8 || (k_scale && v_scale), B200_ERR_INVALID,
             "reshape_and_cache_flash: fp8 cache needs k_scale and v_scale");
  B200_CHECK(kv_scale_stride == 0 || kv_scale_stride == 1, B200_ERR_INVALID,
             "reshape_and_cache_flash: k_scale and v_scale must be of shape [1] or [num_heads]");
#define B200_RCF(T)                                                                               \
  return fp8 ? launch_flash<T, true>(key, value, key_cache, value_cache, slot_mapping, num_tokens, \
                                     num_heads, head_size, block_size, key_stride, value_stride,  \
                                     block_stride, page_stride, head_stride, k_scale, v_scale,    \
                                     kv_scale_stride, stream)                                     \
             : launch_flash<T, false>(key, value, key_cache, value_cache, slot_mapping,           \
                                      num_tokens, num_heads, head_size, block_size, key_stride,   \
                                      value_stride, block_stride, page_stride, head_stride,       \
                                      k_scale, v_scale, kv_scale_stride, stream)
  switch (in_dtype) {
    case B200_BF16: B200_RCF(__nv_bfloat16);
    case B200_F16: B200_RCF(__half);
    case B200_F32: B200_RCF(float);
    default:
      set_last_error("reshape_and_cache_flash: unsupported input type of kv cache: %d", in_dtype);
      return B200_ERR_INVALID;
  }
#undef B200_RCF
}

extern "C" int b200_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                      void* value_cache, const int64_t* slot_mapping,
                                      int64_t num_tokens, int num_heads, int head_size,
                                      int block_size, int x, int64_t key_stride,
                                      int64_t value_stride, int in_dtype, int cache_dtype,
                                      const float* k_scale, const float* v_scale,
                                      b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && num_heads > 0 && head_size > 0 && block_size > 0 && x > 0 &&
                 head_size % x == 0,
             B200_ERR_INVALID, "reshape_and_cache: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(key && value && key_cache && value_cache && slot_mapping, B200_ERR_INVALID,
             "reshape_and_cache: null pointer");
  const bool fp8 = cache_dtype == B200_FP8_E4M3;
  B200_CHECK(fp8 || cache_dtype == in_dtype, B200_ERR_INVALID,
             "reshape_and_cache: cache dtype must equal input dtype or be fp8_e4m3");
  B200_CHECK(!fp8 || (k_scale && v_scale), B200_ERR_INVALID, "reshape_and_cache: fp8 needs scales");
  dim3 grid((unsigned)num_tokens);
#define B200_RCL(T)                                                                              \
  if (fp8)                                                                                       \
    reshape_and_cache_legacy_kernel<T, true><<<grid, 256, 0, stream>>>(                           \
        (const T*)key, (const T*)value, key_cache, value_cache, slot_mapping, num_heads,         \
        head_size, block_size, x, key_stride, value_stride, k_scale, v_scale);                   \
  else                                                                                           \
    reshape_and_cache_legacy_kernel<T, false><<<grid, 256, 0, stream>>>(                          \
        (const T*)key, (const T*)value, key_cache, value_cache, slot_mapping, num_heads,         \
        head_size, block_size, x, key_stride, value_stride, k_scale, v_scale);                   \
  break
  switch (in_dtype) {
    case B200_BF16: B200_RCL(__nv_bfloat16);
    case B200_F16: B200_RCL(__half);
    case B200_F32: B200_RCL(float);
    default:
      set_last_error("reshape_and_cache: unsupported input dtype %d", in_dtype);
      return B200_ERR_INVALID;
  }
#undef B200_RCL
  B200_LAUNCH_CHECK();
  return B200_OK;
}
